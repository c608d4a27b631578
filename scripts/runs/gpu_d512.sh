# full GPU test suite, then the D = 512 sweep points next to the D = 256 headline
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/gpu_tests.log
tail -8 gpurun_out/gpu_tests.log
for wl in sweep_p64k_d256 sweep_p64k_d512 sweep_p256k_d512 sweep_p64k_d512_t1; do
  timeout 300 python bench.py --workload $wl --path f16 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$wl.log 2>&1
  python - <<PY
import json
for ln in open("gpurun_out/bench_$wl.log"):
    if ln.startswith("{"):
        j=json.loads(ln); print("$wl", "ms/step %.3f kernel_ms %.3f TF %.1f frac %.3f e2e %.3g clocks %s %s"%(j["ms_per_step"], j["roofline"]["kernel_ms"], j["roofline"]["achieved"], j["roofline"]["frac"], j["e2e"]["value"], j["clocks"]["sm_mhz"], j["clocks"]["reasons"]))
PY
  tail -2 gpurun_out/bench_$wl.log | cut -c1-300
done
