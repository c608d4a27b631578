mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_default.log 2>&1; grep '^{' gpurun_out/bench_default.log | cut -c1-900
for wl in sweep_p64k_d128 sweep_p64k_d64 sweep_p256k_d256 sweep_p64k_d256_t1; do
  timeout 300 python bench.py --workload $wl --path f16 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/b_$wl.log 2>&1
  python - gpurun_out/b_$wl.log $wl <<'PY'
import json,sys
for ln in open(sys.argv[1]):
    if ln.startswith("{"):
        j=json.loads(ln); print(sys.argv[2], "value %.4g ms/step %.3f kernel_ms %.3f TF %.1f frac %.3f e2e %.4g"%(j["value"], j["ms_per_step"], j["roofline"]["kernel_ms"], j["roofline"]["achieved"], j["roofline"]["frac"], j["e2e"]["value"]))
PY
done
