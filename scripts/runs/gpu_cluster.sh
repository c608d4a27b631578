mkdir -p gpurun_out
for c in 2 4; do
HOP_TC_CLUSTER=$c timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "tensor_core or folded or forward_host" > gpurun_out/tc_tests_c$c.log 2>&1; echo "cluster $c pytest rc=$?"; tail -2 gpurun_out/tc_tests_c$c.log
done
for c in 1 2 4; do
for wl in sweep_p64k_d256 sweep_p256k_d256 sweep_p64k_d128; do
HOP_TC_CLUSTER=$c timeout 300 python bench.py --workload $wl --path f16 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c${c}_$wl.log 2>&1
python - <<PY
import json
for ln in open("gpurun_out/bench_c${c}_$wl.log"):
    if ln.startswith("{"):
        j=json.loads(ln); print("cluster $c $wl", "ms/step %.3f kernel_ms %.3f TF %.1f frac %.3f e2e %.3g clocks %s"%(j["ms_per_step"], j["roofline"]["kernel_ms"], j["roofline"]["achieved"], j["roofline"]["frac"], j["e2e"]["value"], j["clocks"]["sm_mhz"]))
PY
done
done
