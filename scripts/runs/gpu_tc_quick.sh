mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -s -k "tensor_core or folded" > gpurun_out/tc_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/tc_tests.log
grep -E 'rel-L2|passed|failed|rc=|Error|error' gpurun_out/tc_tests.log | tail -40
for wl in ${WLS:-sweep_p64k_d256 sweep_p64k_d128 sweep_p64k_d64 sweep_p64k_d256_t1 sweep_p256k_d256}; do
timeout 300 python bench.py --workload $wl --path f16 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$wl.log 2>&1; echo "rc=$?" >> gpurun_out/bench_$wl.log
python - <<PY
import json
for ln in open("gpurun_out/bench_$wl.log"):
    if ln.startswith("{"):
        j=json.loads(ln); print("$wl", "ms/step %.3f kernel_ms %.3f TF %.1f frac %.3f e2e %.3g launches %d clocks %s"%(j["ms_per_step"], j["roofline"]["kernel_ms"], j["roofline"]["achieved"], j["roofline"]["frac"], j["e2e"]["value"], j["gpu_launches"], j["clocks"]))
    elif "rc=" in ln and "rc=0" not in ln: print("$wl", ln)
PY
done
