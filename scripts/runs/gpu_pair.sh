mkdir -p gpurun_out
for pr in 1; do
HOP_TC_PAIR=$pr timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -s -k "tensor_core or folded or forward_host" > gpurun_out/tc_tests_pair$pr.log 2>&1; echo "pair $pr pytest rc=$?"; grep -E 'rel-L2|passed|failed|Error|error' gpurun_out/tc_tests_pair$pr.log | tail -22
done
for pr in 1 0; do
for wl in sweep_p64k_d256 sweep_p256k_d256 sweep_p64k_d128 sweep_p64k_d64 sweep_p64k_d256_t1; do
HOP_TC_PAIR=$pr timeout 300 python bench.py --workload $wl --path f16 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_pair${pr}_$wl.log 2>&1
python - <<PY
import json
for ln in open("gpurun_out/bench_pair${pr}_$wl.log"):
    if ln.startswith("{"):
        j=json.loads(ln); print("pair $pr $wl", "ms/step %.3f kernel_ms %.3f TF %.1f frac %.3f e2e %.3g clocks %s %s"%(j["ms_per_step"], j["roofline"]["kernel_ms"], j["roofline"]["achieved"], j["roofline"]["frac"], j["e2e"]["value"], j["clocks"]["sm_mhz"], j["clocks"]["reasons"]))
PY
tail -2 gpurun_out/bench_pair${pr}_$wl.log | grep -i -E 'error|Traceback' 
done
done
