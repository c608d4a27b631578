#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_parity_gpu.py tests/test_backward_tc_gpu.py tests/test_chain_gpu.py -m gpu -q -s > gpurun_out/r02j_tests.txt 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|^FAILED|Error" gpurun_out/r02j_tests.txt | tail -8
timeout 600 python scripts/bench_backward.py > gpurun_out/r02j_backward.jsonl 2> gpurun_out/r02j_backward.err; echo "bwd bench rc=$?"
cut -c1-330 gpurun_out/r02j_backward.jsonl; tail -3 gpurun_out/r02j_backward.err
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-sustained > gpurun_out/r02j_bench.json 2> gpurun_out/r02j_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02j_bench.json").read().strip().splitlines()[-1])
print("value %.2fM frac %.3f prep %.3f ms (%.0f GB/s) bwd %.2f ms frac %.3f fwd_incl_prep %.2f"%(d["value"]/1e6,d["roofline"]["frac"],d["prep"]["ms"],d["prep"]["gbs"],d["backward"]["ms_backward"],d["backward"]["frac"],d["backward"]["ms_forward_incl_prep"]))
PY
BWD="python scripts/bench_backward.py --points=16384,65536,256"
timeout 300 $BWD > gpurun_out/r02j_bwd_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02j_bwd_launches.csv $BWD > gpurun_out/r02j_bwd_launches_ncu.log 2>&1
echo "ncu rc=$?"
