#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/probes/bwd_determinism.py > gpurun_out/r02h_bwd_determinism.jsonl 2>&1; cut -c1-330 gpurun_out/r02h_bwd_determinism.jsonl | tail -7
for i in 1 2; do
timeout 900 python -m pytest tests/test_backward_tc_gpu.py -m gpu -q -s > gpurun_out/r02h_bwd_tests_$i.txt 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|bad" gpurun_out/r02h_bwd_tests_$i.txt | tail -3
done
timeout 600 python scripts/bench_backward.py > gpurun_out/r02h_backward.jsonl 2> gpurun_out/r02h_backward.err; echo "bwd bench rc=$?"
cat gpurun_out/r02h_backward.jsonl | cut -c1-330; tail -3 gpurun_out/r02h_backward.err
BWD="python scripts/bench_backward.py --points=16384,65536,256"
timeout 300 $BWD > gpurun_out/r02h_bwd_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02h_bwd_launches.csv $BWD > gpurun_out/r02h_bwd_launches_ncu.log 2>&1
echo "ncu rc=$?"
