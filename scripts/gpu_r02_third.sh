#!/bin/bash
# round 2, third GPU pass (one GPU): whole -m gpu suite, smoke, bench N = 1, backward bench, launch list under ncu
mkdir -p gpurun_out
timeout 2700 python -m pytest tests -m gpu -q -rs --durations=15 -s > gpurun_out/r02c_tests.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r02c_tests.txt
grep -E "passed|failed|rc=" gpurun_out/r02c_tests.txt | tail -4
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02c_smoke.txt 2>&1; echo "smoke rc=$?"
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r02c_bench.json 2> gpurun_out/r02c_bench.err; echo "bench rc=$?"
timeout 600 python scripts/bench_backward.py > gpurun_out/r02c_backward.jsonl 2> gpurun_out/r02c_backward.err; echo "bwd bench rc=$?"
BENCH="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-sustained --no-parity --no-backward"
timeout 600 $BENCH > gpurun_out/r02c_launches_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02c_bench_launches.csv $BENCH > gpurun_out/r02c_launches_ncu.log 2>&1
echo "ncu launches rc=$?"
ls -la gpurun_out | tail -12
