#!/bin/bash
# round 2, GPU pass: parity tests, tensor-core backward tests, smoke, bench (N = 1), launch list of the bench
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -rs --durations=15 --deselect tests/test_backward_tc_gpu.py > gpurun_out/r02_tests.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_tests.txt
tail -5 gpurun_out/r02_tests.txt
timeout 900 python -m pytest tests/test_backward_tc_gpu.py -m gpu -q -s > gpurun_out/r02_bwd_tc.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_bwd_tc.txt
tail -5 gpurun_out/r02_bwd_tc.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02_smoke.txt 2>&1; echo "smoke rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench.json 2> gpurun_out/r02_bench.err; echo "bench rc=$?"
BENCH="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-sustained --no-parity"
$BENCH > gpurun_out/r02_launches_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_bench_launches.csv $BENCH > gpurun_out/r02_launches_ncu.log 2>&1
echo "ncu rc=$?"
