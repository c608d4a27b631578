#!/bin/bash
# round 2, GPU pass: parity tests, tensor-core backward tests, smoke, bench (N = 1), backward bench, ncu evidence
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -rs --durations=15 --deselect tests/test_backward_tc_gpu.py > gpurun_out/r02_tests.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_tests.txt
tail -5 gpurun_out/r02_tests.txt
timeout 900 python -m pytest tests/test_backward_tc_gpu.py -m gpu -q -s > gpurun_out/r02_bwd_tc.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_bwd_tc.txt
tail -5 gpurun_out/r02_bwd_tc.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02_smoke.txt 2>&1; echo "smoke rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench.json 2> gpurun_out/r02_bench.err; echo "bench rc=$?"
timeout 600 python scripts/bench_backward.py > gpurun_out/r02_backward.jsonl 2> gpurun_out/r02_backward.err; echo "bwd bench rc=$?"
BENCH="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-sustained --no-parity --no-backward"
$BENCH > gpurun_out/r02_launches_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_bench_launches.csv $BENCH > gpurun_out/r02_launches_ncu.log 2>&1
echo "ncu launches rc=$?"
$BENCH > gpurun_out/r02_full_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:hop_fwd_tc_kernel -s 12 -c 1 -o gpurun_out/r02_fwd_full $BENCH > gpurun_out/r02_full_ncu.log 2>&1
echo "ncu fwd full rc=$?"
BWD="python scripts/bench_backward.py --points=16384,65536,256"
$BWD > gpurun_out/r02_bwd_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_bwd_launches.csv $BWD > gpurun_out/r02_bwd_launches_ncu.log 2>&1
echo "ncu bwd launches rc=$?"
$BWD > gpurun_out/r02_bwd_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:hop_bwd_tc_kernel -s 6 -c 3 -o gpurun_out/r02_bwd_full $BWD > gpurun_out/r02_bwd_full_ncu.log 2>&1
echo "ncu bwd full rc=$?"
ls -la gpurun_out | tail -20
