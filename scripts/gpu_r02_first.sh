#!/bin/bash
# round 2, first GPU pass: parity tests, tensor-core backward tests, smoke, bench (N = 1)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/r02_first_smi.txt 2>&1
timeout 2400 python -m pytest tests -m gpu -x -q -rs --durations=15 --deselect tests/test_backward_tc_gpu.py > gpurun_out/r02_first_tests.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_first_tests.txt
tail -5 gpurun_out/r02_first_tests.txt
timeout 900 python -m pytest tests/test_backward_tc_gpu.py -m gpu -q -s > gpurun_out/r02_first_bwd_tc.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_first_bwd_tc.txt
tail -5 gpurun_out/r02_first_bwd_tc.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02_first_smoke.txt 2>&1; echo "smoke rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r02_first_bench.json 2> gpurun_out/r02_first_bench.err; echo "bench rc=$?"
tail -c 600 gpurun_out/r02_first_bench.err
