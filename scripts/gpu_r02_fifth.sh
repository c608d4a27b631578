#!/bin/bash
# round 2, fifth GPU pass (one GPU): the bottleneck chain -- tests, timing at configs 1 / 2 / 3, launch list
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_chain_gpu.py -m gpu -q -s > gpurun_out/r02e_chain_tests.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r02e_chain_tests.txt
grep -E "passed|failed|rc=" gpurun_out/r02e_chain_tests.txt | tail -3
timeout 900 python scripts/bench_model_configs.py > gpurun_out/r02e_model_configs.jsonl 2> gpurun_out/r02e_model_configs.err; echo "configs rc=$?"
tail -3 gpurun_out/r02e_model_configs.err
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -q -s -k "preparation or diffuse" > gpurun_out/r02e_tests.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r02e_tests.txt
grep -E "passed|failed|rc=" gpurun_out/r02e_tests.txt | tail -3
