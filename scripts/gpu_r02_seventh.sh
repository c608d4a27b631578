#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_chain_gpu.py -m gpu -q > gpurun_out/r02g_chain_tests.txt 2>&1; echo "pytest rc=$?"
tail -2 gpurun_out/r02g_chain_tests.txt
timeout 600 python scripts/probes/chain_time.py all 0,2,4 > gpurun_out/r02g_chain_time.jsonl 2> gpurun_out/r02g_chain_time.err; echo "probe rc=$?"
cat gpurun_out/r02g_chain_time.jsonl; tail -3 gpurun_out/r02g_chain_time.err
timeout 900 python scripts/bench_model_configs.py > gpurun_out/r02g_model_configs.jsonl 2> gpurun_out/r02g_model_configs.err; echo "configs rc=$?"
