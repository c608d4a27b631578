#!/bin/bash
# chain kernel: tests, timing, one full ncu capture at config 3
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_chain_gpu.py -m gpu -q > gpurun_out/r02f_chain_tests.txt 2>&1; echo "pytest rc=$?"
tail -2 gpurun_out/r02f_chain_tests.txt
timeout 600 python scripts/bench_model_configs.py > gpurun_out/r02f_model_configs.jsonl 2> gpurun_out/r02f_model_configs.err; echo "configs rc=$?"
timeout 300 python scripts/probes/chain_time.py ffhq64 > gpurun_out/r02f_chain_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:hop_chain_kernel -s 20 -c 1 -o gpurun_out/r02f_chain_full python scripts/probes/chain_time.py ffhq64 > gpurun_out/r02f_chain_ncu.log 2>&1
echo "ncu rc=$?"
