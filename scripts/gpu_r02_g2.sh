#!/bin/bash
# round 2, 2-GPU pass on the final tree: sharded tests, bench at N = 2 (sharded sub-record), other-width sweep points
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_sharded_gpu.py -m gpu -q -rs > gpurun_out/r02_g2_sharded_tests.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_g2_sharded_tests.txt; tail -4 gpurun_out/r02_g2_sharded_tests.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 1200 $TR bench.py --gpus 2 --steps 10 --warmup 3 --no-backward > gpurun_out/r02_g2_bench.json 2> gpurun_out/r02_g2_bench.err; echo "bench2 rc=$?"
tail -c 300 gpurun_out/r02_g2_bench.err
timeout 1500 python scripts/sweep_config5.py "--points=65536,1024,0;262144,1024,0;65536,1024,1;65536,128,0;262144,128,0;65536,64,0;65536,128,3;65536,64,1" > gpurun_out/r02z_sweep_other_widths.jsonl 2> gpurun_out/r02z_sweep_other.err; echo "sweep rc=$?"
