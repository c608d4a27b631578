#!/bin/bash
# round 2, 2-GPU pass: sharded tests, re-run of the single-GPU cases fixed since, bench at N = 2 (sharded sub-record),
# host copy ceiling, DDP training driver
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_sharded_gpu.py -m gpu -q -rs > gpurun_out/r02_g2_sharded_tests.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_g2_sharded_tests.txt; tail -4 gpurun_out/r02_g2_sharded_tests.txt
timeout 1500 python -m pytest tests/test_parity_gpu.py tests/test_backward_tc_gpu.py -m gpu -q -s -k "stop_rule or d64_T2 or p256k_d256_T3 or p1m_d256_T3 or backward" > gpurun_out/r02_g2_fixed_tests.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_g2_fixed_tests.txt; tail -4 gpurun_out/r02_g2_fixed_tests.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 1200 $TR bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r02_g2_bench.json 2> gpurun_out/r02_g2_bench.err; echo "bench2 rc=$?"
tail -c 400 gpurun_out/r02_g2_bench.err
timeout 300 $TR scripts/host_copy_ceiling.py > gpurun_out/r02_g2_host_ceiling.json 2>&1; echo "ceiling rc=$?"
timeout 600 $TR scripts/train_ddp.py --config ffhq64 --epochs 2 --steps-per-epoch 20 --out gpurun_out/train_ddp > gpurun_out/r02_g2_train_ddp.json 2> gpurun_out/r02_g2_train_ddp.err; echo "train rc=$?"
rm -rf gpurun_out/train_ddp
