#!/bin/bash
# round 2, 8-GPU pass: sharded tests at 2 / 4 / 8 ranks, bench at N = 1, 2, 4, 8 (sharded sub-record at N > 1), host copy
# ceiling at 8 ranks, DDP training driver (config 3) at 4 and 8 GPUs
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_sharded_gpu.py -m gpu -q -rs > gpurun_out/r02_g8_sharded_tests.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_g8_sharded_tests.txt; tail -4 gpurun_out/r02_g8_sharded_tests.txt
for G in 8 4 2; do
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 2951$G"
  timeout 1200 $TR bench.py --gpus $G --steps 10 --warmup 3 --no-backward > gpurun_out/r02_g${G}_bench.json 2> gpurun_out/r02_g${G}_bench.err; echo "bench$G rc=$?"
  tail -c 300 gpurun_out/r02_g${G}_bench.err
done
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r02_g1_bench.json 2> gpurun_out/r02_g1_bench.err; echo "bench1 rc=$?"
TR8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531"
TR4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29532"
timeout 300 $TR8 scripts/host_copy_ceiling.py > gpurun_out/r02_g8_host_ceiling.json 2>&1; echo "ceiling8 rc=$?"
timeout 300 $TR4 scripts/host_copy_ceiling.py > gpurun_out/r02_g4_host_ceiling.json 2>&1; echo "ceiling4 rc=$?"
timeout 600 $TR8 scripts/train_ddp.py --config ffhq64 --epochs 2 --steps-per-epoch 20 --out gpurun_out/train_ddp8 > gpurun_out/r02_g8_train_ddp.json 2> gpurun_out/r02_g8_train_ddp.err; echo "train8 rc=$?"
timeout 600 $TR4 scripts/train_ddp.py --config ffhq64 --epochs 2 --steps-per-epoch 20 --out gpurun_out/train_ddp4 > gpurun_out/r02_g4_train_ddp.json 2> gpurun_out/r02_g4_train_ddp.err; echo "train4 rc=$?"
timeout 600 python scripts/train_ddp.py --config ffhq64 --epochs 2 --steps-per-epoch 20 --out gpurun_out/train_ddp1 > gpurun_out/r02_g1_train_ddp.json 2> gpurun_out/r02_g1_train_ddp.err; echo "train1 rc=$?"
rm -rf gpurun_out/train_ddp8 gpurun_out/train_ddp4 gpurun_out/train_ddp1
ls -la gpurun_out | tail -20
