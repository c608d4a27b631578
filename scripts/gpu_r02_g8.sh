#!/bin/bash
# round 2, 8-GPU pass on the final tree: sharded tests at 2 / 4 / 8 ranks, bench at N = 8 and 4 (sharded sub-record),
# config 3 (FFHQ-64 under DDP) weak scaling at 8 and strong scaling (global batch 256) at 1 / 2 / 4 / 8 GPUs
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_sharded_gpu.py -m gpu -q -rs > gpurun_out/r02_g8_sharded_tests.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r02_g8_sharded_tests.txt; tail -3 gpurun_out/r02_g8_sharded_tests.txt
for G in 8 4; do
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 2951$G"
  timeout 1200 $TR bench.py --gpus $G --steps 10 --warmup 3 --no-backward > gpurun_out/r02_final_g${G}_bench.json 2> gpurun_out/r02_final_g${G}_bench.err; echo "bench$G rc=$?"
done
: > gpurun_out/r02_train_ddp_strong.jsonl
for G in 8 4 2; do
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 2953$G"
  timeout 600 $TR scripts/train_ddp.py --config ffhq64 --global-batch 256 --epochs 2 --steps-per-epoch 20 --out gpurun_out/tddp$G 2> gpurun_out/r02_train_strong_$G.err | grep '"config"' >> gpurun_out/r02_train_ddp_strong.jsonl; echo "strong$G rc=$?"
done
timeout 600 python scripts/train_ddp.py --config ffhq64 --global-batch 256 --epochs 2 --steps-per-epoch 20 --out gpurun_out/tddp1 2> gpurun_out/r02_train_strong_1.err | grep '"config"' >> gpurun_out/r02_train_ddp_strong.jsonl; echo "strong1 rc=$?"
TR8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541"
timeout 600 $TR8 scripts/train_ddp.py --config ffhq64 --epochs 2 --steps-per-epoch 20 --out gpurun_out/tddp8w 2> gpurun_out/r02_train_weak_8.err | grep '"config"' > gpurun_out/r02_train_ddp_weak8.jsonl; echo "weak8 rc=$?"
rm -rf gpurun_out/tddp*
cut -c1-220 gpurun_out/r02_train_ddp_strong.jsonl
