#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/r02_train_ddp_g8.jsonl
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518"
for mode in "--global-batch 256" "--batch 32"; do
  timeout 300 $TR scripts/train_ddp.py --config ffhq64 $mode --epochs 2 --steps-per-epoch 20 --out gpurun_out/t8 2>/dev/null | grep '"config"' >> gpurun_out/r02_train_ddp_g8.jsonl
done
rm -rf gpurun_out/t8
python -c "
import json
for l in open('gpurun_out/r02_train_ddp_g8.jsonl'):
    d=json.loads(l); print(d['scaling'], d['n_gpus'], d['batch_per_gpu'], 'ms %.2f img/s %.0f'%(d['ms_per_step'], d['images_per_s']))"
