#!/bin/bash
# round 2, last 2-GPU pass: whole -m gpu suite + smoke on the final tree, config 3 weak / strong at 1 and 2 GPUs
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q -rs > gpurun_out/r02_last_tests.txt 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/r02_last_tests.txt
timeout 400 python __graft_entry__.py smoke > gpurun_out/r02_last_smoke.txt 2>&1; echo "smoke rc=$?"; tail -6 gpurun_out/r02_last_smoke.txt
: > gpurun_out/r02_train_ddp_g12.jsonl
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
for mode in "--global-batch 256" "--batch 32"; do
  timeout 400 $TR scripts/train_ddp.py --config ffhq64 $mode --epochs 2 --steps-per-epoch 20 --out gpurun_out/t2 2>/dev/null | grep '"config"' >> gpurun_out/r02_train_ddp_g12.jsonl
  timeout 400 python scripts/train_ddp.py --config ffhq64 $mode --epochs 2 --steps-per-epoch 20 --out gpurun_out/t1 2>/dev/null | grep '"config"' >> gpurun_out/r02_train_ddp_g12.jsonl
done
rm -rf gpurun_out/t1 gpurun_out/t2
python -c "
import json
for l in open('gpurun_out/r02_train_ddp_g12.jsonl'):
    d=json.loads(l); print(d['scaling'], d['n_gpus'], d['batch_per_gpu'], 'ms %.2f img/s %.0f'%(d['ms_per_step'], d['images_per_s']))"
