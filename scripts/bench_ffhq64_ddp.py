"""BASELINE config 3: FFHQ-64 Hop-VAE, synthetic 64x64 batch, forward + backward, batch-parallel (torchrun, one
rank per GPU, DistributedDataParallel; per-GPU batch 32 -> 9,248 query rows per GPU and layer, weak scaling).

  ours   the unchanged model (tests/hopvae_harness.py mirrors HopVAE.py) with `hflayers.HopfieldLayer` = sm_100a kernels
  eager  the same model with the oracle restatement of hflayers run on the SAME GPUs with PyTorch's stock kernels
         (what the reference does on a CUDA device; TF32 off)

Reports ms per training step (forward, loss as main.py:32-33, backward incl. DDP's gradient all-reduce), images/s
and retrieved latents/s (3 Hopfield layers x 289 rows per image), max over ranks, CUDA events.
    python -m torch.distributed.run --nproc-per-node G scripts/bench_ffhq64_ddp.py [steps]"""
import json
import os
import sys

import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch.nn.parallel import DistributedDataParallel as DDP

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(rank)
dev = torch.device("cuda", rank)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
import hflayers  # noqa: E402
import hopvae_harness as H  # noqa: E402
from oracle import hflayers_oracle as O  # noqa: E402  (baseline only)

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
B = 32
x = (torch.rand(B, 3, 64, 64, generator=torch.Generator().manual_seed(1265 + rank)) - 0.5).to(dev)


def build(cls):
    torch.manual_seed(1265)
    m = H.HopVAEHarness(H.FFHQ_64, cls).to(dev).train()
    return DDP(m, device_ids=[rank], find_unused_parameters=True) if world > 1 else m   # post_vq_conv is unused (as upstream)


def step_fn(model, opt):
    def f():
        opt.zero_grad(set_to_none=True)
        recon, aux = model(x)
        (F.mse_loss(recon, x) + aux).backward()
        opt.step()
    return f


def timed(fn):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / steps], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


res = {"config": "ffhq64_fwdbwd_batch_parallel", "n_gpus": world, "batch_per_gpu": B, "rows_per_layer_per_gpu": B * 289}
for tag, cls in (("ours", hflayers.HopfieldLayer), ("eager", O.HopfieldLayer)):
    model = build(cls)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    ms = timed(step_fn(model, opt))
    res[f"ms_step_{tag}"] = ms
    res[f"images_per_s_{tag}"] = world * B / (ms * 1e-3)
    res[f"latents_per_s_{tag}"] = world * B * 289 * 3 / (ms * 1e-3)
    del model, opt
if rank == 0:
    print(json.dumps(res), flush=True)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
