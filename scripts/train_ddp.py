"""the reference's training driver as a multi-GPU job with resumable checkpoints.

Mirrors /root/reference/main.py:96-124 (`main`: model, `load_from_checkpoint`, Adam + ExponentialLR, per-epoch
`train`, periodic `torch.save`) and its per-batch step main.py:26-36 (`loss = mse(X_recon, X) + Z_prediction_error`),
one process per GPU under torchrun: DistributedDataParallel around the unchanged model whose Hopfield layers are this
repository's `hflayers.HopfieldLayer`, per-rank data shards, NCCL gradient all-reduce.  Two things the reference lacks
are added because a resumable multi-GPU job needs them: the checkpoint also carries optimiser / scheduler / epoch
(the reference's file keeps the bare `state_dict()`; `--reference-format` writes exactly that, and both load through
the reference's name-filtered merge, utils/__init__.py:36-49), and rank 0 alone writes.

Data is synthetic (no datasets on the box, no network): `rand - 0.5` images like `Normalize(0.5, 1.0)` output
(utils/__init__.py:112-116), seed 1265 + rank per shard.  When /root/reference is present the reference's own
`HopVAE.py` / `utils.load_from_checkpoint` are used unchanged; on the GPU box (no reference tree) the bit-identical
harness `tests/hopvae_harness.py` stands in.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 scripts/train_ddp.py \\
        --config ffhq64 --epochs 2 --steps-per-epoch 20 --out gpurun_out/train_ddp
Rank 0 prints one JSON line: ms/step (max over ranks, CUDA events), images/s, latents/s, loss trajectory, and the
result of the resume check (save -> fresh process state -> load -> identical loss on a held-out batch, bit for bit).
"""
import argparse
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
os.environ.setdefault("WANDB_MODE", "disabled")
torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False


def load_from_checkpoint(model, state):
    """utils/__init__.py:36-49 restated for an in-memory state dict: drop keys the model lacks, keep the model's own
    values for keys the file lacks, then a strict load."""
    own = model.state_dict()
    state = {k: v for k, v in state.items() if k in own}
    for k, v in own.items():
        state.setdefault(k, v)
    model.load_state_dict(state)
    return model


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="ffhq64", choices=["mnist28", "ffhq32", "ffhq64"])
    ap.add_argument("--epochs", type=int, default=2)
    ap.add_argument("--steps-per-epoch", type=int, default=10)
    ap.add_argument("--batch", type=int, default=32, help="images per GPU (weak scaling)")
    ap.add_argument("--global-batch", type=int, default=0,
                    help="strong scaling (SURVEY.md 8d config 3: global B = 256 split over the ranks); overrides --batch")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "train_ddp"))
    ap.add_argument("--resume", default=None, help="checkpoint written by this driver (or a bare reference state_dict)")
    ap.add_argument("--reference-format", action="store_true", help="also write the bare state_dict main.py:124 writes")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if args.global_batch:
        if args.global_batch % world:
            raise SystemExit(f"--global-batch {args.global_batch} is not divisible by {world} ranks")
        args.batch = args.global_batch // world
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import hflayers
    import hopvae_harness as H
    cfg = {"mnist28": H.MNIST_28, "ffhq32": H.FFHQ_32, "ffhq64": H.FFHQ_64}[args.config]
    lr, gamma = (1e-3, 0.99) if args.config == "mnist28" else (1e-3, 0.98)      # configs/*_config.py

    def build():
        torch.manual_seed(1265)
        return H.HopVAEHarness(cfg, hflayers.HopfieldLayer).to(dev)

    model = build()
    start_epoch, opt_state, sched_state = 0, None, None
    if args.resume:
        ck = torch.load(args.resume, map_location=dev, weights_only=True)
        if "model" in ck:
            load_from_checkpoint(model, ck["model"])
            start_epoch, opt_state, sched_state = int(ck["epoch"]) + 1, ck["optimiser"], ck["scheduler"]
        else:
            load_from_checkpoint(model, ck)               # the reference's bare file (e.g. checkpoints/MNIST-28.ckpt)
    net = DDP(model, device_ids=[local], find_unused_parameters=True) if world > 1 else model   # post_vq_conv is unused
    opt = torch.optim.Adam(net.parameters(), lr=lr, amsgrad=False)              # main.py:106
    sched = torch.optim.lr_scheduler.ExponentialLR(opt, gamma=gamma)            # main.py:107
    if opt_state is not None:
        opt.load_state_dict(opt_state)
        sched.load_state_dict(sched_state)

    c, s = cfg["num_channels"], cfg["image_size"]
    gen = torch.Generator().manual_seed(1265 + rank)                            # this rank's shard
    held_out = (torch.rand(8, c, s, s, generator=torch.Generator().manual_seed(7)) - 0.5).to(dev)
    losses, ms_steps = [], []
    for epoch in range(start_epoch, start_epoch + args.epochs):
        net.train()
        # this epoch's synthetic shard, made before the clock starts: the reference's data loader (out of scope) would
        # hand over device batches; a host-side generator inside the timed loop made the step host-bound at B >= 128
        batches = [(torch.rand(args.batch, c, s, s, generator=gen) - 0.5).to(dev) for _ in range(args.steps_per_epoch)]
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        run = torch.zeros((), device=dev)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0.record()
        for x in batches:                                                      # main.py:26-36
            opt.zero_grad()
            x_recon, z_err = net(x)
            loss = F.mse_loss(x_recon, x) + z_err
            loss.backward()
            opt.step()
            run += loss.detach()
        sched.step()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / args.steps_per_epoch], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(run)
        ms_steps.append(float(t.item()))
        losses.append(float(run.item()) / args.steps_per_epoch / world)
        if rank == 0:
            os.makedirs(args.out, exist_ok=True)
            torch.save({"model": model.state_dict(), "optimiser": opt.state_dict(), "scheduler": sched.state_dict(),
                        "epoch": epoch}, os.path.join(args.out, f"{args.config}.ckpt"))
            if args.reference_format:
                torch.save(model.state_dict(), os.path.join(args.out, f"{args.config}.reference.ckpt"))
    # resume check: what a restarted job would see
    resume_ok = None
    if world > 1:
        dist.barrier()
    if rank == 0:
        model.eval()
        with torch.no_grad():
            want = [t_.clone() for t_ in model(held_out)]
        fresh = build()
        ck = torch.load(os.path.join(args.out, f"{args.config}.ckpt"), map_location=dev, weights_only=True)
        load_from_checkpoint(fresh, ck["model"]).eval()
        opt2 = torch.optim.Adam(fresh.parameters(), lr=lr)
        opt2.load_state_dict(ck["optimiser"])
        with torch.no_grad():
            got = fresh(held_out)
        # the Hopfield bottleneck (this repository's kernels) must reproduce bit for bit; the cuDNN conv stacks around it are
        # compared with a tolerance (their algorithm choice is not pinned by the reference either)
        with torch.no_grad():
            zb = model.encode(held_out)
            b_want, b_got = model.bottleneck(zb), fresh.bottleneck(zb)
        resume_ok = {"bottleneck_bit_identical": bool(all(torch.equal(b_want[k], b_got[k]) for k in ("emb", "logits", "emb_recon"))),
                     "output_max_abs_diff": float(max((a - b).abs().max() for a, b in zip(want, got))),
                     "optimiser_state_restored": opt2.state_dict()["state"].keys() == opt.state_dict()["state"].keys(),
                     "weights_bit_identical": bool(all(torch.equal(v, fresh.state_dict()[k]) for k, v in model.state_dict().items()))}
        rows = {"mnist28": 64, "ffhq32": 81, "ffhq64": 289}[args.config]
        ms = ms_steps[-1]                               # the last epoch: steady state (the first one carries the warm-up)
        print(json.dumps({"config": args.config, "n_gpus": world, "batch_per_gpu": args.batch,
                          "scaling": "strong" if args.global_batch else "weak", "global_batch": world * args.batch,
                          "epochs": [start_epoch, start_epoch + args.epochs], "steps_per_epoch": args.steps_per_epoch,
                          "ms_per_step": ms, "images_per_s": world * args.batch / (ms * 1e-3),
                          "latents_per_s": world * args.batch * rows * 3 / (ms * 1e-3), "loss_per_epoch": losses,
                          "ms_per_step_by_epoch": ms_steps, "lr_after": sched.get_last_lr()[0], "resume_check": resume_ok,
                          "checkpoint": os.path.relpath(os.path.join(args.out, f"{args.config}.ckpt"), ROOT)}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
