"""Host-side ceiling of the e2e (host-buffer) number at N ranks: every rank copies a pinned 16.8 MB buffer to its GPU and
another one back, one cudaMemcpyAsync per call on two streams, nothing else -- what `forward_host` cannot beat.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 scripts/host_copy_ceiling.py
Rank 0 prints one JSON line: aggregate GB/s per direction, and the latents/s that ceiling allows at D = 256."""
import json
import os
import time

import torch
import torch.distributed as dist

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
n, d = 16384, 256
hin = torch.randn(n, d).pin_memory()
hout = torch.empty(n, d).pin_memory()
din, dout = torch.empty(n, d, device=dev), torch.randn(n, d, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def step():
    with torch.cuda.stream(s1):
        din.copy_(hin, non_blocking=True)
    with torch.cuda.stream(s2):
        hout.copy_(dout, non_blocking=True)


for _ in range(20):
    step()
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
t0 = time.perf_counter()
reps = 0
while time.perf_counter() - t0 < 3.0:
    for _ in range(50):
        step()
    s1.synchronize(); s2.synchronize()
    reps += 50
torch.cuda.synchronize()
dt = time.perf_counter() - t0
rate = torch.tensor([reps * hin.numel() * 4 / dt / 1e9], device=dev, dtype=torch.float64)
if world > 1:
    lo = rate.clone(); dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(rate, op=dist.ReduceOp.SUM)
else:
    lo = rate
if rank == 0:
    agg = float(rate.item())
    print(json.dumps({"ranks": world, "bytes_per_copy": hin.numel() * 4, "aggregate_gbs_per_direction": agg,
                      "slowest_rank_gbs": float(lo.item()),
                      "e2e_ceiling_latents_per_s_d256": agg * 1e9 / (d * 4),
                      "method": "pinned cudaMemcpyAsync H2D + D2H concurrently, one copy per call, 3 s"}), flush=True)
if world > 1:
    dist.destroy_process_group()
