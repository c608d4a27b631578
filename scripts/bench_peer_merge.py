"""Micro-benchmark of the pattern-sharded exchange alone (torchrun, one rank per GPU): the fused peer-memory
kernel (`hop_sm100_peer_merge`) against the NCCL form (all_gather of the row statistics + merge kernel +
all_reduce).  CUDA events on the launching stream, max over ranks.
    python -m torch.distributed.run --nproc-per-node G scripts/bench_peer_merge.py [N] [W]"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dev = torch.device("cuda", rank)
dist.init_process_group("nccl", device_id=dev)
import hflayers  # noqa: E402

impl = hflayers._impl
N = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
W = int(sys.argv[2]) if len(sys.argv) > 2 else 256
g = torch.Generator(device=dev).manual_seed(rank)
o = torch.randn(N, W, device=dev, generator=g)
lse = torch.randn(N, device=dev, generator=g)
rowmax = lse - 1.0
arg = torch.randint(0, 1000, (N,), device=dev, dtype=torch.int32, generator=g)


def timed(fn, iters=30, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    ts = []
    for _ in range(iters):
        dist.barrier()   # ranks start together (the barrier's own kernel is outside the events)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    t = torch.tensor([sorted(ts)[len(ts) // 2]], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()) * 1e3


ex = impl.sharded.PeerExchange(N, W, None, dev)
so, sstats, sarg = ex.local_slots()
so.copy_(o); sstats[0].copy_(lse); sstats[1].copy_(rowmax); sarg.copy_(arg)
start, count = impl.batch_parallel_rows(N, world, rank)
res = {"n": N, "w": W, "world": world}
res["peer_all_us"] = timed(lambda: ex.merge(start, count, True))
res["peer_scatter_us"] = timed(lambda: ex.merge(start, count, False))
got = ex.merge(start, count, True)[0].clone()


class _L:  # minimal stand-in to reuse PatternShardedHopfield._exchange
    pass


sh = impl.PatternShardedHopfield.__new__(impl.PatternShardedHopfield)
sh.group, sh.world, sh.rank, sh.backend, sh.scatter_rows = None, world, rank, impl.sharded.CudaBackend(), False
oc = o.clone()


def coll():
    oc.copy_(o)
    return sh._exchange(oc, lse, rowmax, arg)


res["collective_us_incl_16MB_copy"] = timed(coll)
res["copy_us"] = timed(lambda: oc.copy_(o))
want = coll()[0]
res["max_abs_diff"] = float((got - want).abs().max())
bytes_pull = count * W * 4 * (world - 1) + 0.0
res["peer_pull_GBs"] = bytes_pull / (res["peer_scatter_us"] * 1e-6) / 1e9
if rank == 0:
    print(json.dumps(res), flush=True)
dist.barrier()
dist.destroy_process_group()
