"""In-situ kernel times of the graph-captured training step of the three-layer bottleneck (CUPTI through
torch.profiler: no ncu serialisation, warm caches).  Usage: python scripts/prof_train_graph.py [MNIST_28|FFHQ_32|FFHQ_64]"""
import os
import sys
from collections import defaultdict

import torch
from torch.profiler import ProfilerActivity, profile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import hflayers  # noqa: E402
import hopvae_harness as H  # noqa: E402

cfgname = sys.argv[1] if len(sys.argv) > 1 else "MNIST_28"
dev = torch.device("cuda", 0)
torch.manual_seed(1265)
cfg = getattr(H, cfgname)
dut = H.HopVAEHarness(cfg, hflayers.HopfieldLayer).to(dev).train()
R = cfg["representation_dim"]
zz = torch.randn(32, R * R, cfg["embedding_dim"], device=dev).requires_grad_(True)


def step():
    dut.zero_grad(set_to_none=True)
    out = dut.bottleneck(zz)
    (out["emb_recon"].square().mean() + out["emb_recon_loss"]).backward()


s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(3):
        step()
torch.cuda.current_stream().wait_stream(s)
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    step()
for _ in range(5):
    g.replay()
torch.cuda.synchronize()
REPS = 10
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(REPS):
        g.replay()
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
ev.sort(key=lambda e: e.time_range.start)
tot = defaultdict(float)
cnt = defaultdict(int)
for e in ev:
    tot[e.name] += e.time_range.elapsed_us()
    cnt[e.name] += 1
span = (ev[-1].time_range.end - ev[0].time_range.start) / REPS
busy = sum(tot.values()) / REPS
print(f"{cfgname}: {len(ev) // REPS} kernels per step, span {span:.1f} us, kernel-busy {busy:.1f} us")
for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:24]:
    print(f"{v / REPS:8.1f} us  x{cnt[k] // REPS:<3d} {k[:110]}")
# the sequence of one step with gaps
one = ev[: len(ev) // REPS]
print("--- sequence (start offset us, duration us, name)")
t0 = one[0].time_range.start
for e in one:
    print(f"{e.time_range.start - t0:8.1f} {e.time_range.elapsed_us():7.1f}  {e.name[:90]}")
