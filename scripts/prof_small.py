"""Kernel-level and host-level breakdown of the three-layer bottleneck at the MNIST-28 shape (N = 2048, P = 512,
D = 64/64/3): run plain for a cProfile of the host path, or under `ncu --metrics gpu__time_duration.sum` for
the launch list.  Usage: python scripts/prof_small.py [fwd|fwdbwd] [iters]"""
import cProfile
import io
import os
import pstats
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import hflayers  # noqa: E402
import hopvae_harness as H  # noqa: E402

mode = sys.argv[1] if len(sys.argv) > 1 else "fwd"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
cfgname = sys.argv[3] if len(sys.argv) > 3 else "MNIST_28"
dev = torch.device("cuda", 0)
torch.manual_seed(1265)
cfg = getattr(H, cfgname)
dut = H.HopVAEHarness(cfg, hflayers.HopfieldLayer).to(dev)
R = cfg["representation_dim"]
z = torch.randn(32, R * R, cfg["embedding_dim"], device=dev)


def fwd():
    with torch.no_grad():
        dut.bottleneck(z)


zz = z.clone().requires_grad_(True)


def fwdbwd():
    dut.zero_grad(set_to_none=True)
    out = dut.bottleneck(zz)
    (out["emb_recon"].square().mean() + out["emb_recon_loss"]).backward()


if mode == "fwd":
    dut.eval()
    fn = fwd
else:
    dut.train()
    fn = fwdbwd
for _ in range(3):
    fn()
torch.cuda.synchronize()
torch.cuda.nvtx.range_push("measured")
pr = cProfile.Profile()
pr.enable()
for _ in range(iters):
    fn()
pr.disable()
torch.cuda.synchronize()
torch.cuda.nvtx.range_pop()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(35)
print(s.getvalue()[:6000])
