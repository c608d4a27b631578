"""One tensor-core training step (for compute-sanitizer)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import hflayers
D, P, N = 256, 4096, 8192
dut = hflayers.HopfieldLayer(input_size=D, quantity=P, stored_pattern_as_static=True, state_pattern_as_static=True).cuda()
with torch.no_grad():
    dut.lookup_weights.copy_(torch.randn(1, P, D))
dut.kernel_path = "f16"
dut.tune = dict(bwd_tails_simt=("--simt-tails" in sys.argv))
x = torch.randn(N, D, device="cuda", requires_grad=True)
w = torch.randn(N, D, device="cuda")
(dut(x) * w).sum().backward()
torch.cuda.synchronize()
print("ok", float(x.grad.norm()))
