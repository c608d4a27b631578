"""Run-to-run differences of every gradient of one tensor-core training step (same inputs)."""
import os, sys, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import hflayers
from oracle import hflayers_oracle as O
from util import randomize_, rel_l2
simt = '--simt-tails' in sys.argv
for (D, P, N) in [(256, 65536, 2048), (256, 4096, 8192), (512, 4096, 8192), (128, 8192, 4096), (256, 5000, 7001), (256, 65536, 4096)]:
    kw = dict(input_size=D, quantity=P, stored_pattern_as_static=True, state_pattern_as_static=True)
    ref = randomize_(O.HopfieldLayer(**kw), 95, 1.0)
    g = torch.Generator().manual_seed(96)
    with torch.no_grad():
        ref.lookup_weights.copy_(torch.randn(1, P, D, generator=g))
    dut = hflayers.HopfieldLayer(**kw); dut.load_state_dict(ref.state_dict()); dut = dut.cuda(); dut.kernel_path = "f16"
    x = torch.randn(N, D, generator=g).cuda(); w = torch.randn(N, D, generator=g).cuda()
    def run():
        dut.tune = dict(bwd_tails_simt=simt)
        dut.zero_grad(set_to_none=True)
        xd = x.clone().requires_grad_(True)
        (dut(xd) * w).sum().backward(); torch.cuda.synchronize()
        out = {k.replace("hopfield.", "").replace("association_core.", ""): v.grad.clone() for k, v in dut.named_parameters()}
        out["d_x"] = xd.grad.clone()
        return out
    base = run()
    worst = {k: 0.0 for k in base}
    for i in range(8):
        a = run()
        for k in base:
            worst[k] = max(worst[k], float(rel_l2(a[k], base[k])))
    print(json.dumps({"simt_tails": simt, "D": D, "P": P, "N": N, **{k: float(f"{v:.2e}") for k, v in worst.items()}}), flush=True)
