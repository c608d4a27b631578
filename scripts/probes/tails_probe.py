"""Run-to-run differences of the tensor-core backward's gradients (same inputs, fp32 tails): which of dq / dK^ / dV^ moves."""
import os, sys, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import hflayers
from oracle import hflayers_oracle as O
from util import randomize_, rel_l2
for (D, P, N) in [(256, 5000, 7001), (256, 4096, 8192), (256, 65536, 2048), (128, 8192, 4096)]:
    kw = dict(input_size=D, quantity=P, stored_pattern_as_static=True, state_pattern_as_static=True)
    ref = randomize_(O.HopfieldLayer(**kw), 95, 1.0)
    g = torch.Generator().manual_seed(96)
    with torch.no_grad():
        ref.lookup_weights.copy_(torch.randn(1, P, D, generator=g))
    dut = hflayers.HopfieldLayer(**kw); dut.load_state_dict(ref.state_dict()); dut = dut.cuda(); dut.kernel_path = "f16"
    x = torch.randn(N, D, generator=g).cuda(); w = torch.randn(N, D, generator=g).cuda()
    def run():
        dut.tune = dict(bwd_tails_simt=True)
        dut.zero_grad(set_to_none=True)
        xd = x.clone().requires_grad_(True)
        (dut(xd) * w).sum().backward(); torch.cuda.synchronize()
        out = {k: v.grad.clone() for k, v in dut.named_parameters()}
        out["d_x"] = xd.grad.clone()
        return out
    base = run()
    keys = {"d_x": "dq", "hopfield.norm_stored_pattern.weight": "dK", "hopfield.association_core.in_proj_bias": "dV",
            "hopfield.association_core.out_proj.weight": "fwd"}
    worst = {v: 0.0 for v in keys.values()}
    for i in range(8):
        a = run()
        for k, tag in keys.items():
            worst[tag] = max(worst[tag], float(rel_l2(a[k], base[k])))
    print(json.dumps({"D": D, "P": P, "N": N, "worst_run_to_run": worst}), flush=True)
