"""Where do two runs of the same tensor-core backward differ?  (rows of d lookup_weights = pattern rows of dV^ / dK^)"""
import os, sys, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import hflayers
from oracle import hflayers_oracle as O
from util import randomize_, rel_l2
for (D, P, N) in [(256, 65536, 2048)]:
    kw = dict(input_size=D, quantity=P, stored_pattern_as_static=True, state_pattern_as_static=True)
    ref = randomize_(O.HopfieldLayer(**kw), 95, 1.0)
    g = torch.Generator().manual_seed(96)
    with torch.no_grad():
        ref.lookup_weights.copy_(torch.randn(1, P, D, generator=g))
    dut = hflayers.HopfieldLayer(**kw); dut.load_state_dict(ref.state_dict()); dut = dut.cuda(); dut.kernel_path = "f16"
    x = torch.randn(N, D, generator=g).cuda(); w = torch.randn(N, D, generator=g).cuda()
    def run():
        dut.tune = dict(bwd_tails_simt=('--simt-tails' in sys.argv))
        dut.zero_grad(set_to_none=True)
        xd = x.clone().requires_grad_(True)
        (dut(xd) * w).sum().backward(); torch.cuda.synchronize()
        return dut.lookup_weights.grad[0].clone()
    base = run()
    for i in range(6):
        a = run()
        d = (a - base).norm(dim=1) / base.norm(dim=1).clamp_min(1e-30)
        bad = (d > 1e-5).nonzero().flatten()
        top = torch.topk(d, 8)
        print(json.dumps({"D": D, "P": P, "N": N, "rel_l2": float(rel_l2(a, base)), "rows_over_1e-5": int(bad.numel()),
                          "first_bad": bad[:12].tolist(), "last_bad": bad[-4:].tolist(),
                          "bad_mod64": sorted(set((bad % 64).tolist())), "bad_tiles64": sorted(set((bad // 64).tolist()))[:40], "top_rows": top.indices.tolist(), "top_vals": [round(float(v), 5) for v in top.values]}), flush=True)
