"""Does the fused forward give the same output twice?  (all rows, several shapes)"""
import os, sys, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench, hflayers
dev = torch.device("cuda", 0)
for (P, D, T, N) in [(65536, 256, 0, 16384), (65536, 128, 0, 16384), (65536, 64, 0, 16384), (65536, 512, 0, 16384),
                     (65536, 1024, 0, 4096), (65536, 256, 1, 16384), (262144, 256, 0, 16384), (262144, 256, 3, 8192)]:
    w, wv, wo = bench.make_state(P, D, T)
    layer = bench.build_layer(hflayers.HopfieldLayer, P, D, T, w, wv, wo).to(dev)
    layer.kernel_path = "f16"
    xh, _ = bench.make_queries(w, N)
    x = xh.to(dev)
    with torch.no_grad():
        base = layer(x).clone()
        worst, bad_rows, blocks = 0.0, 0, set()
        for i in range(6):
            y = layer(x)
            d = (y - base).norm(dim=1) / base.norm(dim=1).clamp_min(1e-30)
            worst = max(worst, float(d.max()))
            bad = (d > 1e-5).nonzero().flatten()
            bad_rows = max(bad_rows, int(bad.numel()))
            blocks |= set((bad // 16 % 8).tolist())
    print(json.dumps({"P": P, "D": D, "T": T, "N": N, "worst_row_rel_diff": worst, "max_rows_over_1e-5": bad_rows,
                      "row_block_of_16_within_128": sorted(blocks)}), flush=True)
    del layer, x
    torch.cuda.empty_cache()
