"""Which launch forms send rows to the fp32 repair on config 5's recipe (expected: none)."""
import os, sys, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench, hflayers
ops = hflayers._impl.ops
dev = torch.device("cuda", 0)
P, D = 262144, 256
N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
for T, tune in [(0, {}), (0, {"splits": 16}), (0, {"splits": 64}), (1, {}), (2, {}), (3, {}), (3, {"plain_updates": True}),
                (3, {"splits": 16}), (3, {"splits": 32})]:
    w, wv, wo = bench.make_state(P, D, T)
    layer = bench.build_layer(hflayers.HopfieldLayer, P, D, T, w, wv, wo).to(dev)
    layer.kernel_path = "f16"
    layer.tune = dict(tune)
    xh, planted = bench.make_queries(w, N)
    ops.count_repairs = True
    with torch.no_grad():
        layer(xh.to(dev))
        idx = layer.last_argmax.cpu().long()
    ops.count_repairs = False
    print(json.dumps({"T": T, "tune": tune, "N": N, "rescored": ops.last_repair_rows,
                      "planted_found": int((idx == planted).sum())}), flush=True)
    del layer
    torch.cuda.empty_cache()
