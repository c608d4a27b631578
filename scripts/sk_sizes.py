"""Device time of the fused forward (P = 64k, D = 256, T = 0) at several row counts, plain grid vs stream-K."""
import os, sys, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench, hflayers
dev = torch.device("cuda", 0)
p, d = 65536, 256
w, wv, wo = bench.make_state(p, d, 0)
layer = bench.build_layer(hflayers.HopfieldLayer, p, d, 0, w, wv, wo).to(dev)
layer.kernel_path = "f16"
xh, _ = bench.make_queries(w, 16384)
x = xh.to(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for n in (1024, 2048, 4096, 6144, 8192, 16384):
    row = {"rows": n}
    for sk in ("0", "1"):
        os.environ["HOP_TC_STREAMK"] = sk
        xs = x[:n].contiguous()
        with torch.no_grad():
            for _ in range(3):
                layer(xs)
            ts = []
            for _ in range(7):
                flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); layer(xs); e1.record(); torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
        row["ms_streamk" + sk] = sorted(ts)[3]
    print(json.dumps(row), flush=True)
