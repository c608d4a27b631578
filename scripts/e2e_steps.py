"""Per-step wall/device time of HopfieldLayer.forward_host for a few chunk schedules (diagnostic)."""
import os, sys, json, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench, hflayers
dev = torch.device("cuda", 0)
p, d, n = 65536, 256, 16384
w, wv, wo = bench.make_state(p, d, 0)
layer = bench.build_layer(hflayers.HopfieldLayer, p, d, 0, w, wv, wo).to(dev)
layer.kernel_path = "f16"
xh, _ = bench.make_queries(w, n)
xh = xh.contiguous().pin_memory()
out = torch.empty((n, d)).pin_memory()
SCHEDS = [(1, 3, 3, 1), (1, 2, 4, 1), (1, 6, 1), 4]
if len(sys.argv) > 1:
    SCHEDS = [tuple(float(v) for v in a.split(',')) for a in sys.argv[1:]]
for sched in SCHEDS:
    ts, hs = [], []
    for i in range(12):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(); layer.forward_host(xh, out, chunks=sched); e1.record()
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        ts.append(round(e0.elapsed_time(e1), 3)); hs.append(round((t1 - t0) * 1e3, 3))
    print(json.dumps({"chunks": sched, "device_ms_median": sorted(ts[2:])[len(ts[2:]) // 2], "device_ms": ts, "host_enqueue_ms": hs}), flush=True)
