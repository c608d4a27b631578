"""Debug: per-tile timeline of CTA (0,0) of the fused forward (needs `python hop-vae_b200/build.py --trace`).
    HOP_SM100_LIB=hop-vae_b200/libhop_sm100_trace.so python scripts/tc_timeline.py [workload]
"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import hflayers  # noqa: E402

impl = hflayers._impl
wl = sys.argv[1] if len(sys.argv) > 1 else "sweep_p64k_d256"
n, p, d, t = bench.WORKLOADS[wl]
w, wv, wo = bench.make_state(p, d, t)
layer = bench.build_layer(hflayers.HopfieldLayer, p, d, t, w, wv, wo).cuda()
layer.kernel_path = "f16"
x, _ = bench.make_queries(w, n)
x = x.cuda()
with torch.no_grad():
    for _ in range(3):
        layer(x)
    torch.cuda.synchronize()
lib = impl.cabi.load()
R, T, E = 9, 64, 8
buf = np.zeros(R * T * E, dtype=np.int64)
rc = lib.hop_sm100_debug_trace(buf.ctypes.data_as(C.POINTER(C.c_longlong)), buf.size)
assert rc == 0, rc
tr = buf.reshape(R, T, E)
t0 = tr[0, 0, 0]
names = {0: ["j", "qk_issue0", "qk_issued", "p_full", "pv_issued"],
         1: ["wait_s", "got_s", "ldtm", "pairbar", "rescale", "exp", "arrive"]}
print("MMA thread (cycles rel. to first event; deltas)")
for j in range(0, 40):
    a = tr[0, j, :4] - t0
    print(f"  tile {j:2d}: qk_issue_start {a[0]:7d} issued +{a[1]-a[0]:5d}  wait_pfull +{a[2]-a[1]:5d}  pv_issue +{a[3]-a[2]:5d}   period {tr[0,j,0]-tr[0,j-1,0] if j else 0}")
print("softmax warps: per tile, cycles from S-ready (MMA commit seen by the earliest warp) to each warp's p_full arrive")
for j in range(20, 40):
    got = tr[1:9, j, 1]
    arr = tr[1:9, j, 6]
    base = got.min()
    print(f"  tile {j:2d}: got_s-min {got - base}  arrive-min_got {arr - base}  MMA saw p_full at {tr[0, j, 2] - base}")
for role in range(1, 9):
    print(f"softmax warp {role + 1} (quarter {(role + 1) & 3}, half {(role - 1) >> 2}) lane 0")
    for j in range(24, 34):
        a = tr[role, j, :7] - t0
        dl = np.diff(a)
        print(f"  tile {j:2d}: start {a[0]:7d}  waitS {dl[0]:5d} ldtm {dl[1]:4d} max+bar {dl[2]:5d} resc {dl[3]:5d} exp {dl[4]:5d} st+arr {dl[5]:4d}  period {tr[role,j,0]-tr[role,j-1,0] if j else 0}")
