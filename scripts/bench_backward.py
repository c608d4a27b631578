"""Training step of one Hopfield layer at BASELINE config 5's sizes on one B200: forward + backward through the layer's
public API (`loss.backward()`, main.py:35), tcgen05 forward and tcgen05 backward core (hop_bwd_tc.cu).

Per point: device ms of the forward, of the backward, and -- by differencing against a step whose gradient stops at
the association core's inputs -- nothing fancier than CUDA events; the per-kernel split comes from the ncu launch list
(`profiles/r02_bwd_launches.csv`).  Algorithmic FLOPs: F_fwd = 2 N P (D + Dv), F_bwd = 2 N P (3 D + 2 Dv);
fractions are against the measured bf16 peak (MEASURED_PEAKS.json).  Gradient parity of the same code path is
tests/test_backward_tc_gpu.py.
    python scripts/bench_backward.py [--points=N,P,D;...] > profiles/r02_backward.jsonl"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import hflayers  # noqa: E402

dev = torch.device("cuda", 0)
points = [(16384, 65536, 256), (16384, 65536, 512), (16384, 65536, 128), (8192, 262144, 256)]
for a in sys.argv[1:]:
    if a.startswith("--points="):
        points = [tuple(int(v) for v in t.split(",")) for t in a.split("=", 1)[1].split(";")]
steps = 5
peak = bench.peaks()["bf16_tflops"]
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
impl = hflayers._impl

for (n, p, d) in points:
    w, wv, wo = bench.make_state(p, d, 0)
    layer = bench.build_layer(hflayers.HopfieldLayer, p, d, 0, w, wv, wo).to(dev).train()
    layer.kernel_path = "f16"
    xh, _ = bench.make_queries(w, n)
    x = xh.to(dev).requires_grad_(True)
    gout = torch.randn(n, d, device=dev)
    path = impl.ops.resolve_train_path(layer.spec(), n, p)

    def fwd():
        return layer(x)

    for _ in range(2):
        layer.zero_grad(set_to_none=True); x.grad = None
        fwd().backward(gout)
    torch.cuda.synchronize()
    t_f, t_b, launches = [], [], []
    for _ in range(steps):
        layer.zero_grad(set_to_none=True); x.grad = None
        flush.zero_()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        l0 = impl.cabi.launch_count()
        e[0].record(); y = fwd(); e[1].record(); y.backward(gout); e[2].record()
        torch.cuda.synchronize()
        launches.append(impl.cabi.launch_count() - l0)
        t_f.append(e[0].elapsed_time(e[1])); t_b.append(e[1].elapsed_time(e[2]))
    ms_f, ms_b = sorted(t_f)[steps // 2], sorted(t_b)[steps // 2]
    f_fwd, f_bwd = 2.0 * n * p * 2 * d, 2.0 * n * p * 5 * d
    print(json.dumps({
        "rows": n, "patterns": p, "dim": d, "train_path": {1: "fp32_simt", 2: "tc_f16", 3: "tc_bf16"}[path],
        "ms_forward_incl_prep": ms_f, "ms_backward": ms_b, "launches_per_step": launches[-1],
        "bwd_algorithmic_tflops": f_bwd / (ms_b * 1e-3) / 1e12, "bwd_frac_of_measured_peak": f_bwd / (ms_b * 1e-3) / 1e12 / peak,
        "step_algorithmic_tflops": (f_fwd + f_bwd) / ((ms_f + ms_b) * 1e-3) / 1e12,
        "note": "backward = everything loss.backward() enqueues for the layer: stats / packs, the three tcgen05 core "
                "launches (8 executed contraction units for 5 algorithmic), the consistency fix, and the fp32 LayerNorm / "
                "projection / weight-gradient job kernels; forward includes the per-step pattern preparation"}), flush=True)
    del layer, x, w, wv, wo
    torch.cuda.empty_cache()
