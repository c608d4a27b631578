"""BASELINE config 5 on one GPU: the synthetic retrieval sweep P in {64k, 256k, 1M} x D in {256, 512} x T in {0, 1, 3}
(D = 1024 points through --points=P,1024,T;...), N = 16,384 query rows, f16 operands / f32 accumulate.

Per point: device time of `HopfieldLayer.forward` (CUDA events, L2 flushed between steps), algorithmic TFLOP/s
(2 N P [(2T+1) D + Dv]) against the measured bf16 peak, and PARITY at full size against the
chunked fp32/fp64 CPU restatement on a 256-row query subsample (rel-L2 of the layer output; nearest-pattern index on
every row the fp32 oracle itself decides -- gap > 1e-4 nats at T = 0, where the layer's exact-index path applies; with
retrieve updates the iterated state carries the earlier passes' 16-bit rounding and the bar is the 2e-2 of that noise),
plus how many of the 16,384 rows failed the on-chip top-1 / top-2 certificate and were re-scored in fp32.
    python scripts/sweep_config5.py [--quick] > profiles/rNN_sweep_config5.jsonl"""
import json
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import hflayers  # noqa: E402

dev = torch.device("cuda", 0)
N = 16384
SUB = 256
quick = "--quick" in sys.argv
points = [(p, d, t) for d in (256, 512) for p in (65536, 262144, 1048576) for t in (0, 1, 3)]
if quick:
    points = [(65536, 256, 0), (65536, 512, 1)]
for a in sys.argv[1:]:
    if a.startswith("--points="):     # e.g. --points=262144,256,1;262144,256,3
        points = [tuple(int(v) for v in t.split(",")) for t in a.split("=", 1)[1].split(";")]
peak = bench.peaks()["bf16_tflops"]
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


sys.path.insert(0, os.path.join(ROOT, "tests"))
from util import chunked_oracle as cpu_oracle  # noqa: E402  (the checker)


for (p, d, t) in points:
    w, wv, wo = bench.make_state(p, d, t)
    layer = bench.build_layer(hflayers.HopfieldLayer, p, d, t, w, wv, wo).to(dev)
    layer.kernel_path = "f16"
    xh, _ = bench.make_queries(w, N)
    x = xh.to(dev)
    with torch.no_grad():
        for _ in range(3):
            y = layer(x)
        torch.cuda.synchronize()
        ts = []
        for _ in range(5):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); y = layer(x); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ms = sorted(ts)[len(ts) // 2]
        got, got_arg = y[:SUB].cpu(), layer.last_argmax[:SUB].cpu().long()
        ops = hflayers._impl.ops
        ops.count_repairs = True
        layer(x)
        ops.count_repairs = False
    want, arg, gap = cpu_oracle(xh[:SUB], w, wv, wo, t, 1.0 / math.sqrt(d))
    rel = float((got - want).norm() / want.norm())
    decided = gap > (1e-4 if t == 0 else 2e-2)
    fl = bench.fwd_flops(N, p, d, d, t)
    res = {"patterns": p, "dim": d, "update_steps": t, "rows": N, "ms": ms, "latents_per_s": N / (ms * 1e-3),
           "algorithmic_tflops": fl / (ms * 1e-3) / 1e12, "frac_of_measured_peak": fl / (ms * 1e-3) / 1e12 / peak,
           "parity_rows": SUB, "rel_l2": rel, "argmax_decided_rows": int(decided.sum()),
           "argmax_mismatch": int((got_arg[decided] != arg[decided]).sum()), "rows_rescored_fp32": ops.last_repair_rows,
           "rows_failed_certificate": ops.last_flagged_rows}
    if ops.last_repair_rows > N // 100:
        # a diffuse retrieval (the iteration left the planted pattern's basin): the exact index costs an fp32 score pass
        # over the flagged rows; the tensor-core pass alone, for the split of the two
        layer.exact_argmax = False
        with torch.no_grad():
            layer(x)
            torch.cuda.synchronize()
            ts = []
            for _ in range(5):
                flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); layer(x); e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
        ms2 = sorted(ts)[len(ts) // 2]
        res["ms_without_exact_index"] = ms2
        res["frac_without_exact_index"] = fl / (ms2 * 1e-3) / 1e12 / peak
    print(json.dumps(res), flush=True)
    del layer, x, w, wv, wo
    torch.cuda.empty_cache()
