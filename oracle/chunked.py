"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py) -- full-size checker for BASELINE config 5.

The synthetic sweep runs N = 16,384 queries against up to 1M patterns: the N x P score matrix of the
eager oracle (`hflayers_oracle.py`) does not fit a test's time budget there.  This restates the SAME arithmetic --
hflayers `hopfield_core_forward` as used at /root/reference/HopVAE.py:149-170 (LN_k / LN_q / LN_v, V-projection,
beta-scaled scores, softmax, T retrieve updates without re-normalisation, read-out, out-projection)
-- with a P-chunked online softmax on a row subsample: fp32 contractions like the reference's `torch.bmm`, fp64 merge
of the chunks.  Pinned to `hflayers_oracle.HopfieldLayer` in tests/test_oracle_pins.py.  Importers: tests/ and the
`parity` block of bench.py (the checker, outside every timed region).
"""
import math

import torch


def chunked_oracle(x, w, wv, wo, t, beta, chunk=65536):
    """Full-size checker for the config-5 sweep: hflayers semantics restated with a P-chunked
    online softmax (fp32 contractions, fp64 merge); LN affine = identity, bias-free V / out projections.
    Returns (layer output, nearest-pattern index of the last pass, its top-1/top-2 score gap)."""
    ln = torch.nn.functional.layer_norm
    d = w.shape[1]
    k = ln(w, (d,))
    v = ln(w, (d,)) @ wv.t()
    q = beta * ln(x, (d,))
    arg = None
    for it in range(t + 1):
        last = it == t
        b = v if last else k
        m = torch.full((x.shape[0],), -math.inf, dtype=torch.float64)
        l = torch.zeros(x.shape[0], dtype=torch.float64)
        acc = torch.zeros(x.shape[0], b.shape[1], dtype=torch.float64)
        best = torch.full((x.shape[0],), -math.inf)
        second = torch.full((x.shape[0],), -math.inf)
        arg = torch.zeros(x.shape[0], dtype=torch.long)
        for c0 in range(0, w.shape[0], chunk):
            s = q @ k[c0:c0 + chunk].t()
            top = s.topk(2, dim=-1)
            upd = top.values[:, 0] > best
            second = torch.where(upd, torch.maximum(best, top.values[:, 1]), torch.maximum(second, top.values[:, 0]))
            arg = torch.where(upd, top.indices[:, 0] + c0, arg)
            best = torch.where(upd, top.values[:, 0], best)
            s = s.double()
            mn = torch.maximum(m, s.max(-1).values)
            pexp = torch.exp(s - mn[:, None])
            corr = torch.exp(m - mn)
            l = l * corr + pexp.sum(-1)
            acc = acc * corr[:, None] + pexp @ b[c0:c0 + chunk].double()
            m = mn
        r = (acc / l[:, None]).float()
        if not last:
            q = beta * r
    return r @ wo.t(), arg, best - second
