"""Shared helpers for the parity tests."""
import math
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

# Parity bars of BASELINE.json's north_star
REL_L2_BAR = 1e-3          # retrieved latents / layer outputs
RECON_MSE_REL_BAR = 5e-3   # reconstruction MSE, relative


def rel_l2(a, b) -> float:
    a = torch.as_tensor(a, dtype=torch.float64).cpu()
    b = torch.as_tensor(b, dtype=torch.float64).cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-300))


def load_golden():
    return np.load(os.path.join(GOLDEN, "mnist28_golden.npz"))


def load_sample_golden():
    """BASELINE config 4: the reference's own HopVAE.sample() run in the build container
    (tests/golden/make_golden_sample.py)."""
    return np.load(os.path.join(GOLDEN, "pixelcnn_sample_golden.npz"))


def load_ckpt():
    return torch.load(os.path.join(GOLDEN, "MNIST-28.ckpt"), map_location="cpu", weights_only=True)


def layer_state(sd, prefix):
    return {k[len(prefix) + 1:]: v for k, v in sd.items() if k.startswith(prefix + ".")}


def randomize_(layer, seed: int, scale: float = 1.0):
    """Give every parameter of a HopfieldLayer O(1) non-trivial values (LN affine != identity)."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, p in layer.named_parameters():
            if "norm_" in name and name.endswith("weight"):
                p.copy_(1.0 + 0.3 * torch.randn(p.shape, generator=g))
            elif "norm_" in name and name.endswith("bias"):
                p.copy_(0.2 * torch.randn(p.shape, generator=g))
            elif name.endswith("bias"):
                p.copy_(0.1 * torch.randn(p.shape, generator=g))
            elif "lookup" in name:
                p.copy_(scale * torch.randn(p.shape, generator=g))
            else:
                p.copy_(torch.randn(p.shape, generator=g) / (p.shape[-1] ** 0.5))
    return layer


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
