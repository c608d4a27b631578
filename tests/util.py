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


from oracle.chunked import chunked_oracle  # noqa: E402,F401  (full-size checker, shared with bench.py's parity block)
