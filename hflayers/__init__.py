"""`hflayers` -- the reference-facing package name.

joejenb/Hop-VAE does `from hflayers import HopfieldLayer` (/root/reference/HopVAE.py:7).
Putting this repository's root on `sys.path` makes that import resolve to the
B200-native layer: same constructor, forward signature and state-dict layout,
CUDA-only (sm_100a), no CPU fallback.  See INTEGRATION.md.
"""
from ._loader import load_package

_impl = load_package()
HopfieldLayer = _impl.HopfieldLayer
bottleneck_forward = _impl.bottleneck_forward      # HopVAE.forward's three retrievals + glue as one launch (evaluation)

__all__ = ["HopfieldLayer", "bottleneck_forward"]
