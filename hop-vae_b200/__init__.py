"""hop-vae_b200 -- B200-native (sm_100a) Hopfield retrieval for joejenb/Hop-VAE.

Holds only what the hot path needs: `csrc/` (hand-written CUDA kernels + the
C-ABI of include/hop_sm100.h), the ctypes binding, the PyTorch custom ops and
the host-side mirror of the reference's operator interface (`HopfieldLayer`).

The directory name has a hyphen (it is the name the build contract fixes), so
it is imported through `hflayers` (the reference-facing package) or
`load_package()` in `hflayers/_loader.py`, under the module name `hop_vae_b200`.
"""
from . import cabi
from .layer import HopfieldLayer
from . import ops, sharded, chain
from .chain import bottleneck_forward
from .sharded import PatternShardedHopfield, shard_range, batch_parallel_rows
from .ops import LayerSpec, HopfieldAssociation, prepare_patterns_op, retrieve_op, retrieve_backward_op, resolve_path

__all__ = ["HopfieldLayer", "LayerSpec", "HopfieldAssociation", "prepare_patterns_op", "retrieve_op",
           "retrieve_backward_op", "resolve_path", "cabi", "ops", "sharded", "PatternShardedHopfield", "shard_range",
           "batch_parallel_rows", "chain", "bottleneck_forward"]
