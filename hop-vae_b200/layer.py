"""`HopfieldLayer` -- host-side mirror of the reference's operator interface.

Same constructor surface, forward signature and state-dict layout as the
third-party `hflayers.HopfieldLayer` that joejenb/Hop-VAE imports
(/root/reference/HopVAE.py:7, constructed at HopVAE.py:149-170, called at
HopVAE.py:193,215,216,234,255,257,271), so `HopVAE.py`, `main.py` and
`checkpoints/MNIST-28.ckpt` work unchanged.  The arithmetic runs in the sm_100a
kernels behind `include/hop_sm100.h`; this file only holds parameters, builds
the op description and caches the prepared pattern operands.

State-dict keys (pinned by the shipped checkpoint):
    lookup_weights                                   (1, quantity, input_size)
    hopfield.association_core.in_proj_weight / _bias  (rows, input_size) / (rows)
    hopfield.association_core.out_proj.weight / .bias
    hopfield.norm_stored_pattern.weight / .bias
    hopfield.norm_state_pattern.weight / .bias
    hopfield.norm_pattern_projection.weight / .bias
"""
from __future__ import annotations

import math
import os
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn as nn
from torch import Tensor

from . import cabi
from .ops import (LayerSpec, HopfieldAssociation, prepare_patterns, prepare_patterns_op, retrieve, retrieve_op,
                  resolve_path, _eager, _PATTERN_SIDE, _QUERY_SIDE, _PARAM_ORDER)

_TUNE_NAMES = {"streamk": (cabi.TUNE_STREAMK_ON, cabi.TUNE_STREAMK_OFF), "multi_launch": (cabi.TUNE_MULTI_ON, cabi.TUNE_MULTI_OFF),
               "pair": (cabi.TUNE_PAIR_ON, cabi.TUNE_PAIR_OFF), "rows128": (cabi.TUNE_ROWS128, 0),
               "plain_updates": (cabi.TUNE_PLAIN_UPDATES, 0), "bwd_simt": (cabi.TUNE_BWD_SIMT, 0),
               "prep_simt": (cabi.TUNE_PREP_SIMT, 0), "bwd_tails_simt": (cabi.TUNE_BWD_TAILS_SIMT, 0)}

_NOT_BUILT = dict(num_heads=1, normalize_hopfield_space=False, normalize_hopfield_space_affine=False,
                  pattern_projection_as_connected=False, association_activation=None, dropout=0.0,
                  concat_bias_pattern=False, add_zero_association=False, batch_first=True)


class _AssociationCore(nn.Module):
    """Parameter container named like upstream `HopfieldCore` (in_proj_* and out_proj)."""

    def __init__(self, d_in: int, d_assoc: int, d_v: int, d_out: int, q_static: bool, k_static: bool,
                 v_static: bool, bias: bool, out_projection: bool):
        super().__init__()
        self.rows_q = 0 if q_static else d_assoc
        self.rows_k = 0 if k_static else d_assoc
        self.rows_v = 0 if v_static else d_v
        rows = self.rows_q + self.rows_k + self.rows_v
        if rows:
            self.in_proj_weight = nn.Parameter(torch.empty(rows, d_in))
            self.in_proj_bias = nn.Parameter(torch.empty(rows)) if bias else None
        else:
            self.register_parameter("in_proj_weight", None)
            self.register_parameter("in_proj_bias", None)
        self.out_proj = nn.Linear(d_v, d_out, bias=bias) if out_projection else None
        self.reset_parameters()

    def reset_parameters(self) -> None:
        if self.in_proj_weight is not None:
            nn.init.normal_(self.in_proj_weight, mean=0.0, std=0.02)
        if self.in_proj_bias is not None:
            nn.init.zeros_(self.in_proj_bias)
        if self.out_proj is not None:
            nn.init.normal_(self.out_proj.weight, mean=0.0, std=0.02)
            if self.out_proj.bias is not None:
                nn.init.zeros_(self.out_proj.bias)

    def split(self) -> Dict[str, Optional[Tensor]]:
        """Views of the q / k / v slices of in_proj (order q, k, v; static ones absent)."""
        out: Dict[str, Optional[Tensor]] = dict(wq=None, bq=None, wk=None, bk=None, wv=None, bv=None)
        o = 0
        total = self.rows_q + self.rows_k + self.rows_v
        for tag, rows in (("q", self.rows_q), ("k", self.rows_k), ("v", self.rows_v)):
            if rows == total and rows:
                # the only projection (Hop-VAE's layers: q and k static): hand out the parameter itself -- a slice
                # would add a SliceBackward (zero-fill + copy of the full gradient) per parameter and step
                out["w" + tag], out["b" + tag] = self.in_proj_weight, self.in_proj_bias
            elif rows:
                out["w" + tag] = self.in_proj_weight[o:o + rows]
                out["b" + tag] = None if self.in_proj_bias is None else self.in_proj_bias[o:o + rows]
                o += rows
        return out


class _Hopfield(nn.Module):
    """Container named like upstream `Hopfield` (three LayerNorms + association core)."""

    def __init__(self, d_in, d_assoc, d_v, d_out, norms, statics, bias, out_projection):
        super().__init__()
        (n_k, a_k, n_q, a_q, n_v, a_v) = norms
        (k_static, q_static, v_static) = statics
        self.association_core = _AssociationCore(d_in, d_assoc, d_v, d_out, q_static, k_static, v_static, bias,
                                                 out_projection)
        self.norm_stored_pattern = nn.LayerNorm(d_in, elementwise_affine=a_k) if n_k else None
        self.norm_state_pattern = nn.LayerNorm(d_in, elementwise_affine=a_q) if n_q else None
        self.norm_pattern_projection = nn.LayerNorm(d_in, elementwise_affine=a_v) if n_v else None


def host_chunk_bounds(n: int, chunks=None) -> List[Tuple[int, int]]:
    """Row pieces [(begin, end), ...] of `forward_host`'s copy / kernel pipeline: contiguous, covering [0, n), every
    cut on a multiple of 128 rows (the fused kernel's row tile).  `chunks` = number of equal pieces, or a sequence
    of weights (pieces proportional to them); default: a 1:3:3:1 ramp for n >= 8192 -- small first / last pieces
    keep the un-overlapped lead-in copy and lead-out copy short -- two pieces for n >= 2048, else one."""
    if n <= 0:
        return []
    if chunks is None:
        chunks = (1, 3, 3, 1) if n >= 8192 else (2 if n >= 2048 else 1)
    if isinstance(chunks, int):
        rows = max(128, (-(-n // max(chunks, 1)) + 127) // 128 * 128)
        return [(a, min(a + rows, n)) for a in range(0, n, rows)]
    weights = [float(w) for w in chunks]
    if not weights or min(weights) < 0 or sum(weights) <= 0:
        raise ValueError("forward_host: chunk weights must be non-negative with a positive sum")
    total = sum(weights)
    cuts, acc = [0], 0.0
    for w in weights:
        acc += w
        cuts.append(max(cuts[-1], min(n, int(round(n * acc / total / 128.0)) * 128)))
    cuts[-1] = n
    return [(a, b) for a, b in zip(cuts[:-1], cuts[1:]) if b > a]


class HopfieldLayer(nn.Module):
    """Drop-in `hflayers.HopfieldLayer` whose CUDA path is the sm_100a library.

    forward(input): (B, L, input_size) or (N, input_size), float32, CUDA ->
    (B, L, output_size or input_size).  After a call, `last_argmax` holds the
    nearest-pattern index per query row (int32, flattened rows).

    `kernel_path` ("auto" | "fp32" | "f16" | "bf16", or env HOP_SM100_PATH at construction) selects the kernel
    family; "auto" keeps the reference's own sizes and every layer without LayerNorm'd score operands on the exact
    fp32 kernels and sends large N x P problems to tcgen05 (fp16-encoded operands, fp32 accumulate).  A training
    step runs forward and backward on one family (`hop_sm100_select_train_path`).

    `update_steps_max` / `update_steps_eps` follow upstream: at most T retrieve updates, ending early when
    max over the batch of ||A_t - A_(t-1)||_F <= eps (checked behind update t >= 1; `update_steps_max < 0` iterates
    until then, at most 64 updates).  The stop rule is evaluated by the fp32 kernels; the 16-bit tensor-core paths
    cannot resolve 1e-4 changes of the association and run exactly T updates.

    `exact_argmax` (default True): on the tensor-core paths rows whose nearest pattern the 16-bit scores cannot
    decide are re-scored in fp32, so `last_argmax` equals the fp32 reference's index.

    `tune` (dict, tests / experiments): scheduling overrides that travel in the op descriptor, e.g.
    `layer.tune = dict(splits=3)`, `dict(streamk=True)`, `dict(rows128=True)`.
    """

    def __init__(self, input_size: int, hidden_size: Optional[int] = None, output_size: Optional[int] = None,
                 pattern_size: Optional[int] = None, num_heads: int = 1, scaling: Optional[float] = None,
                 update_steps_max: int = 0, update_steps_eps: float = 1e-4,
                 lookup_weights_as_separated: bool = False, lookup_targets_as_trainable: bool = True,
                 normalize_stored_pattern: bool = True, normalize_stored_pattern_affine: bool = True,
                 normalize_state_pattern: bool = True, normalize_state_pattern_affine: bool = True,
                 normalize_pattern_projection: bool = True, normalize_pattern_projection_affine: bool = True,
                 normalize_hopfield_space: bool = False, normalize_hopfield_space_affine: bool = False,
                 stored_pattern_as_static: bool = False, state_pattern_as_static: bool = False,
                 pattern_projection_as_static: bool = False, pattern_projection_as_connected: bool = False,
                 stored_pattern_size: Optional[int] = None, pattern_projection_size: Optional[int] = None,
                 batch_first: bool = True, association_activation: Optional[str] = None, dropout: float = 0.0,
                 input_bias: bool = True, concat_bias_pattern: bool = False, add_zero_association: bool = False,
                 disable_out_projection: bool = False, quantity: int = 1, trainable: bool = True):
        super().__init__()
        given = dict(num_heads=num_heads, normalize_hopfield_space=normalize_hopfield_space,
                     normalize_hopfield_space_affine=normalize_hopfield_space_affine,
                     pattern_projection_as_connected=pattern_projection_as_connected,
                     association_activation=association_activation, dropout=dropout,
                     concat_bias_pattern=concat_bias_pattern, add_zero_association=add_zero_association,
                     batch_first=batch_first)
        for name, default in _NOT_BUILT.items():
            if given[name] != default:
                raise NotImplementedError(f"HopfieldLayer({name}={given[name]!r}) is not built: Hop-VAE "
                                          f"(HopVAE.py:149-170) leaves it at {default!r}")
        if stored_pattern_size not in (None, input_size) or pattern_projection_size not in (None, input_size):
            raise NotImplementedError("stored_pattern_size / pattern_projection_size other than input_size")
        if update_steps_max is None:
            raise ValueError("update_steps_max must be an integer (upstream: < 0 iterates until convergence)")
        if update_steps_max < 0 and not update_steps_eps > 0:
            raise ValueError("update_steps_max < 0 (iterate until convergence) needs update_steps_eps > 0")
        d_in = int(input_size)
        d_assoc = d_in if hidden_size is None else int(hidden_size)
        d_v = d_assoc if pattern_size is None else int(pattern_size)
        d_out = d_in if output_size is None else int(output_size)
        if (stored_pattern_as_static or state_pattern_as_static) and d_assoc != d_in:
            raise ValueError("static stored/state patterns live in the input space: hidden_size must equal input_size")
        if pattern_projection_as_static and d_v != d_in:
            raise ValueError("static pattern projection needs pattern_size == input_size")
        self.input_size, self.hidden_size, self.pattern_size, self.output_size = d_in, d_assoc, d_v, d_out
        self.quantity = int(quantity)
        self.scaling = scaling
        self.update_steps_max = int(update_steps_max)
        self.update_steps_eps = float(update_steps_eps)
        self.hopfield = _Hopfield(
            d_in, d_assoc, d_v, d_out,
            (normalize_stored_pattern, normalize_stored_pattern_affine, normalize_state_pattern,
             normalize_state_pattern_affine, normalize_pattern_projection, normalize_pattern_projection_affine),
            (stored_pattern_as_static, state_pattern_as_static, pattern_projection_as_static),
            input_bias, not disable_out_projection)
        self.lookup_weights = nn.Parameter(torch.empty(1, self.quantity, d_in), requires_grad=trainable)
        if lookup_weights_as_separated:
            self.lookup_targets = nn.Parameter(torch.empty(1, self.quantity, d_in),
                                               requires_grad=lookup_targets_as_trainable)
        else:
            self.register_parameter("lookup_targets", None)
        flags = 0
        flags |= cabi.F_NORM_STATE if normalize_state_pattern else 0
        flags |= cabi.F_NORM_STORED if normalize_stored_pattern else 0
        flags |= cabi.F_NORM_PROJ if normalize_pattern_projection else 0
        flags |= 0 if state_pattern_as_static else cabi.F_PROJ_STATE
        flags |= 0 if stored_pattern_as_static else cabi.F_PROJ_STORED
        flags |= 0 if pattern_projection_as_static else cabi.F_PROJ_PATTERN
        flags |= 0 if disable_out_projection else cabi.F_OUT_PROJ
        self._flags = flags
        self.kernel_path = os.environ.get("HOP_SM100_PATH", "auto")
        self.exact_argmax = True
        self.tune: Dict[str, object] = {}
        self.last_argmax: Optional[Tensor] = None
        self.last_lse: Optional[Tensor] = None
        self._pack_cache = None
        self._cache_key = None
        self._cache = None
        self._paths: Dict[tuple, int] = {}
        self.reset_parameters()

    # ---------------------------------------------------------------- parameters
    def reset_parameters(self) -> None:
        nn.init.normal_(self.lookup_weights, mean=0.0, std=0.02)
        if self.lookup_targets is not None:
            nn.init.normal_(self.lookup_targets, mean=0.0, std=0.02)

    @property
    def beta(self) -> float:
        return float(self.scaling) if self.scaling is not None else 1.0 / math.sqrt(self.hidden_size)

    def spec(self) -> LayerSpec:
        path = cabi.PATH_NAMES.get(self.kernel_path)
        if path is None:
            raise ValueError(f"kernel_path must be one of {sorted(cabi.PATH_NAMES)}, got {self.kernel_path!r}")
        tune_flags = 0
        for name, v in self.tune.items():
            if name == "splits":
                continue
            if name not in _TUNE_NAMES:
                raise ValueError(f"unknown tuning override {name!r} (one of {sorted(_TUNE_NAMES) + ['splits']})")
            on, off = _TUNE_NAMES[name]
            tune_flags |= on if v else off
        return LayerSpec(flags=self._flags | (cabi.F_EXACT_ARGMAX if self.exact_argmax else 0), path=path,
                         d_in=self.input_size, d_assoc=self.hidden_size, d_v=self.pattern_size,
                         d_out=self.output_size, update_steps=self.update_steps_max, beta=self.beta, ln_eps=1e-5,
                         update_eps=max(self.update_steps_eps, 0.0), tune_splits=int(self.tune.get("splits", 0) or 0),
                         tune_flags=tune_flags)

    def param_pack(self) -> Dict[str, Optional[Tensor]]:
        """The module's tensors under the hop_params names.  Under no_grad the dict is cached (rebuilt when a
        parameter's storage moved, e.g. after `.to(device)` -- the q/k/v entries are views of `in_proj_*`); with
        autograd on it is rebuilt every call so that the views belong to the current graph."""
        if torch.is_grad_enabled():
            return self._build_pack()
        core0 = self.hopfield.association_core
        stamp = (self.lookup_weights.data_ptr(),
                 0 if core0.in_proj_weight is None else core0.in_proj_weight.data_ptr(),
                 0 if core0.in_proj_bias is None else core0.in_proj_bias.data_ptr(),
                 0 if self.lookup_targets is None else self.lookup_targets.data_ptr())
        if self._pack_cache is not None and self._pack_cache[0] == stamp:
            return self._pack_cache[1]
        pack = self._build_pack()
        self.__dict__["_pack_cache"] = (stamp, pack)
        return pack

    def _build_pack(self) -> Dict[str, Optional[Tensor]]:
        h = self.hopfield
        core = h.association_core
        ln = lambda m: (None, None) if m is None else (m.weight, m.bias)
        pack: Dict[str, Optional[Tensor]] = dict(
            # squeeze, not [0]: its backward is a view; SelectBackward zero-fills and copies the full gradient
            lookup=self.lookup_weights.squeeze(0),
            lookup_targets=None if self.lookup_targets is None else self.lookup_targets.squeeze(0))
        pack["ln_state_w"], pack["ln_state_b"] = ln(h.norm_state_pattern)
        pack["ln_stored_w"], pack["ln_stored_b"] = ln(h.norm_stored_pattern)
        pack["ln_proj_w"], pack["ln_proj_b"] = ln(h.norm_pattern_projection)
        pack.update(core.split())
        pack["wo"] = None if core.out_proj is None else core.out_proj.weight
        pack["bo"] = None if core.out_proj is None else core.out_proj.bias
        return pack

    # ------------------------------------------------------------------- caching
    def invalidate(self) -> None:
        """Drop the prepared pattern operands (K^ / V^ packs) and the captured `forward_host` graphs.  The cache is
        keyed on (data_ptr, _version) of every pattern-side tensor, which in-place updates through `.data`
        (old-style optimizers, EMA, clipping) do not bump: call this after such a write.  `train()`, `eval()` and
        `load_state_dict()` call it themselves."""
        self.__dict__["_cache"] = None
        self.__dict__["_cache_key"] = None
        self.__dict__["_pack_cache"] = None
        st = self.__dict__.get("_host_state")
        if st is not None:
            st["graphs"].clear()

    def train(self, mode: bool = True):
        self.invalidate()
        return super().train(mode)

    def _load_from_state_dict(self, *args, **kwargs):
        self.invalidate()
        return super()._load_from_state_dict(*args, **kwargs)

    def _prepared(self, pack, spec: LayerSpec, path: int):
        """Prepared pattern operands, recomputed only when a pattern-side tensor changed."""
        key = (path, spec.update_steps, spec.flags, spec.tune_flags) + tuple(
            (None if pack[k] is None else (pack[k].data_ptr(), pack[k]._version, pack[k].device))
            for k in _PATTERN_SIDE)
        if key != self._cache_key:
            with torch.no_grad():
                self._cache = (prepare_patterns if _eager() else prepare_patterns_op)(
                    *[None if pack[k] is None else pack[k].detach().contiguous() for k in _PATTERN_SIDE],
                    spec.flags, path, spec.d_assoc, spec.d_v, spec.update_steps, spec.ln_eps, spec.tune_flags)
            self._cache_key = key
        return self._cache

    # ------------------------------------------------------------------- forward
    def forward(self, input: Tensor, stored_pattern_padding_mask: Optional[Tensor] = None,
                association_mask: Optional[Tensor] = None) -> Tensor:
        if stored_pattern_padding_mask is not None or association_mask is not None:
            raise NotImplementedError("masks are not used by Hop-VAE and are not built")
        if not input.is_cuda:
            raise RuntimeError("hflayers.HopfieldLayer (hop_sm100): input is on the CPU; this layer runs only on an "
                               "sm_100a GPU and has no CPU fallback")
        if input.shape[-1] != self.input_size:
            raise ValueError(f"expected last dimension {self.input_size}, got {tuple(input.shape)}")
        lead = input.shape[:-1]
        x = input.reshape(-1, self.input_size)
        if x.dtype != torch.float32:
            raise RuntimeError(f"hop_sm100: float32 input expected, got {x.dtype}")
        x = x.contiguous()
        spec = self.spec()
        pack = self.param_pack()
        needs_grad = torch.is_grad_enabled() and (
            x.requires_grad or any(t is not None and t.requires_grad for t in pack.values()))
        if needs_grad:
            out, argmax, lse = HopfieldAssociation.apply(spec, x, *[None if pack[k] is None else pack[k].contiguous()
                                                                    for k in _PARAM_ORDER])
        else:
            # upstream's stop rule takes the Frobenius norm per batch item: rows of one (L, D) slice belong together
            rows_per_item = int(lead[-1]) if len(lead) >= 2 else 0
            out, argmax, lse = self._infer(x, spec, pack, rows_per_item=rows_per_item)
        self.__dict__["last_lse"] = lse            # (plain attribute: nn.Module.__setattr__ costs ~5 us per set)
        self.__dict__["last_argmax"] = argmax
        return out.reshape(*lead, out.shape[-1])

    def _path_for(self, spec: LayerSpec, n_rows: int) -> int:
        pkey = (spec, n_rows, self.quantity)
        path = self._paths.get(pkey)
        if path is None:
            path = self._paths[pkey] = resolve_path(spec, n_rows, self.quantity)
        return path

    def _infer(self, x: Tensor, spec: LayerSpec, pack, into=None, path: Optional[int] = None,
               rows_per_item: int = 0) -> Tuple[Tensor, Tensor, Tensor]:
        """No-grad retrieval of x (N, Din) -> (out, nearest-pattern index, lse); `into` = caller-owned result
        buffers (out (N,W), stats (2,N), argmax (N) i32); `path` = kernel family already resolved by the caller."""
        if path is None:
            path = self._path_for(spec, x.shape[0])
        pats = self._prepared(pack, spec, path)
        q = [None if pack[k] is None else pack[k].detach().contiguous() for k in _QUERY_SIDE]
        if _eager():
            res = retrieve(x.detach(), pats, *q, spec.flags, path, spec.d_in, spec.d_out, spec.update_steps,
                           spec.beta, spec.ln_eps, 0, False, False, into, spec.update_eps, rows_per_item,
                           spec.tune_splits, spec.tune_flags)
        else:
            res = retrieve_op(x.detach(), pats, *q, spec.flags, path, spec.d_in, spec.d_out, spec.update_steps,
                              spec.beta, spec.ln_eps, 0, False, spec.update_eps, rows_per_item, spec.tune_splits,
                              spec.tune_flags)
        return res[0], res[4], res[3]

    # ------------------------------------------------------- host-buffer entry point
    def forward_host(self, x_host: Tensor, out_host: Optional[Tensor] = None, chunks=None,
                     graph: Optional[bool] = None) -> Tensor:
        """Host-buffer call: `x_host` (N, input_size) pinned fp32 -> `out_host` (N, output_size) pinned fp32.

        The rows are cut into `chunks` pieces (an int: equal pieces; a sequence of weights: proportional pieces;
        multiples of 128 rows); the host->device copy of piece i+1, the fused retrieval of piece i and the
        device->host copy of piece i-1 run on three streams, so the PCIe transfers hide behind the kernels.
        The whole pipeline (copies, kernels, cross-stream dependencies) is captured ONCE per (buffers, N, schedule,
        parameter version) into a CUDA graph and replayed afterwards: one host call per step instead of ~40, so a
        busy host cannot starve the GPU (`graph=False`, or HOP_SM100_HOST_GRAPH=0, keeps the eager enqueue).
        Returns once all work is *enqueued*; the current stream waits for the last device->host copy, so a
        synchronize / event on the current stream covers everything.  Inference only (no autograd).
        `last_argmax` / `last_lse` are views of per-layer staging buffers, valid until the next call.  The kernel
        family is resolved ONCE per call on all N rows and used for every piece."""
        dev = self.lookup_weights.device
        if dev.type != "cuda":
            raise RuntimeError("hflayers.HopfieldLayer (hop_sm100): the layer is on the CPU; no CPU fallback")
        if x_host.is_cuda or x_host.dtype != torch.float32 or x_host.dim() != 2 or x_host.shape[1] != self.input_size:
            raise ValueError("forward_host expects a 2-D float32 host tensor (N, input_size)")
        if not x_host.is_pinned():
            raise ValueError("forward_host needs pinned host memory (asynchronous copies)")
        n = x_host.shape[0]
        width = self.output_size if (self._flags & cabi.F_OUT_PROJ) else self.pattern_size
        own_out = out_host is None
        if own_out:
            out_host = torch.empty((n, width), dtype=torch.float32).pin_memory()
        elif out_host.shape != (n, width) or out_host.dtype != torch.float32 or not out_host.is_pinned():
            raise ValueError("forward_host: out_host must be a pinned float32 (N, output width) tensor")
        bounds = host_chunk_bounds(n, chunks)
        st = self.__dict__.get("_host_state")
        if st is None or st["dev"] != dev or st["rows"] < n or st["width"] != width:
            st = dict(dev=dev, rows=n, width=width, streams=tuple(torch.cuda.Stream(device=dev) for _ in range(3)),
                      x=torch.empty((n, self.input_size), device=dev, dtype=torch.float32),
                      y=torch.empty((n, width), device=dev, dtype=torch.float32),
                      stats=torch.empty((2, n), device=dev, dtype=torch.float32),
                      arg=torch.empty((n,), device=dev, dtype=torch.int32), graphs={})
            self.__dict__["_host_state"] = st
        # (a fresh result buffer per call would mean a fresh capture per call: graphs need caller-owned buffers)
        use_graph = graph if graph is not None else not own_out
        with torch.no_grad():
            spec, pack = self.spec(), self.param_pack()
            path = self._path_for(spec, n)
            if use_graph and _eager():
                # the prepared pattern operands are baked into the captured launches: their cache key is part of ours
                self._prepared(pack, spec, path)
                key = (x_host.data_ptr(), out_host.data_ptr(), n, tuple(bounds), self._cache_key, path)
                g = st["graphs"].get(key)
                if g is None:
                    self._host_pipeline(x_host, out_host, bounds, st, spec, pack, path)   # warm-up: lazy caches, allocator
                    torch.cuda.current_stream(dev).synchronize()
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g):
                        self._host_pipeline(x_host, out_host, bounds, st, spec, pack, path)
                    if len(st["graphs"]) >= 4:
                        st["graphs"].pop(next(iter(st["graphs"])))
                    st["graphs"][key] = g
                g.replay()
            else:
                self._host_pipeline(x_host, out_host, bounds, st, spec, pack, path)
        self.__dict__["last_argmax"] = st["arg"][:n]
        self.__dict__["last_lse"] = st["stats"][0, :n]
        return out_host

    def _host_pipeline(self, x_host: Tensor, out_host: Tensor, bounds, st, spec: LayerSpec, pack, path: int) -> None:
        """Enqueue H2D / retrieval / D2H of every piece on the three side streams (fork from and join to the current
        stream, so the same code runs eagerly and under CUDA-graph capture); results go through staging buffers."""
        dev = st["dev"]
        h2d, comp, d2h = st["streams"]
        cur = torch.cuda.current_stream(dev)
        for s_ in st["streams"]:
            s_.wait_stream(cur)
        for a, b in bounds:
            xs = st["x"][a:b]
            with torch.cuda.stream(h2d):
                xs.copy_(x_host[a:b], non_blocking=True)
                e1 = h2d.record_event()
            with torch.cuda.stream(comp):
                comp.wait_event(e1)
                self._infer(xs, spec, pack, into=(st["y"][a:b], st["stats"][:, a:b], st["arg"][a:b]), path=path)
                e2 = comp.record_event()
            with torch.cuda.stream(d2h):
                d2h.wait_event(e2)
                out_host[a:b].copy_(st["y"][a:b], non_blocking=True)
        cur.wait_stream(h2d)
        cur.wait_stream(comp)
        cur.wait_stream(d2h)

    def extra_repr(self) -> str:
        return (f"input_size={self.input_size}, hidden_size={self.hidden_size}, pattern_size={self.pattern_size}, "
                f"output_size={self.output_size}, quantity={self.quantity}, beta={self.beta:.6g}, "
                f"update_steps_max={self.update_steps_max}, kernel_path={self.kernel_path!r}")
