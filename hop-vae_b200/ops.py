"""PyTorch custom ops over the C-ABI library (`hop_sm100::prepare_patterns`,
`hop_sm100::retrieve`, `hop_sm100::retrieve_backward`) and the autograd
function that ties them into one differentiable Hopfield association.

PyTorch is plumbing here: it owns device memory and the stream; all arithmetic
runs in `libhop_sm100.so`.  CUDA tensors only -- there is no CPU path.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import List, Optional, Tuple

import torch
from torch import Tensor

from . import cabi

_PATTERN_SIDE = ("lookup", "lookup_targets", "ln_stored_w", "ln_stored_b", "ln_proj_w", "ln_proj_b",
                 "wk", "bk", "wv", "bv", "wo")
_QUERY_SIDE = ("ln_state_w", "ln_state_b", "wq", "bq", "wo", "bo")


@dataclass(frozen=True)
class LayerSpec:
    """Static description of one Hopfield layer (everything in hop_desc but N, P)."""
    flags: int
    path: int
    d_in: int
    d_assoc: int
    d_v: int
    d_out: int
    update_steps: int
    beta: float
    ln_eps: float = 1e-5


def _stream(dev: Optional[torch.device] = None) -> int:
    """Raw handle of the current CUDA stream (the launch-bound reference shapes feel every microsecond here)."""
    idx = torch.cuda.current_device() if dev is None or dev.index is None else dev.index
    return torch._C._cuda_getCurrentRawStream(idx)


class _on_device:
    """`with torch.cuda.device(dev)` only when `dev` is not already current (the common case costs nothing)."""
    __slots__ = ("guard",)

    def __init__(self, dev: torch.device):
        self.guard = None if dev.index is None or dev.index == torch.cuda.current_device() else torch.cuda.device(dev)

    def __enter__(self):
        if self.guard is not None:
            self.guard.__enter__()

    def __exit__(self, *exc):
        if self.guard is not None:
            self.guard.__exit__(*exc)
        return False


def _eager() -> bool:
    """Outside torch.compile the plain Python bodies are called directly: the custom-op dispatcher costs ~50 us
    per call, more than the kernels at the reference's own sizes (N ~ 2k-9k rows, P = 512)."""
    return not torch.compiler.is_compiling()


_EMPTY_F32 = {}


def _empty_f32(dev: torch.device) -> Tensor:
    """Shared zero-element placeholder for outputs the inference call does not produce."""
    t = _EMPTY_F32.get(dev)
    if t is None:
        t = _EMPTY_F32[dev] = torch.empty((0,), device=dev, dtype=torch.float32)
    return t


_WS_BYTES = {}


def _workspace_bytes(d: "cabi.HopDesc", key, for_backward: int) -> int:
    # (the two test / tuning knobs that change the forward workspace are part of the key)
    k = (key, for_backward, os.environ.get("HOP_TC_SPLITS"), os.environ.get("HOP_TC_STREAMK"))
    v = _WS_BYTES.get(k)
    if v is None:
        import ctypes as C
        need = C.c_size_t(0)
        cabi.check(cabi.load().hop_sm100_workspace_bytes(C.byref(d), for_backward, C.byref(need)),
                   "hop_sm100_workspace_bytes")
        v = _WS_BYTES[k] = int(need.value)
    return v


def _req_cuda_f32(t: Optional[Tensor], name: str) -> None:
    if t is None:
        return
    if not t.is_cuda:
        raise RuntimeError(f"hop_sm100: `{name}` is on {t.device}; the Hopfield retrieval op is CUDA (sm_100a) only "
                           f"and has no CPU fallback")
    if t.dtype != torch.float32:
        raise RuntimeError(f"hop_sm100: `{name}` must be float32 (got {t.dtype})")
    if not t.is_contiguous():
        raise RuntimeError(f"hop_sm100: `{name}` must be contiguous")


def _desc(flags: int, path: int, n: int, p: int, d_in: int, d_assoc: int, d_v: int, d_out: int, t: int,
          beta: float, ln_eps: float, pattern_offset: int = 0) -> cabi.HopDesc:
    return cabi.make_desc(flags=flags, path=path, n_rows=n, n_patterns=p, d_in=d_in, d_assoc=d_assoc, d_v=d_v,
                          d_out=d_out, update_steps=t, beta=beta, ln_eps=ln_eps, pattern_offset=pattern_offset)


def resolve_path(spec: LayerSpec, n_rows: int, n_patterns: int) -> int:
    import ctypes as C
    d = _desc(spec.flags, spec.path, n_rows, n_patterns, spec.d_in, spec.d_assoc, spec.d_v, max(spec.d_out, 1),
              spec.update_steps, spec.beta, spec.ln_eps)
    out = C.c_int32(0)
    cabi.check(cabi.load().hop_sm100_select_path(C.byref(d), C.byref(out)), "hop_sm100_select_path")
    return int(out.value)


# --------------------------------------------------------------------------- prepare_patterns
def prepare_patterns(lookup: Tensor, lookup_targets: Optional[Tensor], ln_stored_w: Optional[Tensor],
                        ln_stored_b: Optional[Tensor], ln_proj_w: Optional[Tensor], ln_proj_b: Optional[Tensor],
                        wk: Optional[Tensor], bk: Optional[Tensor], wv: Optional[Tensor], bv: Optional[Tensor],
                        wo: Optional[Tensor], flags: int, path: int, d_assoc: int, d_v: int, update_steps: int,
                        ln_eps: float) -> List[Tensor]:
    """lookup (P, Din) -> [khat (P,H), vhat (P,Dv), k16, kt16, vt16, vot16] (16-bit packs empty on the fp32
    path; vot16 = out_proj folded into the read-out operand, only when Dout == H == Dv).
    `path` must already be resolved (not AUTO)."""
    import ctypes as C
    lib = cabi.load()
    for name, t in (("lookup", lookup), ("lookup_targets", lookup_targets), ("ln_stored_w", ln_stored_w),
                    ("ln_stored_b", ln_stored_b), ("ln_proj_w", ln_proj_w), ("ln_proj_b", ln_proj_b), ("wk", wk),
                    ("bk", bk), ("wv", wv), ("bv", bv), ("wo", wo)):
        _req_cuda_f32(t, name)
    P, d_in = lookup.shape
    d_out = wo.shape[0] if (wo is not None and (flags & cabi.F_OUT_PROJ)) else 1
    d = _desc(flags, path, 1, P, d_in, d_assoc, d_v, d_out, update_steps, 1.0, ln_eps)
    dev = lookup.device
    khat = torch.empty((P, d_assoc), device=dev, dtype=torch.float32)
    vhat = torch.empty((P, d_v), device=dev, dtype=torch.float32)
    if path == cabi.PATH_FP32_SIMT:
        none16 = torch.empty((0,), device=dev, dtype=torch.int16)
        packs = [none16, none16, none16, none16]
    else:
        sizes = [C.c_size_t(0) for _ in range(6)]
        cabi.check(lib.hop_sm100_pattern_bytes(C.byref(d), *[C.byref(s) for s in sizes]), "hop_sm100_pattern_bytes")
        packs = [torch.empty((int(s.value) // 2,), device=dev, dtype=torch.int16) for s in sizes[2:]]
    params = cabi.HopParams(lookup=cabi.ptr(lookup), lookup_targets=cabi.ptr(lookup_targets),
                            ln_stored_w=cabi.ptr(ln_stored_w), ln_stored_b=cabi.ptr(ln_stored_b),
                            ln_proj_w=cabi.ptr(ln_proj_w), ln_proj_b=cabi.ptr(ln_proj_b), wk=cabi.ptr(wk),
                            bk=cabi.ptr(bk), wv=cabi.ptr(wv), bv=cabi.ptr(bv), wo=cabi.ptr(wo))
    pats = cabi.HopPatterns(khat=cabi.ptr(khat), vhat=cabi.ptr(vhat),
                            k16=cabi.ptr(packs[0]) if packs[0].numel() else None,
                            kt16=cabi.ptr(packs[1]) if packs[1].numel() else None,
                            vt16=cabi.ptr(packs[2]) if packs[2].numel() else None,
                            vot16=cabi.ptr(packs[3]) if packs[3].numel() else None)
    with _on_device(dev):
        cabi.check(lib.hop_sm100_prepare_patterns(C.byref(d), C.byref(params), C.byref(pats), _stream(dev)),
                   "hop_sm100_prepare_patterns")
    return [khat, vhat] + packs


@torch.library.custom_op("hop_sm100::prepare_patterns", mutates_args=(), device_types="cuda")
def prepare_patterns_op(lookup: Tensor, lookup_targets: Optional[Tensor], ln_stored_w: Optional[Tensor],
                        ln_stored_b: Optional[Tensor], ln_proj_w: Optional[Tensor], ln_proj_b: Optional[Tensor],
                        wk: Optional[Tensor], bk: Optional[Tensor], wv: Optional[Tensor], bv: Optional[Tensor],
                        wo: Optional[Tensor], flags: int, path: int, d_assoc: int, d_v: int, update_steps: int,
                        ln_eps: float) -> List[Tensor]:
    """torch.compile-visible wrapper of `prepare_patterns`."""
    res = prepare_patterns(lookup, lookup_targets, ln_stored_w, ln_stored_b, ln_proj_w, ln_proj_b, wk, bk, wv, bv,
                           wo, flags, path, d_assoc, d_v, update_steps, ln_eps)
    # custom-op outputs may not alias each other (the fp32 path shares one empty pack)
    return res[:2] + [r if r.numel() else torch.empty_like(r) for r in res[2:]]


@prepare_patterns_op.register_fake
def _(lookup, lookup_targets, ln_stored_w, ln_stored_b, ln_proj_w, ln_proj_b, wk, bk, wv, bv, wo, flags, path,
      d_assoc, d_v, update_steps, ln_eps):
    P = lookup.shape[0]
    ppad = (P + 127) // 128 * 128
    tc = path in (cabi.PATH_TC_F16, cabi.PATH_TC_BF16)
    mk = lambda n: lookup.new_empty((n,), dtype=torch.int16)
    fold = (tc and wo is not None and bool(flags & cabi.F_OUT_PROJ) and wo.shape[0] == d_assoc == d_v
            and d_assoc in (64, 128, 256, 512, 1024))
    return [lookup.new_empty((P, d_assoc)), lookup.new_empty((P, d_v)), mk(ppad * d_assoc if tc else 0),
            mk(ppad * d_assoc if tc and update_steps > 0 else 0), mk(ppad * d_v if tc else 0),
            mk(ppad * d_assoc if fold else 0)]


# --------------------------------------------------------------------------- retrieve (forward)
def retrieve(x: Tensor, khat: Tensor, vhat: Tensor, k16: Tensor, kt16: Tensor, vt16: Tensor, vot16: Tensor,
             ln_state_w: Optional[Tensor], ln_state_b: Optional[Tensor], wq: Optional[Tensor],
             bq: Optional[Tensor], wo: Optional[Tensor], bo: Optional[Tensor], flags: int, path: int,
             d_in: int, d_out: int, update_steps: int, beta: float, ln_eps: float,
             pattern_offset: int, want_retrieved: bool = True, want_saved: bool = True,
             into: Optional[Tuple[Tensor, Tensor, Tensor]] = None) -> List[Tensor]:
    """x (N, Din) -> [out (N,Dout|Dv), retrieved (N,Dv), qhat (N,H), lse (N), argmax (N) int32, rowmax (N)].
    With want_retrieved=False and a folded read-out operand (vot16) the tensor-core kernel writes `out`
    directly and `retrieved` is returned empty.  want_saved=False (inference) additionally skips everything only
    the backward needs: `qhat` is returned empty and, with an out-projection, so is `retrieved`.
    `into=(out (N,W), stats (2,N) f32, argmax (N) i32)` writes the results into caller-owned buffers (the
    peer-mapped exchange buffers of the pattern-sharded mode) instead of fresh tensors."""
    import ctypes as C
    lib = cabi.load()
    for name, t in (("x", x), ("khat", khat), ("vhat", vhat), ("ln_state_w", ln_state_w), ("ln_state_b", ln_state_b),
                    ("wq", wq), ("bq", bq), ("wo", wo), ("bo", bo)):
        _req_cuda_f32(t, name)
    N = x.shape[0]
    P, H = khat.shape
    Dv = vhat.shape[1]
    out_proj = bool(flags & cabi.F_OUT_PROJ)
    key = (flags, path, N, P, d_in, H, Dv, d_out if out_proj else 1, update_steps, beta, ln_eps, pattern_offset)
    d = cabi.make_desc(flags=flags, path=path, n_rows=N, n_patterns=P, d_in=d_in, d_assoc=H, d_v=Dv,
                       d_out=d_out if out_proj else 1, update_steps=update_steps, beta=beta, ln_eps=ln_eps,
                       pattern_offset=pattern_offset)
    dev = x.device
    if into is not None:
        out = into[0]
        if out.shape != (N, d_out if out_proj else Dv) or not out.is_contiguous() or out.dtype != torch.float32:
            raise ValueError("hop_sm100: `into[0]` must be a contiguous float32 (N, out width) tensor")
    else:
        out = torch.empty((N, d_out if out_proj else Dv), device=dev, dtype=torch.float32)
    tc = path != cabi.PATH_FP32_SIMT
    folded = out_proj and not want_retrieved and vot16.numel() > 0 and tc
    if not want_saved and not tc:
        want_retrieved = False              # the fp32 kernel's read-out only feeds its fused out-projection
    if folded or (out_proj and not want_retrieved and not tc):
        retrieved = None
    else:
        retrieved = torch.empty((N, Dv), device=dev, dtype=torch.float32) if out_proj else out
    qhat = torch.empty((N, H), device=dev, dtype=torch.float32) if want_saved else None
    stats = into[1] if into is not None else torch.empty((2, N), device=dev, dtype=torch.float32)   # lse, rowmax
    lse, rowmax = stats[0], stats[1]
    argmax = into[2] if into is not None else torch.empty((N,), device=dev, dtype=torch.int32)
    use_ws = bool(flags & cabi.F_PROJ_STATE) or update_steps > 0 or tc
    need = _workspace_bytes(d, key, 0) if use_ws else 0
    ws = torch.empty((need,), device=dev, dtype=torch.uint8) if use_ws else None
    params = cabi.HopParams(ln_state_w=cabi.ptr(ln_state_w), ln_state_b=cabi.ptr(ln_state_b), wq=cabi.ptr(wq),
                            bq=cabi.ptr(bq), wo=cabi.ptr(wo), bo=cabi.ptr(bo))
    pats = cabi.HopPatterns(khat=cabi.ptr(khat), vhat=cabi.ptr(vhat),
                            k16=cabi.ptr(k16) if k16.numel() else None,
                            kt16=cabi.ptr(kt16) if kt16.numel() else None,
                            vt16=cabi.ptr(vt16) if vt16.numel() else None,
                            vot16=cabi.ptr(vot16) if vot16.numel() else None)
    with _on_device(dev):
        cabi.check(lib.hop_sm100_forward(C.byref(d), C.byref(params), C.byref(pats), cabi.ptr(x), cabi.ptr(out),
                                         cabi.ptr(retrieved), cabi.ptr(qhat), cabi.ptr(lse), cabi.ptr(argmax),
                                         cabi.ptr(rowmax), cabi.ptr(ws), need, _stream(dev)), "hop_sm100_forward")
    if qhat is None:
        qhat = _empty_f32(dev)
    if retrieved is None:
        retrieved = _empty_f32(dev) if not want_saved else torch.empty((0, Dv), device=dev, dtype=torch.float32)
    elif not out_proj:
        retrieved = out.clone() if want_saved else out
    return [out, retrieved, qhat, lse, argmax, rowmax]


@torch.library.custom_op("hop_sm100::retrieve", mutates_args=(), device_types="cuda")
def retrieve_op(x: Tensor, khat: Tensor, vhat: Tensor, k16: Tensor, kt16: Tensor, vt16: Tensor, vot16: Tensor,
                ln_state_w: Optional[Tensor], ln_state_b: Optional[Tensor], wq: Optional[Tensor],
                bq: Optional[Tensor], wo: Optional[Tensor], bo: Optional[Tensor], flags: int, path: int,
                d_in: int, d_out: int, update_steps: int, beta: float, ln_eps: float,
                pattern_offset: int, want_retrieved: bool = True) -> List[Tensor]:
    """torch.compile-visible wrapper of `retrieve` (always returns what the backward needs)."""
    res = retrieve(x, khat, vhat, k16, kt16, vt16, vot16, ln_state_w, ln_state_b, wq, bq, wo, bo, flags, path, d_in,
                   d_out, update_steps, beta, ln_eps, pattern_offset, want_retrieved, True)
    # custom-op outputs may not alias each other
    return [res[0], res[1], res[2], res[3].clone(), res[4], res[5].clone()]


@retrieve_op.register_fake
def _(x, khat, vhat, k16, kt16, vt16, vot16, ln_state_w, ln_state_b, wq, bq, wo, bo, flags, path, d_in, d_out,
      update_steps, beta, ln_eps, pattern_offset, want_retrieved=True):
    N = x.shape[0]
    H, Dv = khat.shape[1], vhat.shape[1]
    out_w = d_out if (flags & cabi.F_OUT_PROJ) else Dv
    folded = (bool(flags & cabi.F_OUT_PROJ) and not want_retrieved and vot16.numel() > 0
              and path != cabi.PATH_FP32_SIMT)
    return [x.new_empty((N, out_w)), x.new_empty((0 if folded else N, Dv)), x.new_empty((N, H)), x.new_empty((N,)),
            x.new_empty((N,), dtype=torch.int32), x.new_empty((N,))]


# --------------------------------------------------------------------------- retrieve backward
_GRAD_ORDER = cabi._GRAD_FIELDS  # order of the returned list


def retrieve_backward(grad_out: Tensor, x: Tensor, qhat: Tensor, retrieved: Tensor, lse: Tensor, khat: Tensor,
                      vhat: Tensor, lookup: Tensor, lookup_targets: Optional[Tensor],
                      ln_state_w: Optional[Tensor], ln_state_b: Optional[Tensor], ln_stored_w: Optional[Tensor],
                      ln_stored_b: Optional[Tensor], ln_proj_w: Optional[Tensor], ln_proj_b: Optional[Tensor],
                      wq: Optional[Tensor], bq: Optional[Tensor], wk: Optional[Tensor], bk: Optional[Tensor],
                      wv: Optional[Tensor], bv: Optional[Tensor], wo: Optional[Tensor], bo: Optional[Tensor],
                      flags: int, d_out: int, beta: float, ln_eps: float) -> List[Optional[Tensor]]:
    """Gradients in the order of hop_grads (None where the parameter is absent)."""
    import ctypes as C
    lib = cabi.load()
    _req_cuda_f32(grad_out, "grad_out")
    N, d_in = x.shape
    P, H = khat.shape
    Dv = vhat.shape[1]
    out_proj = bool(flags & cabi.F_OUT_PROJ)
    key = (flags, cabi.PATH_FP32_SIMT, N, P, d_in, H, Dv, d_out if out_proj else 1, 0, beta, ln_eps, 0)
    d = _desc(flags, cabi.PATH_FP32_SIMT, N, P, d_in, H, Dv, d_out if out_proj else 1, 0, beta, ln_eps)
    dev = x.device
    like = lambda t: None if t is None else torch.empty_like(t)
    g = {
        "d_x": torch.empty_like(x), "d_lookup": torch.empty_like(lookup), "d_lookup_targets": like(lookup_targets),
        "d_ln_state_w": like(ln_state_w), "d_ln_state_b": like(ln_state_b), "d_ln_stored_w": like(ln_stored_w),
        "d_ln_stored_b": like(ln_stored_b), "d_ln_proj_w": like(ln_proj_w), "d_ln_proj_b": like(ln_proj_b),
        "d_wq": like(wq), "d_bq": like(bq), "d_wk": like(wk), "d_bk": like(bk), "d_wv": like(wv), "d_bv": like(bv),
        "d_wo": like(wo), "d_bo": like(bo),
    }
    need = _workspace_bytes(d, key, 1)
    ws = torch.empty((need,), device=dev, dtype=torch.uint8)
    params = cabi.HopParams(lookup=cabi.ptr(lookup), lookup_targets=cabi.ptr(lookup_targets),
                            ln_state_w=cabi.ptr(ln_state_w), ln_state_b=cabi.ptr(ln_state_b),
                            ln_stored_w=cabi.ptr(ln_stored_w), ln_stored_b=cabi.ptr(ln_stored_b),
                            ln_proj_w=cabi.ptr(ln_proj_w), ln_proj_b=cabi.ptr(ln_proj_b), wq=cabi.ptr(wq),
                            bq=cabi.ptr(bq), wk=cabi.ptr(wk), bk=cabi.ptr(bk), wv=cabi.ptr(wv), bv=cabi.ptr(bv),
                            wo=cabi.ptr(wo), bo=cabi.ptr(bo))
    pats = cabi.HopPatterns(khat=cabi.ptr(khat), vhat=cabi.ptr(vhat))
    grads = cabi.HopGrads(**{k: (cabi.ptr(v) if v is not None and v.numel() else None) for k, v in g.items()})
    go = grad_out.contiguous()
    with _on_device(dev):
        cabi.check(lib.hop_sm100_backward(C.byref(d), C.byref(params), C.byref(pats), cabi.ptr(x), cabi.ptr(qhat),
                                          cabi.ptr(retrieved), cabi.ptr(lse), cabi.ptr(go), C.byref(grads),
                                          cabi.ptr(ws), need, _stream(dev)), "hop_sm100_backward")
    return [g[k] for k in _GRAD_ORDER]


@torch.library.custom_op("hop_sm100::retrieve_backward", mutates_args=(), device_types="cuda")
def retrieve_backward_op(grad_out: Tensor, x: Tensor, qhat: Tensor, retrieved: Tensor, lse: Tensor, khat: Tensor,
                         vhat: Tensor, lookup: Tensor, lookup_targets: Optional[Tensor],
                         ln_state_w: Optional[Tensor], ln_state_b: Optional[Tensor], ln_stored_w: Optional[Tensor],
                         ln_stored_b: Optional[Tensor], ln_proj_w: Optional[Tensor], ln_proj_b: Optional[Tensor],
                         wq: Optional[Tensor], bq: Optional[Tensor], wk: Optional[Tensor], bk: Optional[Tensor],
                         wv: Optional[Tensor], bv: Optional[Tensor], wo: Optional[Tensor], bo: Optional[Tensor],
                         flags: int, d_out: int, beta: float, ln_eps: float) -> List[Tensor]:
    """torch.compile-visible wrapper of `retrieve_backward` (empty tensors where the parameter is absent)."""
    res = retrieve_backward(grad_out, x, qhat, retrieved, lse, khat, vhat, lookup, lookup_targets, ln_state_w,
                            ln_state_b, ln_stored_w, ln_stored_b, ln_proj_w, ln_proj_b, wq, bq, wk, bk, wv, bv, wo,
                            bo, flags, d_out, beta, ln_eps)
    return [r if r is not None else torch.empty((0,), device=x.device) for r in res]


@retrieve_backward_op.register_fake
def _(grad_out, x, qhat, retrieved, lse, khat, vhat, lookup, lookup_targets, ln_state_w, ln_state_b, ln_stored_w,
      ln_stored_b, ln_proj_w, ln_proj_b, wq, bq, wk, bk, wv, bv, wo, bo, flags, d_out, beta, ln_eps):
    like = lambda t: torch.empty_like(t) if t is not None else x.new_empty((0,))
    return [torch.empty_like(x), torch.empty_like(lookup), like(lookup_targets), like(ln_state_w), like(ln_state_b),
            like(ln_stored_w), like(ln_stored_b), like(ln_proj_w), like(ln_proj_b), like(wq), like(bq), like(wk),
            like(bk), like(wv), like(bv), like(wo), like(bo)]


# --------------------------------------------------------------------------- autograd glue
_PARAM_ORDER = ("lookup", "lookup_targets", "ln_state_w", "ln_state_b", "ln_stored_w", "ln_stored_b", "ln_proj_w",
                "ln_proj_b", "wq", "bq", "wk", "bk", "wv", "bv", "wo", "bo")
_GRAD_OF_PARAM = {"lookup": "d_lookup", "lookup_targets": "d_lookup_targets", "ln_state_w": "d_ln_state_w",
                  "ln_state_b": "d_ln_state_b", "ln_stored_w": "d_ln_stored_w", "ln_stored_b": "d_ln_stored_b",
                  "ln_proj_w": "d_ln_proj_w", "ln_proj_b": "d_ln_proj_b", "wq": "d_wq", "bq": "d_bq", "wk": "d_wk",
                  "bk": "d_bk", "wv": "d_wv", "bv": "d_bv", "wo": "d_wo", "bo": "d_bo"}


class HopfieldAssociation(torch.autograd.Function):
    """Differentiable  x (N,Din), params  ->  out (N,Dout).  Training path: patterns are
    prepared inside (parameters change every step); backward recomputes A flash-style."""

    @staticmethod
    def forward(ctx, spec: LayerSpec, x: Tensor, *params: Optional[Tensor]):
        pd = dict(zip(_PARAM_ORDER, params))
        N, P = x.shape[0], pd["lookup"].shape[0]
        # the backward kernels are fp32: keep forward and backward on the same (exact) path
        path = cabi.PATH_FP32_SIMT
        eager = _eager()
        pats = (prepare_patterns if eager else prepare_patterns_op)(
            *[pd[k] for k in _PATTERN_SIDE], spec.flags, path, spec.d_assoc, spec.d_v, spec.update_steps, spec.ln_eps)
        res = (retrieve if eager else retrieve_op)(
            x, *pats, *[pd[k] for k in _QUERY_SIDE], spec.flags, path, spec.d_in, spec.d_out, spec.update_steps,
            spec.beta, spec.ln_eps, 0)
        out, retrieved, qhat, lse, argmax, _ = res
        ctx.set_materialize_grads(False)
        ctx.spec = spec
        ctx.present = [p is not None for p in params]
        ctx.save_for_backward(x, qhat, retrieved, lse, pats[0], pats[1], *[p for p in params if p is not None])
        ctx.mark_non_differentiable(argmax)
        return out, argmax

    @staticmethod
    def backward(ctx, grad_out, _grad_argmax):
        spec = ctx.spec
        if grad_out is None:
            return (None,) * (2 + len(ctx.present))
        if spec.update_steps != 0:
            raise NotImplementedError("hop_sm100: backward is implemented for update_steps_max == 0 "
                                      "(the setting every Hop-VAE config trains with)")
        saved = list(ctx.saved_tensors)
        x, qhat, retrieved, lse, khat, vhat = saved[:6]
        it = iter(saved[6:])
        pd = {k: (next(it) if present else None) for k, present in zip(_PARAM_ORDER, ctx.present)}
        grads = (retrieve_backward if _eager() else retrieve_backward_op)(
            grad_out.contiguous(), x, qhat, retrieved, lse, khat, vhat, *[pd[k] for k in _PARAM_ORDER], spec.flags,
            spec.d_out, spec.beta, spec.ln_eps)
        gd = dict(zip(_GRAD_ORDER, grads))
        outs = [None, gd["d_x"] if ctx.needs_input_grad[1] else None]
        for i, k in enumerate(_PARAM_ORDER):
            gk = gd[_GRAD_OF_PARAM[k]]
            outs.append(gk if (pd[k] is not None and ctx.needs_input_grad[2 + i] and gk is not None
                               and gk.numel()) else None)
        return tuple(outs)


# --------------------------------------------------------------------------- shard merge / out-projection
def merge_shards(retrieved: Tensor, lse_all: Tensor, rowmax_all: Tensor, argmax_all: Tensor,
                 shard: int) -> Tuple[Tensor, Tensor]:
    """In place: retrieved (N,Dv) *= exp(lse_all[shard] - logsumexp_g lse_all[g]); returns (global lse, global
    nearest-pattern index).  A SUM over shards of the scaled `retrieved` is the global read-out."""
    import ctypes as C
    lib = cabi.load()
    _req_cuda_f32(retrieved, "retrieved")
    _req_cuda_f32(lse_all, "lse_all")
    _req_cuda_f32(rowmax_all, "rowmax_all")
    G, N = lse_all.shape
    lse = torch.empty((N,), device=retrieved.device, dtype=torch.float32)
    arg = torch.empty((N,), device=retrieved.device, dtype=torch.int32)
    if argmax_all.dtype != torch.int32 or not argmax_all.is_contiguous():
        raise RuntimeError("hop_sm100: argmax_all must be contiguous int32")
    with torch.cuda.device(retrieved.device):
        cabi.check(lib.hop_sm100_merge_shards(N, retrieved.shape[1], G, shard, cabi.ptr(lse_all),
                                              cabi.ptr(rowmax_all), cabi.ptr(argmax_all), cabi.ptr(retrieved),
                                              cabi.ptr(lse), cabi.ptr(arg), _stream()), "hop_sm100_merge_shards")
    return lse, arg


def out_projection(retrieved: Tensor, wo: Tensor, bo: Optional[Tensor]) -> Tensor:
    """(N,Dv) . Wo^T + bo -> (N,Dout) with the library's kernel."""
    lib = cabi.load()
    for n, t in (("retrieved", retrieved), ("wo", wo), ("bo", bo)):
        _req_cuda_f32(t, n)
    out = torch.empty((retrieved.shape[0], wo.shape[0]), device=retrieved.device, dtype=torch.float32)
    with torch.cuda.device(retrieved.device):
        cabi.check(lib.hop_sm100_out_projection(retrieved.shape[0], retrieved.shape[1], wo.shape[0],
                                                cabi.ptr(retrieved), cabi.ptr(wo), cabi.ptr(bo), cabi.ptr(out),
                                                _stream()), "hop_sm100_out_projection")
    return out
