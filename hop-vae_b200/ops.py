"""PyTorch custom ops over the C-ABI library (`hop_sm100::prepare_patterns`,
`hop_sm100::retrieve`, `hop_sm100::retrieve_backward`) and the autograd
function that ties them into one differentiable Hopfield association.

PyTorch is plumbing here: it owns device memory and the stream; all arithmetic
runs in `libhop_sm100.so`.  CUDA tensors only -- there is no CPU path.
"""
from __future__ import annotations

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
    update_eps: float = 0.0     # upstream `update_steps_eps` (honoured by the fp32 path; 0 = exactly T updates)
    tune_splits: int = 0        # hop_desc.tune_splits / tune_flags: scheduling overrides for tests and experiments
    tune_flags: int = 0


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
    # (`key` holds every descriptor field the size depends on, the tuning fields included)
    k = (key, for_backward)
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
          beta: float, ln_eps: float, pattern_offset: int = 0, update_eps: float = 0.0, rows_per_item: int = 0,
          tune_splits: int = 0, tune_flags: int = 0) -> cabi.HopDesc:
    return cabi.make_desc(flags=flags, path=path, n_rows=n, n_patterns=p, d_in=d_in, d_assoc=d_assoc, d_v=d_v,
                          d_out=d_out, update_steps=t, beta=beta, ln_eps=ln_eps, pattern_offset=pattern_offset,
                          update_eps=update_eps, rows_per_item=rows_per_item, tune_splits=tune_splits,
                          tune_flags=tune_flags)


def resolve_path(spec: LayerSpec, n_rows: int, n_patterns: int) -> int:
    import ctypes as C
    d = _desc(spec.flags, spec.path, n_rows, n_patterns, spec.d_in, spec.d_assoc, spec.d_v, max(spec.d_out, 1),
              spec.update_steps, spec.beta, spec.ln_eps, 0, spec.update_eps, 0, spec.tune_splits, spec.tune_flags)
    out = C.c_int32(0)
    cabi.check(cabi.load().hop_sm100_select_path(C.byref(d), C.byref(out)), "hop_sm100_select_path")
    return int(out.value)


# --------------------------------------------------------------------------- prepare_patterns
N_PACKS = len(cabi._PATTERN_FIELDS)     # khat, vhat, k16, kt16, vt16, vot16, k16lo, kt16lo, v16, kstats, scratch


def prepare_patterns(lookup: Tensor, lookup_targets: Optional[Tensor], ln_stored_w: Optional[Tensor],
                     ln_stored_b: Optional[Tensor], ln_proj_w: Optional[Tensor], ln_proj_b: Optional[Tensor],
                     wk: Optional[Tensor], bk: Optional[Tensor], wv: Optional[Tensor], bv: Optional[Tensor],
                     wo: Optional[Tensor], flags: int, path: int, d_assoc: int, d_v: int, update_steps: int,
                     ln_eps: float, tune_flags: int = 0) -> List[Tensor]:
    """lookup (P, Din) -> the hop_patterns buffers in the order of `cabi._PATTERN_FIELDS`: khat (P,H), vhat (P,Dv)
    fp32, then the 16-bit packs (int16 storage; empty on the fp32 path or where `hop_sm100_pattern_bytes` says 0:
    vot16 = out_proj folded into the read-out operand when Dout == H == Dv, k16lo / kt16lo = low halves for the
    hi + lo update passes of T >= 2, v16 / kt16 for the tensor-core backward under F_TRAIN) and kstats (4 floats).
    `path` must already be resolved (not AUTO)."""
    import ctypes as C
    lib = cabi.load()
    for name, t in (("lookup", lookup), ("lookup_targets", lookup_targets), ("ln_stored_w", ln_stored_w),
                    ("ln_stored_b", ln_stored_b), ("ln_proj_w", ln_proj_w), ("ln_proj_b", ln_proj_b), ("wk", wk),
                    ("bk", bk), ("wv", wv), ("bv", bv), ("wo", wo)):
        _req_cuda_f32(t, name)
    P, d_in = lookup.shape
    d_out = wo.shape[0] if (wo is not None and (flags & cabi.F_OUT_PROJ)) else 1
    d = _desc(flags, path, 1, P, d_in, d_assoc, d_v, d_out, update_steps, 1.0, ln_eps, 0,
              1.0 if update_steps < 0 else 0.0, 0, 0, tune_flags)      # (T < 0 is only valid together with a stop rule)
    dev = lookup.device
    sizes = cabi.HopPatternSizes()
    cabi.check(lib.hop_sm100_pattern_bytes(C.byref(d), C.byref(sizes)), "hop_sm100_pattern_bytes")
    pats: List[Tensor] = []
    for name in cabi._PATTERN_FIELDS:
        nbytes = int(getattr(sizes, name))
        if name == "khat":
            pats.append(torch.empty((P, d_assoc), device=dev, dtype=torch.float32))
        elif name == "vhat":
            pats.append(torch.empty((P, d_v), device=dev, dtype=torch.float32))
        elif name == "kstats":
            pats.append(torch.empty((nbytes // 4,), device=dev, dtype=torch.float32))
        else:
            pats.append(torch.empty((nbytes // 2,), device=dev, dtype=torch.int16))
    params = cabi.HopParams(lookup=cabi.ptr(lookup), lookup_targets=cabi.ptr(lookup_targets),
                            ln_stored_w=cabi.ptr(ln_stored_w), ln_stored_b=cabi.ptr(ln_stored_b),
                            ln_proj_w=cabi.ptr(ln_proj_w), ln_proj_b=cabi.ptr(ln_proj_b), wk=cabi.ptr(wk),
                            bk=cabi.ptr(bk), wv=cabi.ptr(wv), bv=cabi.ptr(bv), wo=cabi.ptr(wo))
    ps = cabi.patterns_struct(pats)
    with _on_device(dev):
        cabi.check(lib.hop_sm100_prepare_patterns(C.byref(d), C.byref(params), C.byref(ps), _stream(dev)),
                   "hop_sm100_prepare_patterns")
    pats[-1] = pats[-1].new_empty((0,))     # the scratch of the preparation is not kept (stream-ordered free)
    return pats


@torch.library.custom_op("hop_sm100::prepare_patterns", mutates_args=(), device_types="cuda")
def prepare_patterns_op(lookup: Tensor, lookup_targets: Optional[Tensor], ln_stored_w: Optional[Tensor],
                        ln_stored_b: Optional[Tensor], ln_proj_w: Optional[Tensor], ln_proj_b: Optional[Tensor],
                        wk: Optional[Tensor], bk: Optional[Tensor], wv: Optional[Tensor], bv: Optional[Tensor],
                        wo: Optional[Tensor], flags: int, path: int, d_assoc: int, d_v: int, update_steps: int,
                        ln_eps: float, tune_flags: int = 0) -> List[Tensor]:
    """torch.compile-visible wrapper of `prepare_patterns`."""
    return prepare_patterns(lookup, lookup_targets, ln_stored_w, ln_stored_b, ln_proj_w, ln_proj_b, wk, bk, wv, bv,
                            wo, flags, path, d_assoc, d_v, update_steps, ln_eps, tune_flags)


@prepare_patterns_op.register_fake
def _(lookup, lookup_targets, ln_stored_w, ln_stored_b, ln_proj_w, ln_proj_b, wk, bk, wv, bv, wo, flags, path,
      d_assoc, d_v, update_steps, ln_eps, tune_flags=0):
    P = lookup.shape[0]
    ppad = (P + 127) // 128 * 128
    tc = path in (cabi.PATH_TC_F16, cabi.PATH_TC_BF16)
    train = tc and bool(flags & cabi.F_TRAIN)
    precise = (path == cabi.PATH_TC_F16 and update_steps >= 2 and d_assoc <= 256
               and not (tune_flags & cabi.TUNE_PLAIN_UPDATES))
    mk = lambda n: lookup.new_empty((n,), dtype=torch.int16)
    fold = (tc and wo is not None and bool(flags & cabi.F_OUT_PROJ) and wo.shape[0] == d_assoc == d_v
            and d_assoc in (64, 128, 256, 512, 1024))
    return [lookup.new_empty((P, d_assoc)), lookup.new_empty((P, d_v)), mk(ppad * d_assoc if tc else 0),
            mk(ppad * d_assoc if tc and (update_steps > 0 or train) else 0), mk(ppad * d_v if tc else 0),
            mk(ppad * d_assoc if fold else 0), mk(ppad * d_assoc if precise else 0),
            mk(ppad * d_assoc if precise else 0), mk(ppad * d_v if train else 0),
            lookup.new_empty((4 if tc else 0,)), mk(0)]


# --------------------------------------------------------------------------- retrieve (forward)
# Diagnostic of exact_argmax (tests, bench): with `count_repairs` set, every tensor-core forward is followed by a
# (synchronising) read of how many rows failed the top-1 / top-2 certificate and were re-scored in fp32.
count_repairs = False
last_repair_rows = 0          # rows that took the full fp32 re-scoring pass
last_flagged_rows = 0         # rows that failed the on-chip certificate (most are settled by the 64-pattern block check)


def retrieve(x: Tensor, pats: List[Tensor], ln_state_w: Optional[Tensor], ln_state_b: Optional[Tensor],
             wq: Optional[Tensor], bq: Optional[Tensor], wo: Optional[Tensor], bo: Optional[Tensor], flags: int,
             path: int, d_in: int, d_out: int, update_steps: int, beta: float, ln_eps: float, pattern_offset: int,
             want_retrieved: bool = True, want_saved: bool = True,
             into: Optional[Tuple[Tensor, Tensor, Tensor]] = None, update_eps: float = 0.0, rows_per_item: int = 0,
             tune_splits: int = 0, tune_flags: int = 0) -> List[Tensor]:
    """x (N, Din), pats = `prepare_patterns(...)` ->
    [out (N,Dout|Dv), retrieved (N,Dv), qhat (N,H), lse (N), argmax (N) int32, rowmax (N)].
    With want_retrieved=False and a folded read-out operand (vot16) the tensor-core kernel writes `out`
    directly and `retrieved` is returned empty.  want_saved=False (inference) additionally skips everything only
    the backward needs: `qhat` is returned empty and, with an out-projection, so is `retrieved`.
    `into=(out (N,W), stats (2,N) f32, argmax (N) i32)` writes the results into caller-owned buffers (the
    peer-mapped exchange buffers of the pattern-sharded mode) instead of fresh tensors."""
    import ctypes as C
    lib = cabi.load()
    khat, vhat, vot16 = pats[0], pats[1], pats[5]
    for name, t in (("x", x), ("khat", khat), ("vhat", vhat), ("ln_state_w", ln_state_w), ("ln_state_b", ln_state_b),
                    ("wq", wq), ("bq", bq), ("wo", wo), ("bo", bo)):
        _req_cuda_f32(t, name)
    N = x.shape[0]
    P, H = khat.shape
    Dv = vhat.shape[1]
    out_proj = bool(flags & cabi.F_OUT_PROJ)
    key = (flags, path, N, P, d_in, H, Dv, d_out if out_proj else 1, update_steps, beta, ln_eps, pattern_offset,
           update_eps, rows_per_item, tune_splits, tune_flags)
    d = cabi.make_desc(flags=flags, path=path, n_rows=N, n_patterns=P, d_in=d_in, d_assoc=H, d_v=Dv,
                       d_out=d_out if out_proj else 1, update_steps=update_steps, beta=beta, ln_eps=ln_eps,
                       pattern_offset=pattern_offset, update_eps=update_eps, rows_per_item=rows_per_item,
                       tune_splits=tune_splits, tune_flags=tune_flags)
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
    use_ws = bool(flags & cabi.F_PROJ_STATE) or update_steps != 0 or tc
    need = _workspace_bytes(d, key, 0) if use_ws else 0
    ws = torch.empty((need,), device=dev, dtype=torch.uint8) if use_ws else None
    params = cabi.HopParams(ln_state_w=cabi.ptr(ln_state_w), ln_state_b=cabi.ptr(ln_state_b), wq=cabi.ptr(wq),
                            bq=cabi.ptr(bq), wo=cabi.ptr(wo), bo=cabi.ptr(bo))
    ps = cabi.patterns_struct(pats)
    with _on_device(dev):
        cabi.check(lib.hop_sm100_forward(C.byref(d), C.byref(params), C.byref(ps), cabi.ptr(x), cabi.ptr(out),
                                         cabi.ptr(retrieved), cabi.ptr(qhat), cabi.ptr(lse), cabi.ptr(argmax),
                                         cabi.ptr(rowmax), cabi.ptr(ws), need, _stream(dev)), "hop_sm100_forward")
        if count_repairs and tc and (flags & cabi.F_EXACT_ARGMAX):
            global last_repair_rows, last_flagged_rows
            rows, flagged = C.c_int32(0), C.c_int32(0)
            cabi.check(lib.hop_sm100_repair_rows(C.byref(d), cabi.ptr(ws), _stream(dev), C.byref(rows), C.byref(flagged)),
                       "hop_sm100_repair_rows")
            last_repair_rows, last_flagged_rows = int(rows.value), int(flagged.value)
    if qhat is None:
        qhat = _empty_f32(dev)
    if retrieved is None:
        retrieved = _empty_f32(dev) if not want_saved else torch.empty((0, Dv), device=dev, dtype=torch.float32)
    elif not out_proj:
        retrieved = out.clone() if want_saved else out
    return [out, retrieved, qhat, lse, argmax, rowmax]


@torch.library.custom_op("hop_sm100::retrieve", mutates_args=(), device_types="cuda")
def retrieve_op(x: Tensor, pats: List[Tensor], ln_state_w: Optional[Tensor], ln_state_b: Optional[Tensor],
                wq: Optional[Tensor], bq: Optional[Tensor], wo: Optional[Tensor], bo: Optional[Tensor], flags: int,
                path: int, d_in: int, d_out: int, update_steps: int, beta: float, ln_eps: float,
                pattern_offset: int, want_retrieved: bool = True, update_eps: float = 0.0, rows_per_item: int = 0,
                tune_splits: int = 0, tune_flags: int = 0) -> List[Tensor]:
    """torch.compile-visible wrapper of `retrieve` (always returns what the backward needs)."""
    res = retrieve(x, pats, ln_state_w, ln_state_b, wq, bq, wo, bo, flags, path, d_in, d_out, update_steps, beta,
                   ln_eps, pattern_offset, want_retrieved, True, None, update_eps, rows_per_item, tune_splits,
                   tune_flags)
    # custom-op outputs may not alias each other
    return [res[0], res[1], res[2], res[3].clone(), res[4], res[5].clone()]


@retrieve_op.register_fake
def _(x, pats, ln_state_w, ln_state_b, wq, bq, wo, bo, flags, path, d_in, d_out, update_steps, beta, ln_eps,
      pattern_offset, want_retrieved=True, update_eps=0.0, rows_per_item=0, tune_splits=0, tune_flags=0):
    N = x.shape[0]
    H, Dv = pats[0].shape[1], pats[1].shape[1]
    out_w = d_out if (flags & cabi.F_OUT_PROJ) else Dv
    folded = (bool(flags & cabi.F_OUT_PROJ) and not want_retrieved and pats[5].numel() > 0
              and path != cabi.PATH_FP32_SIMT)
    return [x.new_empty((N, out_w)), x.new_empty((0 if folded else N, Dv)), x.new_empty((N, H)), x.new_empty((N,)),
            x.new_empty((N,), dtype=torch.int32), x.new_empty((N,))]


# --------------------------------------------------------------------------- retrieve backward
_GRAD_ORDER = cabi._GRAD_FIELDS  # order of the returned list


def resolve_train_path(spec: LayerSpec, n_rows: int, n_patterns: int) -> int:
    """Kernel family a training step (forward + backward) runs on: the tcgen05 kernels where the tensor-core backward
    applies, else the exact fp32 kernels (H, Dv <= 256); raises NotImplementedError where neither does."""
    import ctypes as C
    d = _desc(spec.flags, spec.path, n_rows, n_patterns, spec.d_in, spec.d_assoc, spec.d_v, max(spec.d_out, 1),
              spec.update_steps, spec.beta, spec.ln_eps, 0, spec.update_eps, 0, spec.tune_splits, spec.tune_flags)
    out = C.c_int32(0)
    cabi.check(cabi.load().hop_sm100_select_train_path(C.byref(d), C.byref(out)), "hop_sm100_select_train_path")
    return int(out.value)


def retrieve_backward(grad_out: Tensor, x: Tensor, qhat: Tensor, retrieved: Tensor, lse: Tensor, argmax: Tensor,
                      pats: List[Tensor], lookup: Tensor, lookup_targets: Optional[Tensor],
                      ln_state_w: Optional[Tensor], ln_state_b: Optional[Tensor], ln_stored_w: Optional[Tensor],
                      ln_stored_b: Optional[Tensor], ln_proj_w: Optional[Tensor], ln_proj_b: Optional[Tensor],
                      wq: Optional[Tensor], bq: Optional[Tensor], wk: Optional[Tensor], bk: Optional[Tensor],
                      wv: Optional[Tensor], bv: Optional[Tensor], wo: Optional[Tensor], bo: Optional[Tensor],
                      flags: int, path: int, d_out: int, beta: float, ln_eps: float,
                      tune_flags: int = 0) -> List[Optional[Tensor]]:
    """Gradients in the order of hop_grads (None where the parameter is absent).  `path` = the path the forward of
    this step ran on (`resolve_train_path`): it selects the tcgen05 or the fp32 association core."""
    import ctypes as C
    lib = cabi.load()
    _req_cuda_f32(grad_out, "grad_out")
    khat, vhat = pats[0], pats[1]
    N, d_in = x.shape
    P, H = khat.shape
    Dv = vhat.shape[1]
    out_proj = bool(flags & cabi.F_OUT_PROJ)
    key = (flags, path, N, P, d_in, H, Dv, d_out if out_proj else 1, 0, beta, ln_eps, 0, 0.0, 0, 0, tune_flags)
    d = _desc(flags, path, N, P, d_in, H, Dv, d_out if out_proj else 1, 0, beta, ln_eps, 0, 0.0, 0, 0, tune_flags)
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
    ps = cabi.patterns_struct(pats)
    grads = cabi.HopGrads(**{k: (cabi.ptr(v) if v is not None and v.numel() else None) for k, v in g.items()})
    go = grad_out.contiguous()
    with _on_device(dev):
        cabi.check(lib.hop_sm100_backward(C.byref(d), C.byref(params), C.byref(ps), cabi.ptr(x), cabi.ptr(qhat),
                                          cabi.ptr(retrieved), cabi.ptr(lse), cabi.ptr(argmax), cabi.ptr(go),
                                          C.byref(grads),
                                          cabi.ptr(ws), need, _stream(dev)), "hop_sm100_backward")
    return [g[k] for k in _GRAD_ORDER]


@torch.library.custom_op("hop_sm100::retrieve_backward", mutates_args=(), device_types="cuda")
def retrieve_backward_op(grad_out: Tensor, x: Tensor, qhat: Tensor, retrieved: Tensor, lse: Tensor, argmax: Tensor,
                         pats: List[Tensor], lookup: Tensor, lookup_targets: Optional[Tensor],
                         ln_state_w: Optional[Tensor], ln_state_b: Optional[Tensor], ln_stored_w: Optional[Tensor],
                         ln_stored_b: Optional[Tensor], ln_proj_w: Optional[Tensor], ln_proj_b: Optional[Tensor],
                         wq: Optional[Tensor], bq: Optional[Tensor], wk: Optional[Tensor], bk: Optional[Tensor],
                         wv: Optional[Tensor], bv: Optional[Tensor], wo: Optional[Tensor], bo: Optional[Tensor],
                         flags: int, path: int, d_out: int, beta: float, ln_eps: float,
                         tune_flags: int = 0) -> List[Tensor]:
    """torch.compile-visible wrapper of `retrieve_backward` (empty tensors where the parameter is absent)."""
    res = retrieve_backward(grad_out, x, qhat, retrieved, lse, argmax, pats, lookup, lookup_targets, ln_state_w,
                            ln_state_b, ln_stored_w, ln_stored_b, ln_proj_w, ln_proj_b, wq, bq, wk, bk, wv, bv, wo,
                            bo, flags, path, d_out, beta, ln_eps, tune_flags)
    return [r if r is not None else torch.empty((0,), device=x.device) for r in res]


@retrieve_backward_op.register_fake
def _(grad_out, x, qhat, retrieved, lse, argmax, pats, lookup, lookup_targets, ln_state_w, ln_state_b, ln_stored_w,
      ln_stored_b, ln_proj_w, ln_proj_b, wq, bq, wk, bk, wv, bv, wo, bo, flags, path, d_out, beta, ln_eps,
      tune_flags=0):
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

_TRAIN_PATHS = {}


class HopfieldAssociation(torch.autograd.Function):
    """Differentiable  x (N,Din), params  ->  out (N,Dout).  Training path: patterns are
    prepared inside (parameters change every step); backward recomputes A flash-style.
    Returns (out, nearest-pattern index, lse)."""

    @staticmethod
    def forward(ctx, spec: LayerSpec, x: Tensor, *params: Optional[Tensor]):
        pd = dict(zip(_PARAM_ORDER, params))
        N, P = x.shape[0], pd["lookup"].shape[0]
        if spec.update_steps != 0:
            raise NotImplementedError("hop_sm100: the backward is built for update_steps_max == 0 (the setting every "
                                      "Hop-VAE config trains with); wrap inference in torch.no_grad()")
        # forward and backward of one step run on the same kernel family: tcgen05 where the tensor-core backward
        # applies (large N x P, D in {128, 256, 512}), else the exact fp32 kernels
        pkey = (spec, N, P)
        path = _TRAIN_PATHS.get(pkey)
        if path is None:
            path = _TRAIN_PATHS[pkey] = resolve_train_path(spec, N, P)
        flags = spec.flags | (cabi.F_TRAIN if path != cabi.PATH_FP32_SIMT else 0)
        eager = _eager()
        pats = (prepare_patterns if eager else prepare_patterns_op)(
            *[pd[k] for k in _PATTERN_SIDE], flags, path, spec.d_assoc, spec.d_v, spec.update_steps, spec.ln_eps,
            spec.tune_flags)
        res = (retrieve if eager else retrieve_op)(
            x, pats, *[pd[k] for k in _QUERY_SIDE], flags, path, spec.d_in, spec.d_out, spec.update_steps,
            spec.beta, spec.ln_eps, 0, True)
        out, retrieved, qhat, lse, argmax, _ = res
        ctx.set_materialize_grads(False)
        ctx.spec = spec
        ctx.path, ctx.flags = path, flags
        ctx.present = [p is not None for p in params]
        ctx.save_for_backward(x, qhat, retrieved, lse, argmax, *pats, *[p for p in params if p is not None])
        ctx.mark_non_differentiable(argmax, lse)
        return out, argmax, lse

    @staticmethod
    def backward(ctx, grad_out, _grad_argmax, _grad_lse):
        spec = ctx.spec
        if grad_out is None:
            return (None,) * (2 + len(ctx.present))
        saved = list(ctx.saved_tensors)
        x, qhat, retrieved, lse, argmax = saved[:5]
        pats = saved[5:5 + N_PACKS]
        it = iter(saved[5 + N_PACKS:])
        pd = {k: (next(it) if present else None) for k, present in zip(_PARAM_ORDER, ctx.present)}
        grads = (retrieve_backward if _eager() else retrieve_backward_op)(
            grad_out.contiguous(), x, qhat, retrieved, lse, argmax, pats, *[pd[k] for k in _PARAM_ORDER], ctx.flags,
            ctx.path, spec.d_out, spec.beta, spec.ln_eps, spec.tune_flags)
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
