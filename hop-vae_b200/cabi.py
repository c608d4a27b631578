"""ctypes binding of `include/hop_sm100.h` (the C-ABI drop-in boundary).

The library is loaded from the in-tree `libhop_sm100.so`.  There is no CPU
fallback: if the library is missing `load()` raises, and every wrapper raises
`RuntimeError` with the library's own message on a non-zero status.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("HOP_SM100_LIB") or os.path.join(HERE, "libhop_sm100.so")   # env: debug builds only

ABI_VERSION = 3

# hop_desc.flags
F_NORM_STATE, F_NORM_STORED, F_NORM_PROJ = 1 << 0, 1 << 1, 1 << 2
F_PROJ_STATE, F_PROJ_STORED, F_PROJ_PATTERN, F_OUT_PROJ = 1 << 3, 1 << 4, 1 << 5, 1 << 6
F_EXACT_ARGMAX, F_TRAIN = 1 << 7, 1 << 8
# hop_desc.tune_flags (tests / experiments; 0 = the library's heuristics)
TUNE_STREAMK_ON, TUNE_STREAMK_OFF, TUNE_ROWS128, TUNE_MULTI_ON, TUNE_MULTI_OFF = 1 << 0, 1 << 1, 1 << 2, 1 << 3, 1 << 4
TUNE_PAIR_ON, TUNE_PAIR_OFF, TUNE_PLAIN_UPDATES, TUNE_BWD_SIMT, TUNE_PREP_SIMT = 1 << 5, 1 << 6, 1 << 7, 1 << 8, 1 << 9
TUNE_BWD_TAILS_SIMT = 1 << 10
# hop_desc.path
PATH_AUTO, PATH_FP32_SIMT, PATH_TC_F16, PATH_TC_BF16 = 0, 1, 2, 3
PATH_NAMES = {"auto": PATH_AUTO, "fp32": PATH_FP32_SIMT, "f16": PATH_TC_F16, "bf16": PATH_TC_BF16}

STATUS_NAMES = {1: "HOP_ERR_BAD_ARG", 2: "HOP_ERR_BAD_SHAPE", 3: "HOP_ERR_BAD_ALIGN", 4: "HOP_ERR_BAD_ARCH",
                5: "HOP_ERR_CUDA", 6: "HOP_ERR_WORKSPACE", 7: "HOP_ERR_UNSUPPORTED"}

_fp = C.c_void_p  # device pointers travel as integers


class HopDesc(C.Structure):
    _fields_ = [("abi_version", C.c_uint32), ("flags", C.c_uint32), ("path", C.c_int32),
                ("n_rows", C.c_int32), ("n_patterns", C.c_int32), ("d_in", C.c_int32),
                ("d_assoc", C.c_int32), ("d_v", C.c_int32), ("d_out", C.c_int32),
                ("update_steps", C.c_int32), ("pattern_offset", C.c_int32),
                ("beta", C.c_float), ("ln_eps", C.c_float), ("update_eps", C.c_float),
                ("rows_per_item", C.c_int32), ("tune_splits", C.c_int32), ("tune_flags", C.c_uint32)]


_PARAM_FIELDS = ["lookup", "lookup_targets", "ln_state_w", "ln_state_b", "ln_stored_w", "ln_stored_b",
                 "ln_proj_w", "ln_proj_b", "wq", "bq", "wk", "bk", "wv", "bv", "wo", "bo"]
_PATTERN_FIELDS = ["khat", "vhat", "k16", "kt16", "vt16", "vot16", "k16lo", "kt16lo", "v16", "kstats", "scratch"]
_GRAD_FIELDS = ["d_x", "d_lookup", "d_lookup_targets", "d_ln_state_w", "d_ln_state_b", "d_ln_stored_w",
                "d_ln_stored_b", "d_ln_proj_w", "d_ln_proj_b", "d_wq", "d_bq", "d_wk", "d_bk", "d_wv", "d_bv",
                "d_wo", "d_bo"]


class HopParams(C.Structure):
    _fields_ = [(n, _fp) for n in _PARAM_FIELDS]


class HopPatterns(C.Structure):
    _fields_ = [(n, _fp) for n in _PATTERN_FIELDS]


class HopPatternSizes(C.Structure):
    _fields_ = [(n, C.c_size_t) for n in _PATTERN_FIELDS]


class HopGrads(C.Structure):
    _fields_ = [(n, _fp) for n in _GRAD_FIELDS]


MAX_PEERS = 16
PEER_FLAG_WORDS = 64


class HopPeerMerge(C.Structure):
    _fields_ = ([("n_rows", C.c_int32), ("d_v", C.c_int32), ("n_ranks", C.c_int32), ("rank", C.c_int32),
                 ("row_begin", C.c_int32), ("row_count", C.c_int32), ("push_all", C.c_int32), ("epoch", C.c_uint32)]
                + [(n, _fp * MAX_PEERS) for n in ("o", "lse", "rowmax", "argmax", "out", "lse_out", "argmax_out",
                                                  "flags")])


EXPORTS = ["hop_sm100_peer_merge", "hop_sm100_enable_peer_access", "hop_sm100_ipc_export", "hop_sm100_ipc_import",
           "hop_sm100_ipc_close", "hop_sm100_select_train_path", "hop_sm100_repair_rows", "hop_sm100_bottleneck_forward",
           "hop_sm100_bottleneck_workspace_bytes", "hop_sm100_bottleneck_forward_static",
           "hop_sm100_bottleneck_static_max_rows", "hop_sm100_version", "hop_sm100_last_error", "hop_sm100_select_path", "hop_sm100_pattern_bytes",
           "hop_sm100_workspace_bytes", "hop_sm100_prepare_patterns", "hop_sm100_forward", "hop_sm100_backward",
           "hop_sm100_merge_shards", "hop_sm100_out_projection", "hop_sm100_launch_count"]

_lib: Optional[C.CDLL] = None


def load() -> C.CDLL:
    """Load the C-ABI library (once).  Fails loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: the sm_100a Hopfield kernels are not built. Run "
            f"`python hop-vae_b200/build.py` (or `__graft_entry__.build()`); there is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    P = C.POINTER
    lib.hop_sm100_version.restype = C.c_uint32
    lib.hop_sm100_version.argtypes = []
    lib.hop_sm100_last_error.restype = C.c_char_p
    lib.hop_sm100_last_error.argtypes = []
    lib.hop_sm100_launch_count.restype = C.c_uint64
    lib.hop_sm100_launch_count.argtypes = []
    lib.hop_sm100_select_path.restype = C.c_int
    lib.hop_sm100_select_path.argtypes = [P(HopDesc), P(C.c_int32)]
    lib.hop_sm100_bottleneck_workspace_bytes.restype = C.c_int
    lib.hop_sm100_bottleneck_workspace_bytes.argtypes = [C.c_int32, P(C.c_size_t)]
    lib.hop_sm100_bottleneck_forward.restype = C.c_int
    lib.hop_sm100_bottleneck_forward.argtypes = [P(HopDesc), P(HopParams), P(HopPatterns), C.c_int32] + [C.c_void_p] * 7 + [
        C.c_size_t, C.c_void_p]
    lib.hop_sm100_bottleneck_forward_static.restype = C.c_int
    lib.hop_sm100_bottleneck_forward_static.argtypes = [P(HopDesc), P(HopParams), P(HopPatterns), C.c_int32] + [C.c_void_p] * 7
    lib.hop_sm100_bottleneck_static_max_rows.restype = C.c_int
    lib.hop_sm100_bottleneck_static_max_rows.argtypes = []
    lib.hop_sm100_repair_rows.restype = C.c_int
    lib.hop_sm100_repair_rows.argtypes = [P(HopDesc), C.c_void_p, C.c_void_p, P(C.c_int32), P(C.c_int32)]
    lib.hop_sm100_select_train_path.restype = C.c_int
    lib.hop_sm100_select_train_path.argtypes = [P(HopDesc), P(C.c_int32)]
    lib.hop_sm100_pattern_bytes.restype = C.c_int
    lib.hop_sm100_pattern_bytes.argtypes = [P(HopDesc), P(HopPatternSizes)]
    lib.hop_sm100_workspace_bytes.restype = C.c_int
    lib.hop_sm100_workspace_bytes.argtypes = [P(HopDesc), C.c_int, P(C.c_size_t)]
    lib.hop_sm100_prepare_patterns.restype = C.c_int
    lib.hop_sm100_prepare_patterns.argtypes = [P(HopDesc), P(HopParams), P(HopPatterns), _fp]
    lib.hop_sm100_forward.restype = C.c_int
    lib.hop_sm100_forward.argtypes = [P(HopDesc), P(HopParams), P(HopPatterns), _fp, _fp, _fp, _fp, _fp, _fp, _fp,
                                      _fp, C.c_size_t, _fp]
    lib.hop_sm100_backward.restype = C.c_int
    lib.hop_sm100_backward.argtypes = [P(HopDesc), P(HopParams), P(HopPatterns), _fp, _fp, _fp, _fp, _fp, _fp,
                                       P(HopGrads), _fp, C.c_size_t, _fp]
    lib.hop_sm100_merge_shards.restype = C.c_int
    lib.hop_sm100_merge_shards.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_int32, _fp, _fp, _fp, _fp, _fp, _fp,
                                           _fp]
    lib.hop_sm100_out_projection.restype = C.c_int
    lib.hop_sm100_out_projection.argtypes = [C.c_int32, C.c_int32, C.c_int32, _fp, _fp, _fp, _fp, _fp]
    lib.hop_sm100_peer_merge.restype = C.c_int
    lib.hop_sm100_peer_merge.argtypes = [P(HopPeerMerge), _fp]
    lib.hop_sm100_enable_peer_access.restype = C.c_int
    lib.hop_sm100_enable_peer_access.argtypes = [C.c_int32]
    lib.hop_sm100_ipc_export.restype = C.c_int
    lib.hop_sm100_ipc_export.argtypes = [_fp, C.c_char_p, P(C.c_uint64)]
    lib.hop_sm100_ipc_import.restype = C.c_int
    lib.hop_sm100_ipc_import.argtypes = [C.c_char_p, C.c_uint64, P(_fp)]
    lib.hop_sm100_ipc_close.restype = C.c_int
    lib.hop_sm100_ipc_close.argtypes = [_fp, C.c_uint64]
    got = lib.hop_sm100_version()
    if got != ABI_VERSION:
        raise RuntimeError(f"libhop_sm100.so ABI version {got} != binding {ABI_VERSION}; rebuild the library")
    _lib = lib
    return lib


def check(status: int, what: str) -> None:
    if status != 0:
        msg = load().hop_sm100_last_error().decode("utf-8", "replace")
        name = STATUS_NAMES.get(status, str(status))
        if status == 7:
            raise NotImplementedError(f"{what}: {name}: {msg}")
        raise RuntimeError(f"{what}: {name}: {msg}")


def launch_count() -> int:
    return int(load().hop_sm100_launch_count())


def make_desc(*, flags: int, path: int, n_rows: int, n_patterns: int, d_in: int, d_assoc: int, d_v: int,
              d_out: int, update_steps: int, beta: float, ln_eps: float = 1e-5, pattern_offset: int = 0,
              update_eps: float = 0.0, rows_per_item: int = 0, tune_splits: int = 0, tune_flags: int = 0) -> HopDesc:
    return HopDesc(ABI_VERSION, flags, path, n_rows, n_patterns, d_in, d_assoc, d_v, d_out, update_steps,
                   pattern_offset, beta, ln_eps, update_eps, rows_per_item, tune_splits, tune_flags)


def patterns_struct(pats) -> "HopPatterns":
    """hop_patterns from the tensors of `ops.prepare_patterns` (order of _PATTERN_FIELDS; empty tensor -> NULL)."""
    return HopPatterns(**{n: (t.data_ptr() if t is not None and t.numel() else None)
                          for n, t in zip(_PATTERN_FIELDS, pats)})


def ptr(t) -> Optional[int]:
    """Device pointer of a tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()
