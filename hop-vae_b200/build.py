"""In-tree build of the C-ABI library `libhop_sm100.so` (sm_100a only).

    python hop-vae_b200/build.py            # incremental
    python hop-vae_b200/build.py --force

nvcc cross-compiles without a GPU; the resulting .so is git-ignored but travels
to the GPU box with the gpurun snapshot.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libhop_sm100.so")
SOURCES = ["hop_api.cu", "hop_prep.cu", "hop_fwd_simt.cu", "hop_bwd_simt.cu", "hop_merge.cu", "hop_fwd_tc.cu",
           "hop_repair.cu", "hop_bwd_tc.cu", "hop_gemm_tc.cu", "hop_chain.cu"]
HEADERS = [os.path.join(CSRC, "hop_common.cuh"), os.path.join(CSRC, "hop_tc_ptx.cuh"),
           os.path.join(HERE, "..", "include", "hop_sm100.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC"]


def _nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: the sm_100a library cannot be built")
    return exe


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False, trace: bool = False, tag: str = "", defines=()) -> str:
    """trace=True builds the debug variant libhop_sm100_trace.so (-DHOP_TC_TRACE: per-tile clock64 timeline of
    one CTA of the fused forward; never loaded by the product, see scripts/tc_timeline.py)."""
    # `tag` + `defines`: tuning variants (libhop_sm100_<tag>.so), loaded through HOP_SM100_LIB by scripts only
    suffix = ("_trace" if trace else "") + (("_" + tag) if tag else "")
    obj_dir = OBJ + suffix
    lib = LIB.replace(".so", suffix + ".so") if suffix else LIB
    os.makedirs(obj_dir, exist_ok=True)
    nvcc = _nvcc()
    objs, jobs = [], []
    flags = NVCC_FLAGS + (["-DHOP_TC_TRACE"] if trace else []) + ["-D" + d for d in defines]
    for src in SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(obj_dir, src.replace(".cu", ".o"))
        objs.append(o)
        if force or _stale(o, [s] + HEADERS):
            jobs.append([nvcc] + flags + ["-c", s, "-o", o])

    def run(cmd):
        if verbose:
            print(" ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)

    if jobs:
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as ex:
            list(ex.map(run, jobs))
    if force or jobs or _stale(lib, objs):
        run([nvcc, "-shared", "-o", lib] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"])
    return lib


if __name__ == "__main__":
    _tag = next((a.split("=", 1)[1] for a in sys.argv if a.startswith("--tag=")), "")
    _defs = [a[2:] for a in sys.argv if a.startswith("-D")]
    print(build(force="--force" in sys.argv, verbose="--quiet" not in sys.argv, trace="--trace" in sys.argv, tag=_tag,
                defines=_defs))
