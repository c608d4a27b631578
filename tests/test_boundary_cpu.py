"""CPU: the C-ABI library loads and exports every symbol include/hop_sm100.h declares,
the ctypes structs match the C layout (checked against a gcc-compiled probe), and the
reference-facing `hflayers.HopfieldLayer` mirrors the upstream interface.  No compute."""
import ctypes as C
import inspect
import os
import re
import subprocess
import sys

import pytest
import torch

from util import ROOT, load_ckpt, layer_state

HEADER = os.path.join(ROOT, "include", "hop_sm100.h")


@pytest.fixture(scope="module")
def impl():
    sys.path.insert(0, ROOT)
    if not os.path.exists(os.path.join(ROOT, "hop-vae_b200", "libhop_sm100.so")):
        import __graft_entry__
        __graft_entry__.build()
    import hflayers
    return hflayers._impl


def test_library_exports_every_declared_symbol(impl):
    text = open(HEADER).read()
    declared = sorted(set(re.findall(r"\b(hop_sm100_\w+)\s*\(", text)))
    assert len(declared) >= 11
    lib = impl.cabi.load()
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/hop_sm100.h but not exported"
    assert sorted(impl.cabi.EXPORTS) == declared
    assert lib.hop_sm100_version() == impl.cabi.ABI_VERSION == 2


def test_struct_layouts_match_c(impl, tmp_path):
    src = tmp_path / "probe.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "hop_sm100.h"\n'
                   'int main(void){printf("%zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(hop_desc), sizeof(hop_params),'
                   ' sizeof(hop_patterns), sizeof(hop_grads), offsetof(hop_desc, beta), offsetof(hop_desc, n_patterns),'
                   ' offsetof(hop_params, wo), offsetof(hop_grads, d_wo)); return 0;}\n')
    exe = tmp_path / "probe"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    got = [int(v) for v in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()]
    cb = impl.cabi
    want = [C.sizeof(cb.HopDesc), C.sizeof(cb.HopParams), C.sizeof(cb.HopPatterns), C.sizeof(cb.HopGrads),
            cb.HopDesc.beta.offset, cb.HopDesc.n_patterns.offset, cb.HopParams.wo.offset, cb.HopGrads.d_wo.offset]
    assert got == want


def test_error_reporting_without_gpu(impl):
    """Argument validation happens before any CUDA call: status code + thread-local message."""
    cb = impl.cabi
    lib = cb.load()
    d = cb.make_desc(flags=0, path=0, n_rows=4, n_patterns=0, d_in=8, d_assoc=8, d_v=8, d_out=8, update_steps=0,
                     beta=0.3)
    out = C.c_int32(0)
    st = lib.hop_sm100_select_path(C.byref(d), C.byref(out))
    assert st == 2 and b"non-positive" in lib.hop_sm100_last_error()
    d.n_patterns = 16
    d.abi_version = 99
    assert lib.hop_sm100_select_path(C.byref(d), C.byref(out)) == 1
    with pytest.raises(RuntimeError, match="HOP_ERR_BAD_ARG"):
        cb.check(1, "probe")
    d.abi_version = cb.ABI_VERSION
    assert lib.hop_sm100_select_path(C.byref(d), C.byref(out)) == 0 and out.value == cb.PATH_FP32_SIMT
    # the reference's own shapes stay on the exact fp32 kernels; the big sweep goes to tcgen05
    big = cb.make_desc(flags=0, path=0, n_rows=16384, n_patterns=65536, d_in=256, d_assoc=256, d_v=256, d_out=256,
                       update_steps=0, beta=1 / 16)
    assert lib.hop_sm100_select_path(C.byref(big), C.byref(out)) == 0
    assert out.value in (cb.PATH_TC_F16, cb.PATH_FP32_SIMT)
    sizes = [C.c_size_t(0) for _ in range(6)]
    assert lib.hop_sm100_pattern_bytes(C.byref(big), *[C.byref(s) for s in sizes]) == 0
    assert sizes[0].value == 65536 * 256 * 4
    assert sizes[5].value == 0                      # no HOP_F_OUT_PROJ -> nothing to fold
    big.flags = cb.F_OUT_PROJ
    assert lib.hop_sm100_pattern_bytes(C.byref(big), *[C.byref(s) for s in sizes]) == 0
    assert sizes[5].value == (65536 * 256 * 2 if out.value == cb.PATH_TC_F16 else 0)


def test_wide_widths_select_the_tensor_path_and_size_their_workspace(impl, monkeypatch):
    """Host logic of BASELINE config 5's wide widths, no GPU needed (the SM count falls back to 148): D = 512 and
    1024 resolve to the tcgen05 path (the fp32 kernels stop at 256), unsupported widths are refused there, and the
    forward workspace covers the state buffers plus the split partials under both kernel forms."""
    cb = impl.cabi
    lib = cb.load()
    out = C.c_int32(0)
    for d_ in (512, 1024):
        d = cb.make_desc(flags=0, path=0, n_rows=300, n_patterns=2048, d_in=d_, d_assoc=d_, d_v=d_, d_out=d_,
                         update_steps=0, beta=d_ ** -0.5)
        assert lib.hop_sm100_select_path(C.byref(d), C.byref(out)) == 0 and out.value == cb.PATH_TC_F16
        sizes = [C.c_size_t(0) for _ in range(6)]
        assert lib.hop_sm100_pattern_bytes(C.byref(d), *[C.byref(s) for s in sizes]) == 0
        assert sizes[2].value == 2048 * d_ * 2 and sizes[4].value == 2048 * d_ * 2 and sizes[3].value == 0
        state = 3 * ((300 * d_ * 4 + 255) // 256 * 256)          # two state buffers + one read-out buffer
        got = {}
        for halfm in ("1", "0"):
            monkeypatch.setenv("HOP_TC_HALFM", halfm)
            ws = C.c_size_t(0)
            assert lib.hop_sm100_workspace_bytes(C.byref(d), 0, C.byref(ws)) == 0
            assert ws.value >= state and (ws.value - state) % 256 == 0
            got[halfm] = ws.value - state
        # split partials: (splits x N) rows of D floats + three per-row vectors, for some split count >= 1
        for extra in got.values():
            if extra:
                per_split = (300 * d_ * 4 + 255) // 256 * 256
                assert extra >= 2 * per_split
    bad = cb.make_desc(flags=0, path=cb.PATH_TC_F16, n_rows=300, n_patterns=2048, d_in=768, d_assoc=768, d_v=768,
                       d_out=768, update_steps=0, beta=0.05)
    assert lib.hop_sm100_select_path(C.byref(bad), C.byref(out)) != 0
    assert b"1024" in lib.hop_sm100_last_error()


def test_constructor_signature_mirrors_upstream(impl):
    from oracle import hflayers_oracle as O
    ours = inspect.signature(impl.HopfieldLayer.__init__)
    ref = inspect.signature(O.HopfieldLayer.__init__)
    assert list(ours.parameters) == list(ref.parameters)
    for name, p in ref.parameters.items():
        assert ours.parameters[name].default == p.default, name
    fwd = list(inspect.signature(impl.HopfieldLayer.forward).parameters)
    assert fwd[:4] == ["self", "input", "stored_pattern_padding_mask", "association_mask"]


@pytest.mark.parametrize("name,kw", [
    ("hopfield", dict(input_size=64)),
    ("embedding_to_index", dict(input_size=64, output_size=3)),
    ("index_to_embedding", dict(input_size=3, output_size=64)),
])
def test_state_dict_contract(impl, name, kw):
    """HopVAE.py:149-170 construction + utils/__init__.py:36-49 loader: same names, same shapes, nn.Parameters."""
    import hflayers
    layer = hflayers.HopfieldLayer(quantity=512, stored_pattern_as_static=True, state_pattern_as_static=True, **kw)
    sd = layer_state(load_ckpt(), name)
    res = layer.load_state_dict(sd, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    assert all(isinstance(p, torch.nn.Parameter) for p in layer.parameters())
    assert set(layer.state_dict()) == set(sd)
    assert abs(layer.beta - (kw["input_size"] ** -0.5)) < 1e-7


def test_no_cpu_fallback_and_unsupported_options(impl):
    import hflayers
    layer = hflayers.HopfieldLayer(input_size=8, quantity=4, stored_pattern_as_static=True,
                                   state_pattern_as_static=True)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        layer(torch.randn(2, 3, 8))
    for bad in (dict(num_heads=2), dict(dropout=0.1), dict(normalize_hopfield_space=True),
                dict(add_zero_association=True), dict(update_steps_max=-1)):
        with pytest.raises(NotImplementedError):
            hflayers.HopfieldLayer(input_size=8, **bad)
    with pytest.raises(NotImplementedError):
        layer.forward(torch.randn(1, 1, 8), association_mask=torch.zeros(1, 4, dtype=torch.bool))


def test_product_path_never_imports_oracle():
    """The shipped packages must not reference oracle/ (only tests, smoke() and bench's CPU legs may)."""
    for pkg in ("hop-vae_b200", "hflayers"):
        for dirpath, _, files in os.walk(os.path.join(ROOT, pkg)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h")):
                    text = open(os.path.join(dirpath, f)).read()
                    assert not re.search(r"^\s*(from|import)\s+oracle\b", text, re.M), (dirpath, f)


def test_host_pipeline_pieces_cover_the_rows(impl):
    """forward_host's row pieces: contiguous, cover [0, n), every inner cut on the fused kernel's 128-row tile."""
    import random
    bounds_of = impl.layer.host_chunk_bounds
    rng = random.Random(5)
    cases = [(n, c) for n in (1, 127, 128, 129, 2047, 2048, 8192, 16384, 16385, 100000)
             for c in (None, 1, 2, 3, 7, (1, 3, 3, 1), (1, 2, 4, 1), (0, 1, 0), (5, 1))]
    cases += [(rng.randrange(1, 70000), tuple(rng.random() for _ in range(rng.randrange(1, 9)))) for _ in range(200)]
    for n, chunks in cases:
        b = bounds_of(n, chunks)
        assert b and b[0][0] == 0 and b[-1][1] == n, (n, chunks, b)
        for (a0, a1), (b0, _) in zip(b[:-1], b[1:]):
            assert a1 == b0 and a1 % 128 == 0
        assert all(hi > lo for lo, hi in b)
    assert bounds_of(16384, None) == [(0, 2048), (2048, 8192), (8192, 14336), (14336, 16384)]
    assert bounds_of(0, 4) == []
    with pytest.raises(ValueError):
        bounds_of(1000, (0, 0))
