"""CPU: pins the oracle (oracle/) against the reference-generated golden vectors,
an independent PyTorch attention implementation and the fp64 restatement."""
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F

import hopvae_harness as H
from oracle import hflayers_oracle as O
from oracle.hopfield_f64 import hopfield_forward_f64
from util import layer_state, load_ckpt, load_golden, load_sample_golden, rel_l2, randomize_, chunked_oracle


@pytest.fixture(scope="module")
def model():
    m = H.HopVAEHarness(H.MNIST_28, O.HopfieldLayer)
    res = m.load_state_dict(load_ckpt(), strict=True)   # structural pin: every key, every shape
    assert not res.missing_keys and not res.unexpected_keys
    return m.eval()


def test_checkpoint_key_contract():
    sd = load_ckpt()
    hop = {k: tuple(v.shape) for k, v in sd.items() if k.split(".")[0] in
           ("hopfield", "embedding_to_index", "index_to_embedding")}
    assert len(hop) == 33 and len(sd) == 61
    assert hop["hopfield.lookup_weights"] == (1, 512, 64)
    assert hop["embedding_to_index.hopfield.association_core.out_proj.weight"] == (3, 64)
    assert hop["index_to_embedding.hopfield.association_core.in_proj_weight"] == (3, 3)


def test_trained_behaviour_known_answers(model):
    """beta = 1/sqrt(D), T = 0 reproduces the trained reconstruction; anything else is 10^2x worse."""
    g = load_golden()
    x = torch.from_numpy(g["strokes_x"])
    assert torch.equal(x, H.stroke_batch(32))
    with torch.no_grad():
        xr, loss = model(x)
    assert abs(float(loss) - 1.08e-5) < 2e-7
    assert abs(float(F.mse_loss(xr, x)) - 5.19e-4) < 2e-6
    np.testing.assert_allclose(xr.numpy(), g["strokes_x_recon"], rtol=0, atol=2e-5)
    layers = (model.hopfield, model.embedding_to_index, model.index_to_embedding)
    try:
        for t, lo in ((1, 0.3), (3, 0.5)):
            for l in layers:
                l.hopfield.update_steps_max = t
            with torch.no_grad():
                xr, _ = model(x)
            assert float(F.mse_loss(xr, x)) > lo          # B.4: 0.398 and 0.595
        for l in layers:
            l.hopfield.update_steps_max = 0
            l.hopfield.scaling = 1.0
        with torch.no_grad():
            xr, _ = model(x)
        assert float(F.mse_loss(xr, x)) > 0.5             # B.4: 1.089
    finally:
        for l in layers:
            l.hopfield.update_steps_max = 0
            l.hopfield.scaling = None


@pytest.mark.parametrize("tag", ["strokes", "randn"])
def test_golden_bottleneck(model, tag):
    g = load_golden()
    with torch.no_grad():
        b = model.bottleneck(torch.from_numpy(g[f"{tag}_z"]))
    for k in ("emb", "logits", "emb_recon"):
        assert rel_l2(b[k], g[f"{tag}_{k}"]) < 2e-6, k
    assert torch.equal(b["idx"], torch.from_numpy(g[f"{tag}_idx"]))
    for name in ("hopfield", "embedding_to_index", "index_to_embedding"):
        layer = getattr(model, name)
        inp = {"hopfield": g[f"{tag}_z"], "embedding_to_index": g[f"{tag}_emb"],
               "index_to_embedding": g[f"{tag}_idx"]}[name]
        with torch.no_grad():
            _, _, scores, _ = layer(torch.from_numpy(inp), return_details=True)
        am = scores.argmax(-1).reshape(-1).numpy()
        assert (am == g[f"{tag}_argmax_{name}"]).mean() > 0.999   # identical up to fp32 thread-count ties


def _operands(layer, x):
    hp = layer.hopfield
    W = layer.lookup_weights[0]
    q = hp.norm_state_pattern(x)
    k = hp.norm_stored_pattern(W)
    v = F.linear(hp.norm_pattern_projection(W), hp.association_core.in_proj_weight, hp.association_core.in_proj_bias)
    return q, k, v


@pytest.mark.parametrize("d,dout", [(64, None), (64, 3), (3, 64)])
def test_equals_sdpa(d, dout):
    """T = 0 association == F.scaled_dot_product_attention (independent code path on the box)."""
    torch.manual_seed(0)
    layer = randomize_(O.HopfieldLayer(input_size=d, output_size=dout, quantity=96, stored_pattern_as_static=True,
                                       state_pattern_as_static=True), 1)
    x = torch.randn(4, 10, d)
    with torch.no_grad():
        q, k, v = _operands(layer, x)
        ref = F.scaled_dot_product_attention(q, k.expand(4, -1, -1), v.expand(4, -1, -1), scale=1 / math.sqrt(d))
        ref = layer.hopfield.association_core.out_proj(ref)
        got = layer(x)
    assert rel_l2(got, ref) < 2e-6


@pytest.mark.parametrize("T", [0, 1, 3])
def test_equals_fp64_restatement(T):
    layer = randomize_(O.HopfieldLayer(input_size=32, output_size=8, quantity=200, update_steps_max=T,
                                       update_steps_eps=0.0, stored_pattern_as_static=True,
                                       state_pattern_as_static=True), 2)
    x = torch.randn(3, 7, 32, generator=torch.Generator().manual_seed(3))
    hp, core = layer.hopfield, layer.hopfield.association_core
    n = lambda t: t.detach().double().numpy()
    ln = lambda m: (n(m.weight), n(m.bias))
    ref = hopfield_forward_f64(n(x.reshape(-1, 32)), n(layer.lookup_weights[0]), ln(hp.norm_state_pattern),
                               ln(hp.norm_stored_pattern), ln(hp.norm_pattern_projection), n(core.in_proj_weight),
                               n(core.in_proj_bias), n(core.out_proj.weight), n(core.out_proj.bias),
                               1 / math.sqrt(32), update_steps=T)
    with torch.no_grad():
        got = layer(x).reshape(-1, 8)
    assert rel_l2(got, ref["out"]) < 5e-6


def test_faithful_equals_dedup():
    """Expanding the lookup matrix over the batch (upstream) changes nothing but the cost."""
    layer = randomize_(O.HopfieldLayer(input_size=16, quantity=40, stored_pattern_as_static=True,
                                       state_pattern_as_static=True), 4)
    x = torch.randn(5, 6, 16)
    with torch.no_grad():
        a = layer(x)
        layer.faithful = False
        b = layer(x)
    assert rel_l2(a, b) < 1e-6


def test_stop_criterion_is_bounded_by_eps():
    """Early stop (upstream eps rule) vs the fixed-T semantics the kernels implement differ by O(eps)."""
    kw = dict(input_size=16, quantity=64, stored_pattern_as_static=True, state_pattern_as_static=True)
    a = randomize_(O.HopfieldLayer(update_steps_max=8, update_steps_eps=1e-4, **kw), 5)
    b = randomize_(O.HopfieldLayer(update_steps_max=8, update_steps_eps=0.0, **kw), 5)
    x = torch.randn(2, 9, 16, generator=torch.Generator().manual_seed(6))
    with torch.no_grad():
        ya, yb = a(x), b(x)
    assert rel_l2(ya, yb) < 1e-3


def test_projected_patterns_shapes():
    layer = O.HopfieldLayer(input_size=12, hidden_size=20, pattern_size=24, output_size=5, quantity=30)
    assert layer.hopfield.association_core.in_proj_weight.shape == (20 + 20 + 24, 12)
    assert layer(torch.randn(2, 3, 12)).shape == (2, 3, 5)


@pytest.mark.parametrize("tag", ["prior", "batch"])
def test_golden_sample_path(model, tag):
    """BASELINE config 4 (HopVAE.sample, HopVAE.py:187-200): prior indices -> index_to_embedding -> decoder,
    against the outputs of the reference itself (fixture generated by tests/golden/make_golden_sample.py)."""
    g = load_sample_golden()
    idx = torch.from_numpy(g[f"idx_{tag}"])
    with torch.no_grad():
        emb, x = model.sample_from_indices(idx)
    assert torch.equal(emb, torch.from_numpy(g[f"emb_{tag}"]))
    assert torch.equal(x, torch.from_numpy(g[f"x_{tag}"]))


@pytest.mark.parametrize("T", [0, 2])
def test_chunked_checker_equals_oracle(T):
    """The P-chunked full-size checker of the config-5 sweep (util.chunked_oracle) is the same function as the
    restated layer (small size, several chunks)."""
    import bench
    w, wv, wo = bench.make_state(2048, 64, T)
    x, _ = bench.make_queries(w, 24)
    layer = bench.build_layer(O.HopfieldLayer, 2048, 64, T, w, wv, wo)
    with torch.no_grad():
        want, _, scores, _ = layer(x[None], return_details=True)
    got, arg, gap = chunked_oracle(x, w, wv, wo, T, 1.0 / 8.0, chunk=512)
    assert rel_l2(got, want[0]) < 2e-6
    s = scores.reshape(24, 2048)
    assert torch.equal(arg, s.argmax(-1))
    top2 = s.topk(2, -1).values
    assert torch.allclose(gap, top2[:, 0] - top2[:, 1], atol=1e-5)
