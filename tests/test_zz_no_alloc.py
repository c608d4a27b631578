"""GPU (-m gpu), collected last on purpose: `hop_sm100_bottleneck_forward_static`, the allocation-free variant of the
bottleneck launch (scratch in a file-scope __device__ array of the library; include/hop_sm100.h), against the existing
`hop_sm100_bottleneck_forward` (caller workspace) on the same inputs -- two calls in a row, a second shape in between so
that the shared scratch carries another call's contents, and the size limit.  Same kernel, same arithmetic: the bar is the
fp32 kernels' 1e-5 (the outputs are expected to be bit-identical; the loss is summed in tile order on both sides).
The entry point has no counterpart in the reference."""
import ctypes as C

import pytest
import torch

import hopvae_harness as H
from util import rel_l2

pytestmark = pytest.mark.gpu
FP32_BAR = 1e-5


@pytest.fixture(scope="module")
def hf():
    import hflayers
    hflayers._impl.cabi.load()
    return hflayers


def _model(hf, cfg):
    torch.manual_seed(1265)
    return H.HopVAEHarness(cfg, hf.HopfieldLayer).cuda().eval()


def _both(hf, m, z):
    args = (m.hopfield, m.embedding_to_index, m.index_to_embedding, z, m.cfg.num_levels)
    with torch.no_grad():
        want = hf.bottleneck_forward(*args)
        got = hf.bottleneck_forward(*args, static_scratch=True)
    torch.cuda.synchronize()
    return got, want


def test_static_scratch_bottleneck_matches_workspace_form_twice_in_a_row(hf):
    m = _model(hf, H.FFHQ_64)
    small = _model(hf, H.MNIST_28)
    g = torch.Generator().manual_seed(11)
    z1 = torch.randn(32, 289, 64, generator=g).cuda()           # config 3: 9,248 latent rows
    z2 = torch.randn(3, 64, 64, generator=g).cuda()             # a ragged, smaller call that reuses the same scratch
    for call, (model, z) in enumerate([(m, z1), (m, z1), (small, z2), (m, z1)]):
        got, want = _both(hf, model, z)
        for k in ("emb", "logits", "emb_recon"):
            r = rel_l2(got[k], want[k])
            assert r < FP32_BAR, (call, k, r)
        assert torch.equal(got["idx_q"], want["idx_q"]), call
        lg, lw = float(got["emb_recon_loss"]), float(want["emb_recon_loss"])
        print(f"call {call}: rows {z.shape[0] * z.shape[1]}, loss {lg:.8e} vs {lw:.8e}")
        assert abs(lg - lw) <= 1e-6 * abs(lw)


def test_static_scratch_bottleneck_rejects_sizes_beyond_its_array(hf):
    cabi = hf._impl.cabi
    lib = cabi.load()
    limit = int(lib.hop_sm100_bottleneck_static_max_rows())
    assert limit == 262144
    m = _model(hf, H.MNIST_28)
    z = torch.zeros(limit + 64, 64, device="cuda")
    with torch.no_grad(), pytest.raises(Exception, match="maximum"):
        hf.bottleneck_forward(m.hopfield, m.embedding_to_index, m.index_to_embedding, z, 512, static_scratch=True)
    with torch.no_grad():                                        # the workspace form still takes it
        out = hf.bottleneck_forward(m.hopfield, m.embedding_to_index, m.index_to_embedding, z, 512)
    assert out["emb"].shape == (limit + 64, 64)
