"""GPU (-m gpu): Hop-VAE's bottleneck as one launch (hop_chain.cu) against the oracle, the committed
goldens and this library's own three-layer path.

What is compared: every tensor `HopVAE.forward` produces between encoder and decoder (HopVAE.py:255-272).  The bar is
the fp32 kernels' (1e-5 on latents / logits).  The quantised levels `round(sigmoid(logits) * 511)` sit behind a
discontinuity: a logit within fp32 rounding of a level boundary may land on the neighbouring level on either side of
ANY two fp32 implementations (the three-layer path differs from the CPU oracle in the same way, test_parity_gpu.py:
`> 0.9995 identical`); rows with such a flip are compared through their own levels.
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

import hopvae_harness as H
from oracle import hflayers_oracle as O
from util import load_ckpt, load_golden, rel_l2

pytestmark = pytest.mark.gpu
FP32_BAR = 1e-5


@pytest.fixture(scope="module")
def hf():
    import hflayers
    hflayers._impl.cabi.load()
    return hflayers


def _pair(hf, cfg, seed=1265):
    torch.manual_seed(seed)
    ref = H.HopVAEHarness(cfg, O.HopfieldLayer).eval()
    dut = H.HopVAEHarness(cfg, hf.HopfieldLayer)
    dut.load_state_dict(ref.state_dict(), strict=True)
    return ref, dut.cuda().eval()


def _compare(got, want, levels, tag):
    """got: fused outputs (CUDA); want: reference dict (CPU or CUDA tensors / numpy)."""
    w = {k: (torch.from_numpy(v) if isinstance(v, np.ndarray) else v).detach().cpu().float() for k, v in want.items()}
    g = {k: v.detach().cpu().float() for k, v in got.items()}
    for k in ("emb", "logits"):
        r = rel_l2(g[k], w[k])
        print(f"{tag} {k}: rel-L2 {r:.3e}")
        assert r < FP32_BAR, k
    lv_g, lv_w = torch.round(g["idx"] * levels), torch.round(w["idx"] * levels)
    same_rows = (lv_g == lv_w).reshape(-1, lv_g.shape[-1]).all(-1)
    frac = float(same_rows.float().mean())
    print(f"{tag}: identical quantised levels on {frac:.6f} of {same_rows.numel()} rows")
    assert frac > 0.999
    assert float((lv_g - lv_w).abs().max()) <= 1.0          # a flip moves one level, never more
    rg = g["emb_recon"].reshape(-1, g["emb_recon"].shape[-1])[same_rows]
    rw = w["emb_recon"].reshape(-1, w["emb_recon"].shape[-1])[same_rows]
    r = rel_l2(rg, rw)
    print(f"{tag} emb_recon (rows with identical levels): rel-L2 {r:.3e}")
    assert r < FP32_BAR
    lg, lw = float(g["emb_recon_loss"]), float(w["emb_recon_loss"])
    print(f"{tag} emb_recon_loss {lg:.6e} vs {lw:.6e}")
    assert abs(lg - lw) <= 5e-3 * abs(lw) + 1e-9            # BASELINE: <= 0.5 % on the reconstruction error


@pytest.mark.parametrize("name,cfg,batch", [("mnist28", H.MNIST_28, 32), ("ffhq32", H.FFHQ_32, 32), ("ffhq64", H.FFHQ_64, 32),
                                            ("ragged", H.MNIST_28, 3)])
def test_bottleneck_chain_matches_oracle_and_three_launch_path(hf, name, cfg, batch):
    """BASELINE configs 1 / 2 / 3 (random init under seed 1265; 2,048 / 2,592 / 9,248 latent rows) and a ragged batch:
    one launch against the CPU oracle and against the library's own three launches, with the launch count checked."""
    ref, dut = _pair(hf, cfg)
    c, s = cfg["num_channels"], cfg["image_size"]
    x = torch.rand(batch, c, s, s, generator=torch.Generator().manual_seed(7)) - 0.5
    cabi = hf._impl.cabi
    with torch.no_grad():
        z = ref.encode(x)
        want = ref.bottleneck(z)
        zc = z.cuda()
        three = dut.bottleneck(zc)
        dut.bottleneck_fused(zc)                               # (pattern preparation cached)
        torch.cuda.synchronize()
        l0 = cabi.launch_count()
        got = dut.bottleneck_fused(zc)
        torch.cuda.synchronize()
        launches = cabi.launch_count() - l0
    assert launches == 1, f"{launches} kernel launches for the fused bottleneck"
    levels = cfg["num_levels"] - 1
    _compare(got, want, levels, f"{name} vs oracle")
    _compare(got, three, levels, f"{name} vs three launches")
    # deterministic: the squared error is summed in CTA order
    with torch.no_grad():
        again = dut.bottleneck_fused(zc)
    assert torch.equal(again["emb_recon_loss"], got["emb_recon_loss"]) and torch.equal(again["emb"], got["emb"])


@pytest.mark.parametrize("tag", ["strokes", "randn"])
def test_bottleneck_chain_on_checkpoint_goldens(hf, tag):
    """The shipped MNIST-28 checkpoint on the committed golden vectors (tests/golden/mnist28_golden.npz: oracle outputs
    routed through the reference's own HopVAE.py glue, see test_parity_gpu.py)."""
    sd = load_ckpt()
    dut = H.HopVAEHarness(H.MNIST_28, hf.HopfieldLayer)
    dut.load_state_dict(sd, strict=True)
    dut = dut.cuda().eval()
    g = load_golden()
    with torch.no_grad():
        got = dut.bottleneck_fused(torch.from_numpy(g[f"{tag}_z"]).cuda())
        x_recon = dut.decode(got["emb"])
    want = {k: g[f"{tag}_{k}"] for k in ("emb", "logits", "idx", "emb_recon", "emb_recon_loss")}
    _compare(got, want, 511, f"ckpt {tag}")
    x = torch.from_numpy(g[f"{tag}_x"])
    mse, want_mse = float(F.mse_loss(x_recon.cpu(), x)), float(g[f"{tag}_recon_mse"])
    assert abs(mse - want_mse) <= 5e-3 * want_mse


def test_bottleneck_chain_refuses_training_and_foreign_shapes(hf):
    ref, dut = _pair(hf, H.MNIST_28)
    z = torch.randn(2, 64, 64, device="cuda")
    with pytest.raises(RuntimeError, match="evaluation path"):
        dut.bottleneck_fused(z)                                # autograd is on and the layers have trainable parameters
    wide = hf.HopfieldLayer(input_size=128, quantity=64, stored_pattern_as_static=True, state_pattern_as_static=True).cuda()
    with torch.no_grad(), pytest.raises(Exception):
        hf.bottleneck_forward(wide, dut.embedding_to_index, dut.index_to_embedding, torch.randn(4, 128, device="cuda"), 512)


def test_channels_last_conv_stacks_hand_the_latents_over_without_a_copy(hf):
    """SURVEY.md 8(f)-2.  With the conv stacks in `channels_last` (one `.to(memory_format=...)` by the caller, cuDNN's
    preferred layout on this GPU anyway) `pre_vq_conv`'s output already IS the (B, H, W, C) order the Hopfield layer
    reads: the reference's `z.permute(0, 2, 3, 1).contiguous()` (HopVAE.py:252) returns a view of the same storage
    instead of launching a transpose copy, and the layer consumes it as is -- same latents as the NCHW run."""
    ref, dut = _pair(hf, H.FFHQ_64)
    x = (torch.rand(8, 3, 64, 64, generator=torch.Generator().manual_seed(3)) - 0.5).cuda()
    with torch.no_grad():
        z_nchw = dut.pre_vq_conv(dut.encoder(x))
        want = dut.bottleneck(dut.encode(x))
        dut_cl = dut.to(memory_format=torch.channels_last)
        z_cl = dut_cl.pre_vq_conv(dut_cl.encoder(x.contiguous(memory_format=torch.channels_last)))
        assert z_cl.is_contiguous(memory_format=torch.channels_last)
        handed = z_cl.permute(0, 2, 3, 1)
        assert handed.is_contiguous()                                  # the reference's .contiguous() is a no-op here ...
        assert handed.contiguous().data_ptr() == z_cl.data_ptr()       # ... it returns the same storage: no copy kernel
        r = dut.cfg.representation_dim
        got = dut_cl.bottleneck_fused(handed.contiguous().view(-1, r * r, dut.cfg.embedding_dim))
    assert rel_l2(z_cl, z_nchw) < 1e-4            # cuDNN picks another algorithm for the other layout
    assert rel_l2(got["emb"], want["emb"]) < 1e-4
