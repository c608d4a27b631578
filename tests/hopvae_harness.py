"""Test harness: the CALLER of the hot path.

A from-scratch restatement of the Hop-VAE model shell around the three
Hopfield layers -- conv encoder, 1x1 pre-conv, bottleneck glue, conv decoder --
with the reference's state-dict names so `checkpoints/MNIST-28.ckpt` loads
strictly.  `/root/reference` does not exist on the GPU box, so GPU tests and the
bench cannot import the reference's own HopVAE.py; `tests/golden/make_golden.py`
checks (in the build container, where the reference is mounted) that this
harness reproduces the reference's forward bit for bit given the same
`HopfieldLayer` class.

Follows /root/reference/HopVAE.py:11-126 (conv stacks), :128-185 (wiring),
:248-294 (forward) and /root/reference/utils/__init__.py:51-55 (straight-through
round).  The Hopfield layer class is injected: the oracle's for the CPU arm, the
sm_100a one for the GPU arm.
"""
from __future__ import annotations

from types import SimpleNamespace

import torch
import torch.nn as nn
import torch.nn.functional as F

MNIST_28 = dict(image_size=28, num_hiddens=64, num_residual_layers=2, num_residual_hiddens=32, embedding_dim=64,
                num_embeddings=512, num_channels=1, representation_dim=8, num_levels=512, index_dim=3)
# ffhq configs as shipped lack index_dim / prior; the harness supplies index_dim=3.
FFHQ_32 = dict(image_size=32, num_hiddens=128, num_residual_layers=2, num_residual_hiddens=32, embedding_dim=64,
               num_embeddings=512, num_channels=3, representation_dim=9, num_levels=512, index_dim=3)
FFHQ_64 = dict(FFHQ_32, image_size=64, representation_dim=17)


def round_ste(x: torch.Tensor) -> torch.Tensor:
    """round() forward, identity backward."""
    return torch.round(x).detach() + (x - x.detach())


class _ResBlock(nn.Module):
    def __init__(self, ch: int, mid: int):
        super().__init__()
        self._block = nn.Sequential(nn.ReLU(True), nn.Conv2d(ch, mid, 3, 1, 1, bias=False), nn.ReLU(True),
                                    nn.Conv2d(mid, ch, 1, 1, bias=False))

    def forward(self, x):
        return x + self._block(x)


class _ResStack(nn.Module):
    def __init__(self, ch: int, n: int, mid: int):
        super().__init__()
        self._layers = nn.ModuleList([_ResBlock(ch, mid) for _ in range(n)])

    def forward(self, x):
        for layer in self._layers:
            x = layer(x)
        return F.relu(x)


class _Encoder(nn.Module):
    def __init__(self, c_in: int, ch: int, n_res: int, mid: int):
        super().__init__()
        self.conv_1 = nn.Conv2d(c_in, ch // 2, 4, 2, 1)
        self.conv_2 = nn.Conv2d(ch // 2, ch, 4, 2, 1)
        self.conv_3 = nn.Conv2d(ch, ch, 4, 1, 2)
        self.conv_4 = nn.Conv2d(ch, ch, 3, 1, 1)
        self.residual_stack = _ResStack(ch, n_res, mid)

    def forward(self, x):
        x = F.relu(self.conv_1(x))
        x = F.relu(self.conv_2(x))
        x = F.relu(self.conv_3(x))
        return self.residual_stack(self.conv_4(x))


class _Decoder(nn.Module):
    def __init__(self, c_in: int, c_out: int, ch: int, n_res: int, mid: int):
        super().__init__()
        self.conv_1 = nn.Conv2d(c_in, ch, 3, 1, 1)
        self.residual_stack = _ResStack(ch, n_res, mid)
        self.conv_trans_1 = nn.ConvTranspose2d(ch, ch // 2, 4, 1, 2)
        self.conv_trans_2 = nn.ConvTranspose2d(ch // 2, ch // 2, 4, 2, 1)
        self.conv_trans_3 = nn.ConvTranspose2d(ch // 2, c_out, 4, 2, 1)

    def forward(self, x):
        x = self.residual_stack(self.conv_1(x))
        x = F.relu(self.conv_trans_1(x))
        x = F.relu(self.conv_trans_2(x))
        return self.conv_trans_3(x)


class HopVAEHarness(nn.Module):
    """encoder -> hopfield -> embedding_to_index -> sigmoid/round -> index_to_embedding -> decoder."""

    def __init__(self, cfg: dict, layer_cls):
        super().__init__()
        c = SimpleNamespace(**cfg)
        self.cfg = c
        self.encoder = _Encoder(c.num_channels, c.num_hiddens, c.num_residual_layers, c.num_residual_hiddens)
        self.pre_vq_conv = nn.Conv2d(c.num_hiddens, c.embedding_dim, 1, 1)
        static = dict(stored_pattern_as_static=True, state_pattern_as_static=True)
        self.hopfield = layer_cls(input_size=c.embedding_dim, quantity=c.num_embeddings, **static)
        self.embedding_to_index = layer_cls(input_size=c.embedding_dim, output_size=c.index_dim,
                                            quantity=c.num_embeddings, **static)
        self.index_to_embedding = layer_cls(input_size=c.index_dim, output_size=c.embedding_dim,
                                            quantity=c.num_embeddings, **static)
        self.post_vq_conv = nn.Conv2d(c.index_dim, c.index_dim, 1, 1)   # constructed, never called (as upstream)
        self.decoder = _Decoder(c.embedding_dim, c.num_channels, c.num_hiddens, c.num_residual_layers,
                                c.num_residual_hiddens)

    # -- pieces, so tests can feed the bottleneck directly ---------------------------------
    def encode(self, x):
        z = self.pre_vq_conv(self.encoder(x))
        r = self.cfg.representation_dim
        return z.permute(0, 2, 3, 1).contiguous().view(-1, r * r, self.cfg.embedding_dim)

    def bottleneck(self, z):
        """z (B, L, D) -> dict of every intermediate of HopVAE.forward's bottleneck."""
        levels = self.cfg.num_levels - 1
        emb = self.hopfield(z)
        logits = self.embedding_to_index(emb)
        idx_q = round_ste(torch.sigmoid(logits) * levels)
        idx = idx_q / levels
        emb_recon = self.index_to_embedding(idx)
        return dict(emb=emb, logits=logits, idx=idx, emb_recon=emb_recon,
                    emb_recon_loss=F.mse_loss(emb_recon, emb))

    def bottleneck_fused(self, z):
        """The same bottleneck as ONE launch (this repository's `hflayers.bottleneck_forward`); the
        patch a maintainer applies to HopVAE.forward for evaluation is in INTEGRATION.md."""
        import hflayers
        out = hflayers.bottleneck_forward(self.hopfield, self.embedding_to_index, self.index_to_embedding, z,
                                          self.cfg.num_levels, want_idx=True)
        return dict(emb=out["emb"], logits=out["logits"], idx=out["idx"], emb_recon=out["emb_recon"],
                    emb_recon_loss=out["emb_recon_loss"])

    def decode(self, emb):
        r = self.cfg.representation_dim
        e = emb.view(-1, r, r, self.cfg.embedding_dim).permute(0, 3, 1, 2).contiguous()
        return self.decoder(e)

    def sample_from_indices(self, idx):
        """`HopVAE.sample()` after the prior (HopVAE.py:188-200): idx (S, index_dim, R, R) integer levels ->
        (z_embeddings (S, R*R, D), x_sample)."""
        r = self.cfg.representation_dim
        z = idx.type(torch.int64) / (self.cfg.num_levels - 1)
        z = z.permute(0, 2, 3, 1).contiguous().view(-1, r * r, self.cfg.index_dim)
        emb = self.index_to_embedding(z)
        return emb, self.decode(emb)

    def forward(self, x):
        b = self.bottleneck(self.encode(x))
        return self.decode(b["emb"]), b["emb_recon_loss"]


def stroke_batch(batch: int = 32) -> torch.Tensor:
    """Deterministic MNIST-like input of SURVEY.md B.4 (normalised background -0.4242)."""
    x = torch.full((batch, 1, 28, 28), -0.4242)
    g = torch.Generator().manual_seed(0)
    for b in range(batch):
        i, j = torch.randint(4, 20, (2,), generator=g).tolist()
        x[b, 0, i:i + 8, j:j + 3] = 2.8
        x[b, 0, i + 3:i + 5, j - 2:j + 6] = 2.0
    return x
