"""Generates the committed golden fixtures from the REFERENCE itself.

Run in the build container only (it imports /root/reference, which is not on
the GPU box):

    python tests/golden/make_golden.py

What it does
  1. imports the reference's own HopVAE.py / utils / configs with
     `sys.modules['hflayers']` = the oracle restatement (upstream `hflayers` is
     not installable offline -- SURVEY.md section 0.2);
  2. loads checkpoints/MNIST-28.ckpt strictly and reproduces the trained-behaviour
     known answers of SURVEY.md B.4 (asserted here);
  3. checks that tests/hopvae_harness.py (the GPU-box stand-in for HopVAE.py)
     is bit-identical to the reference's forward;
  4. writes tests/golden/mnist28_golden.npz (inputs, every bottleneck
     intermediate, per-layer nearest-pattern indices, losses) and copies the
     checkpoint (a data file, not source) to tests/golden/MNIST-28.ckpt.
"""
import hashlib
import importlib
import os
import shutil
import sys

import numpy as np
import torch
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, REF)

from oracle import hflayers_oracle  # noqa: E402

sys.modules["hflayers"] = hflayers_oracle
ref_hopvae = importlib.import_module("HopVAE")
from utils import MakeConfig  # noqa: E402  (reference utils)
from configs.mnist_28_config import config as ref_config  # noqa: E402

import hopvae_harness as H  # noqa: E402


def layer_details(layer, x):
    """Per-layer oracle outputs incl. nearest-pattern index of the last pass."""
    out, xi, scores, _ = layer(x, return_details=True)
    return out, scores.argmax(dim=-1).reshape(-1).to(torch.int32)


def run(model, x):
    with torch.no_grad():
        z = model.pre_vq_conv(model.encoder(x))
        r = model.representation_dim
        z = z.permute(0, 2, 3, 1).contiguous().view(-1, r * r, model.embedding_dim)
        emb, am0 = layer_details(model.hopfield, z)
        logits, am1 = layer_details(model.embedding_to_index, emb)
        idx = torch.round(torch.sigmoid(logits) * (model.num_levels - 1)) / (model.num_levels - 1)
        emb_recon, am2 = layer_details(model.index_to_embedding, idx)
        x_recon, loss = model(x)
        assert torch.equal(loss, F.mse_loss(emb_recon, emb))
    return dict(x=x, z=z, emb=emb, logits=logits, idx=idx, emb_recon=emb_recon, x_recon=x_recon,
                emb_recon_loss=loss, recon_mse=F.mse_loss(x_recon, x), argmax_hopfield=am0,
                argmax_embedding_to_index=am1, argmax_index_to_embedding=am2)


def main():
    torch.set_num_threads(1)  # deterministic reductions
    cfg = MakeConfig(dict(ref_config))
    torch.manual_seed(1265)
    model = ref_hopvae.HopVAE(cfg, torch.device("cpu"))
    ckpt = os.path.join(REF, "checkpoints", "MNIST-28.ckpt")
    sd = torch.load(ckpt, map_location="cpu", weights_only=True)
    print(model.load_state_dict(sd, strict=True))
    model.eval()

    strokes = run(model, H.stroke_batch(32))
    print("strokes: emb_recon_loss %.6e recon_mse %.6e" % (strokes["emb_recon_loss"], strokes["recon_mse"]))
    assert abs(float(strokes["emb_recon_loss"]) - 1.08e-5) < 2e-7, "B.4 known answer (embedding recon loss)"
    assert abs(float(strokes["recon_mse"]) - 5.19e-4) < 2e-6, "B.4 known answer (reconstruction MSE)"

    g = torch.Generator().manual_seed(1265)
    randn = run(model, torch.randn(8, 1, 28, 28, generator=g))
    print("randn:   emb_recon_loss %.6e recon_mse %.6e" % (randn["emb_recon_loss"], randn["recon_mse"]))

    # the harness must be the reference, bit for bit, given the same layer class
    hm = H.HopVAEHarness(H.MNIST_28, hflayers_oracle.HopfieldLayer)
    print(hm.load_state_dict(sd, strict=True))
    hm.eval()
    with torch.no_grad():
        for name, ref in (("strokes", strokes), ("randn", randn)):
            xr, loss = hm(ref["x"])
            assert torch.equal(xr, ref["x_recon"]) and torch.equal(loss, ref["emb_recon_loss"]), name
            b = hm.bottleneck(hm.encode(ref["x"]))
            for k in ("emb", "logits", "idx", "emb_recon"):
                assert torch.equal(b[k], ref[k]), (name, k)
    print("harness == reference HopVAE.forward (bit-exact)")

    out = {}
    for tag, d in (("strokes", strokes), ("randn", randn)):
        for k, v in d.items():
            a = v.detach().cpu().numpy()
            if a.ndim == 3:  # (B, L, D) -> keep shape
                pass
            out[f"{tag}_{k}"] = a
    path = os.path.join(HERE, "mnist28_golden.npz")
    np.savez_compressed(path, **out)
    shutil.copyfile(ckpt, os.path.join(HERE, "MNIST-28.ckpt"))
    for f in ("mnist28_golden.npz", "MNIST-28.ckpt"):
        h = hashlib.sha256(open(os.path.join(HERE, f), "rb").read()).hexdigest()
        print(f, os.path.getsize(os.path.join(HERE, f)), "bytes sha256", h)


if __name__ == "__main__":
    main()
