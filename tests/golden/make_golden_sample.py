"""Golden fixture for BASELINE config 4 (PixelCNN prior sampling -> index_to_embedding -> decoder), generated
from the REFERENCE itself.  Build container only (imports /root/reference):

    python tests/golden/make_golden_sample.py

The reference's own `HopVAE.sample()` (HopVAE.py:187-200) is run with configs/pixelcnn_mnist_28_config.py, the
Hopfield / conv weights of checkpoints/MNIST-28.ckpt (the checkpoint holds no `prior.*` keys, so the PixelCNN
prior is random-init under manual_seed(1265)) and `sys.modules['hflayers']` = the oracle
restatement.  Written to tests/golden/pixelcnn_sample_golden.npz:
  idx_prior (1,3,8,8) int64   what prior.sample() returned (192 sequential batch-1 forwards)
  emb_prior (1,64,64) f32     index_to_embedding output;  x_prior (1,1,28,28) f32  decoder output
  idx_batch (16,3,8,8) int64  synthetic batch (uniform levels, generator 1265), emb_batch, x_batch likewise
"""
import importlib
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, REF)

from oracle import hflayers_oracle  # noqa: E402

sys.modules["hflayers"] = hflayers_oracle
ref_hopvae = importlib.import_module("HopVAE")
from utils import MakeConfig  # noqa: E402
from configs.pixelcnn_mnist_28_config import config as ref_config  # noqa: E402

import hopvae_harness as H  # noqa: E402

torch.manual_seed(1265)
cfg = MakeConfig(ref_config)
dev = torch.device("cpu")
model = ref_hopvae.HopVAE(cfg, dev).eval()
sd = torch.load(os.path.join(REF, "checkpoints", "MNIST-28.ckpt"), map_location="cpu", weights_only=True)
missing, unexpected = model.load_state_dict(sd, strict=False)
assert not unexpected and all(k.startswith("prior.") for k in missing), (missing, unexpected)


def after_prior(idx):
    """HopVAE.sample() from the prior's indices on (HopVAE.py:188-200)."""
    with torch.no_grad():
        z = idx.type(torch.int64) / (model.num_levels - 1)
        z = z.permute(0, 2, 3, 1).contiguous().view(-1, model.representation_dim ** 2, model.index_dim)
        emb = model.index_to_embedding(z)
        e = emb.view(-1, model.representation_dim, model.representation_dim, model.embedding_dim)
        return emb, model.decoder(e.permute(0, 3, 1, 2).contiguous())


torch.manual_seed(1265)
with torch.no_grad():
    idx_prior = model.prior.sample()
torch.manual_seed(1265)
with torch.no_grad():
    x_ref = model.sample()            # the reference's own call: same seed -> same indices
emb_prior, x_prior = after_prior(idx_prior)
assert torch.equal(x_ref, x_prior), "after_prior() must reproduce HopVAE.sample() bit for bit"

g = torch.Generator().manual_seed(1265)
idx_batch = torch.randint(0, cfg.num_levels, (16, cfg.index_dim, cfg.representation_dim, cfg.representation_dim),
                          generator=g)
emb_batch, x_batch = after_prior(idx_batch)

# the GPU-box stand-in must agree with the reference bit for bit on CPU
hs = H.HopVAEHarness(H.MNIST_28, hflayers_oracle.HopfieldLayer).eval()
hs.load_state_dict({k: v for k, v in sd.items()}, strict=True)
with torch.no_grad():
    e2, x2 = hs.sample_from_indices(idx_batch)
assert torch.equal(e2, emb_batch) and torch.equal(x2, x_batch)

out = os.path.join(HERE, "pixelcnn_sample_golden.npz")
np.savez_compressed(out, idx_prior=idx_prior.numpy(), emb_prior=emb_prior.numpy(), x_prior=x_prior.numpy(),
                    idx_batch=idx_batch.numpy(), emb_batch=emb_batch.numpy(), x_batch=x_batch.numpy())
print("wrote", out, os.path.getsize(out), "bytes; prior indices", idx_prior.flatten()[:8].tolist())
