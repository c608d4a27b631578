"""Hop-VAE's bottleneck as one launch (csrc/hop_chain.cu).

`bottleneck_forward(hopfield, embedding_to_index, index_to_embedding, z, num_levels)` computes what
`HopVAE.forward` does between the encoder and the decoder (/root/reference/HopVAE.py:255-272):

    z_embeddings        = hopfield(z)
    z_indices           = sigmoid(embedding_to_index(z_embeddings))
    z_indices_quantised = straight_through_round(z_indices * (num_levels - 1))      utils/__init__.py:51-55
    z_embeddings_recon  = index_to_embedding(z_indices_quantised / (num_levels - 1))
    embedding_recon_loss = F.mse_loss(z_embeddings_recon, z_embeddings)

with the three `HopfieldLayer`s of this package, in ONE kernel launch (45 eager launches in the reference, 3 with this
package's layers called one by one).  Evaluation only: under autograd (training) call the layers as the reference does.
"""
from __future__ import annotations

from typing import Dict

import torch
from torch import Tensor

from . import cabi
from .ops import _desc, _on_device, _stream, _req_cuda_f32

_QUERY_KEYS = ("ln_state_w", "ln_state_b", "wq", "bq", "wo", "bo")
_ws_cache: Dict[tuple, Tensor] = {}


def bottleneck_forward(hopfield, embedding_to_index, index_to_embedding, z: Tensor, num_levels: int,
                       want_logits: bool = True, want_recon: bool = True, want_idx: bool = False,
                       static_scratch: bool = False) -> Dict[str, Tensor]:
    """z (..., D) float32 CUDA -> dict(emb, logits, idx_q, emb_recon, emb_recon_loss[, idx]) with the reference's shapes
    (leading dimensions of z kept).  `idx_q` are the integer levels the prior is fitted on (HopVAE.py:279-283);
    `idx` = idx_q / (num_levels - 1) costs one more (elementwise) launch and is only built when asked for."""
    import ctypes as C
    layers = (hopfield, embedding_to_index, index_to_embedding)
    if torch.is_grad_enabled() and (z.requires_grad or any(p.requires_grad for l in layers for p in l.parameters())):
        raise RuntimeError("hop_sm100: bottleneck_forward is an evaluation path (no autograd); call it under "
                           "torch.no_grad(), or call the three layers one by one for training")
    _req_cuda_f32(z, "z")
    lead = z.shape[:-1]
    x = z.reshape(-1, z.shape[-1])
    n = x.shape[0]
    lib = cabi.load()
    descs = (cabi.HopDesc * 3)()
    params = (cabi.HopParams * 3)()
    pats = (cabi.HopPatterns * 3)()
    keep = []
    for i, layer in enumerate(layers):
        spec, pack = layer.spec(), layer.param_pack()
        prepared = layer._prepared(pack, spec, cabi.PATH_FP32_SIMT)      # fp32 K^ / V^ (cached per parameter version)
        keep.append((pack, prepared))
        descs[i] = _desc(spec.flags, cabi.PATH_FP32_SIMT, n, layer.quantity, spec.d_in, spec.d_assoc, spec.d_v,
                         max(spec.d_out, 1), spec.update_steps, spec.beta, spec.ln_eps, tune_splits=spec.tune_splits)
        q = {k: (None if pack[k] is None else pack[k].detach().contiguous()) for k in _QUERY_KEYS}
        keep.append(q)
        params[i] = cabi.HopParams(**{k: cabi.ptr(v) for k, v in q.items()})
        pats[i] = cabi.patterns_struct(prepared)
    dev = x.device
    d_emb, d_idx = layers[0].output_size, layers[1].output_size
    emb = torch.empty((n, d_emb), device=dev, dtype=torch.float32)
    logits = torch.empty((n, d_idx), device=dev, dtype=torch.float32) if want_logits else None
    idx_q = torch.empty((n, d_idx), device=dev, dtype=torch.float32)
    recon = torch.empty((n, d_emb), device=dev, dtype=torch.float32) if want_recon else None
    mse = torch.empty((), device=dev, dtype=torch.float32)
    if static_scratch:
        # the library's built-in scratch (hop_sm100_bottleneck_forward_static): no workspace, nothing allocated by the
        # call; rows <= hop_sm100_bottleneck_static_max_rows(), calls on one GPU must not overlap
        with _on_device(dev):
            cabi.check(lib.hop_sm100_bottleneck_forward_static(descs, params, pats, int(num_levels), cabi.ptr(x),
                                                               cabi.ptr(emb), cabi.ptr(logits), cabi.ptr(idx_q),
                                                               cabi.ptr(recon), cabi.ptr(mse), _stream(dev)),
                       "hop_sm100_bottleneck_forward_static")
        return _results(emb, logits, idx_q, recon, mse, lead, d_emb, d_idx, num_levels, want_idx)
    need = C.c_size_t(0)
    cabi.check(lib.hop_sm100_bottleneck_workspace_bytes(n, C.byref(need)), "hop_sm100_bottleneck_workspace_bytes")
    key = (dev, int(need.value))
    ws = _ws_cache.get(key)
    if ws is None:
        ws = _ws_cache[key] = torch.zeros((int(need.value),), device=dev, dtype=torch.uint8)
    with _on_device(dev):
        cabi.check(lib.hop_sm100_bottleneck_forward(descs, params, pats, int(num_levels), cabi.ptr(x), cabi.ptr(emb),
                                                    cabi.ptr(logits), cabi.ptr(idx_q), cabi.ptr(recon), cabi.ptr(mse),
                                                    cabi.ptr(ws), int(need.value), _stream(dev)),
                   "hop_sm100_bottleneck_forward")
    return _results(emb, logits, idx_q, recon, mse, lead, d_emb, d_idx, num_levels, want_idx)


def _results(emb, logits, idx_q, recon, mse, lead, d_emb, d_idx, num_levels, want_idx):
    out = dict(emb=emb.reshape(*lead, d_emb), idx_q=idx_q.reshape(*lead, d_idx), emb_recon_loss=mse)
    if want_idx:
        out["idx"] = (idx_q / float(num_levels - 1)).reshape(*lead, d_idx)
    if logits is not None:
        out["logits"] = logits.reshape(*lead, d_idx)
    if recon is not None:
        out["emb_recon"] = recon.reshape(*lead, d_emb)
    return out
