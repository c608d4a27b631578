"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py) -- fp64 NumPy restatement
of one ``HopfieldLayer.forward`` in its de-duplicated form, used to budget
tolerances (how far is the fp32 oracle itself from exact arithmetic?) and to
derive analytic gradients for the backward parity tests.

Follows SURVEY.md A.3 (restating hflayers/functional.py hopfield_core_forward)
for the configuration built at /root/reference/HopVAE.py:149-170.
"""
from __future__ import annotations

import numpy as np


def layer_norm(x: np.ndarray, w=None, b=None, eps: float = 1e-5) -> np.ndarray:
    x = np.asarray(x, dtype=np.float64)
    mu = x.mean(axis=-1, keepdims=True)
    var = ((x - mu) ** 2).mean(axis=-1, keepdims=True)      # biased, like nn.LayerNorm
    y = (x - mu) / np.sqrt(var + eps)
    if w is not None:
        y = y * np.asarray(w, dtype=np.float64)
    if b is not None:
        y = y + np.asarray(b, dtype=np.float64)
    return y


def softmax_rows(s: np.ndarray) -> np.ndarray:
    m = s.max(axis=-1, keepdims=True)
    e = np.exp(s - m)
    return e / e.sum(axis=-1, keepdims=True)


def hopfield_forward_f64(x, W, ln_q, ln_k, ln_v, Wv, bv, Wo, bo, beta, update_steps: int = 0,
                         Wq=None, bq=None, Wk=None, bk=None, chunk: int = 4096):
    """x: (N, D_in) queries; W: (P, D_in) lookup patterns; ln_*: (weight, bias) or None.

    Returns dict(out, retrieved, lse, argmax) with the fixed-T semantics
    (``update_steps`` extra retrieve passes, no eps stop).  Row-chunked so the
    N x P score block stays small.
    """
    f = np.float64
    x = np.asarray(x, f); W = np.asarray(W, f)
    K = layer_norm(W, *ln_k) if ln_k is not None else W
    Vin = layer_norm(W, *ln_v) if ln_v is not None else W
    if Wk is not None:
        K = K @ np.asarray(Wk, f).T + (0 if bk is None else np.asarray(bk, f))
    V = Vin @ np.asarray(Wv, f).T + (0 if bv is None else np.asarray(bv, f)) if Wv is not None else Vin
    N = x.shape[0]
    retrieved = np.empty((N, V.shape[1]), f)
    lse = np.empty(N, f)
    arg = np.empty(N, np.int64)
    for lo in range(0, N, chunk):
        xs = x[lo:lo + chunk]
        Q = layer_norm(xs, *ln_q) if ln_q is not None else xs
        if Wq is not None:
            Q = Q @ np.asarray(Wq, f).T + (0 if bq is None else np.asarray(bq, f))
        q = beta * Q
        for _ in range(update_steps):
            A = softmax_rows(q @ K.T)
            q = beta * (A @ K)
        S = q @ K.T
        m = S.max(axis=-1, keepdims=True)
        e = np.exp(S - m)
        l = e.sum(axis=-1, keepdims=True)
        retrieved[lo:lo + chunk] = (e / l) @ V
        lse[lo:lo + chunk] = (m + np.log(l))[:, 0]
        arg[lo:lo + chunk] = S.argmax(axis=-1)
    out = retrieved
    if Wo is not None:
        out = retrieved @ np.asarray(Wo, f).T + (0 if bo is None else np.asarray(bo, f))
    return dict(out=out, retrieved=retrieved, lse=lse, argmax=arg)
