"""CPU restatement (numpy) of SpatialTemporalEmbedding -- TEST INFRASTRUCTURE, never imported by the product.

Follows /root/reference/models/cross_modal_fusion.py:130-164 line by line; pinned by tests/golden/st_embed.npz, which
tests/golden/make_golden_st_embed.py produced by running the unmodified reference class.
"""
from __future__ import annotations

import math

import numpy as np


def table_rows(pos: np.ndarray, R: int) -> np.ndarray:
    """(pos * R).long() (:145-146): fp32 product, truncation toward zero; negative indices address from the end as
    PyTorch advanced indexing does; outside [-R, R) the reference raises IndexError -> IndexError here too."""
    s = (pos.astype(np.float32) * np.float32(R)).astype(np.float32)
    if not np.all(np.isfinite(s)):
        raise IndexError("non-finite spatial position")
    i = np.trunc(s).astype(np.int64)
    if np.any(i < -R) or np.any(i >= R):
        raise IndexError("index out of range")
    return np.where(i < 0, i + R, i)


def gelu(x: np.ndarray) -> np.ndarray:
    erf = np.vectorize(math.erf)
    return (0.5 * x * (1.0 + erf(x / math.sqrt(2.0)))).astype(x.dtype)


def gelu_grad(x: np.ndarray) -> np.ndarray:
    erf = np.vectorize(math.erf)
    return 0.5 * (1.0 + erf(x / math.sqrt(2.0))) + x * np.exp(-0.5 * x * x) / math.sqrt(2.0 * math.pi)


def forward(tokens, pos, t, ex, ey, w1, b1, w2, b2, mod):
    """tokens (n, D) fp32; pos (n, 2) | None; t (n,) | None; ex/ey (R, D/4); w1/b1/b2 (D/2,); w2 (D/2, D/2); mod (D,) | None."""
    n, D = tokens.shape
    Dh = D // 2
    emb = np.zeros_like(tokens)                                              # :140
    if pos is not None:
        xi, yi = table_rows(pos[:, 0], ex.shape[0]), table_rows(pos[:, 1], ey.shape[0])
        emb[:, :Dh] += np.concatenate([ex[xi], ey[yi]], axis=-1)            # :148-152
    if t is not None:
        pre = (t[:, None].astype(np.float32) * w1[None, :] + b1[None, :]).astype(np.float32)
        act = gelu(pre.astype(np.float64)).astype(np.float32)
        emb[:, Dh:] += (act.astype(np.float64) @ w2.astype(np.float64).T + b2).astype(np.float32)   # :155-158
    if mod is not None:
        emb = (emb + mod[None, :]).astype(np.float32)                         # :161-162
    return (tokens + emb).astype(np.float32)                                  # :164


def backward(g, pos, t, ex_shape, w1, b1, w2, has_mod):
    """fp64 gradients of the parameters for grad_out g (n, D): dict of arrays."""
    n, D = g.shape
    Dh, Dq = D // 2, D // 4
    g = g.astype(np.float64)
    out = {}
    if pos is not None:
        R = ex_shape[0]
        dex, dey = np.zeros(ex_shape, np.float64), np.zeros(ex_shape, np.float64)
        np.add.at(dex, table_rows(pos[:, 0], R), g[:, :Dq])
        np.add.at(dey, table_rows(pos[:, 1], R), g[:, Dq:Dh])
        out["spatial_embed_x"], out["spatial_embed_y"] = dex, dey
    if t is not None:
        gt = g[:, Dh:]
        pre = t[:, None].astype(np.float64) * w1[None, :].astype(np.float64) + b1[None, :]
        act = gelu(pre)
        out["w2"] = gt.T @ act
        out["b2"] = gt.sum(0)
        dpre = (gt @ w2.astype(np.float64)) * gelu_grad(pre)
        out["w1"] = (dpre * t[:, None]).sum(0)
        out["b1"] = dpre.sum(0)
    if has_mod:
        out["mod"] = g.sum(0)
    return out
