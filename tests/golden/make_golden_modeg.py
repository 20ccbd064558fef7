#!/usr/bin/env python
"""Known-answer vectors for the decomposed / interpolated mode ("Mode G") -> tests/golden/modeg_kat.npz

The reference holds no implementation of this mode (SPECIFICATIONS.MD:1068-1200 is text only; the upstream
Grid4D hash encoder is un-vendored and un-pinned), so these vectors cannot come from running the reference.
They FREEZE the stated math of oracle/grid4d_oracle.c's header on purpose: a few points x {dense level, hashed
level} x D = 3 (plus one xyt-style column selection), every value computed three times -- by the C oracle, by
the numpy restatement in this file and by the torch restatement oracle/grid4d_torch.py -- and asserted equal
before the file is written.  One value is derived by hand below (HAND_*), so a reader can check the semantics
with pencil and paper:

    level A: R = 4, T = 512, dense ((4+1)^3 = 125 <= 512): idx = x + 5 y + 25 z
    p0 = (0.375, 0.5625, 0.8125):  s = (1.5, 2.25, 3.25), cell g = (1, 2, 3), w = (0.5, 0.25, 0.25)
    corners (bit0 = +x, bit1 = +y, bit2 = +z): 86 87 91 92 111 112 116 117
    weights:  .28125 .28125 .09375 .09375 .09375 .09375 .03125 .03125      (sum = 1)
    table[i][0] = ((7 i) mod 64 - 32) / 64  ->  -6 1 29 -28 -23 -16 12 19  (/64)
    feature 0 = (.28125 (-5) + .09375 (1) + .09375 (-39) + .03125 (31)) / 64 = -4 / 64 = -0.0625

Run from the repo root:  python tests/golden/make_golden_modeg.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

PRIMES = np.array([1, 2654435761, 805459861, 3674653429], dtype=np.int64)   # models/hash_encoding.py:64

F = 2
# (name, R, T, dims, dense)
LEVELS = [
    ("A_dense_r4", 4, 512, (0, 1, 2), True),
    ("B_hash_r64", 64, 1024, (0, 1, 2), False),
    ("C_hash_r16_xyt", 16, 2048, (0, 1, 3), False),        # 17^3 = 4913 > 2048: hashed; columns x, y, t
    ("D_dense_r8_yzt", 8, 1024, (1, 2, 3), True),          # 9^3 = 729 <= 1024: dense; columns y, z, t
    ("E_hash_prime_T", 32, 1009, (0, 2, 3), False),        # non power-of-two table: int64 floor-mod path
]
POINTS = np.array([
    [0.375, 0.5625, 0.8125, 0.25],          # exact at R = 4 (the hand-derived point); on the lattice at R = 64
    [0.1171875, 0.90625, 0.03125, 0.5],     # = (7.5, 58, 2) / 64: half-way in x at R = 64
    [-0.25, 1.25, 0.5, -0.0],               # outside [0,1]: negative cell, cell > R; indices wrap by floor-mod
    [0.999, 0.001, 0.5, 0.7],               # not representable: generic fp32 rounding of c * R
    [0.0, 1.0, 0.99999994, 1e-30],          # edges: 0, 1, largest fp32 below 1, tiny
    [0.3, 0.6, 0.9, 0.123456],
], dtype=np.float32)
HAND_FEATURE = -0.0625                       # level A, point 0, feature 0
HAND_CORNERS = [86, 87, 91, 92, 111, 112, 116, 117]
HAND_WEIGHTS = [.28125, .28125, .09375, .09375, .09375, .09375, .03125, .03125]


def kat_table(T: int, level: int) -> np.ndarray:
    """(T, F) fp32, multiples of 1/64 in [-0.5, 0.5): exact in fp32 and (times 2^12) in fp16."""
    i = np.arange(T, dtype=np.int64)[:, None]
    f = np.arange(F, dtype=np.int64)[None, :]
    return (((7 * i + 3 * f + 11 * level) % 64 - 32) / 64.0).astype(np.float32)


def numpy_level(coords, R, T, dims, dense):
    """The stated math (grid4d_oracle.c header) in numpy: -> idx (8, n) int64, w (8, n) fp32."""
    c = coords[:, list(dims)].astype(np.float32)
    s = (c * np.float32(R)).astype(np.float32)
    fl = np.floor(s)
    g = fl.astype(np.int64)
    w = (s - fl).astype(np.float32)
    D = len(dims)
    idx, wk = [], []
    for k in range(1 << D):
        off = np.array([(k >> d) & 1 for d in range(D)], dtype=np.int64)
        cell = g + off
        with np.errstate(over="ignore"):
            if dense:
                h = sum(cell[:, d] * np.int64((R + 1) ** d) for d in range(D))
            else:
                h = np.zeros(len(c), dtype=np.int64)
                for d in range(D):
                    h ^= cell[:, d] * PRIMES[d]
        idx.append(np.mod(h, np.int64(T)))
        u = None
        for d in range(D):
            ud = w[:, d] if (k >> d) & 1 else (np.float32(1.0) - w[:, d]).astype(np.float32)
            u = ud if u is None else (u * ud).astype(np.float32)
        wk.append(u)
    return np.stack(idx), np.stack(wk)


def main():
    import torch
    from oracle.grid4d_oracle import COracle, Level
    from oracle.grid4d_torch import grid4d_forward, level_indices_weights

    orc = COracle()
    n = len(POINTS)
    levels, tabs, off = [], [], 0
    for l, (name, R, T, dims, dense) in enumerate(LEVELS):
        levels.append(Level(T=T, R=R, dims=dims, out_col=l * F, table_off=off, dense=dense))
        tabs.append(kat_table(T, l))
        off += T
    tables = np.concatenate(tabs, 0)
    gout = ((np.arange(n * len(LEVELS) * F, dtype=np.float32).reshape(n, -1) % 7) - 3.0) / 4.0   # exact quarters

    want_idx = orc.hash_indices(levels, POINTS, True)
    want_out = orc.forward(levels, POINTS, tables, F, True)
    want_grad = orc.backward(levels, POINTS, gout, off, F, True)

    out = {"points": POINTS, "grad_out": gout, "n_levels": np.array([len(LEVELS)]), "F": np.array([F]),
           "features": want_out}
    tpts = torch.from_numpy(POINTS)
    for l, (name, R, T, dims, dense) in enumerate(LEVELS):
        idx_np, w_np = numpy_level(POINTS, R, T, dims, dense)
        assert np.array_equal(idx_np, want_idx[l]), name                           # C == numpy
        idx_t, w_t = level_indices_weights(tpts, dims, R, T, dense)
        assert np.array_equal(idx_t.numpy().T, idx_np), name                       # torch == numpy
        assert np.array_equal(w_t.numpy().T, w_np), name
        feat_np = np.zeros((n, F), dtype=np.float32)
        for k in range(8):                                                          # ascending corners, fmaf in the C oracle
            feat_np = feat_np + w_np[k][:, None].astype(np.float64) * tabs[l][idx_np[k]].astype(np.float64)
        assert np.abs(feat_np - want_out[:, l * F:(l + 1) * F]).max() <= 2e-7, name
        out[f"meta_{l}"] = np.array([R, T, int(dense), *dims], dtype=np.int64)
        out[f"idx_{l}"] = idx_np
        out[f"w_{l}"] = w_np
        g = want_grad[levels[l].table_off:levels[l].table_off + T]
        rows = np.nonzero(np.abs(g).sum(1))[0]
        out[f"grad_rows_{l}"] = rows
        out[f"grad_vals_{l}"] = g[rows]
        # numpy scatter restatement of the backward
        g_np = np.zeros((T, F), dtype=np.float64)
        for k in range(8):
            np.add.at(g_np, idx_np[k], w_np[k][:, None].astype(np.float64) * gout[:, l * F:(l + 1) * F].astype(np.float64))
        assert np.abs(g_np - g).max() <= 1e-12, name
    t_out = grid4d_forward(tpts, [torch.from_numpy(t) for t in tabs],
                           [(dims, R, T, dense) for (_, R, T, dims, dense) in LEVELS], True).numpy()
    assert np.abs(t_out - want_out).max() <= 2e-7
    # the hand-derived anchor
    assert out["idx_0"][:, 0].tolist() == HAND_CORNERS
    assert out["w_0"][:, 0].tolist() == HAND_WEIGHTS
    assert float(want_out[0, 0]) == HAND_FEATURE, float(want_out[0, 0])
    path = os.path.join(HERE, "modeg_kat.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
