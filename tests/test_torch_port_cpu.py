"""oracle/grid4d_torch.py -- the op-by-op PyTorch port that bench.py times as the CPU baseline and that the GPU tests
run on the device as a large-size oracle -- checked on CPU against the reference's golden vectors (Mode R) and against
the C oracle (decomposed / interpolated mode: two independent restatements of the same stated math).  CPU only.
"""
import os

import numpy as np
import pytest
import torch

from _synth import coords_mixed, synth_table, uniform01
from oracle.grid4d_oracle import make_levels
from oracle.grid4d_torch import RefHashEncoding, geo2xyzt_norm, grid4d_forward

ENC_CASES = ["enc_d3_l16_t19", "enc_d1_l8_t17", "enc_d2_l5_f4_npow2", "enc_d3_l3_f1_prime", "enc_d3_l2_f3_b1"]


@pytest.mark.parametrize("name", ENC_CASES)
def test_port_reproduces_reference_golden(golden_dir, name):
    g = np.load(os.path.join(golden_dir, name + ".npz"), allow_pickle=False)
    L, F, T, D, seed = (int(v) for v in g["meta"])
    enc = RefHashEncoding(L, F, [int(r) for r in g["resolutions"]], T, D)
    with torch.no_grad():
        for l, emb in enumerate(enc.hash_tables):
            emb.weight.copy_(torch.from_numpy(synth_table(T, F, 1000 * seed + l)))
    x = torch.from_numpy(g["coords"])
    out = enc(x)
    assert np.array_equal(out.detach().numpy(), g["features"])                 # same ATen ops: bit exact
    out.backward(torch.from_numpy(g["grad_out"]))
    for l, emb in enumerate(enc.hash_tables):
        dense = np.zeros((T, F), dtype=np.float32)
        dense[g[f"grad_rows_{l}"]] = g[f"grad_vals_{l}"]
        scale = max(np.abs(dense).max(), 1e-30)
        assert np.abs(emb.weight.grad.numpy() - dense).max() <= 1e-6 * scale


@pytest.mark.parametrize("dense_fallback", [False, True])
def test_port_matches_c_oracle_in_decomposed_mode(oracle, dense_fallback):
    grids = [(0, 1, 2), (0, 1, 3), (1, 2, 3), (0, 2, 3)]                        # SPECIFICATIONS.MD:1174-1176
    res = [[4, 9, 16, 50, 128, 700]] * 4
    T, F, n = 4096, 2, 600
    levels = make_levels(grids, res, [T] * 4, F, dense_fallback)
    tabs = [synth_table(T, F, 40 + i, lo=-1.0, hi=1.0) for i in range(len(levels))]
    coords = coords_mixed(n, 4, seed=91, res_hint=16)
    coords = np.clip(coords, 0.0, np.float32(1.0) - np.float32(2 ** -24)).astype(np.float32)
    want = oracle.forward(levels, coords, np.concatenate(tabs, 0), F, linear=True)
    specs = [(lv.dims, lv.R, lv.T, lv.dense) for lv in levels]
    tt = [torch.from_numpy(t).requires_grad_(True) for t in tabs]
    got = grid4d_forward(torch.from_numpy(coords), tt, specs, linear=True)
    assert got.shape == want.shape
    assert np.abs(got.detach().numpy() - want).max() <= 1e-5 * np.abs(want).max()
    gout = (uniform01(n * want.shape[1], 92).reshape(want.shape) * 2 - 1).astype(np.float32)
    got.backward(torch.from_numpy(gout))
    grads = oracle.backward(levels, coords, gout, len(levels) * T, F, linear=True)
    for i, t in enumerate(tt):
        ref = grads[i * T:(i + 1) * T]
        assert np.abs(t.grad.numpy() - ref).max() <= 1e-4 * max(np.abs(ref).max(), 1e-30)


def test_port_geo_prologue_matches_c_oracle(oracle, golden_dir):
    g = np.load(os.path.join(golden_dir, "geo2xyz.npz"), allow_pickle=False)
    geo = g["florida_geo"]
    t = np.linspace(-1.0, 11.0, geo.shape[0])
    bmin, span = g["florida_bbox_min"], g["florida_bbox_span"]
    want, ecef = oracle.geo2xyzt_norm(geo, t, bmin, span, (0.0, 10.0), want_ecef=True)
    got, xyz = geo2xyzt_norm(torch.from_numpy(geo), torch.from_numpy(t), bmin, span, (0.0, 10.0))
    assert np.abs(xyz.numpy() - g["florida_ecef"]).max() <= 1e-6
    assert np.abs(xyz.numpy() - ecef).max() <= 1e-6
    assert np.abs(got.numpy() - want).max() <= 1.2e-7
