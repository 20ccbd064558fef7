"""Randomised cross-check of the CPU oracle against the UNMODIFIED reference, run live.

Only where the reference tree exists (the CPU container: /root/reference); skipped elsewhere -- the GPU box never
reads it.  The committed fixtures under tests/golden/ pin fixed cases; this widens the pin to a few hundred random
geometries per run (dimension, resolution, table size incl. non powers of two and primes, feature count, coordinates
far outside [0, 1]) so that a restatement error that the fixed cases happen to miss would still show up.
CPU only, seconds.
"""
import os
import sys

import numpy as np
import pytest
import torch

from _synth import coords_mixed, synth_table, uniform01
from oracle.grid4d_oracle import Level, numpy_hash_coords

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.exists(os.path.join(REF, "models", "hash_encoding.py")),
                                reason="reference tree not present (GPU box)")


@pytest.fixture(scope="module")
def ref_cls():
    sys.path.insert(0, REF)
    try:
        from models.hash_encoding import MultiResolutionHashEncoding      # reference, unmodified
    finally:
        sys.path.remove(REF)
    return MultiResolutionHashEncoding


def _table_sizes(rng):
    pool = [2 ** k for k in range(4, 21)] + [7, 97, 1000, 50000, 65537, 1000003, 3 * 2 ** 15, 2 ** 19 - 1, 2 ** 19 + 1]
    return int(pool[rng.integers(len(pool))])


def test_hash_indices_random_geometries(oracle, ref_cls):
    rng = np.random.default_rng(20261104)
    for k in range(160):
        D = int(rng.integers(1, 5))
        T = _table_sizes(rng)
        R = int(2 ** rng.uniform(0, 20)) if rng.random() < 0.8 else int(rng.integers(1, 3000))
        n = 192
        coords = coords_mixed(n, D, seed=1000 + k, res_hint=max(2, min(R, 4096)))
        if k % 4 == 0:                                            # far outside the unit cube, both signs
            coords = (coords * np.float32(rng.choice([10.0, 1e3, 1e5]))).astype(np.float32)
        enc = ref_cls(1, 2, [R], T, D)
        want = enc.hash_coords(torch.from_numpy(coords), R).numpy()
        assert want.min() >= 0 and want.max() < T
        got_c = oracle.hash_indices([Level(T=T, R=R, dims=tuple(range(D)), out_col=0)], coords, False)[0][0]
        assert np.array_equal(got_c, want), f"C oracle: case {k} D={D} T={T} R={R}"
        assert np.array_equal(numpy_hash_coords(coords, R, T), want), f"numpy oracle: case {k} D={D} T={T} R={R}"


def test_forward_backward_random_geometries(oracle, ref_cls):
    rng = np.random.default_rng(7)
    for k in range(24):
        D = int(rng.integers(1, 5))
        L = int(rng.integers(1, 5))
        F = int(rng.choice([1, 2, 3, 4, 8]))
        T = int(rng.choice([64, 1000, 4096, 50000, 2 ** 16]))
        res = [int(2 ** rng.uniform(0, 14)) for _ in range(L)]
        n = 300
        enc = ref_cls(L, F, res, T, D)
        tabs = [synth_table(T, F, 500 * k + l, lo=-1.0, hi=1.0) for l in range(L)]
        with torch.no_grad():
            for emb, t in zip(enc.hash_tables, tabs):
                emb.weight.copy_(torch.from_numpy(t))
        coords = coords_mixed(n, D, seed=77 + k, res_hint=max(2, min(res[0], 4096)))
        x = torch.from_numpy(coords).requires_grad_(True)
        out = enc(x)
        gout = (uniform01(n * L * F, 900 + k).reshape(n, L * F) * 2 - 1).astype(np.float32)
        out.backward(torch.from_numpy(gout))

        levels = [Level(T=T, R=r, dims=tuple(range(D)), out_col=l * F, table_off=l * T) for l, r in enumerate(res)]
        tables = np.concatenate(tabs, 0)
        got = oracle.forward(levels, coords, tables, F, linear=False)
        assert np.array_equal(got, out.detach().numpy()), f"forward: case {k} D={D} L={L} F={F} T={T} res={res}"
        grads = oracle.backward(levels, coords, gout, L * T, F, linear=False)
        for l, emb in enumerate(enc.hash_tables):
            want = emb.weight.grad.numpy().astype(np.float64)
            scale = max(np.abs(want).max(), 1e-30)
            assert np.abs(grads[l * T:(l + 1) * T] - want).max() <= 1e-5 * scale, f"backward: case {k} level {l}"
        assert x.grad is None                                       # graph cut at floor().long(), hash_encoding.py:61


def test_grid4d_encoder_features_and_masks_random(oracle):
    """Reference Grid4DEncoder (models/encoders.py:54-85) feature half with random row masks, three seeds."""
    sys.path.insert(0, REF)
    try:
        from models.configs import DeepEarthConfig                            # reference
        from models.encoders import Grid4DEncoder                             # reference
    finally:
        sys.path.remove(REF)
    T, F = 2 ** 12, 2
    for seed in (1, 2, 3):
        cfg = DeepEarthConfig(hidden_dim=16, n_heads=4, n_layers=1, hash_table_size=T)
        enc = Grid4DEncoder(cfg)
        levels, tabs, off = [], [], 0
        for sub, dims, col0, mid, s0 in ((enc.spatial_encoder, (0, 1, 2), 0, 1, 100), (enc.temporal_encoder, (3,), 32, 2, 200)):
            for l, emb in enumerate(sub.hash_tables):
                t = synth_table(sub.hash_table_size, F, 1000 * seed + s0 + l, lo=-1.0, hi=1.0)
                with torch.no_grad():
                    emb.weight.copy_(torch.from_numpy(t))
                levels.append(Level(T=sub.hash_table_size, R=int(sub.resolutions[l]), dims=dims, out_col=col0 + l * F,
                                    table_off=off, mask_id=mid))
                tabs.append(t)
                off += sub.hash_table_size
        n = 400
        xyzt = coords_mixed(n, 4, seed=300 + seed, res_hint=16)
        sm, tm = uniform01(n, 310 + seed) > 0.4, uniform01(n, 320 + seed) > 0.6
        captured = {}
        h = enc.projection.register_forward_hook(lambda m, i, o: captured.__setitem__("feat", i[0].detach().clone()))
        enc(torch.from_numpy(xyzt), spatial_mask=torch.from_numpy(sm), temporal_mask=torch.from_numpy(tm))
        h.remove()
        got = oracle.forward(levels, xyzt, np.concatenate(tabs, 0), F, linear=False, mask_a=sm, mask_b=tm)
        assert np.array_equal(got, captured["feat"].numpy()), f"seed {seed}"


def test_geo2xyz_random_points(oracle):
    """Reference GeospatialConverter.geodetic_to_xyz + xyz_to_norm (geo2xyz.py:146-160, :327-332) on random points
    incl. poles and the date line: ECEF within 1e-6 m, normalised fp32 within one ulp of [0,1]."""
    sys.path.insert(0, REF)
    try:
        from geospatial.geo2xyz import GeospatialConverter                    # reference
    finally:
        sys.path.remove(REF)
    for seed in (11, 12, 13):
        n = 1000
        geo = np.stack([uniform01(n, seed) * 180 - 90, uniform01(n, seed + 50) * 360 - 180,
                        uniform01(n, seed + 90) * 9500 - 450], 1)
        geo[:4] = [[90, 0, 0], [-90, 12, 100], [0, 180, 0], [0, -180, 5]]
        conv = GeospatialConverter()
        xyz, _ = conv.geodetic_to_xyz(torch.from_numpy(geo))
        norm = conv.xyz_to_norm(xyz).numpy()
        bmin, span = conv.bbox.min_point.numpy(), conv.bbox.span.numpy()
        got, ecef = oracle.geo2xyzt_norm(geo, None, bmin, span, (0.0, 1.0), want_ecef=True)
        assert np.abs(ecef - xyz.numpy()).max() <= 1e-6
        assert np.abs(got[:, :3] - norm.astype(np.float32)).max() <= 1.2e-7
        assert (got[:, 3] == 0).all()
