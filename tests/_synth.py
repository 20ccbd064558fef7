"""Deterministic synthetic inputs shared by the golden generator and the tests.

Nothing here depends on a library RNG stream: tables and coordinates come from a splitmix64
counter hash, so the fixtures under tests/golden/ can be re-derived bit for bit anywhere.
"""
from __future__ import annotations

import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x: np.ndarray) -> np.ndarray:
    x = x.astype(np.uint64)
    with np.errstate(over="ignore"):
        x = (x + np.uint64(0x9E3779B97F4A7C15)) & _M64
        z = x
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return z ^ (z >> np.uint64(31))


def uniform01(n: int, seed: int) -> np.ndarray:
    """n doubles in [0,1) from counter-mode splitmix64."""
    ctr = np.arange(n, dtype=np.uint64) + (np.uint64(seed) << np.uint64(40))
    return (splitmix64(ctr) >> np.uint64(11)).astype(np.float64) * (1.0 / (1 << 53))


def synth_table(T: int, F: int, seed: int, lo: float = -1e-4, hi: float = 1e-4) -> np.ndarray:
    """(T,F) fp32 table in [lo,hi): the reference init range (models/hash_encoding.py:44-45)."""
    u = uniform01(T * F, seed).reshape(T, F)
    return (lo + (hi - lo) * u).astype(np.float32)


def normal(n: int, seed: int) -> np.ndarray:
    """Box-Muller standard normals (fp64)."""
    u1 = np.maximum(uniform01(n, seed), 1e-300)
    u2 = uniform01(n, seed + 7919)
    return np.sqrt(-2.0 * np.log(u1)) * np.cos(2.0 * np.pi * u2)


def coords_mixed(n: int, D: int, seed: int, res_hint: int = 16) -> np.ndarray:
    """(n,D) fp32 coordinates mixing every regime the reference accepts:
    uniform [0,1), standard normal (negative / >1, tests/test_intergrated_model.py:255), the
    SF-bay generator's negative-x box (tests/test_data_generator.py:99-115), exact cell edges
    k/res, 0, 1, -0.0 and tiny magnitudes."""
    q = n // 4
    a = uniform01(q * D, seed).reshape(q, D)
    b = normal(q * D, seed + 1).reshape(q, D)
    c = uniform01(q * D, seed + 2).reshape(q, D) * 0.1
    c[:, 0] = -0.68 + 0.01 * c[:, 0]
    if D > 1:
        c[:, 1] = 0.42 + 0.01 * c[:, 1]
    rem = n - 3 * q
    k = (splitmix64(np.arange(rem * D, dtype=np.uint64) + np.uint64(seed + 3)) % np.uint64(4 * res_hint + 1))
    d = (k.astype(np.float64) - 2 * res_hint).reshape(rem, D) / res_hint  # edges in [-2,2]
    out = np.concatenate([a, b, c, d], 0).astype(np.float32)
    specials = np.array([0.0, -0.0, 1.0, -1.0, 1e-30, -1e-30, 0.5, 0.99999994], dtype=np.float32)
    for i, s in enumerate(specials[: min(len(specials), n)]):
        out[-1 - i, :] = s
    return out


def florida_geo(n: int, seed: int, n_clusters: int = 64):
    """Florida-clustered geodetic points + timestamps.

    K cluster centres uniform in the Central-Florida box (dashboard/dataset_config.json:55-60),
    Zipf(1.0) cluster weights, lat/lon = centre + N(0, 0.1 deg)
    (training/setup_deepearth_data.py:111-112), alt ~ U(0,50) m, t ~ U(2010-01-01, 2025-12-31)
    as POSIX seconds.  Returns geo (n,3) fp64 [lat, lon, alt] and t (n,) fp64.
    """
    N, S, E, W = 28.978, 28.033, -80.902, -81.934
    cu = uniform01(2 * n_clusters, seed).reshape(n_clusters, 2)
    clat = S + (N - S) * cu[:, 0]
    clon = W + (E - W) * cu[:, 1]
    wts = 1.0 / np.arange(1, n_clusters + 1)
    cdf = np.cumsum(wts / wts.sum())
    which = np.minimum(np.searchsorted(cdf, uniform01(n, seed + 11)), n_clusters - 1)
    lat = clat[which] + 0.1 * normal(n, seed + 12)
    lon = clon[which] + 0.1 * normal(n, seed + 13)
    alt = 50.0 * uniform01(n, seed + 14)
    t0, t1 = 1262304000.0, 1767139200.0  # 2010-01-01, 2025-12-31 UTC
    t = t0 + (t1 - t0) * uniform01(n, seed + 15)
    return np.stack([lat, lon, alt], 1), t


FLORIDA_BBOX_GEO = dict(north=28.978, south=28.033, east=-80.902, west=-81.934, alt_min=0.0, alt_max=50.0)
FLORIDA_T_RANGE = (1262304000.0, 1767139200.0)
