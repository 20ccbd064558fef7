"""Pin the CPU oracle against fixtures produced by the unmodified reference
(tests/golden/make_golden.py).  CPU only."""
import os

import numpy as np
import pytest

from _synth import synth_table
from oracle.grid4d_oracle import (Level, numpy_backward_nearest, numpy_forward_nearest,
                                  numpy_hash_coords)

ENC_CASES = ["enc_d3_l16_t19", "enc_d4_l16_t19", "enc_d1_l8_t17", "enc_d2_l5_f4_npow2",
             "enc_d3_l3_f1_prime", "enc_d3_l4_f8", "enc_d3_l2_f3_b1"]


def load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name + ".npz"), allow_pickle=False)


def enc_levels(meta, res):
    L, F, T, D, seed = (int(v) for v in meta)
    levels = [Level(T=T, R=int(r), dims=tuple(range(D)), out_col=l * F, table_off=l * T)
              for l, r in enumerate(res)]
    tables = np.concatenate([synth_table(T, F, 1000 * seed + l) for l in range(L)], 0)
    return levels, tables, (L, F, T, D)


def test_hash_known_answers(oracle, golden_dir):
    g = load(golden_dir, "hash_kat")
    # SURVEY.md section 8c probe values
    lv16 = [Level(T=2 ** 19, R=16, dims=(0, 1, 2), out_col=0)]
    assert oracle.hash_indices(lv16, g["kat_pts"], False)[0][0].tolist() == [322876, 94504, 41480]
    lvf = [Level(T=2 ** 19, R=524288, dims=(0, 1, 2), out_col=0)]
    assert oracle.hash_indices(lvf, g["kat_pts"], False)[0][0].tolist() == [22798, 262144, 262144]
    lv4 = [Level(T=2 ** 19, R=1024, dims=(0, 1, 2, 3), out_col=0)]
    assert oracle.hash_indices(lv4, g["kat4_pts"], False)[0][0].tolist() == [427944, 70136]
    for k in range(int(g["n_cases"][0])):
        D, T, R = (int(v) for v in g[f"c{k}_meta"])
        coords, want = g[f"c{k}_coords"], g[f"c{k}_idx"]
        got_c = oracle.hash_indices([Level(T=T, R=R, dims=tuple(range(D)), out_col=0)], coords, False)[0][0]
        got_np = numpy_hash_coords(coords, R, T)
        assert np.array_equal(got_c, want), f"C oracle case {k} D={D} T={T} R={R}"
        assert np.array_equal(got_np, want), f"numpy oracle case {k} D={D} T={T} R={R}"


@pytest.mark.parametrize("name", ENC_CASES)
def test_encoder_forward_backward_vs_reference(oracle, golden_dir, name):
    g = load(golden_dir, name)
    levels, tables, (L, F, T, D) = enc_levels(g["meta"], g["resolutions"])
    coords = g["coords"]
    idx = oracle.hash_indices(levels, coords, False)
    for l in range(L):
        assert np.array_equal(idx[l][0], g["indices"][l]), f"level {l} indices"
    out = oracle.forward(levels, coords, tables, F, linear=False)
    assert np.array_equal(out, g["features"])  # pure gather: bit exact
    per_level = [tables[l * T:(l + 1) * T] for l in range(L)]
    assert np.array_equal(numpy_forward_nearest(coords, per_level, g["resolutions"]), g["features"])
    grads = oracle.backward(levels, coords, g["grad_out"], L * T, F, linear=False)
    np_grads = numpy_backward_nearest(coords, g["grad_out"], g["resolutions"], T, F)
    for l in range(L):
        dense = np.zeros((T, F), dtype=np.float64)
        dense[g[f"grad_rows_{l}"]] = g[f"grad_vals_{l}"]
        got = grads[l * T:(l + 1) * T]
        scale = max(np.abs(dense).max(), 1e-30)
        assert np.abs(got - dense).max() <= 1e-5 * scale   # fp32 vs fp64 accumulation order
        assert np.abs(np_grads[l] - got).max() <= 1e-12 * scale


def test_grid4d_encoder_features_with_masks(oracle, golden_dir):
    """Feature half of the reference Grid4DEncoder incl. row masks (models/encoders.py:68-82)."""
    g = load(golden_dir, "grid4d_encoder")
    T, F = 2 ** 14, 2
    levels, tabs = [], []
    off = 0
    for l in range(16):
        levels.append(Level(T=T, R=2 ** (4 + l), dims=(0, 1, 2), out_col=l * F, table_off=off, mask_id=1))
        tabs.append(synth_table(T, F, 5000 + l))
        off += T
    for l in range(8):
        levels.append(Level(T=T // 4, R=2 ** (2 + l), dims=(3,), out_col=32 + l * F, table_off=off, mask_id=2))
        tabs.append(synth_table(T // 4, F, 6000 + l))
        off += T // 4
    tables = np.concatenate(tabs, 0)
    out = oracle.forward(levels, g["xyzt"], tables, F, linear=False)
    assert np.array_equal(out, g["features_nomask"])
    out_m = oracle.forward(levels, g["xyzt"], tables, F, linear=False,
                           mask_a=g["spatial_mask"], mask_b=g["temporal_mask"])
    assert np.array_equal(out_m, g["features"])
    assert (out_m[~g["spatial_mask"], :32] == 0).all() and (out_m[~g["temporal_mask"], 32:] == 0).all()


def test_linear_mode_reduces_to_reference(oracle, golden_dir):
    """Stated Mode-G math vs pinned pieces: corner 0 of a hashed level is the reference index;
    every corner index equals the reference hash of the shifted integer cell; weights sum to 1;
    a constant table interpolates to that constant."""
    g = load(golden_dir, "enc_d3_l16_t19")
    levels, tables, (L, F, T, D) = enc_levels(g["meta"], g["resolutions"])
    coords = g["coords"]
    lin = oracle.hash_indices(levels, coords, True)
    for l in range(L):
        assert np.array_equal(lin[l][0], g["indices"][l])
        R = int(g["resolutions"][l])
        cell = np.floor(coords * np.float32(R)).astype(np.float32)
        for k in range(8):
            off = np.array([(k >> d) & 1 for d in range(3)], dtype=np.float32)
            # reference hash of the integer cell: hash_coords(cell + off, resolution=1)
            want = numpy_hash_coords(cell + off, 1, T)
            ok = np.abs(cell).max(axis=1) < 2 ** 23     # cell+off exactly representable in fp32
            assert np.array_equal(lin[l][k][ok], want[ok])
    const = np.full_like(tables, 0.25)
    out = oracle.forward(levels, coords, const, F, linear=True)
    assert np.abs(out - 0.25).max() < 1e-6


def test_dense_fallback_indexing(oracle):
    """Dense levels: idx = x + y(R+1) + z(R+1)^2 for in-range coords, floor-mod T outside."""
    R, T = 16, 2 ** 19
    lv = [Level(T=T, R=R, dims=(0, 1, 2), out_col=0, dense=True)]
    pts = np.array([[0, 0, 0], [0.5, 0.25, 0.75], [0.999, 0.999, 0.999], [-0.1, 0.0, 0.0]], dtype=np.float32)
    idx = oracle.hash_indices(lv, pts, True)[0]
    assert idx[0].tolist()[:3] == [0, 8 + 4 * 17 + 12 * 289, 15 + 15 * 17 + 15 * 289]
    assert idx[7][2] == 16 + 16 * 17 + 16 * 289          # far corner of the last cell < (R+1)^3
    assert idx[0][3] == T - 2                             # cell (-2,0,0) -> -2 mod T
    assert idx.min() >= 0 and idx.max() < T


def test_geo2xyz_vs_reference(oracle, golden_dir):
    g = load(golden_dir, "geo2xyz")
    for name in ("landmarks", "florida", "global"):
        xyzt, ecef = oracle.geo2xyzt_norm(g[f"{name}_geo"], None, g[f"{name}_bbox_min"],
                                          g[f"{name}_bbox_span"], (0.0, 1.0), want_ecef=True)
        assert np.abs(ecef - g[f"{name}_ecef"]).max() <= 2e-9     # metres; libm vs ATen sin/cos
        want = g[f"{name}_norm"].astype(np.float32)
        assert np.abs(xyzt[:, :3] - want).max() <= 1.2e-7
    geo = g["degenerate_geo"]
    xyzt, ecef = oracle.geo2xyzt_norm(geo, None, [0, 0, 0], [1, 1, 1], (0, 1), want_ecef=True)
    x2 = oracle.geo2xyzt_norm(geo, None, ecef.min(0), ecef.max(0) - ecef.min(0), (0, 1))
    assert np.array_equal(x2[:, :3], g["degenerate_norm"].astype(np.float32))   # all 0.5


def scenario_tolerance(span):
    """|d norm| allowed per axis: fp32 rounding of a value in [0, 1] plus the fp64 ECEF ulp noise (libm / CUDA vs ATen
    sin, cos: <= 4e-9 m) amplified by a small box -- the mm-precision scenarios have spans of centimetres."""
    span = np.abs(np.asarray(span, np.float64))
    return 6e-8 + 4e-9 / np.maximum(span, 1e-9)


def test_geo2xyz_reference_scenario_groups(oracle, golden_dir):
    """The reference's own geodetic scenario groups (geospatial/tests/geospatial_tests.py:67-215), captured by running
    its test routine (tests/golden/make_golden_geo_scenarios.py): ECEF and bbox-normalised coordinates per group."""
    g = load(golden_dir, "geo_scenarios")
    n = len(g["names"])
    assert n >= 30
    for i in range(n):
        geo, lo, span = g[f"s{i}_geo"], g[f"s{i}_bbox_min"], g[f"s{i}_bbox_span"]
        xyzt, ecef = oracle.geo2xyzt_norm(geo, None, lo, span, (0.0, 1.0), want_ecef=True)
        assert np.abs(ecef - g[f"s{i}_ecef"]).max() <= 4e-9, (g["names"][i], np.abs(ecef - g[f"s{i}_ecef"]).max())
        err = np.abs(xyzt[:, :3].astype(np.float64) - g[f"s{i}_norm"])
        assert (err <= scenario_tolerance(span)[None, :]).all(), (g["names"][i], err.max(0), span)


def kat_levels(g):
    """Level list + tables of tests/golden/modeg_kat.npz (make_golden_modeg.py: table[i][f] = ((7i+3f+11l)%64-32)/64)."""
    F = int(g["F"][0])
    levels, tabs, off = [], [], 0
    for l in range(int(g["n_levels"][0])):
        R, T, dense, *dims = (int(v) for v in g[f"meta_{l}"])
        levels.append(Level(T=T, R=R, dims=tuple(dims), out_col=l * F, table_off=off, dense=bool(dense)))
        i = np.arange(T, dtype=np.int64)[:, None]
        f = np.arange(F, dtype=np.int64)[None, :]
        tabs.append((((7 * i + 3 * f + 11 * l) % 64 - 32) / 64.0).astype(np.float32))
        off += T
    return levels, tabs, F


def test_modeg_known_answers(oracle, golden_dir):
    """The interpolated mode has no reference implementation: its stated math (grid4d_oracle.c header) is frozen by
    committed known-answer vectors -- dense, hashed and prime-T levels, in- and out-of-range points -- and by one
    value derived by hand (make_golden_modeg.py header): level A, point 0 -> corners, weights and feature."""
    g = load(golden_dir, "modeg_kat")
    levels, tabs, F = kat_levels(g)
    pts = g["points"]
    assert g["idx_0"][:, 0].tolist() == [86, 87, 91, 92, 111, 112, 116, 117]
    assert g["w_0"][:, 0].tolist() == [.28125, .28125, .09375, .09375, .09375, .09375, .03125, .03125]
    assert float(g["features"][0, 0]) == -0.0625
    idx = oracle.hash_indices(levels, pts, True)
    for l in range(len(levels)):
        assert np.array_equal(idx[l], g[f"idx_{l}"]), l
    tables = np.concatenate(tabs, 0)
    out = oracle.forward(levels, pts, tables, F, True)
    assert np.array_equal(out, g["features"])
    # fp16 working tables hold these values exactly (multiples of 1/64 times 2^12): same features
    half = (tables * np.float32(4096.0)).astype(np.float16)
    assert np.array_equal(half.astype(np.float32), tables * np.float32(4096.0))
    assert np.abs(oracle.forward(levels, pts, half, F, True, out_scale=1.0 / 4096.0) - g["features"]).max() <= 1e-7
    grad = oracle.backward(levels, pts, g["grad_out"], sum(lv.T for lv in levels), F, True)
    for l, lv in enumerate(levels):
        want = np.zeros((lv.T, F))
        want[g[f"grad_rows_{l}"]] = g[f"grad_vals_{l}"]
        assert np.abs(grad[lv.table_off:lv.table_off + lv.T] - want).max() <= 1e-12, l
    # the torch restatement (the device-resident checker of the full-size GPU tests) reproduces them too
    import torch
    from oracle.grid4d_torch import grid4d_forward, level_indices_weights
    tp = torch.from_numpy(pts)
    for l, lv in enumerate(levels):
        it, wt = level_indices_weights(tp, lv.dims, lv.R, lv.T, lv.dense)
        assert np.array_equal(it.numpy().T, g[f"idx_{l}"]) and np.array_equal(wt.numpy().T, g[f"w_{l}"]), l
    t_out = grid4d_forward(tp, [torch.from_numpy(t) for t in tabs], [(lv.dims, lv.R, lv.T, lv.dense) for lv in levels], True)
    assert np.abs(t_out.numpy() - g["features"]).max() <= 2e-7
