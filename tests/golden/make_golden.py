"""Generate tests/golden/*.npz by running the UNMODIFIED reference from /root/reference.

Run in the CPU container only (the reference tree does not exist on the GPU box):
    python tests/golden/make_golden.py
The fixtures pin the oracle (tests/test_oracle_golden.py) and, through it, the CUDA path.
Inputs come from tests/_synth.py (counter-hash, no RNG-stream dependence), so only coordinates
and reference OUTPUTS are stored; tables are re-synthesised by the tests.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, "/root/reference")

from _synth import coords_mixed, synth_table, uniform01, florida_geo  # noqa: E402
from models.hash_encoding import MultiResolutionHashEncoding  # noqa: E402  (reference)
from models.encoders import Grid4DEncoder  # noqa: E402  (reference)
from models.configs import DeepEarthConfig  # noqa: E402  (reference)
from geospatial.geo2xyz import GeospatialConverter  # noqa: E402  (reference)

torch.manual_seed(0)


def load_tables(enc: MultiResolutionHashEncoding, seed0: int):
    for l, emb in enumerate(enc.hash_tables):
        tab = synth_table(enc.hash_table_size, enc.n_features_per_level, seed0 + l)
        with torch.no_grad():
            emb.weight.copy_(torch.from_numpy(tab))


def sparse_rows(g: torch.Tensor):
    rows = torch.nonzero(g.abs().sum(-1)).flatten()
    return rows.numpy().astype(np.int64), g[rows].numpy()


def encoder_case(name, L, F, res, T, D, n, seed):
    enc = MultiResolutionHashEncoding(L, F, res, T, D)
    load_tables(enc, seed0=1000 * seed)
    coords = coords_mixed(n, D, seed, res_hint=res[0])
    x = torch.from_numpy(coords)
    idx = np.stack([enc.hash_coords(x, r).numpy() for r in res[:L]])
    out = enc(x)
    gout = torch.from_numpy((uniform01(n * L * F, seed + 77).reshape(n, L * F) * 2 - 1).astype(np.float32))
    out.backward(gout)
    payload = dict(coords=coords, indices=idx, features=out.detach().numpy(), grad_out=gout.numpy(),
                   meta=np.array([L, F, T, D, seed], dtype=np.int64), resolutions=np.array(res[:L], dtype=np.int64))
    for l, emb in enumerate(enc.hash_tables):
        r, v = sparse_rows(emb.weight.grad)
        payload[f"grad_rows_{l}"] = r
        payload[f"grad_vals_{l}"] = v
    np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **payload)
    print(name, "features", out.shape, "checksum", float(out.double().sum()))


def hash_kat():
    """Known-answer hash indices incl. the SURVEY.md section 8c probes."""
    enc3 = MultiResolutionHashEncoding(1, 2, [16], 2 ** 19, 3)
    pts = torch.tensor([[-0.3, 0.7, -1.2], [0.5, 0.5, 0.5], [1.5, -2.0, 3.0]])
    a = enc3.hash_coords(pts, 16).numpy()
    b = enc3.hash_coords(pts, 524288).numpy()
    enc4 = MultiResolutionHashEncoding(1, 2, [1024], 2 ** 19, 4)
    c = enc4.hash_coords(torch.tensor([[.1, .2, .3, .4], [-.1, -.2, -.3, -.4]]), 1024).numpy()
    assert a.tolist() == [322876, 94504, 41480] and b.tolist() == [22798, 262144, 262144]
    assert c.tolist() == [427944, 70136]
    cases = {}
    k = 0
    for D in (1, 2, 3, 4):
        for T in (2 ** 19, 2 ** 17, 1000003, 50000, 7):
            for R in (1, 4, 16, 1000, 1024, 524288):
                enc = MultiResolutionHashEncoding(1, 2, [R], T, D)
                coords = coords_mixed(256, D, 31 * k + 5, res_hint=max(R, 2) if R <= 1024 else 16)
                cases[f"c{k}_coords"] = coords
                cases[f"c{k}_idx"] = enc.hash_coords(torch.from_numpy(coords), R).numpy()
                cases[f"c{k}_meta"] = np.array([D, T, R], dtype=np.int64)
                k += 1
    cases["n_cases"] = np.array([k], dtype=np.int64)
    cases["kat_pts"] = pts.numpy()
    cases["kat_r16"] = a
    cases["kat_r524288"] = b
    cases["kat4_pts"] = np.array([[.1, .2, .3, .4], [-.1, -.2, -.3, -.4]], dtype=np.float32)
    cases["kat4_r1024"] = c
    np.savez_compressed(os.path.join(HERE, "hash_kat.npz"), **cases)
    print("hash_kat", k, "cases")


def grid4d_encoder_case():
    """Reference Grid4DEncoder (models/encoders.py:12-85) with masks, small hidden_dim."""
    cfg = DeepEarthConfig(hidden_dim=32, n_heads=4, n_layers=1, hash_table_size=2 ** 14)
    enc = Grid4DEncoder(cfg)
    load_tables(enc.spatial_encoder, seed0=5000)
    load_tables(enc.temporal_encoder, seed0=6000)
    # projection weights: deterministic synthetic
    with torch.no_grad():
        for i, p in enumerate(enc.projection.parameters()):
            p.copy_(torch.from_numpy((uniform01(p.numel(), 900 + i).reshape(p.shape) * 0.4 - 0.2).astype(np.float32)))
    n = 512
    xyzt = coords_mixed(n, 4, 42, res_hint=16)
    sm = (uniform01(n, 43) > 0.3)
    tm = (uniform01(n, 44) > 0.5)
    captured = {}
    enc.projection.register_forward_hook(lambda m, i, o: captured.__setitem__("feat", i[0].detach().clone()))
    x = torch.from_numpy(xyzt)
    emb_nomask = enc(x).detach().numpy()
    feat_nomask = captured["feat"].numpy()
    emb = enc(x, spatial_mask=torch.from_numpy(sm), temporal_mask=torch.from_numpy(tm))
    feat = captured["feat"].numpy()
    gout = torch.from_numpy((uniform01(n * 32, 45).reshape(n, 32) * 2 - 1).astype(np.float32))
    emb.backward(gout)
    payload = dict(xyzt=xyzt, spatial_mask=sm, temporal_mask=tm, features_nomask=feat_nomask,
                   embedding_nomask=emb_nomask, features=feat, embedding=emb.detach().numpy(),
                   grad_out=gout.numpy(), state_keys=np.array(list(enc.state_dict().keys())))
    for nm, sub in (("s", enc.spatial_encoder), ("t", enc.temporal_encoder)):
        for l, e in enumerate(sub.hash_tables):
            r, v = sparse_rows(e.weight.grad)
            payload[f"grad_rows_{nm}{l}"] = r
            payload[f"grad_vals_{nm}{l}"] = v
    for i, p in enumerate(enc.projection.parameters()):
        payload[f"proj_grad_{i}"] = p.grad.numpy()
    np.savez_compressed(os.path.join(HERE, "grid4d_encoder.npz"), **payload)
    print("grid4d_encoder", emb.shape, list(enc.state_dict().keys())[:3])


def geo_case():
    """GeospatialConverter.geodetic_to_xyz + xyz_to_norm (geo2xyz.py:128-160, :327-332)."""
    landmarks = np.array([  # geospatial/tests/geospatial_tests.py:364-381 style landmarks
        [27.9881, 86.9250, 8848.86], [31.5590, 35.4732, -430.5], [0.0, 0.0, 0.0], [90.0, 0.0, 0.0],
        [-90.0, 0.0, 0.0], [0.0, 180.0, 0.0], [0.0, -180.0, 0.0], [48.8584, 2.2945, 330.0],
        [40.6892, -74.0445, 93.0], [-33.8568, 151.2153, 65.0], [28.5, -81.4, 25.0]], dtype=np.float64)
    fl, t = florida_geo(2048, 7)
    glob = np.stack([uniform01(512, 1) * 180 - 90, uniform01(512, 2) * 360 - 180, uniform01(512, 3) * 9000 - 400], 1)
    payload = {}
    for name, geo in (("landmarks", landmarks), ("florida", fl), ("global", glob)):
        conv = GeospatialConverter()
        xyz, _ = conv.geodetic_to_xyz(torch.from_numpy(geo))
        norm = conv.xyz_to_norm(xyz)
        bb = conv.bbox
        payload[f"{name}_geo"] = geo
        payload[f"{name}_ecef"] = xyz.numpy()
        payload[f"{name}_norm"] = norm.numpy()
        payload[f"{name}_norm_dtype"] = np.array([str(norm.dtype)])
        payload[f"{name}_bbox_min"] = bb.min_point.numpy()
        payload[f"{name}_bbox_span"] = bb.span.numpy()
    payload["florida_t"] = t
    # degenerate span: every point identical -> 0.5 (utils.py:54-77)
    conv = GeospatialConverter()
    same = np.repeat(landmarks[-1:], 4, 0)
    xyz, _ = conv.geodetic_to_xyz(torch.from_numpy(same))
    payload["degenerate_geo"] = same
    payload["degenerate_norm"] = conv.xyz_to_norm(xyz).numpy()
    np.savez_compressed(os.path.join(HERE, "geo2xyz.npz"), **payload)
    print("geo2xyz ok; florida norm dtype", payload["florida_norm_dtype"])


if __name__ == "__main__":
    hash_kat()
    # BASELINE.json config 1 geometry (spatial encoder), default Grid4DEncoder halves, odd shapes
    encoder_case("enc_d3_l16_t19", 16, 2, [2 ** i for i in range(4, 20)], 2 ** 19, 3, 1024, 1)
    encoder_case("enc_d4_l16_t19", 16, 2, [2 ** i for i in range(4, 20)], 2 ** 19, 4, 1024, 2)
    encoder_case("enc_d1_l8_t17", 8, 2, [2 ** i for i in range(2, 10)], 2 ** 17, 1, 1024, 3)
    encoder_case("enc_d2_l5_f4_npow2", 5, 4, [3, 10, 33, 100, 1000], 1000, 2, 777, 4)
    encoder_case("enc_d3_l3_f1_prime", 3, 1, [7, 50, 300], 65521, 3, 513, 5)
    encoder_case("enc_d3_l4_f8", 4, 8, [8, 16, 64, 512], 4096, 3, 300, 6)
    encoder_case("enc_d3_l2_f3_b1", 2, 3, [16, 128], 2 ** 10, 3, 1, 7)
    grid4d_encoder_case()
    geo_case()
