"""Parity of the CUDA path (through the nn.Module / C-ABI boundary) against the CPU oracle and the
reference's golden vectors.  Needs a GPU: run with ``-m gpu`` on the B200 box.

Tolerances (BASELINE.json north_star): hash indices bit-exact; forward features 1e-5 relative with
fp32 tables (nearest mode is a pure gather and must be bit-exact), 1e-2 with fp16 tables; table
gradients 1e-4 relative (atomic ordering).  "Relative" is taken against max|reference| of the tensor
compared, because individual interpolated features can cancel to ~0.
"""
import os

import numpy as np
import pytest
import torch

from _synth import coords_mixed, synth_table, uniform01

pytestmark = pytest.mark.gpu

FWD_TOL_F32 = 1e-5
FWD_TOL_F16 = 1e-2
GRAD_TOL = 1e-4


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")


def rel_err(got, want):
    want = np.asarray(want, dtype=np.float64)
    return float(np.abs(np.asarray(got, dtype=np.float64) - want).max() / max(np.abs(want).max(), 1e-30))


def load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name + ".npz"), allow_pickle=False)


def make_encoder(de, L, F, res, T, D, seed, dev, **kw):
    enc = de.MultiResolutionHashEncoding(L, F, list(res), T, D, **kw).to(dev)
    tabs = [synth_table(T, F, 1000 * seed + l) for l in range(L)]
    with torch.no_grad():
        for emb, t in zip(enc.hash_tables, tabs):
            emb.weight.copy_(torch.from_numpy(t))
    return enc, np.concatenate(tabs, 0)


def oracle_levels(enc, linear):
    from oracle.grid4d_oracle import Level
    T, F = enc.hash_table_size, enc.n_features_per_level
    return [Level(T=s.table_size, R=s.resolution, dims=s.dims, out_col=s.out_col, table_off=l * T, dense=s.dense)
            for l, s in enumerate(enc.level_specs())]


ENC_CASES = ["enc_d3_l16_t19", "enc_d4_l16_t19", "enc_d1_l8_t17", "enc_d2_l5_f4_npow2",
             "enc_d3_l3_f1_prime", "enc_d3_l4_f8", "enc_d3_l2_f3_b1"]


# ---------------------------------------------------------------------------------------------------
# reference mode (nearest) against golden vectors produced by the unmodified reference
# ---------------------------------------------------------------------------------------------------
def test_hash_known_answers(dev, golden_dir):
    import deepearth_b200 as de
    g = load(golden_dir, "hash_kat")
    enc3 = de.MultiResolutionHashEncoding(1, 2, [16], 2 ** 19, 3).to(dev)
    pts = torch.from_numpy(g["kat_pts"]).to(dev)
    assert enc3.hash_coords(pts, 16).tolist() == [322876, 94504, 41480]          # SURVEY.md 8c probes
    assert enc3.hash_coords(pts, 524288).tolist() == [22798, 262144, 262144]
    enc4 = de.MultiResolutionHashEncoding(1, 2, [1024], 2 ** 19, 4).to(dev)
    assert enc4.hash_coords(torch.from_numpy(g["kat4_pts"]).to(dev), 1024).tolist() == [427944, 70136]
    for k in range(int(g["n_cases"][0])):
        D, T, R = (int(v) for v in g[f"c{k}_meta"])
        enc = de.MultiResolutionHashEncoding(1, 2, [R], T, D)
        got = enc.hash_coords(torch.from_numpy(g[f"c{k}_coords"]).to(dev), R)
        assert got.dtype == torch.int64
        assert np.array_equal(got.cpu().numpy(), g[f"c{k}_idx"]), f"case {k}: D={D} T={T} R={R}"


@pytest.mark.parametrize("name", ENC_CASES)
def test_encoder_vs_reference_golden(dev, golden_dir, name):
    import deepearth_b200 as de
    g = load(golden_dir, name)
    L, F, T, D, seed = (int(v) for v in g["meta"])
    enc, _ = make_encoder(de, L, F, g["resolutions"], T, D, seed, dev)
    coords = torch.from_numpy(g["coords"]).to(dev).requires_grad_(True)
    for l, R in enumerate(g["resolutions"]):
        assert np.array_equal(enc.hash_coords(coords.detach(), int(R)).cpu().numpy(), g["indices"][l])
    out = enc(coords)
    assert out.shape == (coords.shape[0], L * F) and out.dtype == torch.float32
    assert np.array_equal(out.detach().cpu().numpy(), g["features"])            # pure gather: bit exact
    out.backward(torch.from_numpy(g["grad_out"]).to(dev))
    assert coords.grad is None                                                  # hash_encoding.py:61 cuts the graph
    for l, emb in enumerate(enc.hash_tables):
        want = np.zeros((T, F), dtype=np.float32)
        want[g[f"grad_rows_{l}"]] = g[f"grad_vals_{l}"]
        assert emb.weight.grad.shape == (T, F)
        assert rel_err(emb.weight.grad.cpu().numpy(), want) <= GRAD_TOL


def test_grid4d_encoder_vs_reference_golden(dev, golden_dir):
    import deepearth_b200 as de
    g = load(golden_dir, "grid4d_encoder")
    cfg = de.DeepEarthConfig(hidden_dim=32, n_heads=4, n_layers=1, hash_table_size=2 ** 14)
    enc = de.Grid4DEncoder(cfg).to(dev)
    with torch.no_grad():
        for l, e in enumerate(enc.spatial_encoder.hash_tables):
            e.weight.copy_(torch.from_numpy(synth_table(2 ** 14, 2, 5000 + l)))
        for l, e in enumerate(enc.temporal_encoder.hash_tables):
            e.weight.copy_(torch.from_numpy(synth_table(2 ** 12, 2, 6000 + l)))
        for i, p in enumerate(enc.projection.parameters()):
            p.copy_(torch.from_numpy((uniform01(p.numel(), 900 + i).reshape(p.shape) * 0.4 - 0.2).astype(np.float32)))
    xyzt = torch.from_numpy(g["xyzt"]).to(dev)
    sm = torch.from_numpy(g["spatial_mask"]).to(dev)
    tm = torch.from_numpy(g["temporal_mask"]).to(dev)
    assert np.array_equal(enc.encode_features(xyzt).detach().cpu().numpy(), g["features_nomask"])
    assert np.array_equal(enc.encode_features(xyzt, sm, tm).detach().cpu().numpy(), g["features"])
    assert rel_err(enc(xyzt).detach().cpu().numpy(), g["embedding_nomask"]) <= 1e-5
    emb = enc(xyzt, spatial_mask=sm, temporal_mask=tm)
    assert rel_err(emb.detach().cpu().numpy(), g["embedding"]) <= 1e-5
    emb.backward(torch.from_numpy(g["grad_out"]).to(dev))
    for nm, sub, T in (("s", enc.spatial_encoder, 2 ** 14), ("t", enc.temporal_encoder, 2 ** 12)):
        for l, e in enumerate(sub.hash_tables):
            want = np.zeros((T, 2), dtype=np.float32)
            want[g[f"grad_rows_{nm}{l}"]] = g[f"grad_vals_{nm}{l}"]
            assert rel_err(e.weight.grad.cpu().numpy(), want) <= GRAD_TOL, (nm, l)
    for i, p in enumerate(enc.projection.parameters()):
        assert rel_err(p.grad.cpu().numpy(), g[f"proj_grad_{i}"]) <= 1e-4
    # the stand-alone sub-encoders keep the reference call signature
    s = enc.spatial_encoder(xyzt[:, :3])
    t = enc.temporal_encoder(xyzt[:, 3:4])
    assert np.array_equal(torch.cat([s, t], -1).detach().cpu().numpy(), g["features_nomask"])


# ---------------------------------------------------------------------------------------------------
# seeded inputs against the C oracle: both modes, both table dtypes
# ---------------------------------------------------------------------------------------------------
CFGS = [
    # L, F, resolutions, T, D, n, linear, dense
    (16, 2, [2 ** i for i in range(4, 20)], 2 ** 19, 3, 20000, False, False),
    (16, 2, [2 ** i for i in range(4, 20)], 2 ** 19, 3, 20000, True, True),
    (8, 2, [2 ** i for i in range(2, 10)], 2 ** 17, 1, 5000, True, True),
    (6, 4, [5, 16, 47, 128, 500, 4096], 2 ** 12, 2, 7001, True, True),
    (5, 2, [4, 8, 16, 32, 64], 2 ** 16, 4, 3000, True, True),
    (5, 1, [3, 9, 27, 81, 243], 30011, 3, 4099, True, True),      # non power-of-two table: int64 path
    (3, 8, [16, 64, 256], 2 ** 10, 3, 2500, True, False),
    (2, 3, [16, 128], 1000, 3, 1025, True, True),                 # runtime-F path
]


@pytest.mark.parametrize("cfg", CFGS, ids=lambda c: f"L{c[0]}F{c[1]}T{c[3]}D{c[4]}{'lin' if c[6] else 'nn'}")
def test_modes_vs_oracle_fp32(dev, oracle, cfg):
    import deepearth_b200 as de
    from deepearth_b200 import functional as Fn
    L, F, res, T, D, n, linear, dense = cfg
    enc, tables = make_encoder(de, L, F, res, T, D, 11, dev, interpolation="linear" if linear else "nearest",
                               dense_fallback=dense)
    levels = oracle_levels(enc, linear)
    assert any(lv.dense for lv in levels) == (dense and (res[0] + 1) ** D <= T)
    coords_np = coords_mixed(n, D, 5, res_hint=res[0])
    coords = torch.from_numpy(coords_np).to(dev)
    idx = Fn.hash_indices(enc.level_specs(), coords, linear).cpu().numpy()
    want_idx = oracle.hash_indices(levels, coords_np, linear)
    for l in range(L):
        assert np.array_equal(idx[l], want_idx[l]), f"level {l}"            # bit exact, every corner
    out = enc(coords)
    want = oracle.forward(levels, coords_np, tables, F, linear)
    err = rel_err(out.detach().cpu().numpy(), want)
    assert err <= (FWD_TOL_F32 if linear else 0.0), err
    gout_np = (uniform01(n * L * F, 99).reshape(n, L * F) * 2 - 1).astype(np.float32)
    out.backward(torch.from_numpy(gout_np).to(dev))
    want_g = oracle.backward(levels, coords_np, gout_np, L * T, F, linear)
    for l, emb in enumerate(enc.hash_tables):
        assert rel_err(emb.weight.grad.cpu().numpy(), want_g[l * T:(l + 1) * T]) <= GRAD_TOL, f"level {l}"


def test_fp16_tables_vs_oracle(dev, oracle):
    import deepearth_b200 as de
    L, F, T, D, n = 16, 2, 2 ** 19, 3, 20000
    res = [2 ** i for i in range(4, 20)]
    enc, tables = make_encoder(de, L, F, res, T, D, 13, dev, interpolation="linear", dense_fallback=True,
                               table_dtype="float16")
    assert enc.table_scale == 4096.0
    levels = oracle_levels(enc, True)
    coords_np = coords_mixed(n, D, 6)
    out = enc(torch.from_numpy(coords_np).to(dev)).detach().cpu().numpy()
    half_tables = (tables * np.float32(enc.table_scale)).astype(np.float16)
    assert np.array_equal(enc._work.cpu().numpy().reshape(-1, F), half_tables)   # cast kernel: exact
    want_h = oracle.forward(levels, coords_np, half_tables, F, True, out_scale=1.0 / enc.table_scale)
    assert rel_err(out, want_h) <= FWD_TOL_F32              # same fp16 tables: only accumulation order
    want_f = oracle.forward(levels, coords_np, tables, F, True)
    assert rel_err(out, want_f) <= FWD_TOL_F16              # vs fp32 tables: the stated fp16 tolerance
    # unscaled fp16 storage of the 1e-4 init is subnormal -- show why the scale is there
    enc1, _ = make_encoder(de, L, F, res, T, D, 13, dev, interpolation="linear", dense_fallback=True,
                           table_dtype="float16", table_scale=1.0)
    out1 = enc1(torch.from_numpy(coords_np).to(dev)).detach().cpu().numpy()
    assert rel_err(out1, want_f) <= FWD_TOL_F16
    # max error can tie (both hit the same fp16-normal entry); the scale shows in the mean error
    assert np.abs(out - want_f).mean() < np.abs(out1 - want_f).mean()


def test_decomposed_encoder_vs_oracle(dev, oracle):
    """BASELINE config 2 geometry at reduced size: xyz/xyt/yzt/xzt x 16 levels, fp16 tables, fp32 grads."""
    import deepearth_b200 as de
    from oracle.grid4d_oracle import GRID_DIMS, Level
    T, F, L, n = 2 ** 16, 2, 16, 6000
    res = [2 ** i for i in range(4, 20)]
    cfg = de.DeepEarthConfig(hidden_dim=16, n_heads=4, n_layers=1, hash_table_size=T, n_temporal_levels=16,
                             temporal_resolutions=res, grid4d_mode="decomposed", grid4d_table_dtype="float16")
    enc = de.Grid4DEncoder(cfg).to(dev)
    assert enc.feature_dim == 128
    names = ["xyz", "xyt", "yzt", "xzt"]
    levels, tabs = [], []
    for gi, nm in enumerate(names):
        sub = getattr(enc, f"{nm}_encoder")
        for l in range(L):
            t = synth_table(T, F, 7000 + 100 * gi + l)
            tabs.append(t)
            with torch.no_grad():
                sub.hash_tables[l].weight.copy_(torch.from_numpy(t))
            levels.append(Level(T=T, R=res[l], dims=GRID_DIMS[nm], out_col=(gi * L + l) * F,
                                table_off=(gi * L + l) * T, dense=(res[l] + 1) ** 3 <= T))
    tables = np.concatenate(tabs, 0)
    xyzt_np = coords_mixed(n, 4, 21)
    xyzt = torch.from_numpy(xyzt_np).to(dev)
    feats = enc.encode_features(xyzt)
    half = (tables * np.float32(4096.0)).astype(np.float16)
    want = oracle.forward(levels, xyzt_np, half, F, True, out_scale=1.0 / 4096.0)
    assert rel_err(feats.detach().cpu().numpy(), want) <= FWD_TOL_F32
    assert rel_err(feats.detach().cpu().numpy(), oracle.forward(levels, xyzt_np, tables, F, True)) <= FWD_TOL_F16
    gout_np = (uniform01(n * 128, 5).reshape(n, 128) * 2 - 1).astype(np.float32)
    feats.backward(torch.from_numpy(gout_np).to(dev))
    want_g = oracle.backward(levels, xyzt_np, gout_np, 4 * L * T, F, True)
    k = 0
    for nm in names:
        for e in getattr(enc, f"{nm}_encoder").hash_tables:
            assert e.weight.grad.dtype == torch.float32
            assert rel_err(e.weight.grad.cpu().numpy(), want_g[k * T:(k + 1) * T]) <= GRAD_TOL, (nm, k)
            k += 1
    # spec masking: coordinates, not features, are zeroed (SPECIFICATIONS.MD:1159-1164)
    sm = torch.from_numpy(uniform01(n, 8) > 0.5).to(dev)
    masked = xyzt_np.copy()
    masked[~sm.cpu().numpy(), :3] = 0
    want_m = oracle.forward(levels, masked, half, F, True, out_scale=1.0 / 4096.0)
    assert rel_err(enc.encode_features(xyzt, spatial_mask=sm).detach().cpu().numpy(), want_m) <= FWD_TOL_F32


# ---------------------------------------------------------------------------------------------------
# edge cases
# ---------------------------------------------------------------------------------------------------
def test_edge_shapes(dev, oracle):
    import deepearth_b200 as de
    enc, tables = make_encoder(de, 4, 2, [16, 32, 64, 128], 2 ** 12, 3, 3, dev, interpolation="linear")
    empty = enc(torch.empty((0, 3), device=dev))
    assert empty.shape == (0, 8)
    one = enc(torch.tensor([[0.3, -0.2, 1.7]], device=dev))
    assert one.shape == (1, 8)
    # non-contiguous input: a column slice of an xyzt batch is consumed in place (row stride 4)
    xyzt_np = coords_mixed(1000, 4, 9)
    xyzt = torch.from_numpy(xyzt_np).to(dev)
    a = enc(xyzt[:, :3])
    b = enc(xyzt[:, :3].contiguous())
    assert torch.equal(a, b)
    t = xyzt.t().contiguous().t()[:, :3]          # column-major view: needs the copy path
    assert torch.equal(enc(t), b)
    with pytest.raises(TypeError):
        enc(xyzt[:, :3].double())
    with pytest.raises(ValueError):
        enc(xyzt[:, :2])
    # levels_per_group override gives identical results
    enc.levels_per_group = 1
    assert torch.equal(enc(xyzt[:, :3]), b)
    enc.levels_per_group = 2
    assert torch.equal(enc(xyzt[:, :3]), b)
    # no_grad / eval and autocast keep fp32 features
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        c = enc(xyzt[:, :3])
    assert c.dtype == torch.float32 and torch.equal(c, b)


def test_frozen_tables_and_grad_accumulation(dev):
    import deepearth_b200 as de
    enc, _ = make_encoder(de, 3, 2, [16, 32, 64], 2 ** 10, 3, 4, dev)
    enc.hash_tables[1].weight.requires_grad_(False)
    x = torch.rand(500, 3, device=dev)
    enc(x).sum().backward()
    g0 = enc.hash_tables[0].weight.grad.clone()
    assert enc.hash_tables[1].weight.grad is None
    enc(x).sum().backward()                        # second backward accumulates
    assert torch.allclose(enc.hash_tables[0].weight.grad, 2 * g0)
    assert float(g0.sum()) == pytest.approx(500 * 2, rel=1e-6)


def test_training_step_contract_autocast_scaler_clip(dev):
    """One optimiser step the way the reference's trainer drives the encoder (hpc/train_distrbuted.py:284-313:
    autocast, GradScaler, unscale_ + clip_grad_norm_, AdamW) against the torch restatement of the reference
    modules holding the same state dict; also state-dict interchange in both directions and .eval() / no_grad."""
    import deepearth_b200 as de
    from oracle.grid4d_torch import RefGrid4DFeatures
    T = 2 ** 14
    cfg = de.DeepEarthConfig(hidden_dim=64, n_heads=4, n_layers=1, hash_table_size=T)

    class RefEncoder(torch.nn.Module):                       # models/encoders.py:19-85, same attribute names
        def __init__(self):
            super().__init__()
            f = RefGrid4DFeatures(hash_table_size=T)
            self.spatial_encoder, self.temporal_encoder = f.spatial_encoder, f.temporal_encoder
            self._features = [f]
            self.projection = torch.nn.Sequential(torch.nn.Linear(48, 64), torch.nn.LayerNorm(64), torch.nn.GELU(),
                                                  torch.nn.Linear(64, 64))

        def forward(self, xyzt, spatial_mask=None, temporal_mask=None):
            return self.projection(self._features[0](xyzt, spatial_mask, temporal_mask))

    torch.manual_seed(11)
    mine = de.Grid4DEncoder(cfg).to(dev)
    ref = RefEncoder().to(dev)
    with torch.no_grad():
        for p in mine.parameters():
            if p.dim() == 2 and p.shape[1] == 2:
                p.uniform_(-0.5, 0.5)                        # gradients that survive the clip visibly
    assert set(mine.state_dict().keys()) == set(ref.state_dict().keys())
    ref.load_state_dict(mine.state_dict())                   # ours -> reference layout
    mine.load_state_dict(ref.state_dict())                   # and back
    g = torch.Generator(device=dev).manual_seed(12)
    xyzt = torch.randn((20000, 4), device=dev, generator=g) * 0.6 + 0.3     # negative and > 1 coordinates included
    smask = torch.rand(20000, device=dev, generator=g) > 0.2
    target = torch.randn((20000, 64), device=dev, generator=g)

    def step(model):
        opt = torch.optim.AdamW(model.parameters(), lr=1e-2)
        scaler = torch.amp.GradScaler("cuda", init_scale=1024.0)
        with torch.autocast("cuda", dtype=torch.float16):
            loss = torch.nn.functional.mse_loss(model(xyzt, spatial_mask=smask).float(), target)
        scaler.scale(loss).backward()
        scaler.unscale_(opt)
        norm = torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        scaler.step(opt)
        scaler.update()
        return float(loss.detach()), float(norm)

    loss_m, norm_m = step(mine)
    loss_r, norm_r = step(ref)
    assert loss_m == pytest.approx(loss_r, rel=1e-5)
    assert norm_m == pytest.approx(norm_r, rel=1e-4)
    sd_m, sd_r = mine.state_dict(), ref.state_dict()
    for k in sd_r:
        assert float((sd_m[k] - sd_r[k]).abs().max()) <= 1e-4 * max(float(sd_r[k].abs().max()), 1e-6) + 1e-7, k
    mine.eval(), ref.eval()
    with torch.no_grad():
        assert float((mine(xyzt) - ref(xyzt)).abs().max()) <= 1e-4


def test_ddp_wraps_the_encoder(dev):
    """hpc/train_distrbuted.py:499-504 wraps the whole model in DistributedDataParallel: the encoder's table gradients
    (views of one flat buffer) must pass through DDP's hooks and buckets unchanged.  One process, NCCL, 127.0.0.1."""
    import socket
    import torch.distributed as dist
    from torch.nn.parallel import DistributedDataParallel as DDP
    import deepearth_b200 as de
    if dist.is_initialized():
        pytest.skip("a process group already exists in this process")
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    cfg = de.DeepEarthConfig(hidden_dim=64, n_heads=4, n_layers=1, hash_table_size=2 ** 14)
    torch.manual_seed(5)
    plain = de.Grid4DEncoder(cfg).to(dev)
    wrapped = de.Grid4DEncoder(cfg).to(dev)
    wrapped.load_state_dict(plain.state_dict())
    g = torch.Generator(device=dev).manual_seed(6)
    xyzt = torch.rand((5000, 4), device=dev, generator=g)
    target = torch.randn((5000, 64), device=dev, generator=g)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=0, world_size=1)
    try:
        ddp = DDP(wrapped, device_ids=[dev.index if dev.index is not None else 0])
        for _ in range(2):                                   # second pass: gradients accumulate into bucket views
            torch.nn.functional.mse_loss(ddp(xyzt), target).backward()
            torch.nn.functional.mse_loss(plain(xyzt), target).backward()
        torch.cuda.synchronize()
    finally:
        dist.destroy_process_group()
    for (name, a), (_, b) in zip(plain.named_parameters(), wrapped.named_parameters()):
        assert a.grad is not None and b.grad is not None, name
        denom = float(a.grad.abs().max()) or 1.0
        assert float((a.grad - b.grad).abs().max()) <= GRAD_TOL * denom, name


# ---------------------------------------------------------------------------------------------------
# BASELINE.json sizes: size-independent properties + the device-resident torch restatement
# ---------------------------------------------------------------------------------------------------
def test_full_size_properties(dev):
    """Config 2 (4 grids x 16 levels x 2 feat, T=2^19, 2^22 points, fp16 tables): constant tables
    interpolate to the constant, the encode is linear in the tables, and sum(grad_tables) equals
    sum(grad_out) per feature column group (corner weights sum to 1)."""
    import deepearth_b200 as de
    T, L, F, n = 2 ** 19, 16, 2, 2 ** 22
    res = [2 ** i for i in range(4, 20)]
    cfg = de.DeepEarthConfig(hidden_dim=16, n_heads=4, n_layers=1, hash_table_size=T, n_temporal_levels=16,
                             temporal_resolutions=res, grid4d_mode="decomposed", grid4d_table_dtype="float16",
                             grid4d_table_scale=1.0)
    enc = de.Grid4DEncoder(cfg).to(dev)
    g = torch.Generator(device=dev).manual_seed(1234)
    xyzt = torch.rand((n, 4), device=dev, generator=g)
    tabs = [e.weight for sub in enc._encoders for e in sub.hash_tables]
    assert len(tabs) == 64
    with torch.no_grad():
        for p in tabs:
            p.fill_(0.25)
    feats = enc.encode_features(xyzt)
    assert float((feats.detach() - 0.25).abs().max()) <= 1e-6
    del feats
    with torch.no_grad():
        for i, p in enumerate(tabs):
            # multiples of 2^-10 in [-1,1]: exact in fp16 and still exact (and normal) after halving
            p.copy_(torch.round(torch.rand(p.shape, device=dev, generator=g) * 2048 - 1024) / 1024)
    f1 = enc.encode_features(xyzt).detach()
    with torch.no_grad():
        for p in tabs:
            p.mul_(0.5)                                  # exact in fp16
    f2 = enc.encode_features(xyzt).detach()
    assert torch.equal(f2, f1 * 0.5)                     # linear in the tables, exactly for a power of two
    gout = torch.randn((n, 128), device=dev, generator=g)
    feats = enc.encode_features(xyzt)
    feats.backward(gout)
    col_sums = gout.double().sum(0).view(64, F)
    abs_sum = float(gout.abs().double().sum(0).max())
    for k, p in enumerate(tabs):
        got = p.grad.double().sum(0)
        assert float((got - col_sums[k]).abs().max()) <= 1e-5 * abs_sum, k


def test_large_vs_torch_restatement_on_device(dev):
    """2^20 points against oracle/grid4d_torch.py running through ATen on the same GPU."""
    import deepearth_b200 as de
    from oracle.grid4d_torch import grid4d_forward
    T, L, F, n = 2 ** 19, 16, 2, 2 ** 20
    res = [2 ** i for i in range(4, 20)]
    g = torch.Generator(device=dev).manual_seed(7)
    for linear in (False, True):
        enc = de.MultiResolutionHashEncoding(L, F, res, T, 3, interpolation="linear" if linear else "nearest",
                                             dense_fallback=linear).to(dev)
        with torch.no_grad():
            for e in enc.hash_tables:
                e.weight.copy_(torch.rand(e.weight.shape, device=dev, generator=g) * 2e-4 - 1e-4)
        coords = torch.randn((n, 3), device=dev, generator=g) * 0.7
        specs = [((0, 1, 2), s.resolution, T, s.dense) for s in enc.level_specs()]
        tabs = [e.weight for e in enc.hash_tables]
        out = enc(coords)
        want = grid4d_forward(coords, tabs, specs, linear)
        if linear:
            assert float(((out - want).abs().max() / want.abs().max()).detach()) <= FWD_TOL_F32
        else:
            assert torch.equal(out, want)
        gout = torch.randn(out.shape, device=dev, generator=g)
        out.backward(gout)
        mine = [e.weight.grad.clone() for e in enc.hash_tables]
        for e in enc.hash_tables:
            e.weight.grad = None
        want.backward(gout)
        for l, e in enumerate(enc.hash_tables):
            denom = float(e.weight.grad.abs().max())
            assert float((mine[l] - e.weight.grad).abs().max()) <= GRAD_TOL * denom, l


# ---------------------------------------------------------------------------------------------------
# geodetic prologue
# ---------------------------------------------------------------------------------------------------
def test_geo2xyz_reference_scenario_groups_on_device(dev, golden_dir):
    """Prologue kernel on the reference's own scenario groups (geospatial/tests/geospatial_tests.py:67-215): ECEF within
    1e-6 m, normalised coordinates within fp32 rounding (+ the ECEF noise over the box span for the cm-sized boxes), the
    bounding box the kernel folds itself equal to the reference's."""
    import deepearth_b200 as de
    from test_oracle_golden import scenario_tolerance
    g = load(golden_dir, "geo_scenarios")
    for i in range(len(g["names"])):
        geo = torch.from_numpy(g[f"s{i}_geo"]).to(dev)
        conv = de.GeospatialConverter(device="cuda:0")
        xyzt, ecef = conv.geodetic_to_xyzt_norm(geo, update_bbox=True, return_ecef=True)
        assert np.abs(ecef.cpu().numpy() - g[f"s{i}_ecef"]).max() <= 1e-6, g["names"][i]
        assert np.allclose(np.array(conv.bbox.min_point), g[f"s{i}_bbox_min"], rtol=0, atol=1e-8), g["names"][i]
        err = np.abs(xyzt[:, :3].cpu().numpy().astype(np.float64) - g[f"s{i}_norm"])
        assert (err <= scenario_tolerance(g[f"s{i}_bbox_span"])[None, :]).all(), (g["names"][i], err.max(0))
        # the same points as fp32 tensors (how the reference's test builds them) take the _as_fp64 path
        x32 = conv.geodetic_to_xyzt_norm(geo.float())
        if bool((geo.float().double() == geo).all()):
            assert torch.equal(x32, xyzt)


def test_geo2xyz_vs_reference_and_oracle(dev, oracle, golden_dir):
    import deepearth_b200 as de
    g = load(golden_dir, "geo2xyz")
    for name in ("landmarks", "florida", "global"):
        geo = torch.from_numpy(g[f"{name}_geo"]).to(dev)
        conv = de.GeospatialConverter(device="cuda:0")
        xyzt, ecef = conv.geodetic_to_xyzt_norm(geo, update_bbox=True, return_ecef=True)
        assert np.abs(ecef.cpu().numpy() - g[f"{name}_ecef"]).max() <= 1e-6          # metres
        assert np.allclose(np.array(conv.bbox.min_point), g[f"{name}_bbox_min"], rtol=0, atol=1e-6)
        assert np.allclose(np.array(conv.bbox.span), g[f"{name}_bbox_span"], rtol=0, atol=1e-6)
        want = g[f"{name}_norm"].astype(np.float32)
        assert np.abs(xyzt[:, :3].cpu().numpy() - want).max() <= 1.2e-7
        assert float(xyzt[:, 3].abs().max()) == 0.0
        # pre-fitted box + oracle: fp32 outputs identical except where fp64 libm vs CUDA rounding differs
        fixed = de.GeospatialConverter(device="cuda:0", bbox=de.BoundingBox(*g[f"{name}_bbox_min"],
                                       *(g[f"{name}_bbox_min"] + g[f"{name}_bbox_span"])), t_range=(0.0, 10.0))
        tt = torch.linspace(-1, 11, geo.shape[0], dtype=torch.float64, device=dev)
        x2 = fixed.geodetic_to_xyzt_norm(geo, tt).cpu().numpy()
        o2 = oracle.geo2xyzt_norm(g[f"{name}_geo"], tt.cpu().numpy(), g[f"{name}_bbox_min"],
                                  np.array(fixed.bbox.span), (0.0, 10.0))
        assert np.abs(x2 - o2).max() <= 1.2e-7
        assert (x2 != o2).mean() <= 1e-3
    same = torch.from_numpy(g["degenerate_geo"]).to(dev)
    conv = de.GeospatialConverter(device="cuda:0")
    assert np.array_equal(conv.geodetic_to_xyzt_norm(same, update_bbox=True)[:, :3].cpu().numpy(),
                          g["degenerate_norm"].astype(np.float32))                    # all 0.5
    # fp32 geodetic input is up-cast like _as_fp64 (geo2xyz.py:147)
    geo32 = torch.from_numpy(g["florida_geo"]).float().to(dev)
    a = de.GeospatialConverter(device="cuda:0").geodetic_to_xyz(geo32)
    b = de.GeospatialConverter(device="cuda:0").geodetic_to_xyz(geo32.double())
    assert torch.equal(a, b)
    x, y, z = de.geo2xyz(28.5, -81.4, 25.0)
    assert abs(float(x) - g["landmarks_ecef"][-1, 0]) < 1e-6


def test_geo2xyz_statistics_normalisation_and_bound(dev, oracle, golden_dir):
    """Spec pipeline (SPECIFICATIONS.MD:729, :853-854, :294): (ecef - mean) / (2 max std), optionally remapped from
    [-bound, bound] to [0, 1]; ECEF taken from the reference's own converter (golden), statistics as the spec computes
    them (torch .std = unbiased).  One affine map inside the prologue kernel."""
    import deepearth_b200 as de
    g = load(golden_dir, "geo2xyz")
    geo = torch.from_numpy(g["florida_geo"]).to(dev)
    ecef = torch.from_numpy(g["florida_ecef"])
    center = ecef.mean(dim=0)
    scale = float(ecef.std(dim=0).max() * 2.0)
    fit = de.GeospatialConverter(device="cuda:0").fit_statistics(geo)
    assert np.allclose(fit["center"], center.numpy(), rtol=0, atol=1e-5) and abs(fit["scale"] - scale) <= 1e-6 * scale
    want = ((ecef - center) / scale).numpy()
    conv = de.GeospatialConverter.from_statistics(center.tolist(), scale, t_range=(0.0, 4.0))
    tt = torch.linspace(0, 4, geo.shape[0], dtype=torch.float64, device=dev)
    got = conv.geodetic_to_xyzt_norm(geo, tt).cpu().numpy()
    assert np.abs(got[:, :3] - want.astype(np.float32)).max() <= 2.4e-7 * max(1.0, np.abs(want).max())
    assert np.abs(got[:, 3] - (tt.cpu().numpy() / 4.0).astype(np.float32)).max() <= 1.2e-7
    o = oracle.geo2xyzt_norm(g["florida_geo"], tt.cpu().numpy(), center.numpy(), np.full(3, scale), (0.0, 4.0))
    assert np.abs(got - o).max() <= 2.4e-7 and (got != o).mean() <= 1e-3
    bound = 1.6
    conv_b = de.GeospatialConverter.from_statistics(center.tolist(), scale, bound=bound)
    got_b = conv_b.geodetic_to_xyzt_norm(geo).cpu().numpy()[:, :3]
    want_b = (want + bound) / (2 * bound)
    assert np.abs(got_b - want_b.astype(np.float32)).max() <= 1.2e-7
    assert got_b.min() >= 0.0 and got_b.max() <= 1.0
    with pytest.raises(ValueError):
        de.GeospatialConverter.from_statistics(center.tolist(), scale, bound=0.0)
    with pytest.raises(ValueError):
        conv.geodetic_to_xyzt_norm(geo, tt, update_bbox=True)
    # ECEF itself is unaffected by the normalisation the converter carries
    assert np.abs(conv.geodetic_to_xyz(geo).cpu().numpy() - g["florida_ecef"]).max() <= 1e-6


def test_sum_and_probe_helpers(dev):
    from deepearth_b200 import functional as Fn
    x = torch.randn(1_000_003, device=dev)
    assert float(Fn.sum_f32(x)) == pytest.approx(float(x.double().sum()), rel=1e-9, abs=1e-6)


# ---------------------------------------------------------------------------------------------------
# A/B switches: the opt-in paths must reproduce the default lane-pair path
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("table_dtype", ["float32", "float16"])
def test_smem_coarse_and_staged_paths_match_plain_path(dev, oracle, table_dtype):
    """Dense coarse levels served from shared memory (forward) and the TMA-staged backward against the
    plain L2-path kernels (bit-identical forward: same corner order) and against the C oracle."""
    import deepearth_b200 as de
    from deepearth_b200 import _lib
    from oracle.grid4d_oracle import Level
    T, L, F, n = 2 ** 16, 6, 2, 2 ** 16 + 777
    res = [4, 8, 16, 32, 64, 128]                      # (R+1)^3 <= T up to R=32: four dense levels, two hashed
    g = torch.Generator(device=dev).manual_seed(11)
    enc = de.MultiResolutionHashEncoding(L, F, res, T, 3, interpolation="linear", dense_fallback=True,
                                         table_dtype=table_dtype).to(dev)
    with torch.no_grad():
        for e in enc.hash_tables:
            e.weight.copy_(torch.round(torch.rand(e.weight.shape, device=dev, generator=g) * 2048 - 1024) / 2 ** 22)
    coords = torch.rand((n, 3), device=dev, generator=g)
    coords[:4096] = torch.randn((4096, 3), device=dev, generator=g) * 1.5     # out-of-range: index wraps past the live cells
    gout = torch.randn((n, L * F), device=dev, generator=g)

    def run(flags):
        enc.extra_flags = flags
        for e in enc.hash_tables:
            e.weight.grad = None
        out = enc(coords)
        out.backward(gout)
        return out.detach().clone(), torch.cat([e.weight.grad.reshape(-1) for e in enc.hash_tables])

    out_opt, grad_opt = run(_lib.DE_FLAG_STAGED_BWD | _lib.DE_FLAG_SMEM_COARSE)
    out_ref, grad_ref = run(0)
    # the default backward spreads the two coarsest levels over gradient replicas; without a workspace it must agree
    from deepearth_b200 import functional as Fn
    specs = enc.level_specs()
    flat = torch.zeros(L * T * F, device=dev)
    ptrs = [flat.data_ptr() + 4 * l * T * F for l in range(L)]
    assert _lib.load().de_grid4d_backward_workspace_bytes(Fn.pack_levels(specs, None, ptrs), L, F, n, 1) > 0
    Fn.encode_backward(specs, coords, gout, ptrs, F, True, use_replicas=False)
    assert float((flat - grad_ref).abs().max()) <= GRAD_TOL * float(grad_ref.abs().max())
    assert torch.equal(out_opt, out_ref)
    denom = float(grad_ref.abs().max())
    assert float((grad_opt - grad_ref).abs().max()) <= GRAD_TOL * denom

    levels = [Level(T=T, R=R, dims=(0, 1, 2), out_col=l * F, table_off=l * T, dense=(R + 1) ** 3 <= T)
              for l, R in enumerate(res)]
    tables = np.concatenate([e.weight.detach().cpu().numpy() for e in enc.hash_tables], 0)
    if table_dtype == "float16":
        tables = (tables * np.float32(enc.table_scale)).astype(np.float16)
        want = oracle.forward(levels, coords.cpu().numpy(), tables, F, True, out_scale=1.0 / enc.table_scale)
    else:
        want = oracle.forward(levels, coords.cpu().numpy(), tables, F, True)
    err = np.abs(out_opt.cpu().numpy() - want).max() / np.abs(want).max()
    assert err <= FWD_TOL_F32, err
    want_g = oracle.backward(levels, coords.cpu().numpy(), gout.cpu().numpy(), L * T, F, True)
    gerr = np.abs(grad_opt.cpu().numpy().reshape(-1, F) - want_g).max() / np.abs(want_g).max()
    assert gerr <= GRAD_TOL, gerr


@pytest.mark.parametrize("levels_per_group,n_extra", [(2, 0), (4, 1234), (8, 7)])
def test_regrouped_gradient_rows_match_in_place_rows(dev, levels_per_group, n_extra):
    """Backward with the gradient rows re-laid group-major first (regroup_kernel; taken when a level group's slice of
    a row is <= 64 B, i.e. large tables) against the same backward reading the slices in place (itself checked
    against the C oracle by the tests above)."""
    from deepearth_b200 import _lib
    from deepearth_b200 import functional as Fn
    from deepearth_b200.functional import LevelSpec
    T, L, F = 2 ** 18, 32, 2
    n = 2 ** 20 + n_extra
    g = torch.Generator(device=dev).manual_seed(21)
    res = [16 * 2 ** (l % 12) + l for l in range(L)]
    specs = [LevelSpec(table_size=T, resolution=R, dims=(0, 1, 3) if l % 2 else (1, 2, 3), out_col=4 + l * F,
                       dense=(R + 1) ** 3 <= T) for l, R in enumerate(res)]
    coords = torch.rand((n, 4), device=dev, generator=g)
    gout = torch.randn((n, 4 + L * F + 4), device=dev, generator=g)         # rows wider than the level columns, c0 = 4

    def run(flags):
        flat = torch.zeros(L * T * F, device=dev)
        ptrs = [flat.data_ptr() + 4 * l * T * F for l in range(L)]
        Fn.encode_backward(specs, coords, gout, ptrs, F, True, levels_per_group=levels_per_group, extra_flags=flags)
        return flat

    lib = _lib.load()
    flat0 = torch.zeros(1, device=dev)
    lv = Fn.pack_levels(specs, None, [flat0.data_ptr()] * L)
    lin = _lib.DE_INTERP_LINEAR
    with_rg = lib.de_grid4d_backward_workspace_bytes_ex(lv, L, F, n, lin, levels_per_group)
    without = lib.de_grid4d_backward_workspace_bytes_ex(lv, L, F, n, lin | _lib.DE_FLAG_NO_REGROUP, levels_per_group)
    assert with_rg >= without + n * L * F * 4                                 # the regrouped rows are part of the scratch
    got = run(0)
    ref = run(_lib.DE_FLAG_NO_REGROUP)
    denom = float(ref.abs().max())
    assert float((got - ref).abs().max()) <= GRAD_TOL * denom


def test_replica_budget_mixes_replicated_and_plain_levels(dev):
    """Many points per table entry make every level of a group eligible for gradient replicas; only the coarsest ones
    get them (16 MiB of copies per level group, c_abi.cu kRepGroupBytes).  The mixed state -- some levels reduced
    into copies and folded, the rest reduced in place -- must give the same gradients as no replicas at all."""
    from deepearth_b200 import functional as Fn
    from deepearth_b200.functional import LevelSpec
    T, L, F = 2 ** 16, 8, 2                  # 512 KiB of gradients per level -> 16 copies = 7.5 MiB extra each
    n = 2 ** 21 + 5                          # 8 corners * n / T = 256 updates per entry: every level qualifies
    g = torch.Generator(device=dev).manual_seed(33)
    res = [48 * 2 ** l + 1 for l in range(L)]
    specs = [LevelSpec(table_size=T, resolution=R, dims=(0, 1, 2), out_col=l * F, dense=False) for l, R in enumerate(res)]
    coords = torch.rand((n, 4), device=dev, generator=g)
    gout = torch.randn((n, L * F), device=dev, generator=g)

    def run(use_replicas):
        flat = torch.zeros(L * T * F, device=dev)
        ptrs = [flat.data_ptr() + 4 * l * T * F for l in range(L)]
        Fn.encode_backward(specs, coords, gout, ptrs, F, True, use_replicas=use_replicas)
        return flat

    got, ref = run(True), run(False)
    denom = float(ref.abs().max())
    assert denom > 0
    assert float((got - ref).abs().max()) <= GRAD_TOL * denom


def test_fused_head_hands_group_major_gradients_to_the_encoder(dev, monkeypatch):
    """Large tables + fused head: the head's backward kernel writes d(loss)/d(features) group-major and the hash-grid
    backward walks it directly (no regroup pass).  Same gradients as the row-major hand-off, and the path is taken."""
    import deepearth_b200 as de
    from deepearth_b200 import functional as Fn
    res = [2 ** i for i in range(4, 12)]
    cfg = de.DeepEarthConfig(hidden_dim=64, n_heads=4, n_layers=1, hash_table_size=2 ** 21, n_spatial_levels=8,
                             n_temporal_levels=8, spatial_resolutions=res, temporal_resolutions=res,
                             grid4d_mode="decomposed", grid4d_table_dtype="float16", grid4d_fused_head=True)
    torch.manual_seed(7)
    enc = de.Grid4DEncoder(cfg).to(dev)
    n = 2 ** 20 + 3
    g = torch.Generator(device=dev).manual_seed(9)
    xyzt = torch.rand((n, 4), device=dev, generator=g)
    gy = torch.randn((n, 64), device=dev, generator=g)
    widths = []
    real_backward = Fn.encode_backward

    def spy(*a, **k):
        widths.append(k.get("group_width", 0))
        return real_backward(*a, **k)

    monkeypatch.setattr(Fn, "encode_backward", spy)

    def run():
        for p in enc.parameters():
            p.grad = None
        enc(xyzt).backward(gy)
        tables = torch.cat([p.grad.reshape(-1) for k, p in enc.named_parameters() if "hash_tables" in k])
        head = torch.cat([p.grad.reshape(-1) for k, p in enc.named_parameters() if "projection" in k])
        return tables, head

    t_grouped, h_grouped = run()
    assert widths and widths[-1] == 8, widths                     # 16 MiB gradient tables -> 4 levels per group x F = 2
    monkeypatch.setattr(Fn, "backward_group_width", lambda *a, **k: 0)
    t_rows, h_rows = run()
    assert widths[-1] == 0
    assert float((t_grouped - t_rows).abs().max()) <= GRAD_TOL * float(t_rows.abs().max())
    assert float((h_grouped - h_rows).abs().max()) <= 1e-4 * float(h_rows.abs().max())


# ---------------------------------------------------------------------------------------------------
# tensor-core MLP head (tcgen05): Linear -> LayerNorm -> GELU -> Linear, hidden = out = 64
# ---------------------------------------------------------------------------------------------------
def _head_reference(x, seq):
    """fp32 PyTorch evaluation with the kernel's roundings: bf16 inputs / weights / hidden row, fp32 accumulate."""
    mods = list(seq)
    bf = lambda t: t.to(torch.bfloat16).to(torch.float32)
    h = bf(x) @ bf(mods[0].weight).t() + mods[0].bias
    for m in mods[1:-1]:
        h = m(h)
    return bf(h) @ bf(mods[-1].weight).t() + mods[-1].bias


HEAD_BWD_TOL = 8e-3          # B200 (gpurun r02p): worst 3.7e-3 over every shape and gradient; 2x that


@pytest.mark.parametrize("in_dim,layernorm,act", [(128, True, "gelu"), (48, True, "gelu"), (64, False, "relu"),
                                                  (32, False, "none"), (96, True, "relu")])
def test_mlp_head_vs_torch(dev, in_dim, layernorm, act):
    from deepearth_b200.mlp_head import FusedProjection
    torch.manual_seed(3)
    mods = [torch.nn.Linear(in_dim, 64)]
    if layernorm:
        mods.append(torch.nn.LayerNorm(64))
    if act != "none":
        mods.append(torch.nn.GELU() if act == "gelu" else torch.nn.ReLU())
    mods.append(torch.nn.Linear(64, 64))
    seq = torch.nn.Sequential(*mods).to(dev)
    if layernorm:
        with torch.no_grad():
            seq[1].weight.uniform_(0.5, 1.5)
            seq[1].bias.uniform_(-0.2, 0.2)
    head = FusedProjection(seq)
    assert head.supported()
    g = torch.Generator(device=dev).manual_seed(5)
    for n in (1, 127, 128, 129, 5000, 148 * 4 * 128 + 17):      # ragged tails and more tiles than resident CTAs
        x = torch.randn((n, in_dim), device=dev, generator=g)
        with torch.no_grad():
            got = head(x)
            want = _head_reference(x, seq)
            full = seq(x)
        scale = float(want.abs().max())
        assert float((got - want).abs().max()) <= 2e-3 * scale, (n, float((got - want).abs().max()), scale)
        assert float((got - full).abs().max()) <= 3e-2 * scale      # bf16 operands vs the fp32 module
    # training, fp32 recompute backward: forward through the kernel, exact module gradients
    head_r = FusedProjection(seq, fused_backward=False)
    x = torch.randn((777, in_dim), device=dev, generator=g, requires_grad=True)
    y = head_r(x)
    gy = torch.randn(y.shape, device=dev, generator=g)
    y.backward(gy)
    gx, gw = x.grad.clone(), seq[0].weight.grad.clone()
    x.grad = None
    seq.zero_grad()
    seq(x).backward(gy)
    assert torch.allclose(gx, x.grad, rtol=1e-5, atol=1e-6) and torch.allclose(gw, seq[0].weight.grad, rtol=1e-5, atol=1e-6)
    # training, tcgen05 backward kernel (bf16 operands, fp32 accumulation) against autograd through an fp32 model
    # that sees the same bf16-rounded operands in its forward (straight-through rounding): the recomputed hidden
    # row -- and with it every ReLU mask / GELU slope -- is then the kernel's own, up to accumulation order
    ste = lambda t: t + (t.to(torch.bfloat16).to(torch.float32) - t).detach()

    def ref_model(x):
        h = torch.nn.functional.linear(ste(x), ste(seq[0].weight), seq[0].bias)
        for m in list(seq)[1:-1]:
            h = m(h)
        return torch.nn.functional.linear(ste(h), ste(seq[-1].weight), seq[-1].bias)

    worst = {}
    for n in (777, 128 * 148 * 3 + 5):
        x = torch.randn((n, in_dim), device=dev, generator=g, requires_grad=True)
        gy = torch.randn((n, 64), device=dev, generator=g)
        seq.zero_grad()
        ref_model(x).backward(gy)
        want = {"x": x.grad.clone(), **{k: p.grad.clone() for k, p in seq.named_parameters()}}
        x.grad = None
        seq.zero_grad()
        head(x).backward(gy)
        got = {"x": x.grad, **{k: p.grad for k, p in seq.named_parameters()}}
        for k, w in want.items():
            assert got[k] is not None, k
            err = float((got[k] - w).abs().max()) / max(float(w.abs().max()), 1e-20)
            worst[k] = max(worst.get(k, 0.0), err)
            assert err <= HEAD_BWD_TOL, (n, k, err)
    print(f"head backward in_dim={in_dim} ln={layernorm} act={act}:", {k: f"{v:.1e}" for k, v in worst.items()})


def test_grid4d_encoder_fused_head(dev):
    import deepearth_b200 as de
    cfg = de.DeepEarthConfig(hidden_dim=64, n_heads=4, n_layers=1, hash_table_size=2 ** 14, grid4d_fused_head=True)
    enc = de.Grid4DEncoder(cfg).to(dev)
    with torch.no_grad():
        for p in enc.parameters():
            if p.dim() == 2 and p.shape[1] == 2:
                p.uniform_(-1.0, 1.0)
    xyzt = torch.rand((4096 + 5, 4), device=dev)
    with torch.no_grad():
        got = enc(xyzt)
        want = enc.projection(enc.encode_features(xyzt))
    assert enc._fused_head not in (None, False)
    assert float((got - want).abs().max()) <= 3e-2 * float(want.abs().max())
    enc(xyzt).sum().backward()
    assert all(p.grad is not None for p in enc.parameters())


def test_small1d_shared_memory_path_equals_general_kernels(dev):
    """Nearest-cell 1-D levels (the reference's temporal encoder) served from shared memory: features identical to the
    general kernels' (DE_FLAG_NO_SMALL1D), gradients equal up to summation order -- coordinates partly outside [0, 1]
    (those rows take the global path), a temporal row mask, fp32 and fp16 tables, ragged tails."""
    import deepearth_b200 as de
    from deepearth_b200 import _lib
    for n, table_dtype in ((70001, "float32"), (333, "float16"), (1 << 20, "float32")):
        torch.manual_seed(n)
        cfg = de.DeepEarthConfig(hidden_dim=32, n_heads=4, n_layers=1, hash_table_size=2 ** 17, grid4d_table_dtype=table_dtype)
        enc = de.Grid4DEncoder(cfg).to(dev)
        g = torch.Generator(device=dev).manual_seed(7)
        xyzt = torch.rand((n, 4), device=dev, generator=g) * 1.4 - 0.2
        tmask = torch.rand(n, device=dev, generator=g) > 0.25
        gout = torch.randn((n, 48), device=dev, generator=g)
        res = []
        for flags in (0, _lib.DE_FLAG_NO_SMALL1D):
            for e in enc._encoders:
                e.extra_flags = flags
            enc.zero_grad(set_to_none=True)
            f = enc.encode_features(xyzt, temporal_mask=tmask)
            f.backward(gout)
            res.append((f.detach().clone(), [w.grad.clone() for w in enc.temporal_encoder.weights()]))
        assert torch.equal(res[0][0], res[1][0])
        for a, b in zip(res[0][1], res[1][1]):
            assert float((a - b).abs().max()) <= 1e-4 * max(float(b.abs().max()), 1e-30)   # summation order only (B200: 1.2e-5 at 2^20 points)


def test_sort_by_cell_and_warp_aggregated_backward(dev):
    """The north_star's alternative for densely colliding levels: counting sort of the points by coarse cell, then a
    backward whose warps aggregate runs of lanes that share a cell.  perm must be a cell-ordered permutation; the gradients
    must equal the default (replica) path's on Florida-clustered points -- coarse dense and fine hashed levels, ragged tail."""
    from bench import make_geo
    from deepearth_b200 import functional as Fn
    from deepearth_b200.functional import LevelSpec
    n, T, F = (1 << 18) + 37, 2 ** 16, 2
    geo, t, conv = make_geo(n, dev, 5)
    xyzt = conv.geodetic_to_xyzt_norm(geo, t)
    dims = (0, 1, 3)
    perm = Fn.sort_points(xyzt, dims, bits=5)
    assert torch.equal(torch.sort(perm.long()).values, torch.arange(n, device=dev))
    cells = (xyzt[:, list(dims)].clamp(0, 1) * 32).clamp(max=31).floor().long()
    key = (cells[:, 0] * 32 + cells[:, 1]) * 32 + cells[:, 2]
    sk = key[perm.long()]
    assert bool((sk[1:] >= sk[:-1]).all()), "perm is not in cell order"
    res = [16, 32, 64, 1024]
    specs = [LevelSpec(T, R, dims, l * F, dense=Fn.dense_fits(R, 3, T)) for l, R in enumerate(res)]
    g = torch.Generator(device=dev).manual_seed(3)
    gout = torch.randn((n, len(res) * F), device=dev, generator=g)
    a = torch.zeros((len(res), T, F), device=dev)
    b = torch.zeros_like(a)
    Fn.encode_backward(specs, xyzt, gout, [a[l].data_ptr() for l in range(len(res))], F, True)
    Fn.encode_backward_sorted(specs, xyzt, gout, [b[l].data_ptr() for l in range(len(res))], F, perm)
    for l in range(len(res)):
        err = float((a[l] - b[l]).abs().max()) / float(a[l].abs().max())
        assert err <= 1e-4, (res[l], err)
    # any permutation gives the same gradients
    c = torch.zeros_like(a)
    Fn.encode_backward_sorted(specs, xyzt, gout, [c[l].data_ptr() for l in range(len(res))], F,
                              torch.randperm(n, device=dev, generator=g).int())
    assert float((a - c).abs().max()) / float(a.abs().max()) <= 1e-4


def test_graphed_features_match_eager_and_follow_optimiser_steps(dev):
    """Grid4DEncoder.graphed_features: forward + backward as CUDA graphs for a fixed small batch -- same features and table
    gradients as the eager path, new inputs and updated tables are seen on replay."""
    import deepearth_b200 as de
    torch.manual_seed(3)
    res = [2 ** i for i in range(4, 12)]
    cfg = de.DeepEarthConfig(hidden_dim=32, n_heads=4, n_layers=1, hash_table_size=2 ** 15, n_spatial_levels=8,
                             n_temporal_levels=8, spatial_resolutions=res, temporal_resolutions=res,
                             grid4d_mode="decomposed", grid4d_table_dtype="float16")
    enc = de.Grid4DEncoder(cfg).to(dev).train()
    n = 4096
    g = torch.Generator(device=dev).manual_seed(1)
    graphed = enc.graphed_features(torch.rand((n, 4), device=dev, generator=g))
    tables = [w for e in enc._encoders for w in e.weights()]
    for it in range(3):
        xyzt = torch.rand((n, 4), device=dev, generator=g)
        gout = torch.randn((n, enc.feature_dim), device=dev, generator=g)
        enc.zero_grad(set_to_none=True)
        f_eager = enc.encode_features(xyzt)
        f_eager.backward(gout)
        want = [w.grad.clone() for w in tables]
        enc.zero_grad(set_to_none=True)
        f_graph = graphed(xyzt)
        f_graph.backward(gout)
        assert torch.equal(f_graph, f_eager)
        for w, ref in zip(tables, want):
            assert float((w.grad - ref).abs().max()) <= 1e-5 * max(float(ref.abs().max()), 1e-30)
        with torch.no_grad():                                  # an "optimiser step": the next replay must see the new tables
            for w in tables:
                w.add_(torch.randn_like(w) * 1e-4)


@pytest.mark.parametrize("mode", ["reference", "decomposed"])
def test_flat_table_parameter_matches_per_level_layout(dev, mode):
    """config.grid4d_flat_tables: identical features, the flat gradient equals the stacked per-level gradients, optimiser
    step + state-dict round trip through the reference's keys, and the CUDA-graph path on top of it."""
    import deepearth_b200 as de
    kw = dict(hidden_dim=32, n_heads=4, n_layers=1, hash_table_size=2 ** 14)
    if mode == "decomposed":
        res = [2 ** i for i in range(4, 12)]
        kw.update(n_spatial_levels=8, n_temporal_levels=8, spatial_resolutions=res, temporal_resolutions=res,
                  grid4d_mode="decomposed", grid4d_table_dtype="float16")
    torch.manual_seed(11)
    a = de.Grid4DEncoder(de.DeepEarthConfig(**kw)).to(dev).train()
    torch.manual_seed(11)
    b = de.Grid4DEncoder(de.DeepEarthConfig(grid4d_flat_tables=True, **kw)).to(dev).train()
    assert len(list(b.parameters())) == len(a._encoders) + 6
    g = torch.Generator(device=dev).manual_seed(2)
    n = 5000
    xyzt = torch.rand((n, 4), device=dev, generator=g) * 1.2 - 0.1
    gout = torch.randn((n, a.feature_dim), device=dev, generator=g)
    fa = a.encode_features(xyzt)
    fb = b.encode_features(xyzt)
    assert torch.equal(fa, fb)
    fa.backward(gout)
    fb.backward(gout)
    for ea, eb in zip(a._encoders, b._encoders):
        want = torch.stack([w.grad for w in ea.weights()])
        assert eb.tables.grad.shape == want.shape
        assert float((eb.tables.grad - want).abs().max()) <= 1e-5 * max(float(want.abs().max()), 1e-30)
    opt = torch.optim.AdamW(b.parameters(), lr=1e-3)
    opt.step()
    a.load_state_dict(b.state_dict())                          # reference keys in, per-level layout
    with torch.no_grad():
        assert torch.equal(a.encode_features(xyzt), b.encode_features(xyzt))
    # CUDA graphs on top of the flat layout.  Captured on a FRESH module: autograd ties a parameter's AccumulateGrad node to
    # the stream of its first backward, so graphed_features must be called before the module's first eager backward
    torch.manual_seed(11)
    c = de.Grid4DEncoder(de.DeepEarthConfig(grid4d_flat_tables=True, **kw)).to(dev).train()
    c.load_state_dict(b.state_dict())
    graphed = c.graphed_features(xyzt)
    c.zero_grad(set_to_none=True)
    fg = graphed(xyzt)
    fg.backward(gout)
    a.zero_grad(set_to_none=True)
    fe = a.encode_features(xyzt)
    fe.backward(gout)
    assert torch.equal(fg, fe)
    for ea, ec in zip(a._encoders, c._encoders):
        want = torch.stack([w.grad for w in ea.weights()])
        assert float((ec.tables.grad - want).abs().max()) <= 1e-5 * max(float(want.abs().max()), 1e-30)
