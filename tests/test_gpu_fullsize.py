"""Parity of the benchmarked mode AT the benchmarked geometries (BASELINE.json configs[1], [2], [4]).

The decomposed / interpolated encoder ("Mode G") is compared with the device-resident torch restatement
(oracle/grid4d_torch.py, itself pinned on the CPU by tests/test_torch_port_cpu.py and by the committed known-answer
vectors tests/golden/modeg_kat.npz) on the exact level lists bench.py times: 4 grids x 16 levels x F = 2, fp16
working tables, fp32 gradients, T = 2^19 with 2^22 uniform points (c2) and T = 2^22 with 2^22 Florida-clustered
points produced by the geo2xyz prologue kernel (c3) -- replicas, the regroup pass and, for c5, the grouped-head
hand-off all live.  Tolerances are BASELINE.json's: per-corner indices bit-exact, features <= 1e-5 of max|ref|
against the same fp16 tables and <= 1e-2 against the fp32 tables, table gradients <= 1e-4 of max|ref|.
"""
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

FWD_TOL_SAME_TABLES = 1e-5
FWD_TOL_F16 = 1e-2
GRAD_TOL = 1e-4
RES = [2 ** i for i in range(4, 20)]


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")


def decomposed_config(de, T, **kw):
    return de.DeepEarthConfig(hidden_dim=kw.pop("hidden_dim", 64), n_heads=4, n_layers=1, hash_table_size=T,
                              n_spatial_levels=16, n_temporal_levels=16, spatial_resolutions=RES,
                              temporal_resolutions=RES, grid4d_mode="decomposed", grid4d_table_dtype="float16", **kw)


def florida_xyzt(n, dev, seed):
    """Florida-clustered points pushed through the prologue kernel (bench.py's c3 input)."""
    from bench import make_geo
    geo, t, conv = make_geo(n, dev, seed)
    return conv.geodetic_to_xyzt_norm(geo, t)


def level_tables(enc):
    """[(fp32 parameter, fp16 working table view (T, F))] in plan order."""
    out = []
    for sub in enc._encoders:
        for l, emb in enumerate(sub.hash_tables):
            out.append((emb.weight, sub._work[l]))
    return out


def reference_level(xyzt, spec, table_f32):
    """Stated math through ATen on the device: -> (idx (B, 8), w (B, 8), features (B, F) fp32)."""
    from oracle.grid4d_torch import level_indices_weights
    idx, w = level_indices_weights(xyzt, spec.dims, spec.resolution, spec.table_size, spec.dense)
    v = torch.nn.functional.embedding(idx, table_f32)
    return idx, w, (v * w.unsqueeze(-1)).sum(1)


def reference_grad(idx, w, gout_l, T):
    """fp64 scatter-add of w_k * grad_out (what embedding_dense_backward computes, without fp32 ordering noise)."""
    F = gout_l.shape[1]
    g = torch.zeros((T, F), dtype=torch.float64, device=idx.device)
    vals = (w.unsqueeze(-1).double() * gout_l.unsqueeze(1).double()).reshape(-1, F)
    g.index_add_(0, idx.reshape(-1), vals)
    return g


def check_against_restatement(enc, xyzt, feats, gout, table_grads, n_index_points=1 << 18):
    """Per level: indices bit-exact (all 8 corners, first n_index_points points), features and gradients in tolerance."""
    from deepearth_b200 import functional as Fn
    plan = enc._plan()
    F = plan.n_features
    scale = enc._encoders[0].table_scale
    tabs = level_tables(enc)
    sub_pts = xyzt[:n_index_points].contiguous()
    worst = {"fwd_same": 0.0, "fwd_f32": 0.0, "grad": 0.0}
    for g0 in range(0, len(plan.specs), 16):                      # one grid at a time: (16, 8, n) int64 indices
        got_idx = Fn.hash_indices(plan.specs[g0:g0 + 16], sub_pts, linear=True)
        for i in range(g0, g0 + 16):
            spec = plan.specs[i]
            w32, w16 = tabs[i]
            cols = slice(spec.out_col, spec.out_col + F)
            idx, w, f_same = reference_level(xyzt, spec, w16.float())
            assert torch.equal(got_idx[i - g0], idx[:n_index_points].t()), f"level {i}: corner indices differ"
            f_same = f_same * (1.0 / scale)
            denom = float(f_same.abs().max())
            e_same = float((feats[:, cols] - f_same).abs().max()) / denom
            f_f32 = (torch.nn.functional.embedding(idx, w32.detach()) * w.unsqueeze(-1)).sum(1)
            e_f32 = float((feats[:, cols] - f_f32).abs().max()) / float(f_f32.abs().max())
            want_g = reference_grad(idx, w, gout[:, cols], spec.table_size)
            e_g = float((table_grads[i].double() - want_g).abs().max()) / float(want_g.abs().max())
            assert e_same <= FWD_TOL_SAME_TABLES, (i, e_same)
            assert e_f32 <= FWD_TOL_F16, (i, e_f32)
            assert e_g <= GRAD_TOL, (i, e_g)
            worst = {"fwd_same": max(worst["fwd_same"], e_same), "fwd_f32": max(worst["fwd_f32"], e_f32),
                     "grad": max(worst["grad"], e_g)}
            del idx, w, f_same, f_f32, want_g
    return worst


def randomise_tables(enc, dev, seed, lo=-1e-4, hi=1e-4):
    g = torch.Generator(device=dev).manual_seed(seed)
    with torch.no_grad():
        for sub in enc._encoders:
            for emb in sub.hash_tables:
                emb.weight.copy_(torch.rand(emb.weight.shape, device=dev, generator=g) * (hi - lo) + lo)


# ---------------------------------------------------------------------------------------------------
# known answers (frozen stated math) through the C ABI
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("table_dtype", ["float32", "float16"])
def test_modeg_known_answers_on_device(dev, golden_dir, table_dtype):
    from deepearth_b200 import functional as Fn
    from deepearth_b200.functional import LevelSpec
    from test_oracle_golden import kat_levels
    g = np.load(os.path.join(golden_dir, "modeg_kat.npz"))
    levels, tabs, F = kat_levels(g)
    specs = [LevelSpec(table_size=lv.T, resolution=lv.R, dims=tuple(lv.dims), out_col=lv.out_col, dense=lv.dense)
             for lv in levels]
    pts = torch.from_numpy(g["points"]).to(dev)
    for l, sp in enumerate(specs):
        idx = Fn.hash_indices([sp], pts, linear=True)[0].cpu().numpy()
        assert np.array_equal(idx, g[f"idx_{l}"]), l
    dt = torch.float16 if table_dtype == "float16" else torch.float32
    scale = 4096.0 if table_dtype == "float16" else 1.0
    dtabs = [(torch.from_numpy(t).to(dev) * scale).to(dt).contiguous() for t in tabs]
    out = Fn.encode_forward(specs, pts, [t.data_ptr() for t in dtabs], dt, F, len(specs) * F, True, 1.0 / scale)
    assert np.abs(out.cpu().numpy() - g["features"]).max() <= 1e-7          # values are multiples of 2^-6 times fp32 weights
    assert float(out[0, 0]) == -0.0625                                       # the hand-derived value
    flat = torch.zeros(sum(lv.T for lv in levels) * F, device=dev)
    ptrs = [flat.data_ptr() + 4 * lv.table_off * F for lv in levels]
    Fn.encode_backward(specs, pts, torch.from_numpy(g["grad_out"]).to(dev), ptrs, F, True)
    got = flat.view(-1, F).cpu().numpy()
    for l, lv in enumerate(levels):
        want = np.zeros((lv.T, F))
        want[g[f"grad_rows_{l}"]] = g[f"grad_vals_{l}"]
        assert np.abs(got[lv.table_off:lv.table_off + lv.T] - want).max() <= 1e-6, l


# ---------------------------------------------------------------------------------------------------
# configs[1] and configs[2]: the exact level lists at (near) full size
# ---------------------------------------------------------------------------------------------------
def test_c2_geometry_uniform_vs_device_restatement(dev):
    """configs[1]: T = 2^19, 2^22 uniform points (the bench's default line, same seed and generator)."""
    import deepearth_b200 as de
    T, n = 2 ** 19, 2 ** 22
    torch.manual_seed(1234)
    enc = de.Grid4DEncoder(decomposed_config(de, T)).to(dev)
    g = torch.Generator(device=dev).manual_seed(1234)
    xyzt = torch.rand((n, 4), device=dev, generator=g)
    gout = torch.randn((n, 128), device=dev, generator=g)
    feats = enc.encode_features(xyzt)
    assert feats.shape == (n, 128)
    feats.backward(gout)
    grads = [w.grad for w, _ in level_tables(enc)]
    worst = check_against_restatement(enc, xyzt, feats.detach(), gout, grads)
    print("c2 parity (max over 64 levels):", worst)


def test_c3_geometry_florida_vs_device_restatement(dev):
    """configs[2] geometry: T = 2^22 (HBM-resident tables: 4 forward groups, 2 backward groups, regroup pass and
    replicas live), Florida-clustered points from the prologue kernel -- the contended-reduction case."""
    import deepearth_b200 as de
    from deepearth_b200 import functional as Fn
    T, n = 2 ** 22, 2 ** 22
    torch.manual_seed(1234)
    enc = de.Grid4DEncoder(decomposed_config(de, T)).to(dev)
    xyzt = florida_xyzt(n, dev, 1234)
    # cluster centres sit anywhere in the box and spread 0.1 degree: a minority of points falls outside the fitted
    # bbox (normalised coordinates < 0 or > 1), which the kernels must index exactly as the restatement does
    inside = ((xyzt[:, :3] >= 0) & (xyzt[:, :3] <= 1)).all(1).float().mean()
    assert float(inside) >= 0.6 and bool(torch.isfinite(xyzt).all())
    # the clustered inputs really are clustered: a handful of coarse (R = 16) cells holds most of the points
    cells = (xyzt[:, 0] * 16).floor() * 1000 + (xyzt[:, 1] * 16).floor()
    _, counts = torch.unique(cells, return_counts=True)
    assert float(counts.sort(descending=True).values[:32].sum()) >= 0.25 * n      # uniform points: 32 / 4913 of them
    g = torch.Generator(device=dev).manual_seed(77)
    gout = torch.randn((n, 128), device=dev, generator=g)
    plan = enc._plan()
    assert Fn.backward_group_width(plan.specs, n, 2, True) == 4          # 2 levels per group x F: regrouped rows
    feats = enc.encode_features(xyzt)
    feats.backward(gout)
    grads = [w.grad for w, _ in level_tables(enc)]
    worst = check_against_restatement(enc, xyzt, feats.detach(), gout, grads)
    print("c3 parity (max over 64 levels):", worst)


def test_c5_grouped_head_handoff_at_c3_geometry(dev):
    """configs[4] at T = 2^22: prologue -> encode -> fused 64-wide head; the head's backward writes d(loss)/d(features)
    group-major (W = 4) and the hash-grid backward walks it in place.  Table gradients against the fp64 scatter of the
    row-major gradient the same head kernel produces without the hand-off."""
    import deepearth_b200 as de
    T, n = 2 ** 22, 2 ** 21 + 11
    torch.manual_seed(4321)
    enc = de.Grid4DEncoder(decomposed_config(de, T, grid4d_fused_head=True)).to(dev)
    randomise_tables(enc, dev, 5, -1.0, 1.0)                       # features of order 1 so the head sees real inputs
    xyzt = florida_xyzt(n, dev, 99)
    g = torch.Generator(device=dev).manual_seed(3)
    gy = torch.randn((n, 64), device=dev, generator=g)
    out = enc(xyzt)
    assert enc._fused_head not in (None, False)
    out.backward(gy)
    grads = [w.grad.clone() for w, _ in level_tables(enc)]
    with torch.no_grad():
        feats = enc.encode_features(xyzt)
        gx = enc._fused_head.backward_fused(feats, gy, group_width=0)["x"]      # row-major d(loss)/d(features)
    plan = enc._plan()
    tabs = level_tables(enc)
    worst = 0.0
    for i, spec in enumerate(plan.specs):
        idx, w, _ = reference_level(xyzt, spec, tabs[i][1].float())
        want = reference_grad(idx, w, gx[:, spec.out_col:spec.out_col + 2], T)
        err = float((grads[i].double() - want).abs().max()) / float(want.abs().max())
        assert err <= GRAD_TOL, (i, err)
        worst = max(worst, err)
    print("c5 grouped hand-off, table gradients (max over 64 levels):", worst)


# ---------------------------------------------------------------------------------------------------
# the multi-GPU backward (GradSync: level runs + side-stream all-reduce) against the plain backward
# ---------------------------------------------------------------------------------------------------
@pytest.fixture()
def nccl_world1(dev):
    import torch.distributed as dist
    if dist.is_initialized():
        pytest.skip("a process group already exists in this process")
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=0, world_size=1)
    try:
        yield dist
    finally:
        torch.cuda.synchronize()
        dist.destroy_process_group()


@pytest.mark.parametrize("T,chunks", [(2 ** 19, 4), (2 ** 21, 4), (2 ** 19, 3)])
def test_gradsync_backward_equals_plain_backward(dev, nccl_world1, T, chunks):
    """distributed.GradSync.backward is the code the scaling bench times: the level list cut into runs, each run's
    slice of the flat gradient buffer all-reduced on a side stream.  With one rank the all-reduce is the identity, so
    the result must equal one plain de_grid4d_backward over the whole list (row-major gradient rows)."""
    import deepearth_b200 as de
    from deepearth_b200 import functional as Fn
    from deepearth_b200.distributed import GradSync
    n = 2 ** 20 + 13
    torch.manual_seed(8)
    enc = de.Grid4DEncoder(decomposed_config(de, T)).to(dev)
    plan = enc._plan()
    g = torch.Generator(device=dev).manual_seed(2)
    xyzt = torch.rand((n, 4), device=dev, generator=g)
    gout = torch.randn((n, 128), device=dev, generator=g)
    plain = torch.zeros(plan.grad_numel, device=dev)
    Fn.encode_backward(plan.specs, xyzt, gout, [plain.data_ptr() + 4 * o for o in plan.grad_offsets], 2, True)
    for average in (False, True):
        sync = GradSync(average=average, chunks=chunks)
        flat = torch.zeros(plan.grad_numel, device=dev)
        sync.backward(plan.specs, xyzt, gout, flat, plan.grad_offsets, 2, True)
        torch.cuda.synchronize()
        assert float((flat - plain).abs().max()) <= GRAD_TOL * float(plain.abs().max())
    # the module path (enable_grad_sync -> Grid4DFunction.backward -> GradSync) gives the same parameter gradients
    enc.enable_grad_sync(average=True, chunks=chunks)
    enc.encode_features(xyzt).backward(gout)
    torch.cuda.synchronize()
    got = torch.cat([w.grad.reshape(-1) for w, _ in level_tables(enc)])
    assert float((got - plain).abs().max()) <= GRAD_TOL * float(plain.abs().max())


def test_gradsync_grouped_gradient_planes(dev, nccl_world1):
    """The grouped-W variant: the head wrote d(loss)/d(features) as planes g[c // W][p][c % W]; GradSync walks the
    planes run by run.  Against the plain backward of the same gradient in row-major form."""
    import deepearth_b200 as de
    from deepearth_b200 import functional as Fn
    from deepearth_b200.distributed import GradSync
    T, n = 2 ** 21, 2 ** 20 + 5
    torch.manual_seed(9)
    enc = de.Grid4DEncoder(decomposed_config(de, T)).to(dev)
    plan = enc._plan()
    W = Fn.backward_group_width(plan.specs, n, 2, True)
    assert W == 8                                                   # 16 MiB gradient tables: 4 levels per group
    sync = GradSync(average=False, chunks=4)
    assert sync.grouped_ok(plan.specs, n, 2, True, 0, 0, W)
    g = torch.Generator(device=dev).manual_seed(4)
    xyzt = torch.rand((n, 4), device=dev, generator=g)
    gout = torch.randn((n, 128), device=dev, generator=g)
    planes = gout.view(n, 128 // W, W).permute(1, 0, 2).contiguous()          # g[c // W][p][c % W]
    plain = torch.zeros(plan.grad_numel, device=dev)
    Fn.encode_backward(plan.specs, xyzt, gout, [plain.data_ptr() + 4 * o for o in plan.grad_offsets], 2, True)
    flat = torch.zeros(plan.grad_numel, device=dev)
    sync.backward(plan.specs, xyzt, planes.view(n, 128), flat, plan.grad_offsets, 2, True, group_width=W)
    torch.cuda.synchronize()
    assert float((flat - plain).abs().max()) <= GRAD_TOL * float(plain.abs().max())


# ---------------------------------------------------------------------------------------------------
# C-ABI contract: a workspace smaller than asked for
# ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("fraction", [0.5, 0.1, 0.999])
def test_backward_with_partial_workspace(dev, fraction):
    """include/de_grid4d.h: "A smaller workspace is accepted; the call then skips what does not fit."  The replicas
    that ARE granted must be zeroed by the call even though the scratch is smaller than the full request (and holds
    garbage): gradients must equal the no-workspace result."""
    from deepearth_b200 import functional as Fn
    from deepearth_b200.functional import LevelSpec
    T, L, F = 2 ** 14, 6, 2
    n = 2 ** 20 + 3
    res = [4, 8, 12, 16, 24, 200]                       # dense coarse levels: every one of the first five gets replicas
    specs = [LevelSpec(table_size=T, resolution=R, dims=(0, 1, 2), out_col=l * F, dense=(R + 1) ** 3 <= T)
             for l, R in enumerate(res)]
    g = torch.Generator(device=dev).manual_seed(17)
    coords = torch.rand((n, 4), device=dev, generator=g)
    gout = torch.randn((n, L * F), device=dev, generator=g)

    def run(**kw):
        flat = torch.zeros(L * T * F, device=dev)
        ptrs = [flat.data_ptr() + 4 * l * T * F for l in range(L)]
        # poison the allocator's free blocks so a recycled scratch is visibly not zero
        junk = torch.full((64 << 20,), float("nan"), device=dev)
        del junk
        Fn.encode_backward(specs, coords, gout, ptrs, F, True, extra_flags=0x800, **kw)    # NO_REGROUP: replicas only
        return flat

    ref = run(use_replicas=False)
    full = run()
    part = run(workspace_fraction=fraction)
    denom = float(ref.abs().max())
    assert torch.isfinite(part).all()
    assert float((full - ref).abs().max()) <= GRAD_TOL * denom
    assert float((part - ref).abs().max()) <= GRAD_TOL * denom
