"""SpatialTemporalEmbedding: oracle vs the reference's golden vectors on the CPU, the CUDA
module vs the same vectors on the GPU (models/cross_modal_fusion.py:99-165)."""
import os

import numpy as np
import pytest
import torch

from oracle import st_embed_oracle as orc

CASES = ["full", "neg", "spatial_only", "temporal_only", "wide"]
FWD_TOL = 1e-5      # temporal half: fp32 summation order of the second Linear
GRAD_TOL = 1e-4


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "st_embed.npz"))


def unpack(gold, tag):
    spatial, temporal, modality = (bool(v) for v in gold[f"{tag}_flags"])
    p = {k[len(tag) + 7:]: gold[k] for k in gold.files if k.startswith(f"{tag}_param_")}
    g = {k[len(tag) + 6:]: gold[k] for k in gold.files if k.startswith(f"{tag}_grad_")}
    return spatial, temporal, modality, p, g


@pytest.mark.parametrize("tag", CASES)
def test_oracle_matches_reference_golden(gold, tag):
    spatial, temporal, modality, p, g = unpack(gold, tag)
    tokens = gold[f"{tag}_tokens"]
    B, N, D = tokens.shape
    tok = tokens.reshape(-1, D)
    pos = gold[f"{tag}_pos"].reshape(-1, 2) if spatial else None
    t = gold[f"{tag}_t"].reshape(-1) if temporal else None
    ex, ey = p["spatial_embed_x"][0], p["spatial_embed_y"][0]
    w1, b1 = p["temporal_embed.0.weight"][:, 0], p["temporal_embed.0.bias"]
    w2, b2 = p["temporal_embed.2.weight"], p["temporal_embed.2.bias"]
    mod = p["modality_embeds.vision"].reshape(-1) if modality else None
    out = orc.forward(tok, pos, t, ex, ey, w1, b1, w2, b2, mod)
    want = gold[f"{tag}_out"].reshape(-1, D)
    assert np.array_equal(out[:, :D // 2], want[:, :D // 2]), "spatial half must be bit-exact"
    assert np.abs(out - want).max() <= FWD_TOL * np.abs(want).max()
    grads = orc.backward(gold[f"{tag}_gout"].reshape(-1, D), pos, t, ex.shape, w1, b1, w2, modality)
    pairs = []
    if spatial:
        pairs += [("spatial_embed_x", grads["spatial_embed_x"][None]), ("spatial_embed_y", grads["spatial_embed_y"][None])]
    if temporal:
        pairs += [("temporal_embed.0.weight", grads["w1"][:, None]), ("temporal_embed.0.bias", grads["b1"]),
                  ("temporal_embed.2.weight", grads["w2"]), ("temporal_embed.2.bias", grads["b2"])]
    if modality:
        pairs += [("modality_embeds.vision", grads["mod"].reshape(1, 1, -1))]
    for name, got in pairs:
        ref = g[name]
        assert np.abs(got - ref).max() <= GRAD_TOL * max(np.abs(ref).max(), 1e-30), name


def test_oracle_index_rules():
    assert list(orc.table_rows(np.array([0.0, 0.999, -0.0078125, -1.0, -0.5], np.float32), 64)) == [0, 63, 0, 0, 32]
    for bad in (1.0, -1.02, np.nan):
        with pytest.raises(IndexError):
            orc.table_rows(np.array([bad], np.float32), 64)


# ---------------------------------------------------------------------------------------------------
# GPU
# ---------------------------------------------------------------------------------------------------
def build_module(gold, tag, dev):
    import deepearth_b200 as de
    spatial, temporal, modality, p, g = unpack(gold, tag)
    D = gold[f"{tag}_tokens"].shape[-1]
    m = de.SpatialTemporalEmbedding(D)
    m.add_modality("vision")
    m.load_state_dict({k: torch.from_numpy(v) for k, v in p.items()})      # the reference's own state dict
    return m.to(dev), spatial, temporal, modality, g


@pytest.mark.gpu
@pytest.mark.parametrize("tag", CASES)
def test_cuda_module_matches_reference_golden(gold, tag):
    dev = torch.device("cuda:0")
    m, spatial, temporal, modality, g = build_module(gold, tag, dev)
    tokens = torch.from_numpy(gold[f"{tag}_tokens"]).to(dev).requires_grad_(True)
    pos = torch.from_numpy(gold[f"{tag}_pos"]).to(dev) if spatial else None
    t = torch.from_numpy(gold[f"{tag}_t"]).to(dev) if temporal else None
    out = m(tokens, pos, t, "vision" if modality else None)
    want = torch.from_numpy(gold[f"{tag}_out"]).to(dev)
    D = want.shape[-1]
    if temporal:
        assert torch.equal(out[..., :D // 2], want[..., :D // 2]), "gather half must be bit-exact"
        assert float((out - want).abs().max()) <= FWD_TOL * float(want.abs().max())
    else:
        assert torch.equal(out, want)
    gout = torch.from_numpy(gold[f"{tag}_gout"]).to(dev)
    out.backward(gout)
    assert torch.equal(tokens.grad, gout)
    for name, prm in m.named_parameters():
        if name in g:
            ref = torch.from_numpy(g[name]).to(dev)
            err = float((prm.grad - ref).abs().max()) / max(float(ref.abs().max()), 1e-30)
            assert err <= GRAD_TOL, (name, err)
        else:
            assert prm.grad is None or float(prm.grad.abs().max()) == 0.0, name


@pytest.mark.gpu
def test_cuda_module_index_errors_and_wrap(gold):
    import deepearth_b200 as de
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    m = de.SpatialTemporalEmbedding(32).to(dev)
    tok = torch.zeros((1, 3, 32), device=dev)
    pos = torch.tensor([[[0.0, 0.5], [-0.5, -1.0], [0.99, -0.001]]], device=dev)
    out = m(tok, pos)
    ex, ey = m.spatial_embed_x[0], m.spatial_embed_y[0]
    want = torch.stack([torch.cat([ex[0], ey[32]]), torch.cat([ex[32], ey[0]]), torch.cat([ex[63], ey[0]])])
    assert torch.equal(out[0, :, :16], want)
    assert float(out[0, :, 16:].abs().max()) == 0.0
    with pytest.raises(IndexError):
        m(tok, torch.tensor([[[1.0, 0.5], [0.1, 0.1], [0.1, 0.1]]], device=dev))
    with pytest.raises(RuntimeError):
        m(tok.cpu(), pos.cpu())


@pytest.mark.gpu
def test_cuda_module_large_vs_torch_eager(gold):
    """2^15 tokens x D = 512 against the same math through ATen on the device (fp32)."""
    import deepearth_b200 as de
    dev = torch.device("cuda:0")
    torch.manual_seed(5)
    D, B, N = 512, 32, 1024
    m = de.SpatialTemporalEmbedding(D).to(dev)
    m.add_modality("x")
    g = torch.Generator(device=dev).manual_seed(9)
    tok = torch.randn((B, N, D), device=dev, generator=g)
    pos = torch.rand((B, N, 2), device=dev, generator=g) * 0.999
    t = torch.rand((B, N, 1), device=dev, generator=g)
    out = m(tok, pos, t, "x")
    xi, yi = (pos[..., 0] * 64).long(), (pos[..., 1] * 64).long()
    emb = torch.zeros_like(tok)
    emb[..., :D // 2] += torch.cat([m.spatial_embed_x[0][xi], m.spatial_embed_y[0][yi]], -1)
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        emb[..., D // 2:] += m.temporal_embed(t)
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    emb += m.modality_embeds["x"].expand(B, N, -1)
    want = tok + emb
    assert torch.equal(out[..., :D // 2], want[..., :D // 2])
    assert float((out - want).abs().max()) <= FWD_TOL * float(want.abs().max())


@pytest.mark.gpu
@pytest.mark.parametrize("D,n,with_mod", [(128, 1000, True), (256, 64 * 7 + 3, False), (512, 70001, True), (2048, 3000, True),
                                          (1024, 2 * 65536 + 777, True)])      # the last one: three token slabs, ragged tail
def test_tensor_core_mlp_matches_fp32_path_and_fp64(D, n, with_mod):
    """The temporal MLP on the tensor cores (split-bf16 products, csrc/st_embed_tc.cu; the default when D % 128 == 0) against the
    fp32 SIMT path (tensor_core_mlp=False) and against an fp64 evaluation: forward and every parameter gradient, ragged token
    counts (slab tails, k-blocks of the token contraction that are not full), with and without a modality embedding.  Both
    paths must sit at fp32 rounding distance from fp64 -- the split product is not a reduced-precision mode."""
    import deepearth_b200 as de
    dev = torch.device("cuda:0")
    torch.manual_seed(D + n)
    m_tc = de.SpatialTemporalEmbedding(D).to(dev)
    if with_mod:
        m_tc.add_modality("x")
    m_32 = de.SpatialTemporalEmbedding(D, tensor_core_mlp=False).to(dev)
    if with_mod:
        m_32.add_modality("x")
    m_32.load_state_dict(m_tc.state_dict())
    g = torch.Generator(device=dev).manual_seed(n)
    tok = torch.randn((1, n, D), device=dev, generator=g)
    pos = torch.rand((1, n, 2), device=dev, generator=g) * 0.999
    t = torch.rand((1, n, 1), device=dev, generator=g)
    gout = torch.randn((1, n, D), device=dev, generator=g)
    name = "x" if with_mod else None
    outs, grads = [], []
    for m in (m_tc, m_32):
        out = m(tok, pos, t, name)
        out.backward(gout)
        outs.append(out.detach())
        grads.append({k: p.grad.clone() for k, p in m.named_parameters()})
    # fp64 evaluation of the temporal half and its gradients
    lin1, lin2 = m_tc.temporal_embed[0], m_tc.temporal_embed[2]
    w1, b1, w2, b2 = (p.detach().double().requires_grad_(True) for p in (lin1.weight, lin1.bias, lin2.weight, lin2.bias))
    h = torch.nn.functional.gelu(t.double() @ w1.t() + b1)
    emb_t = h @ w2.t() + b2
    if with_mod:
        emb_t = emb_t + m_tc.modality_embeds["x"].detach().double().reshape(-1)[D // 2:]
    want_t = tok.double()[..., D // 2:] + emb_t
    emb_t.backward(gout.double()[..., D // 2:])
    ref = {"temporal_embed.0.weight": w1.grad, "temporal_embed.0.bias": b1.grad, "temporal_embed.2.weight": w2.grad,
           "temporal_embed.2.bias": b2.grad}
    scale = float(want_t.abs().max())
    e_tc = float((outs[0][..., D // 2:].double() - want_t).abs().max()) / scale
    e_32 = float((outs[1][..., D // 2:].double() - want_t).abs().max()) / scale
    assert torch.equal(outs[0][..., :D // 2], outs[1][..., :D // 2])                    # the gather half is the same kernel
    assert e_tc <= 1e-6 and e_32 <= 1e-6, (e_tc, e_32)             # B200: tensor-core 5e-8 .. 2.7e-7, SIMT 7e-8 .. 2.9e-7
    worst = {}
    for k, r in ref.items():
        s = float(r.abs().max())
        worst[k] = (float((grads[0][k].double() - r).abs().max()) / s, float((grads[1][k].double() - r).abs().max()) / s)
        assert worst[k][0] <= 1e-5 and worst[k][1] <= 1e-5, (k, worst[k])       # B200: both paths <= 3e-6 of max
    for k in grads[0]:
        if k not in ref:                                                               # tables, modality: shared kernels
            assert float((grads[0][k] - grads[1][k]).abs().max()) <= 1e-6 * max(float(grads[1][k].abs().max()), 1e-30), k
    print(f"st_embed D={D} n={n}: forward vs fp64 tensor-core {e_tc:.1e} / SIMT {e_32:.1e}; gradients (tc, simt):",
          {k.replace('temporal_embed.', ''): (f"{a:.1e}", f"{b:.1e}") for k, (a, b) in worst.items()})
