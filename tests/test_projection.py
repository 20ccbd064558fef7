"""The projection at its real width (models/encoders.py:47-52, hidden_dim 768 / 2048): the
TMA + tcgen05 GEMM kernel on its own, and the two-kernel forward against the fp32 nn.Sequential."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda:0")


@pytest.mark.parametrize("m,n,k", [(128, 256, 64), (1000, 768, 768), (4096 + 17, 768, 768), (300, 200, 72), (257, 2048, 2048),
                                   (64, 8, 128)])
@pytest.mark.parametrize("out_dtype", [torch.float32, torch.bfloat16])
def test_gemm_bf16_vs_torch(dev, m, n, k, out_dtype):
    """C = A B^T + bias: bf16 operands are exact in the products, so an fp64 product of the SAME bf16 values differs
    only by the kernel's fp32 accumulation (<= 2e-6 of max|C| per K = 2048) and, for bf16 output, the final rounding."""
    from deepearth_b200.mlp_head import gemm_bf16
    g = torch.Generator(device=dev).manual_seed(m * 7 + n)
    a = torch.randn((m, k), device=dev, generator=g).to(torch.bfloat16)
    b = torch.randn((n, k), device=dev, generator=g).to(torch.bfloat16)
    bias = torch.randn(n, device=dev, generator=g)
    c = gemm_bf16(a, b, bias, out_dtype=out_dtype)
    want = a.double() @ b.double().t() + bias.double()
    tol = 2e-5 if out_dtype == torch.float32 else 6e-3
    err = float((c.double() - want).abs().max()) / float(want.abs().max())
    assert err <= tol, err
    c2 = gemm_bf16(a, b, None, out_dtype=out_dtype)
    err2 = float((c2.double() - (want - bias.double())).abs().max()) / float(want.abs().max())
    assert err2 <= tol, err2


def same_rounding_reference(seq, x):
    """fp64 evaluation with the kernel's roundings: bf16 x, W1, W2 and bf16 hidden activations."""
    lin1, ln, _, lin2 = seq
    bf = lambda t: t.detach().to(torch.bfloat16).double()      # noqa: E731
    h = bf(x) @ bf(lin1.weight).t() + lin1.bias.double()
    mu = h.mean(-1, keepdim=True)
    var = ((h - mu) ** 2).mean(-1, keepdim=True)
    v = (h - mu) / torch.sqrt(var + ln.eps) * ln.weight.double() + ln.bias.double()
    a = torch.nn.functional.gelu(v)
    return bf(a.float()) @ bf(lin2.weight).t() + lin2.bias.double(), a


@pytest.mark.parametrize("in_dim,hidden,n", [(48, 768, 1000), (128, 768, 4096 + 5), (128, 2048, 513), (64, 128, 130), (32, 256, 1)])
def test_projection_forward_vs_nn_sequential(dev, in_dim, hidden, n):
    import torch.nn as nn
    from deepearth_b200.mlp_head import FusedProjection, projection_forward
    torch.manual_seed(hidden + in_dim)
    seq = nn.Sequential(nn.Linear(in_dim, hidden), nn.LayerNorm(hidden), nn.GELU(), nn.Linear(hidden, hidden)).to(dev)
    with torch.no_grad():
        seq[1].weight.uniform_(0.5, 1.5)
        seq[1].bias.uniform_(-0.5, 0.5)
        seq[0].bias.uniform_(-1.0, 1.0)
    g = torch.Generator(device=dev).manual_seed(n)
    x = torch.randn((n, in_dim), device=dev, generator=g) * 2.0 + 0.3
    fp = FusedProjection(seq)
    assert fp.wide and fp.supported()
    w1, w2 = fp._weights()
    y, hid, stats = projection_forward(x, w1, seq[0].bias, seq[1].weight, seq[1].bias, seq[1].eps, w2, seq[3].bias, want_stats=True)
    want_same, act = same_rounding_reference(seq, x)
    scale = float(want_same.detach().abs().max())
    # hidden activations: bf16 rounding of the fp64 value computed from the same rounded operands (+ GELU approximation 3e-7)
    herr = float((hid.double() - act).abs().max()) / float(act.abs().max())
    assert herr <= 6e-3, herr                    # one bf16 ulp of the largest activation (B200: 2.1e-3 .. 3.1e-3)
    e_same = float((y.double() - want_same).abs().max()) / scale
    assert e_same <= 1e-3, e_same                # B200: <= 2.6e-4
    with torch.no_grad():
        want = seq(x)
    e_mod = float((y - want).abs().max()) / float(want.abs().max())
    assert e_mod <= 1e-2, e_mod                  # B200: <= 3.7e-3
    # LayerNorm statistics of the pre-activation rows
    h = x.to(torch.bfloat16).double() @ w1.double().t() + seq[0].bias.double()
    assert float((stats[:, 0].double() - h.mean(-1)).abs().max()) <= 1e-4 * max(1.0, float(h.abs().max()))
    rstd = 1.0 / torch.sqrt(h.var(-1, unbiased=False) + seq[1].eps)
    assert float((stats[:, 1].double() / rstd - 1).abs().max()) <= 1e-4
    assert torch.equal(fp(x), y)
    print(f"projection K1={in_dim} H={hidden} n={n}: hidden {herr:.1e}, same-rounding {e_same:.1e}, vs fp32 module {e_mod:.1e}")


@pytest.mark.parametrize("m,n,k", [(768, 768, 4096), (256, 128, 1000), (768, 48, 64 * 37 + 5), (136, 248, 70000)])
def test_gemm_bf16_transposed_accumulate_vs_torch(dev, m, n, k):
    """Weight-gradient form: C (M, N) += A (K, M)^T B (K, N), the contraction (points) split over the CTAs."""
    from deepearth_b200.mlp_head import gemm_bf16
    g = torch.Generator(device=dev).manual_seed(k)
    a = torch.randn((k, m), device=dev, generator=g).to(torch.bfloat16)
    b = torch.randn((k, n), device=dev, generator=g).to(torch.bfloat16)
    want = a.double().t() @ b.double()
    c = gemm_bf16(a, b, transposed=True, accumulate=True)
    err = float((c.double() - want).abs().max()) / float(want.abs().max())
    assert err <= 2e-5, err
    c = gemm_bf16(a, b, transposed=True, accumulate=True, out=c)            # accumulates on top
    err = float((c.double() - 2 * want).abs().max()) / float(want.abs().max())
    assert err <= 4e-5, err
    c1 = gemm_bf16(a, b, transposed=True)                                   # plain store, one K range
    # one fp32 accumulation chain of length K in tensor memory: the bound grows with K (B200: 8e-5 at K = 70000)
    assert float((c1.double() - want).abs().max()) / float(want.abs().max()) <= 2e-5 * max(1.0, k / 4096)


@pytest.mark.parametrize("in_dim,hidden,n", [(48, 768, 1000), (128, 768, 2048 + 5), (128, 2048, 300), (64, 128, 130)])
def test_projection_backward_vs_autograd(dev, in_dim, hidden, n):
    """Gradients against autograd through an fp32 evaluation that shares the kernel's operand roundings (bf16 x, W1, W2;
    straight-through at the roundings).  The kernel additionally rounds the hidden activations, grad_y, da and dpre to
    bf16 (2^-9 relative each): bounds are printed and set at ~2x what a B200 measures."""
    import torch.nn as nn
    from deepearth_b200.mlp_head import FusedProjection
    torch.manual_seed(hidden * 3 + in_dim)
    seq = nn.Sequential(nn.Linear(in_dim, hidden), nn.LayerNorm(hidden), nn.GELU(), nn.Linear(hidden, hidden)).to(dev)
    with torch.no_grad():
        seq[1].weight.uniform_(0.5, 1.5)
        seq[1].bias.uniform_(-0.5, 0.5)
    g = torch.Generator(device=dev).manual_seed(n + 1)
    x = (torch.randn((n, in_dim), device=dev, generator=g) * 1.5).requires_grad_(True)
    gy = torch.randn((n, hidden), device=dev, generator=g)
    fp = FusedProjection(seq)
    assert fp.wide and fp.supported() and fp.fused_backward
    y = fp(x)
    y.backward(gy)
    got = {"x": x.grad.clone()}
    got.update({name: p.grad.clone() for name, p in seq.named_parameters()})
    x.grad = None
    seq.zero_grad(set_to_none=True)

    def ste_bf16(t):
        return t + (t.to(torch.bfloat16).float() - t).detach()
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        h = torch.nn.functional.linear(ste_bf16(x), ste_bf16(seq[0].weight), seq[0].bias)
        a = torch.nn.functional.gelu(seq[1](h))
        y_ref = torch.nn.functional.linear(ste_bf16(a), ste_bf16(seq[3].weight), seq[3].bias)
        y_ref.backward(gy)
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    want = {"x": x.grad}
    want.update({name: p.grad for name, p in seq.named_parameters()})
    worst = {}
    for k in want:
        worst[k] = float((got[k] - want[k]).abs().max()) / max(float(want[k].abs().max()), 1e-30)
    print(f"projection backward K1={in_dim} H={hidden} n={n}:", {k: f"{v:.1e}" for k, v in worst.items()})
    for k, v in worst.items():
        assert v <= 1e-2, (k, v)          # B200: <= 5e-3 on every shape (gpurun r02c)


@pytest.mark.parametrize("in_dim,hidden,n", [(128, 768, 2048 + 5), (48, 256, 300)])
def test_projection_bf16_output_and_bf16_grad_y(dev, in_dim, hidden, n):
    """The autocast form (de_projection_forward_ex / _backward_ex with DE_BF16): y leaves the GEMM rounded once -- equal to
    the fp32 output rounded to bf16 -- and a bf16 grad_y is consumed in place: every gradient equals the one computed from
    grad_y.float() through the fp32 entry point (same operands, same kernels), db2 up to the order of its column sums."""
    import torch.nn as nn
    from deepearth_b200.mlp_head import FusedProjection, projection_backward, projection_forward
    torch.manual_seed(in_dim + hidden)
    seq = nn.Sequential(nn.Linear(in_dim, hidden), nn.LayerNorm(hidden), nn.GELU(), nn.Linear(hidden, hidden)).to(dev)
    g = torch.Generator(device=dev).manual_seed(n)
    x = torch.randn((n, in_dim), device=dev, generator=g)
    gy16 = torch.randn((n, hidden), device=dev, generator=g).to(torch.bfloat16)
    fp = FusedProjection(seq)
    w1, w2 = fp._weights()
    args = (x, w1, seq[0].bias, seq[1].weight, seq[1].bias, seq[1].eps, w2, seq[3].bias)
    y32, hid, stats = projection_forward(*args, want_stats=True)
    y16, hid16, stats16 = projection_forward(*args, want_stats=True, out_dtype=torch.bfloat16)
    assert y16.dtype == torch.bfloat16
    assert torch.equal(y16, y32.to(torch.bfloat16)) and torch.equal(hid, hid16) and torch.equal(stats, stats16)
    bargs = (w1, fp._w1t, fp._w2t, seq[0].bias, seq[1].weight, seq[1].bias, hid, stats)
    want = projection_backward(x, gy16.float(), *bargs)
    got = projection_backward(x, gy16, *bargs)
    names = ["grad_x", "grad_w1", "grad_b1", "grad_gamma", "grad_beta", "grad_w2", "grad_b2"]
    for name, a, b in zip(names, got, want):
        scale = float(b.abs().max())
        if name == "grad_x":
            assert torch.equal(a, b), name                       # no atomics on its path
        else:
            assert float((a - b).abs().max()) <= 2e-4 * scale, (name, float((a - b).abs().max()) / scale)
    # a row pitch that is not the row length: grad_y as a column slice of a wider bf16 buffer
    wide = torch.zeros((n, hidden + 16), device=dev, dtype=torch.bfloat16)
    wide[:, 8:8 + hidden] = gy16
    got2 = projection_backward(x, wide[:, 8:8 + hidden], *bargs)
    assert torch.equal(got2[0], want[0])
    assert float((got2[5] - want[5]).abs().max()) <= 2e-4 * float(want[5].abs().max())


def test_grid4d_encoder_under_autocast_bf16(dev):
    """Under torch.autocast(bfloat16) -- how the reference trains (hpc/train_distrbuted.py:284) -- the encoder returns bf16
    like the reference's nn.Linear would, through the bf16-output GEMM, and its backward takes the bf16 gradient directly."""
    import deepearth_b200 as de
    torch.manual_seed(1)
    cfg = de.DeepEarthConfig(hash_table_size=2 ** 14)
    enc = de.Grid4DEncoder(cfg).to(dev)
    with torch.no_grad():
        for e in enc._encoders:
            for t in e.hash_tables:
                t.weight.uniform_(-1.0, 1.0)
    xyzt = torch.rand((2500, 4), device=dev)
    out32 = enc(xyzt)
    out32.square().mean().backward()
    ref = {k: p.grad.clone() for k, p in enc.named_parameters()}
    enc.zero_grad(set_to_none=True)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        out16 = enc(xyzt)
    assert out16.dtype == torch.bfloat16
    assert torch.equal(out16, out32.detach().to(torch.bfloat16))
    out16.float().square().mean().backward()
    worst = 0.0
    for k, p in enc.named_parameters():
        assert p.grad is not None and p.grad.dtype == torch.float32, k
        scale = float(ref[k].abs().max())
        if scale > 0:
            worst = max(worst, float((p.grad - ref[k]).abs().max()) / scale)
    print(f"autocast bf16 vs fp32 output: worst relative gradient difference {worst:.2e}")
    assert worst <= 1e-2, worst              # y and grad_y rounded to bf16 (2^-9 each); B200: 4.8e-3


def test_grid4d_encoder_default_width_runs_on_our_kernels(dev):
    """Grid4DEncoder(config) with the reference's default hidden_dim = 768 (models/configs.py:44) and grid4d_fused_head:
    forward + backward through the wide projection kernels, checked against the fp32 nn.Sequential on the same features."""
    import deepearth_b200 as de
    torch.manual_seed(0)
    cfg = de.DeepEarthConfig(hash_table_size=2 ** 14, grid4d_fused_head=True)
    assert cfg.hidden_dim == 768
    enc = de.Grid4DEncoder(cfg).to(dev)
    with torch.no_grad():
        for e in enc._encoders:
            for t in e.hash_tables:
                t.weight.uniform_(-1.0, 1.0)
    xyzt = torch.rand((3000, 4), device=dev)
    out = enc(xyzt)
    assert enc._fused_head and enc._fused_head.wide
    with torch.no_grad():
        want = enc.projection(enc.encode_features(xyzt))
    assert float((out.detach() - want).abs().max()) <= 1e-2 * float(want.abs().max())
    out.sum().backward()
    assert all(p.grad is not None and bool(torch.isfinite(p.grad).all()) for p in enc.parameters())


def test_projection_repeats_bit_for_bit_in_fresh_processes():
    """scripts/stress_projection.py in three fresh processes: hidden / grad_x repeat bit for bit and no launch fails.
    (Round 2 found a double tcgen05.dealloc here that trapped in ~1 process of 3 at 2^20 rows -- a fault only a
    repeated, full-size run shows; a fresh process per run keeps a failure from poisoning this one's CUDA context.)"""
    import os
    import subprocess
    import sys
    script = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scripts", "stress_projection.py")
    for _ in range(3):
        r = subprocess.run([sys.executable, script, "--n", str(1 << 20), "--reps", "6"], capture_output=True, text=True,
                           timeout=600)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        assert "mismatches=0" in r.stdout


def test_single_cta_gemm_path_still_matches(dev):
    """DE_GEMM_PAIR=0 (read once per process) selects the single-CTA form of the GEMM kernel (M = 128 tiles, 4-stage ring) that
    the CTA-pair default replaced; it stays in the library as the A/B arm and the fallback, so its parity is kept green: the
    GEMM and projection-backward tests of this file again, in a child process with the switch set."""
    import os
    import subprocess
    import sys
    env = dict(os.environ, DE_GEMM_PAIR="0")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-m", "gpu", "-q", "-x", "-k",
                        "gemm_bf16 or backward_vs_autograd", "-p", "no:cacheprovider"],
                       capture_output=True, text=True, timeout=900, env=env,
                       cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert " passed" in r.stdout and "failed" not in r.stdout
