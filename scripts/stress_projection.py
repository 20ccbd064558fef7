"""Repeat the wide projection's forward + backward on fixed inputs and compare every repetition with the first one.

grad_x and the hidden activations have no atomics on their path: they must repeat bit for bit; the parameter gradients
(split-K / red.global accumulation) must repeat to rounding.  A race in the TMA / tcgen05 / TMEM pipelines shows up here
as a mismatch (or a launch failure) long before a tolerance test notices.

    python scripts/stress_projection.py [--n 131072] [--in-dim 128] [--hidden 768] [--reps 200]
"""
import argparse
import sys
import os

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deepearth_b200 import mlp_head  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1 << 17)
    ap.add_argument("--in-dim", type=int, default=128)
    ap.add_argument("--hidden", type=int, default=768)
    ap.add_argument("--reps", type=int, default=200)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    g = torch.Generator(device="cpu").manual_seed(5)
    n, k1, h = args.n, args.in_dim, args.hidden
    x = torch.randn(n, k1, generator=g).to(dev)
    gy = torch.randn(n, h, generator=g).to(dev)
    w1 = (torch.randn(h, k1, generator=g) / k1 ** 0.5).to(dev).bfloat16()
    w2 = (torch.randn(h, h, generator=g) / h ** 0.5).to(dev).bfloat16()
    b1 = (0.1 * torch.randn(h, generator=g)).to(dev)
    b2 = (0.1 * torch.randn(h, generator=g)).to(dev)
    gamma = (1.0 + 0.1 * torch.randn(h, generator=g)).to(dev)
    beta = (0.1 * torch.randn(h, generator=g)).to(dev)
    w1t, w2t = w1.t().contiguous(), w2.t().contiguous()
    ref = None
    bad = 0
    for rep in range(args.reps):
        y, hid, stats = mlp_head.projection_forward(x, w1, b1, gamma, beta, 1e-5, w2, b2, want_stats=True)
        out = mlp_head.projection_backward(x, gy, w1, w1t, w2t, b1, gamma, beta, hid, stats)
        torch.cuda.synchronize()
        cur = (y, hid, stats) + tuple(out)
        if ref is None:
            ref = [t.clone() for t in cur]
            continue
        names = ["y", "hidden", "stats", "grad_x", "grad_w1", "grad_b1", "grad_gamma", "grad_beta", "grad_w2", "grad_b2"]
        for name, a, b in zip(names, cur, ref):
            if name in ("y", "hidden", "stats", "grad_x"):
                ok = torch.equal(a, b)
            else:
                scale = float(b.abs().max()) + 1e-30
                ok = float((a - b).abs().max()) <= 2e-4 * scale
            if not ok:
                bad += 1
                diff = (a.float() - b.float()).abs()
                print(f"rep {rep}: {name} differs: max |d| {float(diff.max()):.3e} at {int(diff.argmax())}, "
                      f"{int((diff > 0).sum())} elements", flush=True)
    print(f"stress_projection: n={n} in_dim={k1} hidden={h} reps={args.reps} mismatches={bad}")
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
