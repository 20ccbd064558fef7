#!/usr/bin/env python
"""A/B of the head backward kernels on a GPU box: ping-pong kernel (variant 0, or 2 = swapped GEMM-4 LBO/SBO)
against the serial kernel (variant 1) on identical inputs.  Usage: head_bwd_check.py [variant] [in_dim] [n ...]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deepearth_b200.mlp_head import FusedProjection  # noqa: E402

variant = sys.argv[1] if len(sys.argv) > 1 else "0"
in_dim = int(sys.argv[2]) if len(sys.argv) > 2 else 128
ns = [int(v) for v in sys.argv[3:]] or [1, 129, 5000, 148 * 4 * 128 + 17]
dev = torch.device("cuda", 0)
torch.manual_seed(3)
seq = torch.nn.Sequential(torch.nn.Linear(in_dim, 64), torch.nn.LayerNorm(64), torch.nn.GELU(), torch.nn.Linear(64, 64)).to(dev)
head = FusedProjection(seq)
g = torch.Generator(device=dev).manual_seed(5)
names = {id(p): k for k, p in seq.named_parameters()}
for n in ns:
    x = torch.randn((n, in_dim), device=dev, generator=g)
    gy = torch.randn((n, 64), device=dev, generator=g)
    os.environ["DE_HEAD_BWD_VARIANT"] = "1"
    want = head.backward_fused(x, gy)
    torch.cuda.synchronize()
    os.environ["DE_HEAD_BWD_VARIANT"] = variant
    got = head.backward_fused(x, gy)
    torch.cuda.synchronize()
    line = []
    for k, w in want.items():
        err = float((got[k] - w).abs().max()) / max(float(w.abs().max()), 1e-20)
        line.append(f"{names.get(k, k)}={err:.2e}")
    print(f"variant {variant} in_dim {in_dim} n {n}: " + " ".join(line), flush=True)
