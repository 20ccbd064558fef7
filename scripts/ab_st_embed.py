"""SpatialTemporalEmbedding: our kernels against the same math through ATen (the reference's forward, op by op) on the same GPU.

    python scripts/ab_st_embed.py [--tokens 65536] [--dim 512]
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import deepearth_b200 as de  # noqa: E402


def timeit(fn, steps=10, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tokens", type=int, default=65536)
    ap.add_argument("--dim", type=int, default=512)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    D, n = a.dim, a.tokens
    m = de.SpatialTemporalEmbedding(D).to(dev)
    m.add_modality("x")
    g = torch.Generator(device=dev).manual_seed(9)
    tok = torch.randn((n // 256, 256, D), device=dev, generator=g).requires_grad_(True)
    pos = torch.rand((n // 256, 256, 2), device=dev, generator=g) * 0.999
    t = torch.rand((n // 256, 256, 1), device=dev, generator=g)
    gout = torch.randn_like(tok)

    def eager_fwd():                    # models/cross_modal_fusion.py:130-164, op by op
        xi, yi = (pos[..., 0] * 64).long(), (pos[..., 1] * 64).long()
        emb = torch.zeros_like(tok)
        emb[..., :D // 2] += torch.cat([m.spatial_embed_x[0][xi], m.spatial_embed_y[0][yi]], -1)
        emb[..., D // 2:] += m.temporal_embed(t)
        emb = emb + m.modality_embeds["x"].expand(tok.shape[0], tok.shape[1], -1)
        return tok + emb

    def ours_step():
        tok.grad = None
        m.zero_grad(set_to_none=True)
        m(tok, pos, t, "x").backward(gout)

    def eager_step():
        tok.grad = None
        m.zero_grad(set_to_none=True)
        eager_fwd().backward(gout)
    res = {"tokens": n, "dim": D}
    with torch.no_grad():
        res["ours_forward_ms"] = timeit(lambda: m(tok, pos, t, "x"))
        res["eager_forward_ms"] = timeit(eager_fwd)
    res["ours_fwd_bwd_ms"] = timeit(ours_step)
    res["eager_fwd_bwd_ms"] = timeit(eager_step)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
