"""Generate tests/golden/st_embed.npz by running the UNMODIFIED reference SpatialTemporalEmbedding
(/root/reference/models/cross_modal_fusion.py:99-165) on the CPU.  Container only:
    python tests/golden/make_golden_st_embed.py
Stored: the parameters (small), the inputs, the reference output and the reference's autograd gradients.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle.ref_loader import load_reference_models  # noqa: E402

ref = load_reference_models()
assert ref is not None and ref.SpatialTemporalEmbedding is not None, "run oracle/make_ref.py first"


def case(tag, D, B, N, seed, spatial=True, temporal=True, modality=True, negative=False):
    torch.manual_seed(seed)
    m = ref.SpatialTemporalEmbedding(D)
    m.add_modality("vision")
    g = torch.Generator().manual_seed(seed + 1)
    tokens = torch.randn((B, N, D), generator=g, requires_grad=True)
    pos = torch.rand((B, N, 2), generator=g) * 0.999
    if negative:
        pos = pos * 2 - 1                                   # [-1, 1): negative indices wrap to the end of the table
        pos[0, 0, 0], pos[0, 0, 1] = -1.0, -0.0078125       # exact ends: -64 -> row 0, -0.5 -> trunc 0
    t = torch.rand((B, N, 1), generator=g) * 2 - 0.5
    gout = torch.randn((B, N, D), generator=g)
    out = m(tokens, pos if spatial else None, t if temporal else None, "vision" if modality else None)
    out.backward(gout)
    d = {f"{tag}_tokens": tokens.detach().numpy(), f"{tag}_pos": pos.numpy(), f"{tag}_t": t.numpy(),
         f"{tag}_gout": gout.numpy(), f"{tag}_out": out.detach().numpy(),
         f"{tag}_flags": np.array([spatial, temporal, modality], dtype=np.int64)}
    assert torch.equal(tokens.grad, gout)                   # d(out)/d(tokens) is the identity (:164)
    for name, p in m.named_parameters():
        d[f"{tag}_param_{name}"] = p.detach().numpy()
        if p.grad is not None:
            d[f"{tag}_grad_{name}"] = p.grad.numpy()
    print(tag, out.shape, float(out.detach().double().sum()))
    return d


payload = {}
payload.update(case("full", 64, 3, 37, 1))
payload.update(case("neg", 32, 2, 50, 2, negative=True))
payload.update(case("spatial_only", 48, 1, 19, 3, temporal=False, modality=False))
payload.update(case("temporal_only", 40, 2, 21, 4, spatial=False))
payload.update(case("wide", 256, 2, 65, 5))
np.savez_compressed(os.path.join(HERE, "st_embed.npz"), **payload)
print("wrote st_embed.npz")
