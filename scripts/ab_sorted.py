#!/usr/bin/env python
"""A/B on a B200: gradient replicas (default) vs sort-by-cell + warp-aggregated backward (opt-in) on the levels where
collisions are dense -- the coarse dense levels of the four grids -- for uniform and Florida-clustered points.

    python scripts/ab_sorted.py [--log2-points 22] [--T 524288] > gpurun_out/ab_sorted.json
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch  # noqa: E402

from bench import _time_cuda, make_geo  # noqa: E402
from deepearth_b200 import functional as Fn  # noqa: E402
from deepearth_b200.functional import GRID_DIMS, LevelSpec  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--log2-points", type=int, default=22)
ap.add_argument("--T", type=int, default=2 ** 19)
ap.add_argument("--bits", type=int, default=5)
a = ap.parse_args()
dev = torch.device("cuda:0")
n, T, F = 2 ** a.log2_points, a.T, 2
out = {"points": n, "T": T, "bits": a.bits, "cases": []}
for coords_kind in ("uniform", "florida"):
    if coords_kind == "uniform":
        xyzt = torch.rand((n, 4), device=dev)
    else:
        geo, t, conv = make_geo(n, dev, 1234)
        xyzt = conv.geodetic_to_xyzt_norm(geo, t)
    for res in ([16], [16, 32, 64], [16, 32, 64, 128, 256, 512, 1024, 2048]):
        specs_by_grid = {nm: [LevelSpec(T, R, dims, l * F, dense=Fn.dense_fits(R, 3, T)) for l, R in enumerate(res)]
                         for nm, dims in GRID_DIMS.items()}
        gout = torch.randn((n, len(res) * F), device=dev)
        grads = torch.zeros((4, len(res), T, F), device=dev)
        ptrs = {nm: [grads[g, l].data_ptr() for l in range(len(res))] for g, nm in enumerate(GRID_DIMS)}

        def default():
            for nm in GRID_DIMS:
                Fn.encode_backward(specs_by_grid[nm], xyzt, gout, ptrs[nm], F, True)

        perms = {}

        def sort_only():
            for nm, dims in GRID_DIMS.items():
                perms[nm] = Fn.sort_points(xyzt, dims, a.bits)

        def sorted_bwd():
            for nm in GRID_DIMS:
                Fn.encode_backward_sorted(specs_by_grid[nm], xyzt, gout, ptrs[nm], F, perms[nm])
        ms_default = _time_cuda(default)
        ms_sort = _time_cuda(sort_only)
        ms_sorted = _time_cuda(sorted_bwd)
        out["cases"].append({"coords": coords_kind, "resolutions": res, "default_replicas_ms": ms_default, "sort_ms": ms_sort,
                             "sorted_aggregated_ms": ms_sorted, "sorted_total_ms": ms_sort + ms_sorted})
print(json.dumps(out, indent=1))
