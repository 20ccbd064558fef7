#!/usr/bin/env python
"""Per-level forward / backward time of the decomposed encoder (diagnostic; needs a GPU).

    python scripts/level_sweep.py [--workload c2|c3] [--log2-points N]

Times de_grid4d_forward / de_grid4d_backward restricted to one level row at a time (grid xyz and xyt),
then whole grids, so the cost of contention on coarse levels and of HBM-resident fine levels is visible.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch  # noqa: E402

import bench  # noqa: E402
from deepearth_b200 import functional as Fn  # noqa: E402


def time_ms(fn, reps=5):
    for _ in range(2):
        fn()
    best = 1e30
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--log2-points", type=int, default=0)
    ap.add_argument("--tag", default="")
    a = ap.parse_args()
    wl = bench.WORKLOADS[a.workload]
    dev = torch.device("cuda:0")
    n = 2 ** (a.log2_points or wl["log2_points"])
    job = bench.Job(wl, dev, 1234, n)
    rows = []
    L = 16

    def run_subset(ids, name):
        specs = [job.plan.specs[i] for i in ids]
        tp = [job.table_ptrs[i] for i in ids]
        gp = [job.grad_ptrs[i] for i in ids]
        f = time_ms(lambda: Fn.encode_forward(specs, job.coords, tp, job.plan.table_dtype, job.F, job.plan.out_cols,
                                              job.linear, job.out_scale, out=job.out,
                                              levels_per_group=job.plan.levels_per_group))
        b = time_ms(lambda: Fn.encode_backward(specs, job.coords, job.gout, gp, job.F, job.linear,
                                               levels_per_group=job.plan.bwd_levels_per_group))
        rows.append({"subset": name, "levels": len(ids), "fwd_ms": round(f, 4), "bwd_ms": round(b, 4)})
        print(rows[-1], flush=True)

    for g, gname in ((0, "xyz"), (1, "xyt")):
        for l in range(L):
            s = job.plan.specs[g * L + l]
            run_subset([g * L + l], f"{gname}.L{l} R={s.resolution} dense={int(s.dense)}")
    for g, gname in enumerate(["xyz", "xyt", "yzt", "xzt"]):
        run_subset(list(range(g * L, (g + 1) * L)), f"{gname}.all")
    run_subset([g * L + l for g in range(4) for l in range(0, 3)], "coarse(L0-2 x4)")
    run_subset([g * L + l for g in range(4) for l in range(3, 16)], "fine(L3-15 x4)")
    run_subset(list(range(64)), "all")
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", f"level_sweep_{a.workload}{a.tag}.json"), "w") as f:
        json.dump({"workload": a.workload, "points": n, "rows": rows}, f, indent=1)


if __name__ == "__main__":
    main()
