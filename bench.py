#!/usr/bin/env python
"""bench.py -- Grid4D hash-encode forward+backward throughput (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c4|c5|c1r] [--impl reference]

A "step" is one pass of the hot path over one batch of synthetic points: zero the table-gradient
buffer, forward encode (de_grid4d_forward), backward scatter (de_grid4d_backward) and -- at N>1 -- the
NCCL all-reduce of the flat gradient buffer.  Per-GPU work is fixed as N grows (weak scaling): points
shard across ranks, tables are replicated.

Workloads (BASELINE.json configs):
  c2  (default) decomposed xyz/xyt/yzt/xzt x 16 levels x 2 feat, T=2^19, 2^22 uniform points per GPU,
      fp16 working tables, fp32 output and gradients                       [configs[1]]
  c3  same grids with T=2^22 (HBM-resident tables), 2^24 Florida-clustered points run through the
      geo2xyz prologue first                                               [configs[2]]
  c4  the c2 grids on 2^26 points IN TOTAL, sharded over the ranks ("scaling": "strong")   [configs[3]]
  c5  c3 + geo2xyz prologue + 64-wide tcgen05 projection head, forward and backward        [configs[4]]
  c1r the reference's own Grid4DEncoder geometry (xyz 16 levels T=2^19 + t 8 levels T=2^17, nearest,
      fp32 tables), 2^22 points                                            [configs[0] geometry on GPU]

One JSON line on stdout (rank 0).  `value` times the C-ABI calls with device-resident inputs;
`e2e` times the public nn.Module API with pinned-host coordinates copied in and a scalar read back.

Besides the headline workload the default line carries (key "extra"):
  N = 1:  "c3"  -- the north_star's target configuration (2^24 Florida-clustered points, T = 2^22) with its own
                   kernels_ms / roofline / e2e;  "c1r" -- the reference's own geometry (the only mode with a reference
                   implementation);  "eager_cuda_baseline" -- the reference op chain through PyTorch-CUDA eager on the
                   same GPU (the restatement oracle/grid4d_torch.py at the c2 geometry; at the c1r geometry the
                   UNMODIFIED reference classes from oracle/_ref when build() could copy them);
  N > 1:  "c4"  -- configs[3], 2^26 points in total (strong scaling) beside the weak-scaling headline;
          "grad_check" -- max relative error of the all-reduced gradient against the gradient of the union of all
                   ranks' points computed locally (mean over ranks, hpc/train_distrbuted.py:499-504).
--extras none switches them off (development runs); an explicit --workload runs that workload alone.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402
import torch  # noqa: E402

METRIC = "grid4d_hash_encode_fwd_bwd_mpoints_per_s"
UNIT = "Mpoints/s"

WORKLOADS = {
    # name: (mode, T, log2 points/GPU, coords, table dtype)
    "c2": dict(mode="decomposed", T=2 ** 19, log2_points=22, coords="uniform", table_dtype="float16",
               desc="configs[1]: 4 grids (xyz/xyt/yzt/xzt) x 16 levels x 2 feat, T=2^19/level, 2^22 points/GPU, fp16 tables, fp32 grads"),
    "c3": dict(mode="decomposed", T=2 ** 22, log2_points=24, coords="florida", table_dtype="float16",
               desc="configs[2]: 4 grids x 16 levels x 2 feat, T=2^22/level (HBM resident), 2^24 Florida-clustered points/GPU"),
    "c5": dict(mode="decomposed", T=2 ** 22, log2_points=24, coords="florida", table_dtype="float16", head=True,
               desc="configs[4]: geo2xyz prologue + 4 grids x 16 levels x 2 feat (T=2^22) + Linear-LayerNorm-GELU-Linear 64-wide head "
                    "(tcgen05 kernels forward and backward), 2^24 Florida-clustered geodetic points/GPU"),
    "c5h768": dict(mode="decomposed", T=2 ** 19, log2_points=21, coords="florida", table_dtype="float16", head=True, hidden=768,
                   desc="encoder step at the reference's default width: geo2xyz prologue + 4 grids x 16 levels x 2 feat (T=2^19) + "
                        "Linear(128,768)-LayerNorm-GELU-Linear(768,768) projection (models/configs.py:44) on the tcgen05 / TMA kernels, "
                        "forward and backward, 2^21 Florida-clustered geodetic points/GPU"),
    "c4": dict(mode="decomposed", T=2 ** 19, log2_points=26, total=True, coords="uniform", table_dtype="float16",
               desc="configs[3]: 2^26 points in total, sharded over the ranks (2^26 / N per GPU), 4 grids x 16 levels x 2 feat, "
                    "T=2^19/level, replicated tables, all-reduce of the table gradients"),
    "c1r": dict(mode="reference", T=2 ** 19, log2_points=22, coords="uniform", table_dtype="float32",
                desc="reference Grid4DEncoder geometry (xyz L16 T=2^19 + t L8 T=2^17, nearest, fp32 tables), 2^22 points/GPU"),
}


def algorithmic_bytes_per_point(n_level_rows: int, corners: int, F: int, table_bytes: int):
    """fwd = 16 + G*L*(C*F*s_t + F*4); bwd = 16 + G*L*(F*4 + C*F*4)."""
    fwd = 16 + n_level_rows * (corners * F * table_bytes + F * 4)
    bwd = 16 + n_level_rows * (F * 4 + corners * F * 4)
    return fwd, bwd


# L1->L2 request ("packet" = distinct 32 B sector per instruction) peaks measured on B200 by scripts/peaks_gather.py
# (profiles/r01c_peaks_gather.json, L2-resident working set, 2 adjacent lanes on adjacent entries):
# loads 574 G lane-ops/s = 287 G packets/s (1.0 / clk / SM); red.global.add.v2.f32 430 G lane-ops/s = 215 G packets/s.
PEAKS_FILE = os.path.join("profiles", "r01c_peaks_gather.json")


def packet_peaks():
    """Best L2-resident rate of the "2 adjacent lanes" pattern (one packet per lane pair): 4 B entries for the
    forward's fp16 x 2 gathers, 8 B entries for the backward's red.v2.f32 -- read from the committed probe output."""
    with open(os.path.join(ROOT, PEAKS_FILE)) as f:
        rows = json.load(f)["rows"]

    def best(op, eb):
        return max(r["gops"] for r in rows if r["working_set"].startswith("L2_") and r["op"] == op
                   and r["entry_bytes"] == eb and r["pattern"] == "2 adjacent lanes") / 2.0
    return {"forward": best("load", 4), "backward": best("red.add.f32", 8)}


def packets_per_point(n_level_rows: int, corners: int, F: int, entry_bytes: int):
    """Distinct 32 B sectors one point touches per level row when the x-neighbour pair is issued by adjacent lanes
    (DESIGN.md 2.0): C/2 x-pairs, each one sector unless the pair straddles a sector boundary."""
    if corners == 1:
        return float(n_level_rows)
    per_sector = max(1, 32 // (F * entry_bytes))
    return n_level_rows * (corners / 2) * (1.0 + 1.0 / per_sector)


def morton_sort(coords, bits=10):
    """Experiment (--sort-points): order the points along a 4-D Morton curve of their 2^bits cells."""
    q = (coords.clamp(0, 1 - 1e-7) * (1 << bits)).to(torch.int64)
    key = torch.zeros(coords.shape[0], dtype=torch.int64, device=coords.device)
    for b in range(bits):
        for d in range(4):
            key |= ((q[:, d] >> b) & 1) << (4 * b + d)
    return coords[torch.argsort(key)].contiguous()


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md: 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows = []
        self.proc = None
        self.index = index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def mark(self):
        """Host time stamp; samples are attributed to the window between two marks."""
        return time.perf_counter()

    def n_between(self, t0, t1):
        return sum(1 for t, _ in self.rows if t0 <= t <= t1)

    def stop(self, t0=None, t1=None):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.1)
        self.proc.terminate()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, r in self.rows:
            if t0 is not None and not (t0 <= t <= t1):
                continue
            parts = [p.strip() for p in r.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                smax.append(float(parts[1]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------------
# workload construction
# ---------------------------------------------------------------------------------------------------
def build_config(wl):
    import deepearth_b200 as de
    res = [2 ** i for i in range(4, 20)]
    if wl["mode"] == "decomposed":
        return de.DeepEarthConfig(hidden_dim=wl.get("hidden", 64), n_heads=4, n_layers=1, hash_table_size=wl["T"],
                                  n_spatial_levels=16, n_temporal_levels=16, spatial_resolutions=res,
                                  temporal_resolutions=res, grid4d_mode="decomposed",
                                  grid4d_table_dtype=wl["table_dtype"], grid4d_fused_head=bool(wl.get("head")))
    return de.DeepEarthConfig(hidden_dim=64, n_heads=4, n_layers=1, hash_table_size=wl["T"])


def make_geo(n, device, seed):
    """Florida-clustered geodetic points: (geo (n,3) fp64, t (n,) fp64, converter with pre-fitted bbox)."""
    import deepearth_b200 as de
    from _synth import FLORIDA_BBOX_GEO as B, FLORIDA_T_RANGE
    g = torch.Generator(device=device).manual_seed(seed)
    K = 64
    cu = torch.rand((K, 2), device=device, generator=g, dtype=torch.float64)
    clat = B["south"] + (B["north"] - B["south"]) * cu[:, 0]
    clon = B["west"] + (B["east"] - B["west"]) * cu[:, 1]
    w = 1.0 / torch.arange(1, K + 1, device=device, dtype=torch.float64)
    which = torch.multinomial(w / w.sum(), n, replacement=True, generator=g)
    lat = clat[which] + 0.1 * torch.randn(n, device=device, generator=g, dtype=torch.float64)
    lon = clon[which] + 0.1 * torch.randn(n, device=device, generator=g, dtype=torch.float64)
    alt = 50.0 * torch.rand(n, device=device, generator=g, dtype=torch.float64)
    t = FLORIDA_T_RANGE[0] + (FLORIDA_T_RANGE[1] - FLORIDA_T_RANGE[0]) * torch.rand(n, device=device, generator=g, dtype=torch.float64)
    geo = torch.stack([lat, lon, alt], 1).contiguous()
    # bbox = the Central-Florida box at alt 0..50 m, fitted once from its corners (+ edge midpoints)
    conv = de.GeospatialConverter(device=str(device), t_range=FLORIDA_T_RANGE)
    lats = torch.linspace(B["south"], B["north"], 9, dtype=torch.float64, device=device)
    lons = torch.linspace(B["west"], B["east"], 9, dtype=torch.float64, device=device)
    alts = torch.tensor([B["alt_min"], B["alt_max"]], dtype=torch.float64, device=device)
    conv.update_bbox(torch.cartesian_prod(lats, lons, alts))
    return geo, t, conv


def make_coords(wl, n, device, seed):
    if wl["coords"] == "uniform":
        g = torch.Generator(device=device).manual_seed(seed)
        return torch.rand((n, 4), device=device, generator=g)
    geo, t, conv = make_geo(n, device, seed)
    return conv.geodetic_to_xyzt_norm(geo, t)       # the geo2xyz prologue kernel


class Job:
    """Device-resident state of one rank: module, flat gradient buffer, inputs, level plan."""

    def __init__(self, wl, device, seed, n_points):
        import deepearth_b200 as de
        self.de = de
        self.wl = wl
        self.device = device
        self.n = n_points
        torch.manual_seed(1234)                       # identical (replicated) tables on every rank
        self.enc = de.Grid4DEncoder(build_config(wl)).to(device)
        self.plan = self.enc._plan()
        self.weights = [w for e in self.enc._encoders for w in e.weights()]
        self.F = self.plan.n_features
        self.linear = self.plan.linear
        if self.plan.table_dtype == torch.float16:
            self.table_ptrs = [p for e in self.enc._encoders for p in e.working_ptrs()]
        else:
            self.table_ptrs = [w.data_ptr() for w in self.weights]
        self.weight_ptrs = [w.data_ptr() for w in self.weights]
        self.flat_grad = torch.zeros(self.plan.grad_numel, dtype=torch.float32, device=device)
        self.grad_ptrs = [self.flat_grad.data_ptr() + 4 * o for o in self.plan.grad_offsets]
        self.head = None
        self.hid = self.stats = None
        g = torch.Generator(device=device).manual_seed(seed + 1)
        if wl.get("head"):
            # configs[4]: the step starts from geodetic points (prologue kernel inside the timed region) and
            # ends in the 64-wide head; the incoming gradient is d(loss)/d(head output)
            from deepearth_b200.mlp_head import FusedProjection
            self.geo, self.t, self.conv = make_geo(n_points, device, seed)
            self.coords = self.conv.geodetic_to_xyzt_norm(self.geo, self.t)
            self.head = FusedProjection(self.enc.projection)
            assert self.head.supported(), "projection shape not served by the tcgen05 kernels"
            self.head_params = list(self.enc.projection.parameters())
            self.gy = torch.randn((n_points, wl.get("hidden", 64)), device=device, generator=g)
            self.gout = None
        else:
            self.coords, self.gout = Job.inputs(wl, n_points, device, seed, self.plan.out_cols)
        self.out = torch.empty((n_points, self.plan.out_cols), dtype=torch.float32, device=device)
        corners = 8 if self.linear else 1
        tb = 2 if self.plan.table_dtype == torch.float16 else 4
        self.bytes_fwd, self.bytes_bwd = algorithmic_bytes_per_point(len(self.plan.specs), corners, self.F, tb)
        self.packets_fwd = packets_per_point(len(self.plan.specs), corners, self.F, tb)
        self.packets_bwd = packets_per_point(len(self.plan.specs), corners, self.F, 4)
        self.out_scale = 1.0 / self.plan.table_scale if self.plan.table_scale != 1.0 else 1.0
        self.use_replicas = True
        self.head_bwd = "tcgen05"
        self.group_w = 0          # > 0: the head backward writes d(loss)/d(features) group-major for the hash-grid backward

    @staticmethod
    def inputs(wl, n_points, device, seed, out_cols):
        """(coords, grad_out) of the rank whose seed this is: reproducible on any rank (grad_check rebuilds them)."""
        coords = make_coords(wl, n_points, device, seed)
        if wl.get("sort_points"):
            coords = morton_sort(coords)
        g = torch.Generator(device=device).manual_seed(seed + 1)
        return coords, torch.randn((n_points, out_cols), device=device, generator=g)

    def refresh_tables(self):
        """fp32 master tables -> fp16 working tables (one de_tables_to_f16 launch over all 64 levels); no-op for fp32."""
        if self.plan.table_dtype != torch.float16:
            return
        from deepearth_b200 import functional as Fn
        Fn.tables_to_f16(self.plan.specs, self.weight_ptrs, self.table_ptrs, self.F, self.plan.table_scale, self.device)

    def forward(self):
        from deepearth_b200 import functional as Fn
        Fn.encode_forward(self.plan.specs, self.coords, self.table_ptrs, self.plan.table_dtype, self.F,
                          self.plan.out_cols, self.linear, self.out_scale, out=self.out,
                          levels_per_group=self.plan.levels_per_group, extra_flags=self.plan.extra_flags)

    def backward(self):
        from deepearth_b200 import functional as Fn
        Fn.encode_backward(self.plan.specs, self.coords, self.gout, self.grad_ptrs, self.F, self.linear,
                           levels_per_group=self.plan.bwd_levels_per_group, extra_flags=self.plan.extra_flags,
                           use_replicas=self.use_replicas, group_width=self.group_w)

    # ---- configs[4] pieces ---------------------------------------------------------------------
    def prologue(self):
        self.coords = self.conv.geodetic_to_xyzt_norm(self.geo, self.t)

    def head_forward(self):
        if self.head.wide:
            self.y, self.hid, self.stats = self.head.fused_wide(self.out, want_stats=True)
            return
        self.y = self.head.fused(self.out)

    def head_backward(self):
        """Gradients of the head: the tcgen05 backward kernel (default) or, for A/B, a recompute through the fp32
        modules (cuBLAS + eager LayerNorm / GELU), which is what autograd does with fused_backward=False."""
        if self.head_bwd == "tcgen05" and self.head.wide:
            g = self.head.backward_wide(self.out, self.hid, self.stats, self.gy)
            self.gout = g.pop("x")
            self.head_grads = g
            return
        if self.head_bwd == "tcgen05":
            g = self.head.backward_fused(self.out, self.gy, group_width=self.group_w)
            self.gout = g.pop("x")
            self.head_grads = g
            return
        with torch.enable_grad():
            xd = self.out.detach().requires_grad_(True)
            y = self.enc.projection(xd)
            grads = torch.autograd.grad(y, [xd] + self.head_params, self.gy)
        self.gout = grads[0]
        self.head_grads = grads[1:]

    def n_launches_per_step(self):
        # one launch per run of equal-D levels (<= 64) for forward and for backward
        runs, prev = 0, None
        cnt = 0
        for s in self.plan.specs:
            if len(s.dims) != prev or cnt == 64:
                runs += 1
                prev, cnt = len(s.dims), 0
            cnt += 1
        from deepearth_b200 import functional as Fn, _lib
        extra = 0
        # backward helpers: the regroup pass (unless the head wrote the group-major layout itself) and the replica fold
        if not self.group_w and Fn.backward_group_width(self.plan.specs, self.n, self.F, self.linear,
                                                        self.plan.bwd_levels_per_group, self.plan.extra_flags):
            extra += 1
        lin = _lib.DE_INTERP_LINEAR if self.linear else _lib.DE_INTERP_NEAREST
        lv = Fn.pack_levels(self.plan.specs, None, self.grad_ptrs)
        if self.use_replicas and _lib.load().de_grid4d_backward_workspace_bytes_ex(
                lv, len(self.plan.specs), self.F, self.n, lin | _lib.DE_FLAG_NO_REGROUP, self.plan.bwd_levels_per_group):
            extra += runs
        cast = 1 if self.plan.table_dtype == torch.float16 else 0            # de_tables_to_f16 refresh
        head = 0
        if self.head is not None:                 # prologue + head forward + head backward kernels
            head = 1 + (2 + 6 if self.head.wide else 1 + 1)
        return 2 * runs + extra + cast + head


# ---------------------------------------------------------------------------------------------------
# CPU baseline / reference arm.  The reference is PyTorch eager; oracle/grid4d_torch.py issues the same ATen ops
# (and is pinned to the reference's golden vectors by tests/test_torch_port_cpu.py).  At the reference's own
# geometry (c1r) the UNMODIFIED reference classes are used when build() could copy them to oracle/_ref.
# ---------------------------------------------------------------------------------------------------
def reference_step_factory(wl, n, device, threads_note=""):
    """-> (step(), kind, description): one fwd+bwd of the reference op chain on `device` over n uniform points."""
    from oracle.grid4d_torch import RefGrid4DFeatures, grid4d_forward
    g = torch.Generator(device=device).manual_seed(99)
    xyzt = torch.rand((n, 4), generator=g, device=device)
    res = [2 ** i for i in range(4, 20)]
    if wl["mode"] == "decomposed":
        T = wl["T"]
        dims_list = [(0, 1, 2), (0, 1, 3), (1, 2, 3), (0, 2, 3)]
        specs = [(d, R, T, (R + 1) ** 3 <= T) for d in dims_list for R in res]
        tabs = [torch.nn.Parameter((torch.rand((T, 2), generator=g, device=device) * 2e-4 - 1e-4)) for _ in specs]
        gout = torch.randn((n, 128), generator=g, device=device)

        def step():
            for t in tabs:
                t.grad = None
            out = grid4d_forward(xyzt, tabs, specs, linear=True)
            out.backward(gout)
        return step, "port", ("oracle/grid4d_torch.py: the reference's hash_coords op chain per corner + multilinear weights "
                              "(the reference holds no implementation of the decomposed mode), fp32 tables")
    from oracle.ref_loader import load_reference_models
    ref = load_reference_models()
    gout = torch.randn((n, 48), generator=g, device=device)
    if ref is not None:
        torch.manual_seed(1234)
        sp = ref.MultiResolutionHashEncoding(16, 2, res, wl["T"], 3).to(device)
        tp = ref.MultiResolutionHashEncoding(8, 2, [2 ** i for i in range(2, 10)], wl["T"] // 4, 1).to(device)
        params = list(sp.parameters()) + list(tp.parameters())

        def step():
            for q in params:
                q.grad = None
            torch.cat([sp(xyzt[:, :3]), tp(xyzt[:, 3:4])], dim=-1).backward(gout)      # models/encoders.py:68-82
        return step, "reference", "unmodified /root/reference/models/hash_encoding.py classes (oracle/_ref), fp32"
    model = RefGrid4DFeatures(hash_table_size=wl["T"]).to(device)

    def step():
        model.zero_grad(set_to_none=True)
        model(xyzt).backward(gout)
    return step, "port", "oracle/grid4d_torch.py RefGrid4DFeatures (op-by-op port of models/hash_encoding.py), fp32"


def cpu_baseline_run(wl, log2_sample, steps, warmup):
    n = 2 ** log2_sample
    torch.manual_seed(1234)
    threads = torch.get_num_threads()
    step, kind, what = reference_step_factory(wl, n, torch.device("cpu"))
    for _ in range(warmup):
        step()
    times = []
    for _ in range(steps):
        t0 = time.perf_counter()
        step()
        times.append(time.perf_counter() - t0)
    sec = float(np.mean(times))
    return {"value": n / sec / 1e6, "unit": UNIT, "cores": threads, "kind": kind,
            "sample": f"{steps} fwd+bwd steps of 2^{log2_sample} uniform points, same grids/levels/T as the GPU arm, "
                      f"PyTorch eager CPU ops, {threads} threads; {what}",
            "ms_per_step": sec * 1e3}


def eager_cuda_baseline(wl, n, device, steps=3, warmup=2):
    """The reference op chain through PyTorch-CUDA eager on the same GPU: the like-for-like comparator of `value`."""
    step, kind, what = reference_step_factory(wl, n, device)
    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    peak_gib = torch.cuda.max_memory_allocated(device) / 2 ** 30
    return {"value": n / (ms * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": ms, "points": n, "kind": kind,
            "what": what + "; fwd+bwd, CUDA events, same B200", "peak_mem_gib": round(peak_gib, 1)}


def encoder_default_ab(device, log2_points=20):
    """The whole module a user of the reference calls -- Grid4DEncoder(DeepEarthConfig()) with every default (reference
    geometry: 16 x 3-D + 8 x 1-D nearest-cell levels, projection at hidden_dim 768) -- forward + backward over 2^20 points: the
    UNMODIFIED reference classes (oracle/_ref) in PyTorch-CUDA eager against ours with the same state dict, in fp32 and under
    torch.autocast(bfloat16) (the reference's training mode, hpc/train_distrbuted.py:284)."""
    import deepearth_b200 as de
    from oracle.ref_loader import load_reference_models
    ref = load_reference_models()
    if ref is None:
        return {"unavailable": "oracle/_ref not built"}
    n = 2 ** log2_points
    torch.manual_seed(1234)
    enc_ref = ref.Grid4DEncoder(ref.DeepEarthConfig()).to(device)
    enc = de.Grid4DEncoder(de.DeepEarthConfig()).to(device)
    enc.load_state_dict(enc_ref.state_dict())
    g = torch.Generator(device=device).manual_seed(5)
    xyzt = torch.rand((n, 4), device=device, generator=g)
    gout = torch.randn((n, enc.projection[-1].out_features), device=device, generator=g)
    with torch.no_grad():
        a, b = enc_ref(xyzt[:65536]), enc(xyzt[:65536])
        rel = float((a - b).abs().max()) / float(a.abs().max())

    def step(m, amp):
        def run():
            m.zero_grad(set_to_none=True)
            if amp:
                with torch.autocast("cuda", dtype=torch.bfloat16):
                    y = m(xyzt)
                y.backward(gout.to(y.dtype))
            else:
                m(xyzt).backward(gout)
        return run
    out = {"points": n, "hidden_dim": int(enc.projection[0].out_features), "forward_max_rel_diff_vs_reference_fp32": rel}
    for tag, amp in (("fp32", False), ("autocast_bf16", True)):
        ms_ref = _time_cuda(step(enc_ref, amp), steps=3, warmup=1)
        ms_ours = _time_cuda(step(enc, amp), steps=5, warmup=2)
        out[tag] = {"reference_eager_ms": ms_ref, "ours_ms": ms_ours, "speedup": ms_ref / ms_ours,
                    "ours_mpoints_per_s": n / ms_ours / 1e3, "reference_mpoints_per_s": n / ms_ref / 1e3}
    out["what"] = ("Grid4DEncoder(DeepEarthConfig()) forward + backward, same state dict, same batch, same B200: unmodified reference "
                   "classes (oracle/_ref) in eager PyTorch vs deepearth_b200 (hash grids + the 768-wide projection on tcgen05 / TMA)")
    return out


def _time_cuda(fn, steps=5, warmup=2):
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


def field_bench(device, n_side=2048):
    """forward-only evaluation of the c2 grids over a regular n_side x n_side lat/lon grid (2^22 points):
    the patch-ordered field kernel (coordinates generated on the device) against the general forward over the same
    points (coordinate tensor from the prologue kernel) -- identical features, checked here on every run."""
    import deepearth_b200 as de
    from _synth import FLORIDA_BBOX_GEO as B, FLORIDA_T_RANGE
    torch.manual_seed(1234)
    enc = de.Grid4DEncoder(build_config(WORKLOADS["c2"])).to(device).eval()
    _, _, conv = make_geo(16, device, 1)
    lat_range, lon_range = (B["south"], B["north"]), (B["west"], B["east"])
    args = (conv, lat_range, lon_range, n_side, n_side)
    kw = dict(altitude=25.0, time=0.5 * (FLORIDA_T_RANGE[0] + FLORIDA_T_RANGE[1]))
    feats, xyzt = enc.encode_field(*args, return_coords=True, **kw)
    with torch.no_grad():
        same = bool(torch.equal(feats, enc.encode_features(xyzt)))
    del feats
    ms_field = _time_cuda(lambda: enc.encode_field(*args, **kw))
    with torch.no_grad():
        ms_general = _time_cuda(lambda: enc.encode_features(xyzt))
    n = n_side * n_side
    return {"points": n, "grid": [n_side, n_side], "field_forward_ms": ms_field, "general_forward_ms": ms_general,
            "speedup": ms_general / ms_field, "mpoints_per_s": n / ms_field / 1e3, "bit_identical": same, "gpu_launches": 14,
            "what": "Grid4DEncoder.encode_field (de_grid4d_field_forward: 16 x 16-point patches per CTA, a warp = 16 neighbouring "
                    "points of one level) vs Grid4DEncoder.encode_features over the same 2^22 points; c2 grids, fp16 tables, no_grad"}


def _st_embed_case(device, n_tokens, dim, peak):
    """One size of the f3 measurement: our module (tensor-core temporal MLP, and the fp32 SIMT path for A/B) beside the
    reference's forward op by op through ATen on the same GPU (models/cross_modal_fusion.py:130-164: index gathers, cat,
    nn.Sequential with a cuBLAS fp32 SGEMM, adds)."""
    import deepearth_b200 as de
    torch.manual_seed(7)
    m = de.SpatialTemporalEmbedding(dim, check_indices=False).to(device)
    m.add_modality("vision")
    g = torch.Generator(device=device).manual_seed(3)
    tok = torch.randn((n_tokens // 256, 256, dim), device=device, generator=g, requires_grad=True)
    pos = torch.rand((n_tokens // 256, 256, 2), device=device, generator=g) * 0.999
    t = torch.rand((n_tokens // 256, 256, 1), device=device, generator=g)
    gout = torch.randn_like(tok)

    def eager_fwd():
        xi, yi = (pos[..., 0] * 64).long(), (pos[..., 1] * 64).long()
        emb = torch.zeros_like(tok)
        emb[..., :dim // 2] += torch.cat([m.spatial_embed_x[0][xi], m.spatial_embed_y[0][yi]], -1)
        emb[..., dim // 2:] += m.temporal_embed(t)
        emb = emb + m.modality_embeds["vision"].expand(tok.shape[0], tok.shape[1], -1)
        return tok + emb

    def step(fwd):
        def run():
            tok.grad = None
            m.zero_grad(set_to_none=True)
            fwd().backward(gout)
        return run
    ours = lambda: m(tok, pos, t, "vision")        # noqa: E731
    out = {"tokens": n_tokens, "dim": dim}
    with torch.no_grad():
        out["forward_ms"] = _time_cuda(ours)
        out["forward_gather_only_ms"] = _time_cuda(lambda: m(tok, pos, None, "vision"))
        out["eager_aten_forward_ms"] = _time_cuda(eager_fwd, steps=3, warmup=1)
    out["fwd_bwd_ms"] = _time_cuda(step(ours))
    out["eager_aten_fwd_bwd_ms"] = _time_cuda(step(eager_fwd), steps=3, warmup=1)
    m.tensor_core_mlp = False
    with torch.no_grad():
        out["simt_fp32_forward_ms"] = _time_cuda(ours, steps=3, warmup=1)
    out["simt_fp32_fwd_bwd_ms"] = _time_cuda(step(ours), steps=3, warmup=1)
    m.tensor_core_mlp = True
    stream_bytes = n_tokens * dim * 4 * 2
    mlp_ms = max(out["forward_ms"] - out["forward_gather_only_ms"], 1e-3)
    out.update({"gather_only_gbs": stream_bytes / out["forward_gather_only_ms"] / 1e6,
                "gather_only_frac_of_hbm_peak": stream_bytes / out["forward_gather_only_ms"] / 1e6 / peak,
                "forward_gbs": stream_bytes / out["forward_ms"] / 1e6, "forward_frac_of_hbm_peak": stream_bytes / out["forward_ms"] / 1e6 / peak,
                "temporal_mlp_fp32_equivalent_gflops": 2.0 * n_tokens * (dim // 2) ** 2 / (mlp_ms * 1e6),
                "speedup_vs_eager_aten_forward": out["eager_aten_forward_ms"] / out["forward_ms"],
                "speedup_vs_eager_aten_fwd_bwd": out["eager_aten_fwd_bwd_ms"] / out["fwd_bwd_ms"]})
    return out


def st_embed_bench(device, n_tokens=1 << 16, dim=512):
    """SpatialTemporalEmbedding forward + backward; achieved bytes/s of the streamed tensors (tokens in, out
    written) against the HBM copy peak, and the same module through ATen on the same GPU.  Two sizes: the round-1 / round-2
    line (65536 tokens x 512) and the integrated model's width (universal_dim 2048, 2^18 tokens), where the temporal MLP
    dominates."""
    peak, _ = measured_peaks()
    out = _st_embed_case(device, n_tokens, dim, peak)
    out["d2048"] = _st_embed_case(device, 1 << 18, 2048, peak)
    out["gpu_launches"] = 2 * 7 + 7 * 7
    out["what"] = ("deepearth_b200.SpatialTemporalEmbedding (de_st_embedding_forward_tc / _backward_tc): table gathers + modality + "
                   "residual in one streaming pass (tokens read + out written = the streamed bytes); the temporal MLP's second Linear "
                   "as a split-bf16 (3 parts, 6 products) TMA + tcgen05 GEMM at fp32 accuracy; simt_fp32_* = the fp32 tile-GEMM path "
                   "it replaced; eager_aten_* = the reference's forward op by op (cuBLAS SGEMM) on the same B200")
    return out


def projection_bench(device, log2_points=20, in_dim=128, hidden=768):
    """Grid4DEncoder.projection at the reference's default width (models/configs.py:44) on the tcgen05 kernels,
    forward + backward over 2^20 feature rows, beside the same nn.Sequential through cuBLAS (fp32, and bf16 autocast).
    The GEMM fraction is the second Linear alone (2 n H^2 flop) against the measured dense bf16 peak."""
    import torch.nn as nn
    from deepearth_b200.mlp_head import FusedProjection, gemm_bf16
    n = 2 ** log2_points
    torch.manual_seed(11)
    seq = nn.Sequential(nn.Linear(in_dim, hidden), nn.LayerNorm(hidden), nn.GELU(), nn.Linear(hidden, hidden)).to(device)
    g = torch.Generator(device=device).manual_seed(5)
    x = torch.randn((n, in_dim), device=device, generator=g).requires_grad_(True)
    gy = torch.randn((n, hidden), device=device, generator=g)
    fp = FusedProjection(seq)
    assert fp.wide and fp.supported()

    def ours():
        x.grad = None
        seq.zero_grad(set_to_none=True)
        fp(x).backward(gy)

    def torch_fp32():
        x.grad = None
        seq.zero_grad(set_to_none=True)
        seq(x).backward(gy)

    def torch_amp():
        x.grad = None
        seq.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            y = seq(x)
        y.backward(gy.to(y.dtype))
    def ours_amp():                     # the same call sequence as torch_amp: bf16 y, its gradient cast to bf16 by the caller
        x.grad = None
        seq.zero_grad(set_to_none=True)
        y = fp(x, out_dtype=torch.bfloat16)
        y.backward(gy.to(y.dtype))
    with torch.no_grad():
        ms_fwd = _time_cuda(lambda: fp(x))
        ms_fwd16 = _time_cuda(lambda: fp(x, out_dtype=torch.bfloat16))
    ms_step = _time_cuda(ours)
    ms_step16 = _time_cuda(ours_amp)
    hid = torch.randn((n, hidden), device=device, generator=g).to(torch.bfloat16)
    w2 = seq[3].weight.detach().to(torch.bfloat16)
    out = torch.empty((n, hidden), device=device)
    ms_gemm = _time_cuda(lambda: gemm_bf16(hid, w2, seq[3].bias, out=out))
    out16 = torch.empty((n, hidden), device=device, dtype=torch.bfloat16)
    ms_gemm16 = _time_cuda(lambda: gemm_bf16(hid, w2, seq[3].bias, out=out16))
    del hid, out, out16
    ms_t32 = _time_cuda(torch_fp32, steps=3, warmup=1)
    ms_amp = _time_cuda(torch_amp, steps=3, warmup=1)
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except OSError:
        pass
    burst, sustained = float(peaks.get("bf16_tflops", 1590.0)), float(peaks.get("bf16_tflops_sustained", 1400.0))
    tf = 2.0 * n * hidden * hidden / (ms_gemm * 1e9)
    tf16 = 2.0 * n * hidden * hidden / (ms_gemm16 * 1e9)
    hbm, _ = measured_peaks()
    gemm_bytes = n * hidden * (2 + 4)
    return {"points": n, "in_dim": in_dim, "hidden": hidden, "forward_ms": ms_fwd, "fwd_bwd_ms": ms_step,
            "mpoints_per_s_fwd_bwd": n / ms_step / 1e3,
            "bf16_io": {"forward_ms": ms_fwd16, "fwd_bwd_ms": ms_step16, "speedup_vs_cublas_bf16_autocast": ms_amp / ms_step16,
                        "what": "bf16 y out of the GEMM epilogue and bf16 grad_y consumed in place (de_projection_forward_ex / "
                                "_backward_ex): what Grid4DEncoder does under torch.autocast(bfloat16), the reference's training mode"},
            "gemm2": {"ms_fp32_out": ms_gemm, "tflops_fp32_out": tf, "frac_of_bf16_burst": tf / burst, "frac_of_bf16_sustained": tf / sustained,
                      "hbm_gbs_fp32_out": gemm_bytes / ms_gemm / 1e6, "hbm_frac_fp32_out": gemm_bytes / ms_gemm / 1e6 / hbm,
                      "ms_bf16_out": ms_gemm16, "tflops_bf16_out": tf16, "frac_of_bf16_burst_bf16_out": tf16 / burst,
                      "peak_source": "MEASURED_PEAKS.json bf16_tflops / bf16_tflops_sustained" if peaks else "fallback 1590 / 1400",
                      "what": "y (n, H) = a (n, H) bf16 . W2^T + b2: persistent TMA + tcgen05 kernel, CTA pairs (256 x 256 tiles); with fp32 output the "
                              "kernel writes 4 B per 2 H flop (H = 768: 3 KB per 1.18 MFLOP), so HBM binds before the tensor cores do"},
            "torch_cublas_fp32_fwd_bwd_ms": ms_t32, "torch_cublas_bf16_autocast_fwd_bwd_ms": ms_amp,
            "speedup_vs_cublas_fp32": ms_t32 / ms_step, "speedup_vs_cublas_bf16_autocast": ms_amp / ms_step,
            "gpu_launches": 7 * 2 + 7 * 8 + 14,
            "what": "FusedProjection (proj_hidden_kernel + gemm_bf16_tma_kernel forward; cast, da GEMM, proj_dpre_kernel, dW2 / dW1 "
                    "split-K GEMMs, dx GEMM backward) vs the identical nn.Sequential through cuBLAS / ATen"}


def small_batch_bench(device, log2_points=12):
    """The reference's own batch sizes (a few thousand points, tests/run_tests.py:265-324): one c2 training step at 4096
    points through the module API -- eager (host-bound: Python, ctypes and autograd bookkeeping for 64 table parameters),
    with forward + backward replayed as CUDA graphs (Grid4DEncoder.graphed_features), and both again with one flat table
    parameter per grid (config.grid4d_flat_tables: 4 parameters instead of 64, reference state-dict keys kept)."""
    import deepearth_b200 as de
    n = 2 ** log2_points
    out = {"points": n}
    for tag, flat in (("", False), ("_flat", True)):
        wl = dict(WORKLOADS["c2"])
        cfg = build_config(wl)
        cfg.grid4d_flat_tables = flat
        torch.manual_seed(1234)
        enc = de.Grid4DEncoder(cfg).to(device).train()
        params = [p for e in enc._encoders for p in e.autograd_tensors()]
        coords, gout = Job.inputs(wl, n, device, 1234, enc.feature_dim)

        def eager():
            for p in params:
                p.grad = None
            enc.encode_features(coords).backward(gout)
        out["eager" + tag + "_ms_per_step"] = _time_cuda(eager, steps=200, warmup=20)
        graphed = enc.graphed_features(coords)

        def replay():
            for p in params:
                p.grad = None
            graphed(coords).backward(gout)
        out["graphed" + tag + "_ms_per_step"] = _time_cuda(replay, steps=200, warmup=20)
        del enc, graphed, params
        torch.cuda.empty_cache()
    best = min(v for k, v in out.items() if k.endswith("_ms_per_step"))
    out.update({"best_ms_per_step": best, "speedup_vs_eager": out["eager_ms_per_step"] / best, "best_mpoints_per_s": n / best / 1e3,
                "gpu_launches": 4 * 4 * 220,
                "what": "c2 grids, 4096 uniform points, Grid4DEncoder.encode_features + autograd backward, CUDA events over 200 steps; "
                        "graphed = torch.cuda.make_graphed_callables over the same four kernels (fp16 table refresh, forward, gradient "
                        "zeroing, backward); _flat = one (16, T, 2) parameter per grid"})
    return out


def cpu_sample_log2(wl, steps, warmup):
    """Bounded CPU sample: the decomposed mode runs ~0.03 Mpoints/s on 16 host threads, so 2^17 points x 13 steps is
    about a minute; more steps shrink the sample."""
    if wl["mode"] != "decomposed":
        return 18
    log2 = 17 if steps + warmup <= 16 else 16
    if wl["T"] > 2 ** 20:
        log2 -= 1
    return log2


L2_POLICY = "inputs larger than L2 (coords + grad_out + out > 4 GiB per step at the workload's size); no flush"


def describe_config(name, wl, n_points, world=1):
    """The workload, identically worded by both arms (the driver compares the two `config` objects); what is specific to a
    run -- the reference arm's bounded sample, the GPU arm's L2 limits and overlap -- lives outside it."""
    return {"workload": name, "l2_policy": L2_POLICY, "description": wl["desc"], "points_per_gpu": n_points, "coords": wl["coords"],
            "levels": 64 if wl["mode"] == "decomposed" else 24, "features_per_level": 2, "table_size": wl["T"],
            "interpolation": "linear" if wl["mode"] == "decomposed" else "nearest",
            "parallelism": f"points-sharded x{world}, replicated tables, NCCL all-reduce of table grads"}


def run_reference_arm(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    log2_sample = cpu_sample_log2(wl, max(1, args.steps), max(1, args.warmup))
    cb = cpu_baseline_run(wl, log2_sample, max(1, args.steps), max(1, args.warmup))
    cfg = describe_config(args.workload, wl, 2 ** wl["log2_points"], args.gpus)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": cb["ms_per_step"], "higher_is_better": True,
            "scaling": "strong" if wl.get("total") else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": cfg,
            "cpu_baseline": {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------
class Ctx:
    """Process-wide state of the GPU arm: ranks, device, clock sampler."""

    def __init__(self):
        import torch.distributed as dist
        self.dist = dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device; the GPU arm has no CPU fallback (use --impl reference for the CPU port)")
        torch.cuda.set_device(self.local)
        self.device = torch.device("cuda", self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.device)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(self, v: float) -> float:
        if self.world == 1:
            return v
        t = torch.tensor([v], device=self.device, dtype=torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t)


def measure(args, name, wl, ctx, steps, want_clocks=True):
    """Times one workload on this rank (device-resident `value` leg, then the `e2e` leg); rank 0 gets the result dict."""
    dist, world, rank, device = ctx.dist, ctx.world, ctx.rank, ctx.device
    from deepearth_b200 import functional as Fn
    from deepearth_b200 import _lib as _L
    import ctypes as _ct
    _pl2, _pl2max = _ct.c_longlong(0), _ct.c_longlong(0)
    _L.load().de_get_persisting_l2(_ct.byref(_pl2), _ct.byref(_pl2max))
    _gran = _ct.c_int(0)
    _L.load().de_get_l2_fetch_granularity(_ct.byref(_gran))

    n = 2 ** (args.log2_points if args.log2_points else wl["log2_points"])
    if wl.get("total"):
        n //= world                      # strong scaling: the workload fixes the job's points, not the rank's
    job = Job(wl, device, seed=1234 + rank, n_points=n)
    if args.levels_per_group:
        job.plan.levels_per_group = args.levels_per_group
    if args.bwd_levels_per_group:
        job.plan.bwd_levels_per_group = args.bwd_levels_per_group
    job.plan.extra_flags = args.flags
    job.use_replicas = not args.no_replicas
    job.head_bwd = args.head_bwd
    if job.head is not None:
        job.head.fused_backward = args.head_bwd == "tcgen05"
        if args.head_bwd == "tcgen05" and not args.no_grouped_head and not job.head.wide:
            job.group_w = Fn.backward_group_width(job.plan.specs, n, job.F, job.linear, job.plan.bwd_levels_per_group,
                                                  job.plan.extra_flags)

    sync = None
    if world > 1 and not args.no_overlap:
        from deepearth_b200.distributed import GradSync
        sync = GradSync(average=False, chunks=args.sync_chunks)
        job.enc.enable_grad_sync(average=True, chunks=args.sync_chunks)     # the e2e (autograd) path
        if job.group_w and not sync.grouped_ok(job.plan.specs, n, job.F, job.linear, job.plan.bwd_levels_per_group,
                                               job.plan.extra_flags, job.group_w):
            job.group_w = 0

    def step(ev=None):
        job.flat_grad.zero_()
        job.refresh_tables()             # fp32 masters -> fp16 working tables: what a training step does after the optimiser
        if ev:
            ev[0].record()
        if job.head is not None:
            job.prologue()
        job.forward()
        if job.head is not None:
            if ev:
                ev[4].record()
            job.head_forward()
            if ev:
                ev[5].record()
            job.head_backward()
        if ev:
            ev[1].record()
        if sync is not None:
            # backward + all-reduce overlapped (distributed.GradSync): kernels_ms.allreduce is the exposed tail only
            sync.backward(job.plan.specs, job.coords, job.gout, job.flat_grad, job.plan.grad_offsets, job.F, job.linear,
                          levels_per_group=job.plan.bwd_levels_per_group, extra_flags=job.plan.extra_flags,
                          group_width=job.group_w)
            if ev:
                ev[2].record()
                ev[3].record()
        else:
            job.backward()
            if ev:
                ev[2].record()
            if world > 1:
                dist.all_reduce(job.flat_grad)
                if ev:
                    ev[3].record()

    sampler = ClockSampler(ctx.local)
    if rank == 0 and want_clocks:
        sampler.start()                    # nvidia-smi takes ~0.5 s to deliver its first sample
    for _ in range(max(args.warmup, 3)):
        step()
    ctx.barrier()
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(6)] for _ in range(steps)]
    t_start, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ctx.barrier()
    w0 = sampler.mark()
    t_start.record()
    for k in range(steps):
        step(evs[k])
    t_end.record()
    ctx.barrier()
    w1 = sampler.mark()
    clock_window = "timed region"
    # the timed region may be shorter than a few nvidia-smi periods: then every rank keeps the same load
    # running (untimed) until rank 0 has enough samples, and the JSON line says so
    extra = 0
    if rank == 0 and sampler.proc is not None and sampler.n_between(w0, w1) < 4:
        extra = max(1, int(1500.0 / max(t_start.elapsed_time(t_end) / steps, 0.05)))
    if world > 1:
        ex = torch.tensor([extra], device=device, dtype=torch.int64)
        dist.broadcast(ex, 0)
        extra = int(ex)
    if extra:
        clock_window = "timed region + identical untimed steps run straight after it"
        for _ in range(extra):
            step()
        ctx.barrier()
        w1 = sampler.mark()
    clocks = sampler.stop(w0, w1) if (rank == 0 and want_clocks) else None
    if clocks is not None:
        clocks["window"] = clock_window
    total_ms = t_start.elapsed_time(t_end)
    head_ms = None
    if job.head is not None:
        fwd_ms = float(np.mean([e[0].elapsed_time(e[4]) for e in evs]))        # prologue + encode
        head_ms = {"head_forward_tcgen05": float(np.mean([e[4].elapsed_time(e[5]) for e in evs])),
                   f"head_backward_{args.head_bwd}": float(np.mean([e[5].elapsed_time(e[1]) for e in evs]))}
    else:
        fwd_ms = float(np.mean([e[0].elapsed_time(e[1]) for e in evs]))
    fwd_enc_ms = fwd_ms          # c5: includes the ~0.15 ms prologue
    bwd_ms = float(np.mean([e[1].elapsed_time(e[2]) for e in evs]))
    ar_ms = float(np.mean([e[2].elapsed_time(e[3]) for e in evs])) if world > 1 else 0.0
    total_ms = ctx.max_over_ranks(total_ms)
    ms_per_step = total_ms / steps
    value = world * n / (ms_per_step * 1e-3) / 1e6

    # ---- end to end through the public nn.Module API, pinned-host coordinates in, scalar out --------
    if job.head is not None:
        host_geo, host_t = job.geo.cpu().pin_memory(), job.t.cpu().pin_memory()
        h2d_bytes = host_geo.numel() * 8 + host_t.numel() * 8
        params = job.weights + job.head_params
        api = "GeospatialConverter.geodetic_to_xyzt_norm + Grid4DEncoder.forward (tcgen05 head) + autograd backward; pinned host geodetic points -> device every step (copy of step i+1 overlaps step i); output checksum reduced + read back on a side stream during the backward"
        job.gout = job.head_grads = job.y = job.hid = job.stats = None
    else:
        host = job.coords.cpu().pin_memory()
        h2d_bytes = host.numel() * 4
        params = job.weights
        api = "Grid4DEncoder.encode_features + autograd backward; pinned host xyzt -> device every step (copy of step i+1 overlaps step i); output checksum reduced + read back on a side stream during the backward"
    del job.out      # the module allocates its own output
    torch.cuda.empty_cache()

    # Input pipeline of the e2e loop: every step's inputs come from pinned host memory; the copy for step i+1 is
    # issued on a copy stream at the start of step i (what a pinned-memory DataLoader with non_blocking copies
    # does), so K timed steps issue K host->device copies inside the timed region.
    copy_stream = torch.cuda.Stream(device=device)

    def fetch():
        with torch.cuda.stream(copy_stream):
            if job.head is not None:
                bufs = (host_geo.to(device, non_blocking=True), host_t.to(device, non_blocking=True))
            else:
                bufs = (host.to(device, non_blocking=True),)
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        return bufs, ev

    pending = [fetch()]
    # The step's result metric (fp64 sum of the forward output) is reduced and read back on a side stream as soon as
    # the forward has finished, while the backward runs: what a training loop that logs a forward metric does.
    metric_stream = torch.cuda.Stream(device=device)
    metric_host = torch.zeros((), dtype=torch.float64).pin_memory()

    def e2e_step():
        for p in params:
            p.grad = None
        bufs, ev = pending.pop()
        pending.append(fetch())                      # next step's inputs travel while this step computes
        main = torch.cuda.current_stream(device)
        main.wait_event(ev)
        for b in bufs:
            b.record_stream(main)
        if job.head is not None:
            xyzt = job.conv.geodetic_to_xyzt_norm(bufs[0], bufs[1])
            res = job.enc(xyzt)
            gres = job.gy
        else:
            res = job.enc.encode_features(bufs[0])
            gres = job.gout
        fwd_done = torch.cuda.Event()
        fwd_done.record(main)
        with torch.cuda.stream(metric_stream):
            metric_stream.wait_event(fwd_done)
            out_d = res.detach()
            out_d.record_stream(metric_stream)
            metric_host.copy_(Fn.sum_f32(out_d), non_blocking=True)
            metric_done = torch.cuda.Event()
            metric_done.record(metric_stream)
        res.backward(gres)
        if world > 1 and sync is None:
            from deepearth_b200.distributed import allreduce_table_grads
            allreduce_table_grads(job.weights, average=True)
        metric_done.synchronize()         # device -> host read of the step's result
        return float(metric_host)

    for _ in range(3):
        e2e_step()
    ctx.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        checksum = e2e_step()
    e1.record()
    ctx.barrier()
    e2e_ms = ctx.max_over_ranks(e0.elapsed_time(e1))
    e2e_value = world * n / (e2e_ms / steps * 1e-3) / 1e6
    launches = job.n_launches_per_step() * steps
    result = None
    if rank == 0:
        peak, peak_src = measured_peaks()
        pk = packet_peaks()
        dom = "backward" if bwd_ms >= fwd_ms else "forward"
        dom_ms = max(bwd_ms, fwd_ms)
        dom_bytes = (job.bytes_bwd if dom == "backward" else job.bytes_fwd) * n
        achieved = dom_bytes / (dom_ms * 1e-3) / 1e9
        traffic, traffic_src = None, None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            with open(tpath) as f:
                tfile = json.load(f)
            tj = tfile.get(name, {})
            traffic, traffic_src = tj.get(dom), (tfile.get("_source") if tj.get(dom) is not None else None)
        # where the dominant kernel's working set lives: one level group's tables (forward) / gradient tables (backward)
        lvl_bytes = wl["T"] * job.F * (4 if dom == "backward" else (2 if wl["table_dtype"] == "float16" else 4))
        resident = "l2" if lvl_bytes * 16 <= 64 * 2 ** 20 or wl["mode"] != "decomposed" else "hbm"
        cfg = describe_config(name, wl, n, world)
        run_detail = {"l2_fetch_granularity": _gran.value, "head_grad_group_width": job.group_w,
                      "l2_persisting_bytes": _pl2.value, "l2_persisting_max_bytes": _pl2max.value,
                      "allreduce": (f"overlapped with the backward in {args.sync_chunks} level runs" if sync is not None
                                    else ("one collective after the backward" if world > 1 else "none")),
                      "step": "zero grad buffer + fp32->fp16 table refresh + forward + backward" + (" + all-reduce" if world > 1 else "")}
        result = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong" if wl.get("total") else "weak",
            "vs_baseline": None, "dtype": "f16 tables / f32 accumulate" if wl["table_dtype"] == "float16" else "f32",
            "data": "synthetic",
            "config": cfg,
            "run": run_detail,
            "kernels_ms": {"forward": fwd_ms, "backward": bwd_ms, "allreduce": ar_ms, **(head_ms or {})},
            "roofline": {"bound": "hbm", "kernel": f"grid4d_{dom}_kernel", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": peak_src,
                         "working_set": ("one level group's tables stay L2-resident: the unit that binds is the L1->L2 request "
                                         "rate (roofline_packets), HBM carries only the streamed rows" if resident == "l2" else
                                         "tables exceed L2 (HBM-resident): gathers / reductions miss to HBM per level group"),
                         "working_set_level": resident,
                         "algorithmic_bytes_per_point": {"forward": job.bytes_fwd, "backward": job.bytes_bwd}},
            # the bound that actually applies while a level group's tables sit in L2: distinct 32 B
            # sectors requested per second against the probe-measured L1->L2 request rate of the same access shape
            "roofline_packets": {
                "bound": "l2_request_rate", "unit": "Gpackets/s",
                "forward": {"packets_per_point": job.packets_fwd, "achieved": job.packets_fwd * n / (fwd_enc_ms * 1e-3) / 1e9,
                            "peak": pk["forward"],
                            "frac": job.packets_fwd * n / (fwd_enc_ms * 1e-3) / 1e9 / pk["forward"]},
                "backward": {"packets_per_point": job.packets_bwd, "achieved": job.packets_bwd * n / (bwd_ms * 1e-3) / 1e9,
                             "peak": pk["backward"],
                             "frac": job.packets_bwd * n / (bwd_ms * 1e-3) / 1e9 / pk["backward"]},
                "peak_source": PEAKS_FILE + " (scripts/peaks_gather.py on B200; best L2-resident rate, 2 adjacent lanes)"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d_bytes * world),
                    "d2h_bytes_per_step": 8 * world, "ms_per_step": e2e_ms / steps, "checksum": checksum,
                    "api": api},
            "gpu_launches": launches,
            "clocks": clocks,
        }
    # release this workload's device memory before the next one
    job.enc.disable_grad_sync()
    del job, pending, params
    import gc
    gc.collect()
    torch.cuda.empty_cache()
    return result


def grad_check(args, ctx, log2_points=16):
    """N > 1: the all-reduced (mean) table gradient of a small deterministic batch, computed through the overlapped
    GradSync path bench.py times, against the gradient of the UNION of all ranks' points computed locally by one plain
    de_grid4d_backward per rank's batch (rank r's points and gradient rows are rebuilt from seeds 1234 + r)."""
    from deepearth_b200 import functional as Fn
    from deepearth_b200.distributed import GradSync
    wl = dict(WORKLOADS["c2"])
    n = 2 ** log2_points
    job = Job(wl, ctx.device, seed=1234 + ctx.rank, n_points=n)
    sync = GradSync(average=True, chunks=args.sync_chunks)
    flat = torch.zeros_like(job.flat_grad)
    sync.backward(job.plan.specs, job.coords, job.gout, flat, job.plan.grad_offsets, job.F, job.linear)
    ref = torch.zeros_like(job.flat_grad)
    ptrs = [ref.data_ptr() + 4 * o for o in job.plan.grad_offsets]
    for r in range(ctx.world):
        other = Job.inputs(wl, n, ctx.device, 1234 + r, job.plan.out_cols)
        Fn.encode_backward(job.plan.specs, other[0], other[1], ptrs, job.F, job.linear)
    ref.div_(ctx.world)
    torch.cuda.synchronize()
    err = float((flat - ref).abs().max()) / max(float(ref.abs().max()), 1e-30)
    err = ctx.max_over_ranks(err)
    del job
    torch.cuda.empty_cache()
    return {"max_rel_err": err, "tolerance": 1e-4, "ok": bool(err <= 1e-4), "points_per_rank": n, "ranks": ctx.world,
            "what": "GradSync.backward (mean over ranks, overlapped level runs) vs plain backward over the union of all "
                    "ranks' points / world, max over ranks of max|diff| / max|ref| on the flat 256 MiB gradient buffer"}


EXTRA_NAMES = ("c3", "c1r", "field", "st_embed", "projection_h768", "small_batch", "eager_cuda_baseline")


def run_extra_child(name, steps, limit_s):
    """One side measurement in a child process (`bench.py --extra-only NAME`), JSON on its last stdout line."""
    cmd = [sys.executable, os.path.abspath(__file__), "--extra-only", name, "--steps", str(steps)]
    env = dict(os.environ)
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE"):
        env.pop(k, None)
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=limit_s, env=env)
    except subprocess.TimeoutExpired:
        return {"error": f"timed out after {limit_s} s"}
    lines = [ln for ln in r.stdout.strip().splitlines() if ln.startswith("{")]
    if r.returncode != 0 or not lines:
        return {"error": f"exit code {r.returncode}: " + (r.stderr.strip().splitlines() or ["no output"])[-1][:300]}
    try:
        return json.loads(lines[-1])
    except ValueError as e:
        return {"error": f"unparsable output: {e}"}


def run_extra_only(args):
    """Child side of run_extra_child."""
    ctx = Ctx()
    name = args.extra_only
    if name in ("c3", "c1r"):
        r = measure(args, name, dict(WORKLOADS[name]), ctx, max(2, args.steps))
        out = {k: r[k] for k in ("value", "unit", "steps", "ms_per_step", "config", "kernels_ms", "roofline", "roofline_packets",
                                 "e2e", "gpu_launches", "clocks")}
    elif name == "field":
        out = field_bench(ctx.device)
    elif name == "st_embed":
        out = st_embed_bench(ctx.device)
    elif name == "projection_h768":
        out = projection_bench(ctx.device)
    elif name == "small_batch":
        out = small_batch_bench(ctx.device)
    elif name == "eager_cuda_baseline":
        out = {"c2": eager_cuda_baseline(dict(WORKLOADS["c2"]), 2 ** 22, ctx.device)}
        torch.cuda.empty_cache()
        out["c1r"] = eager_cuda_baseline(dict(WORKLOADS["c1r"]), 2 ** 22, ctx.device)
        torch.cuda.empty_cache()
        out["encoder_default"] = encoder_default_ab(ctx.device)
    else:
        raise SystemExit(f"unknown extra {name!r}")
    print(json.dumps(out), flush=True)


def run_gpu_arm(args, wl):
    ctx = Ctx()
    from deepearth_b200 import _lib as _L
    if args.l2_fetch:
        _L.check(_L.load().de_set_l2_fetch_granularity(args.l2_fetch), "de_set_l2_fetch_granularity")
    if args.persist_l2_mb:
        _L.check(_L.load().de_set_persisting_l2(-1 if args.persist_l2_mb < 0 else args.persist_l2_mb << 20), "de_set_persisting_l2")
    extras_on = args.extras == "all" or (args.extras == "auto" and not args.workload_given)
    line = measure(args, args.workload, wl, ctx, args.steps)
    extra = {}
    if extras_on:
        side_steps = max(2, min(args.steps, 5))
        if ctx.world == 1:
            # Each side measurement runs in its own process with a time limit: a fault or a hang in one of them costs that
            # entry (recorded as {"error": ...}), never the headline line.  This process keeps only its CUDA context.
            torch.cuda.empty_cache()
            for nm, limit in (("c3", 240), ("c1r", 120), ("field", 120), ("st_embed", 120), ("projection_h768", 180),
                              ("small_batch", 120), ("eager_cuda_baseline", 240)):
                extra[nm] = run_extra_child(nm, side_steps, limit)
        else:
            r = measure(args, "c4", dict(WORKLOADS["c4"]), ctx, side_steps, want_clocks=False)
            if ctx.rank == 0:
                extra["c4"] = {k: r[k] for k in ("value", "unit", "scaling", "steps", "ms_per_step", "config", "kernels_ms", "e2e")}
            gc_res = grad_check(args, ctx)
            extra["grad_check"] = gc_res
    if ctx.rank == 0:
        if extra:
            line["extra"] = extra
            line["gpu_launches"] += sum(v.get("gpu_launches", 0) for v in extra.values() if isinstance(v, dict))
        if ctx.world == 1 and not args.no_cpu_baseline:
            cb = cpu_baseline_run(wl, cpu_sample_log2(wl, 3, 1) - 1, 3, 1)
            line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample")}
        print(json.dumps(line), flush=True)
    if ctx.world > 1:
        ctx.dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS))
    ap.add_argument("--extras", default="auto", choices=["auto", "all", "none"],
                    help="side measurements in the same line: auto = on for the default workload, off for an explicit --workload")
    ap.add_argument("--extra-only", default=None, choices=EXTRA_NAMES, help=argparse.SUPPRESS)
    ap.add_argument("--log2-points", type=int, default=0, help="override points per GPU (power of two)")
    ap.add_argument("--levels-per-group", type=int, default=0)
    ap.add_argument("--bwd-levels-per-group", type=int, default=0)
    ap.add_argument("--coords", default="", choices=["", "uniform", "florida"], help="override the workload's coordinate distribution")
    ap.add_argument("--no-overlap", action="store_true", help="N>1: one all-reduce after the backward instead of the overlapped slices")
    ap.add_argument("--sync-chunks", type=int, default=4, help="N>1: level runs the backward / all-reduce overlap is cut into")
    ap.add_argument("--head-bwd", default="tcgen05", choices=["tcgen05", "cublas"], help="c5: head backward kernel")
    ap.add_argument("--no-grouped-head", action="store_true", help="c5 A/B: head backward writes row-major gradients (the hash-grid backward regroups them)")
    ap.add_argument("--sort-points", action="store_true", help="experiment: Morton-order the synthetic points before timing")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-replicas", action="store_true", help="backward without gradient replicas for contended coarse levels (A/B)")
    ap.add_argument("--flags", type=lambda v: int(v, 0), default=0, help="DE_FLAG_* opt-in paths (0x100 TMA-staged backward, 0x200 shared-memory coarse forward)")
    ap.add_argument("--persist-l2-mb", type=int, default=0, help="cudaLimitPersistingL2CacheSize in MiB (-1 = device maximum)")
    ap.add_argument("--l2-fetch", type=int, default=0, help="set cudaLimitMaxL2FetchGranularity (32/64/128)")
    args = ap.parse_args()
    args.workload_given = args.workload is not None
    if args.workload is None:
        args.workload = "c2"
    wl = dict(WORKLOADS[args.workload])
    if args.coords:
        wl["coords"] = args.coords
        wl["desc"] += f" [coords overridden: {args.coords}]"
    if args.sort_points:
        wl["sort_points"] = True
        wl["desc"] += " [points Morton-sorted before timing]"
    if args.extra_only:
        run_extra_only(args)
    elif args.impl == "reference":
        run_reference_arm(args, wl)
    else:
        run_gpu_arm(args, wl)


if __name__ == "__main__":
    main()
