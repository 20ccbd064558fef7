"""Grid4DEncoder -- B200 drop-in for /root/reference/models/encoders.py:12-85.

Reference mode (default): ``spatial_encoder`` (xyz, D=3) + ``temporal_encoder`` (t, D=1, T//4), optional
row masks that multiply the *features* (:73-79), concat to (B, 48) and the Linear-LayerNorm-GELU-Linear
``projection`` (:47-52).  Same attribute names, parameter names and state-dict keys; the feature half
is ONE fused launch over all 24 levels writing straight into the concatenated (B, 48) buffer.

Decomposed mode (``config.grid4d_mode = "decomposed"``; SPECIFICATIONS.MD:1068-1200): four 3-D grids
xyz / xyt / yzt / xzt with multilinear interpolation, dense-index fallback on coarse levels and optional
fp16 working tables; masks zero the *coordinates* as the spec does (:1159-1164).
"""
from __future__ import annotations

import copy

from typing import List, Optional

import torch
import torch.nn as nn

from . import functional as Fn
from .functional import GRID_DIMS, EncodePlan, Grid4DFunction
from .hash_encoding import MultiResolutionHashEncoding
from .mlp_head import FusedProjection


def _cfg(config, name, default):
    v = getattr(config, name, default)
    return default if v is None else v


class _GraphedFeatures(nn.Module):
    """The feature half of a Grid4DEncoder as a module of its own (only the hash tables are its parameters): what
    ``Grid4DEncoder.graphed_features`` hands to ``torch.cuda.make_graphed_callables``."""

    def __init__(self, encoder: "Grid4DEncoder"):
        super().__init__()
        self.grids = nn.ModuleList(encoder._encoders)          # shared parameters, not copies
        object.__setattr__(self, "_encoder", encoder)
        self.train(encoder.training)

    def forward(self, xyzt: torch.Tensor) -> torch.Tensor:
        return self._encoder.encode_features(xyzt)


class Grid4DEncoder(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.config = config
        self.mode = _cfg(config, "grid4d_mode", "reference")
        F = config.n_features_per_level
        table_dtype = _cfg(config, "grid4d_table_dtype", "float32")
        table_scale = getattr(config, "grid4d_table_scale", None)
        flat = bool(_cfg(config, "grid4d_flat_tables", False))

        if self.mode == "reference":
            interp = _cfg(config, "grid4d_interpolation", "nearest")
            dense = bool(_cfg(config, "grid4d_dense_fallback", False))
            self.spatial_encoder = MultiResolutionHashEncoding(
                n_levels=config.n_spatial_levels, n_features_per_level=F,
                resolutions=config.spatial_resolutions, hash_table_size=config.hash_table_size, coords_dim=3,
                interpolation=interp, dense_fallback=dense, table_dtype=table_dtype, table_scale=table_scale,
                dims=(0, 1, 2), flat_tables=flat)
            self.temporal_encoder = MultiResolutionHashEncoding(
                n_levels=config.n_temporal_levels, n_features_per_level=F,
                resolutions=config.temporal_resolutions, hash_table_size=config.hash_table_size // 4,
                coords_dim=1, interpolation=interp, dense_fallback=dense, table_dtype=table_dtype,
                table_scale=table_scale, dims=(3,), flat_tables=flat)
            self._encoders = [self.spatial_encoder, self.temporal_encoder]
            self._mask_ids = [1, 2]
        elif self.mode == "decomposed":
            interp = _cfg(config, "grid4d_interpolation", "linear")
            dense = bool(_cfg(config, "grid4d_dense_fallback", True))
            st_res = _cfg(config, "spacetime_resolutions", config.temporal_resolutions)
            kw = dict(n_features_per_level=F, hash_table_size=config.hash_table_size, coords_dim=3,
                      interpolation=interp, dense_fallback=dense, table_dtype=table_dtype, table_scale=table_scale,
                      flat_tables=flat)
            self.xyz_encoder = MultiResolutionHashEncoding(
                n_levels=config.n_spatial_levels, resolutions=config.spatial_resolutions, dims=GRID_DIMS["xyz"], **kw)
            self.xyt_encoder = MultiResolutionHashEncoding(
                n_levels=config.n_temporal_levels, resolutions=st_res, dims=GRID_DIMS["xyt"], **kw)
            self.yzt_encoder = MultiResolutionHashEncoding(
                n_levels=config.n_temporal_levels, resolutions=st_res, dims=GRID_DIMS["yzt"], **kw)
            self.xzt_encoder = MultiResolutionHashEncoding(
                n_levels=config.n_temporal_levels, resolutions=st_res, dims=GRID_DIMS["xzt"], **kw)
            self._encoders = [self.xyz_encoder, self.xyt_encoder, self.yzt_encoder, self.xzt_encoder]
            self._mask_ids = [0, 0, 0, 0]
        else:
            raise ValueError(f"unknown grid4d_mode {self.mode!r}")

        self.feature_dim = sum(e.output_dim for e in self._encoders)
        self._fused_head = None   # FusedProjection | False, resolved on first forward
        self._grad_sync = None
        self.projection = nn.Sequential(
            nn.Linear(self.feature_dim, config.hidden_dim),
            nn.LayerNorm(config.hidden_dim),
            nn.GELU(),
            nn.Linear(config.hidden_dim, config.hidden_dim),
        )

    # ---- fused feature path -------------------------------------------------------------------------
    def _plan(self) -> EncodePlan:
        e0 = self._encoders[0]
        key = (tuple(e._geometry_key() + (e.dims,) for e in self._encoders), tuple(self._mask_ids), e0.interpolation,
               e0.table_dtype, e0.table_scale, e0.levels_per_group, e0.bwd_levels_per_group, e0.extra_flags)
        cache = self.__dict__.setdefault("_plan_cache", {})
        plan = cache.get(key)
        if plan is None:
            specs, col = [], 0
            for enc, mid in zip(self._encoders, self._mask_ids):
                specs += enc.level_specs(out_col0=col, mask_id=mid)
                col += enc.output_dim
            plan = EncodePlan(specs, e0.n_features_per_level, self.feature_dim, e0.interpolation == "linear",
                              e0.table_dtype, e0.table_scale, e0.levels_per_group, e0.bwd_levels_per_group,
                              e0.extra_flags)
            if len(cache) > 16:
                cache.clear()
            cache[key] = plan
        plan = copy.copy(plan)          # per-call fields live on the copy; the level list is shared (and packed once)
        plan.grad_sync = self._grad_sync
        return plan

    def enable_grad_sync(self, group=None, average: bool = True, chunks: int = 4) -> None:
        """Data parallel without DDP around the tables: the backward all-reduces the flat table-gradient
        buffer itself, slice by slice on a side stream, overlapped with the scatter (distributed.GradSync).
        Replaces what DDP does for these parameters in hpc/train_distrbuted.py:499-504."""
        from .distributed import GradSync
        self._grad_sync = GradSync(group, average, chunks)

    def disable_grad_sync(self) -> None:
        self._grad_sync = None

    def encode_features(self, xyzt: torch.Tensor, spatial_mask: Optional[torch.Tensor] = None,
                        temporal_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        """(B,4) -> (B, feature_dim): the concatenated hash features, before the projection."""
        if xyzt.dim() != 2 or xyzt.shape[1] != 4:
            raise ValueError(f"xyzt must be (B, 4), got {tuple(xyzt.shape)}")
        Fn._require_cuda(xyzt, "xyzt")
        weights: List[torch.Tensor] = []
        leaves: List[torch.Tensor] = []
        for enc in self._encoders:
            weights += enc.weights()
            leaves += enc.autograd_tensors()
        Fn._require_cuda(weights[0], "hash table weights")
        ptrs = (self._working_ptrs(weights) if self._encoders[0].table_dtype == torch.float16
                else [w.data_ptr() for w in weights])
        if self.mode == "decomposed" and (spatial_mask is not None or temporal_mask is not None):
            xyzt = xyzt.clone()                       # SPECIFICATIONS.MD:1159-1164: mask coordinates
            if spatial_mask is not None:
                xyzt[~spatial_mask.bool(), :3] = 0
            if temporal_mask is not None:
                xyzt[~temporal_mask.bool(), 3] = 0
            spatial_mask = temporal_mask = None
        return Grid4DFunction.apply(xyzt, spatial_mask, temporal_mask, self._plan(), ptrs, *leaves)

    @torch.no_grad()
    def encode_field(self, converter, lat_range, lon_range, n_lat: int, n_lon: int, altitude: float = 0.0,
                     time: Optional[float] = None, return_coords: bool = False):
        """Hash features over the regular grid ``np.linspace(*lat_range, n_lat)`` x ``np.linspace(*lon_range, n_lon)`` at a
        fixed altitude / time -- the query set of `visualize_embedding_field` (SPECIFICATIONS.MD:1963-2088) -- without a
        coordinate tensor from the host.  -> (n_lat * n_lon, feature_dim), latitude-major; inference only (no autograd).
        Decomposed mode runs the patch-ordered kernel (csrc/grid4d_field_kernels.cu); the reference geometry generates
        the coordinates on the device and runs the general forward.  Either way the features are bit-identical to
        ``encode_features(converter.geodetic_to_xyzt_norm(points))``."""
        import ctypes
        from . import _lib
        weights: List[torch.Tensor] = []
        for enc in self._encoders:
            weights += enc.weights()
        Fn._require_cuda(weights[0], "hash table weights")
        dev = weights[0].device
        e0 = self._encoders[0]
        plan = self._plan()
        n = int(n_lat) * int(n_lon)
        fused = (self.mode == "decomposed" and plan.linear and plan.n_features in (1, 2, 4) and len(plan.specs) <= 64
                 and all(len(s.dims) == 3 for s in plan.specs))
        if not fused:
            xyzt = converter.field_xyzt(lat_range, lon_range, n_lat, n_lon, altitude, time)
            feats = self.encode_features(xyzt)
            return (feats, xyzt) if return_coords else feats
        ptrs = self._working_ptrs(weights) if e0.table_dtype == torch.float16 else [w.data_ptr() for w in weights]
        offset, span = converter._norm_constants()
        desc = converter.field_desc(lat_range, lon_range, n_lat, n_lon, altitude, time)
        out = torch.empty((n, self.feature_dim), dtype=torch.float32, device=dev)
        xyzt = torch.empty((n, 4), dtype=torch.float32, device=dev) if return_coords else None
        dbl = lambda v: (ctypes.c_double * len(v))(*[float(x) for x in v])      # noqa: E731
        dt = _lib.DE_F16 if e0.table_dtype == torch.float16 else _lib.DE_F32
        with torch.cuda.device(dev):
            rc = _lib.load().de_grid4d_field_forward(
                Fn.pack_levels(plan.specs, ptrs), len(plan.specs), ctypes.byref(desc), dbl(offset), dbl(span),
                dbl(converter.t_range) if converter.t_range is not None else None, dt, plan.n_features,
                1.0 / plan.table_scale if plan.table_scale != 1.0 else 1.0, Fn._ptr(out), out.stride(0), Fn._ptr(xyzt),
                _lib.DE_INTERP_LINEAR | plan.extra_flags, plan.levels_per_group, Fn._stream_ptr(dev))
        _lib.check(rc, "de_grid4d_field_forward")
        return (out, xyzt) if return_coords else out

    @torch.no_grad()
    def forward_field(self, converter, lat_range, lon_range, n_lat: int, n_lon: int, altitude: float = 0.0,
                      time: Optional[float] = None) -> torch.Tensor:
        """(n_lat, n_lon, hidden_dim) embedding field: ``encode_field`` followed by the projection."""
        feats = self.encode_field(converter, lat_range, lon_range, n_lat, n_lon, altitude, time)
        return self._project(feats).reshape(int(n_lat), int(n_lon), -1)

    def graphed_features(self, sample_xyzt: torch.Tensor, num_warmup_iters: int = 3):
        """``xyzt -> features`` with forward AND backward captured as CUDA graphs (``torch.cuda.make_graphed_callables``) for
        the batch shape of ``sample_xyzt``: the reference trains and validates on batches of a few thousand points, where a
        step of this encoder is ~0.17 ms of kernels behind ~0.75 ms of Python / ctypes / autograd bookkeeping (64 table
        parameters).  Replays cost two graph launches.  Every kernel of the path is capture-safe (no synchronisation, no
        host-side state, scratch from the caching allocator); the fp32 -> fp16 table refresh is part of the forward graph,
        so optimiser steps between replays are seen.  Usual CUDA-graph rules apply: fixed shape, no masks, gradients are
        written into static buffers (set ``p.grad = None`` or step the optimiser before the next replay), and the call must
        come BEFORE the module's first eager backward (autograd binds a parameter's AccumulateGrad node to the stream of its
        first backward; one bound to the default stream cannot run inside a capture).  With ``grid4d_flat_tables`` a c2 step
        at 4096 points takes 0.18 ms (0.76 ms eager, per-level parameters)."""
        Fn._require_cuda(sample_xyzt, "sample_xyzt")
        feats = _GraphedFeatures(self)
        return torch.cuda.make_graphed_callables(feats, (sample_xyzt,), num_warmup_iters=num_warmup_iters)

    def refresh_tables(self) -> None:
        """Force the fp16 working tables to be re-derived on the next forward.  Needed only after the fp32 parameters were
        rewritten through ``.data`` (``p.data.copy_(ema)`` does not bump ``Parameter._version``, which is what the cache
        keys on outside training); optimiser steps, ``load_state_dict`` and ``.to()`` are tracked automatically."""
        self.__dict__["_work_key_all"] = None
        for enc in self._encoders:
            enc._work_key = None

    def _working_ptrs(self, weights: List[torch.Tensor]) -> List[int]:
        """fp16 working tables of ALL grids, refreshed by ONE de_tables_to_f16 launch when the fp32 parameters may have
        changed (same rule as MultiResolutionHashEncoding.working_ptrs, without one launch + host round per grid)."""
        dev = weights[0].device
        ptrs: List[int] = []
        fresh = True
        for enc in self._encoders:
            had = enc._work is not None and enc._work.device == dev
            ptrs += enc._ensure_work(dev)
            fresh = fresh and had
        stepping = self.training and torch.is_grad_enabled()      # an optimiser is rewriting the tables: always refresh
        key = None if stepping else (dev, tuple(w._version for w in weights), tuple(w.data_ptr() for w in weights))
        if stepping or not fresh or key != self.__dict__.get("_work_key_all"):
            e0 = self._encoders[0]
            with torch.no_grad():
                Fn.tables_to_f16(self._plan().specs, [w.data_ptr() for w in weights], ptrs, e0.n_features_per_level,
                                 e0.table_scale, dev)
            self.__dict__["_work_key_all"] = key
            for enc in self._encoders:
                enc._work_key = None          # a stand-alone call of one grid re-derives its own key
        return ptrs

    def _head_wanted(self) -> bool:
        """config.grid4d_fused_head: True = tcgen05 kernels whenever the shape is supported; None (default) = for the wide
        projection the reference ships (hidden_dim a multiple of 128: 768, 2048); False = the fp32 nn.Sequential."""
        want = getattr(self.config, "grid4d_fused_head", None)
        if want is None:
            return self.config.hidden_dim % 128 == 0
        return bool(want)

    def _project(self, feats: torch.Tensor) -> torch.Tensor:
        # The projection runs on our tcgen05 kernels (bf16 operands, fp32 accumulation): csrc/proj_kernels.cu for the
        # reference's widths, csrc/mlp_head.cu at hidden_dim = 64 (opt-in).  Unsupported shapes and
        # grid4d_fused_head=False use the reference's fp32 nn.Sequential.  Under autocast the 64-wide head falls back (its
        # autograd hand-off is fp32-only); the wide path runs as is and returns the autocast dtype, as nn.Linear would.
        if not self._head_wanted():
            return self.projection(feats)
        if self._fused_head is None:
            try:
                fp = FusedProjection(self.projection)
                self._fused_head = fp if fp.supported() else False
            except TypeError:
                self._fused_head = False
        head = self._fused_head
        if not head:
            return self.projection(feats)
        if torch.is_autocast_enabled():
            if not head.wide:
                return self.projection(feats)
            dt = torch.get_autocast_dtype("cuda")
            with torch.autocast(device_type="cuda", enabled=False):
                if dt == torch.bfloat16:          # y leaves the GEMM in bf16, its gradient arrives in bf16: no cast passes
                    return head(feats.float(), out_dtype=torch.bfloat16)
                y = head(feats.float())
            return y.to(dt)
        return head(feats)

    def forward(self, xyzt: torch.Tensor, spatial_mask: Optional[torch.Tensor] = None,
                temporal_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        feats = self.encode_features(xyzt, spatial_mask, temporal_mask)
        offered = self._offer_grouped_grads(feats)
        try:
            return self._project(feats)
        finally:
            if offered:
                self._fused_head._grouped_next = None      # consumed by _HeadFunction.forward, or never used

    def _offer_grouped_grads(self, feats: torch.Tensor) -> bool:
        """Fused head + large tables: let the head's backward kernel write d(loss)/d(features) in the group-major
        layout the hash-grid backward walks (functional.backward_group_width), instead of rows that the backward
        would first regroup.  The features tensor never leaves forward(), so its gradient has exactly one producer
        and one consumer; both sides share a per-call token."""
        head = self._fused_head
        if head is None and self._head_wanted() and not torch.is_autocast_enabled():
            self._project(feats[:0])                       # resolves self._fused_head
            head = self._fused_head
        fn = feats.grad_fn
        # the group-major hand-off is private to (head backward -> hash-grid backward): anything else that could look at
        # d(loss)/d(features) -- tensor hooks, anomaly mode -- gets the ordinary row-major gradient instead
        if torch.is_anomaly_enabled() or getattr(feats, "_backward_hooks", None):
            return False
        if not head or head.wide or not getattr(head, "fused_backward", False) or fn is None:
            return False
        if torch.is_autocast_enabled() or not hasattr(fn, "plan") or feats.shape[0] < (1 << 20):
            return False
        plan = fn.plan
        W = Fn.backward_group_width(plan.specs, feats.shape[0], plan.n_features, plan.linear,
                                    plan.bwd_levels_per_group, plan.extra_flags)
        if W and self._grad_sync is not None and not self._grad_sync.grouped_ok(
                plan.specs, feats.shape[0], plan.n_features, plan.linear, plan.bwd_levels_per_group, plan.extra_flags, W):
            W = 0                                          # the overlapped backward walks level runs: they must agree on W
        if W:
            plan.grouped_token = object()
            head._grouped_next = (plan.grouped_token, W)
        return bool(W)
