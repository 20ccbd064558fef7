"""MultiResolutionHashEncoding -- B200 drop-in for the reference module of the same name.

Reference contract mirrored here (/root/reference/models/hash_encoding.py):
  * constructor ``(n_levels, n_features_per_level, resolutions, hash_table_size, coords_dim)`` (:19-20),
    attributes ``n_levels n_features_per_level resolutions hash_table_size coords_dim hash_tables`` (:31-41);
  * ``hash_tables`` is an ``nn.ModuleList`` of ``nn.Embedding(T, F)`` initialised U(-1e-4, 1e-4) in the same
    RNG order (:38-45), so state-dict keys are ``hash_tables.{l}.weight`` and the same seed gives the same
    weights as the reference;
  * ``hash_coords(coords, resolution) -> (B,) int64`` (:47-74) and ``forward(coords (B,D)) -> (B, L*F)`` fp32,
    level-major / feature-minor (:76-96); tables get dense fp32 grads, coords get none.
The whole level loop is one CUDA launch (deepearth_b200/csrc/grid4d_kernels.cu).  Opt-in keyword
arguments add multilinear interpolation, the dense-index fallback and fp16 working tables.
"""
from __future__ import annotations

import copy

from typing import List, Optional, Sequence

import torch
import torch.nn as nn

from . import functional as Fn
from .functional import EncodePlan, Grid4DFunction, LevelSpec, dense_fits

_DTYPES = {"float32": torch.float32, "fp32": torch.float32, "float16": torch.float16, "fp16": torch.float16,
           torch.float32: torch.float32, torch.float16: torch.float16}

DEFAULT_F16_TABLE_SCALE = 4096.0  # 2^12: lifts the U(-1e-4,1e-4) init out of the fp16 subnormal range


class _LevelView:
    """`hash_tables[l]` of a flat-table encoder: exposes `.weight` (a (T, F) view of the flat parameter) and the
    nn.Embedding attributes callers read."""

    def __init__(self, owner, level: int):
        self._owner, self._level = owner, level

    @property
    def weight(self) -> torch.Tensor:
        return self._owner.tables[self._level]

    @property
    def num_embeddings(self) -> int:
        return self._owner.hash_table_size

    @property
    def embedding_dim(self) -> int:
        return self._owner.n_features_per_level


class _TableViews:
    def __init__(self, owner):
        self._owner = owner

    def __len__(self) -> int:
        return self._owner.n_levels

    def __getitem__(self, l: int) -> _LevelView:
        if not -len(self) <= l < len(self):
            raise IndexError(l)
        return _LevelView(self._owner, l % len(self))

    def __iter__(self):
        return (_LevelView(self._owner, l) for l in range(len(self)))


class MultiResolutionHashEncoding(nn.Module):
    def __init__(self, n_levels: int, n_features_per_level: int, resolutions: List[int],
                 hash_table_size: int, coords_dim: int, *, interpolation: str = "nearest",
                 dense_fallback: bool = False, table_dtype="float32", table_scale: Optional[float] = None,
                 dims: Optional[Sequence[int]] = None, levels_per_group: int = 0,
                 bwd_levels_per_group: int = 0, flat_tables: bool = False):
        super().__init__()
        if not 1 <= coords_dim <= 4:
            raise ValueError(f"coords_dim must be in 1..4 (the reference has 4 primes, hash_encoding.py:64), got {coords_dim}")
        if len(resolutions) < n_levels:
            # the reference silently emits fewer features here (slice at hash_encoding.py:87) and then
            # breaks downstream shapes; reject it up front
            raise ValueError(f"need at least n_levels={n_levels} resolutions, got {len(resolutions)}")
        if interpolation not in ("nearest", "linear"):
            raise ValueError(f"interpolation must be 'nearest' or 'linear', got {interpolation!r}")
        if table_dtype not in _DTYPES:
            raise ValueError(f"table_dtype must be float32 or float16, got {table_dtype!r}")
        self.n_levels = n_levels
        self.n_features_per_level = n_features_per_level
        self.resolutions = resolutions
        self.hash_table_size = hash_table_size
        self.coords_dim = coords_dim
        self.interpolation = interpolation
        self.dense_fallback = bool(dense_fallback)
        self.table_dtype = _DTYPES[table_dtype]
        if table_scale is None:
            table_scale = DEFAULT_F16_TABLE_SCALE if self.table_dtype == torch.float16 else 1.0
        if self.table_dtype == torch.float32 and table_scale != 1.0:
            raise ValueError("table_scale applies to float16 working tables only")
        self.table_scale = float(table_scale)
        self.dims = tuple(dims) if dims is not None else tuple(range(coords_dim))
        if len(self.dims) != coords_dim:
            raise ValueError("dims must name coords_dim coordinate columns")
        self.levels_per_group = levels_per_group
        self.bwd_levels_per_group = bwd_levels_per_group
        self.extra_flags = 0   # DE_FLAG_* A/B switches (include/de_grid4d.h); 0 = every optimisation on

        # the reference's construction (and RNG consumption): n_levels nn.Embedding, then uniform init (hash_encoding.py:38-45)
        tables = nn.ModuleList([nn.Embedding(hash_table_size, n_features_per_level) for _ in range(n_levels)])
        for table in tables:
            nn.init.uniform_(table.weight, -1e-4, 1e-4)
        self.flat_tables = bool(flat_tables)
        if self.flat_tables:
            # ONE (L, T, F) parameter: autograd, optimisers and DDP see one tensor per encoder instead of L (the per-level
            # bookkeeping is ~0.3 ms of a small-batch step).  `hash_tables[l].weight` stays readable (a view) and state dicts
            # keep the reference's keys `hash_tables.{l}.weight` (_save_to_state_dict / _load_from_state_dict below).
            self.tables = nn.Parameter(torch.stack([t.weight.detach() for t in tables]))
            object.__setattr__(self, "hash_tables", _TableViews(self))
        else:
            self.hash_tables = tables

        self._work: Optional[torch.Tensor] = None   # flat fp16 working copy (L, T, F)
        self._work_key = None

    # ---- geometry ---------------------------------------------------------------------------------
    @property
    def output_dim(self) -> int:
        return self.n_levels * self.n_features_per_level

    def _geometry_key(self):
        return (self.hash_table_size, self.n_levels, self.n_features_per_level, self.dense_fallback,
                tuple(int(r) for r in self.resolutions[:self.n_levels]))

    def level_specs(self, dims: Optional[Sequence[int]] = None, out_col0: int = 0, mask_id: int = 0) -> List[LevelSpec]:
        """Level geometry (cached: callers get the same list object for the same geometry and must not mutate it)."""
        dims = tuple(dims) if dims is not None else self.dims
        key = (dims, out_col0, mask_id, self._geometry_key())
        cache = self.__dict__.setdefault("_spec_cache", {})
        specs = cache.get(key)
        if specs is None:
            F = self.n_features_per_level
            specs = [LevelSpec(table_size=self.hash_table_size, resolution=int(R), dims=dims, out_col=out_col0 + l * F,
                               dense=self.dense_fallback and dense_fits(R, len(dims), self.hash_table_size),
                               mask_id=mask_id)
                     for l, R in enumerate(self.resolutions[:self.n_levels])]
            if len(cache) > 16:
                cache.clear()
            cache[key] = specs
        return specs

    def weights(self) -> List[torch.Tensor]:
        """Per-level (T, F) tensors (parameters, or views of the flat parameter): what the kernels take pointers of."""
        if self.flat_tables:
            return list(self.tables.unbind(0))
        return [t.weight for t in self.hash_tables]

    def autograd_tensors(self) -> List[torch.Tensor]:
        """The leaves autograd differentiates: the L level parameters, or the one flat parameter."""
        return [self.tables] if self.flat_tables else [t.weight for t in self.hash_tables]

    # ---- state dict: the reference's keys in both layouts -----------------------------------------------------
    def _save_to_state_dict(self, destination, prefix, keep_vars):
        if not self.flat_tables:
            return super()._save_to_state_dict(destination, prefix, keep_vars)
        for l in range(self.n_levels):
            t = self.tables[l]
            destination[f"{prefix}hash_tables.{l}.weight"] = t if keep_vars else t.detach()

    def _load_from_state_dict(self, state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys, error_msgs):
        if not self.flat_tables:
            return super()._load_from_state_dict(state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys,
                                                 error_msgs)
        with torch.no_grad():
            for l in range(self.n_levels):
                key = f"{prefix}hash_tables.{l}.weight"
                if key not in state_dict:
                    missing_keys.append(key)
                    continue
                src = state_dict[key]
                if tuple(src.shape) != tuple(self.tables.shape[1:]):
                    error_msgs.append(f"size mismatch for {key}: copying a param with shape {tuple(src.shape)} from checkpoint, "
                                      f"the shape in current model is {tuple(self.tables.shape[1:])}.")
                    continue
                self.tables[l].copy_(src)
        self._work_key = None

    # ---- fp16 working tables ----------------------------------------------------------------------
    def working_ptrs(self) -> Optional[List[int]]:
        """Device pointers the forward kernel reads: the fp32 parameters themselves, or the scaled fp16
        working copy (refreshed whenever the parameters may have changed)."""
        if self.table_dtype == torch.float32:
            return None
        ws = self.weights()
        dev = ws[0].device
        key = (dev, tuple(w._version for w in ws), tuple(w.data_ptr() for w in ws))
        stale = self._work is None or self._work.device != dev or key != self._work_key
        if stale or (self.training and torch.is_grad_enabled()):
            self.refresh_tables()
            self._work_key = key
        per = self.hash_table_size * self.n_features_per_level * 2
        return [self._work.data_ptr() + l * per for l in range(self.n_levels)]

    def _ensure_work(self, dev: torch.device) -> List[int]:
        """Allocate the fp16 working copy if needed; returns its per-level device pointers."""
        if self._work is None or self._work.device != dev:
            self._work = torch.empty((self.n_levels, self.hash_table_size, self.n_features_per_level),
                                     dtype=torch.float16, device=dev)
            self._work_key = None
        per = self.hash_table_size * self.n_features_per_level * 2
        base = self._work.data_ptr()
        return [base + l * per for l in range(self.n_levels)]

    @torch.no_grad()
    def refresh_tables(self) -> None:
        """Re-derive the fp16 working tables from the fp32 parameters: half(w * table_scale)."""
        if self.table_dtype == torch.float32:
            return
        ws = self.weights()
        dev = ws[0].device
        Fn._require_cuda(ws[0], "hash table weights")
        if self._work is None or self._work.device != dev:
            self._work = torch.empty((self.n_levels, self.hash_table_size, self.n_features_per_level),
                                     dtype=torch.float16, device=dev)
        per = self.hash_table_size * self.n_features_per_level * 2
        Fn.tables_to_f16(self.level_specs(), [w.data_ptr() for w in ws],
                         [self._work.data_ptr() + l * per for l in range(self.n_levels)],
                         self.n_features_per_level, self.table_scale, dev)

    def _apply(self, fn, *args, **kwargs):
        self._work = None       # parameters moved / cast: drop the derived copy
        self._work_key = None
        return super()._apply(fn, *args, **kwargs)

    # ---- reference API ----------------------------------------------------------------------------
    def hash_coords(self, coords: torch.Tensor, resolution: int) -> torch.Tensor:
        """(B, D) coords -> (B,) int64 table indices of the floor cell (hash_encoding.py:47-74)."""
        if coords.dim() != 2 or coords.shape[1] != self.coords_dim:
            raise ValueError(f"coords must be (B, {self.coords_dim}), got {tuple(coords.shape)}")
        spec = LevelSpec(self.hash_table_size, int(resolution), tuple(range(self.coords_dim)), 0, False, 0)
        return Fn.hash_indices([spec], coords, linear=False)[0, 0]

    def plan(self, dims: Optional[Sequence[int]] = None) -> EncodePlan:
        dims = tuple(dims) if dims is not None else self.dims
        key = (dims, self._geometry_key(), self.interpolation, self.table_dtype, self.table_scale,
               self.levels_per_group, self.bwd_levels_per_group, self.extra_flags)
        cache = self.__dict__.setdefault("_plan_cache", {})
        plan = cache.get(key)
        if plan is None:
            plan = EncodePlan(self.level_specs(dims), self.n_features_per_level, self.output_dim,
                              self.interpolation == "linear", self.table_dtype, self.table_scale,
                              self.levels_per_group, self.bwd_levels_per_group, self.extra_flags)
            if len(cache) > 16:
                cache.clear()
            cache[key] = plan
        return copy.copy(plan)          # per-call fields (grad_sync, grouped_token) live on the copy

    def forward(self, coords: torch.Tensor) -> torch.Tensor:
        if coords.dim() != 2 or coords.shape[1] < self.coords_dim:
            raise ValueError(f"coords must be (B, {self.coords_dim}), got {tuple(coords.shape)}")
        # stand-alone call with exactly coords_dim columns (the reference signature) reads columns
        # 0..D-1; a wider row (e.g. xyzt) is read through self.dims
        dims = self.dims if coords.shape[1] > max(self.dims) else tuple(range(self.coords_dim))
        Fn._require_cuda(coords, "coords")
        ws = self.weights()
        Fn._require_cuda(ws[0], "hash table weights")
        ptrs = self.working_ptrs() if self.table_dtype == torch.float16 else [w.data_ptr() for w in ws]
        return Grid4DFunction.apply(coords, None, None, self.plan(dims), ptrs, *self.autograd_tensors())

    def extra_repr(self) -> str:
        return (f"levels={self.n_levels}, F={self.n_features_per_level}, T={self.hash_table_size}, D={self.coords_dim}, "
                f"interpolation={self.interpolation}, dense_fallback={self.dense_fallback}, "
                f"table_dtype={self.table_dtype}, table_scale={self.table_scale}")
