"""Tensor-level wrappers over the C ABI (include/de_grid4d.h) and the autograd bridge.

Everything here takes CUDA tensors, passes raw ``data_ptr()``s plus the current stream to
libde_grid4d.so, and raises if the tensors are not on a CUDA device -- there is no CPU path.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib
from ._lib import DeLevelDesc

GRID_DIMS = {  # SPECIFICATIONS.MD:1168-1183: column selection and concat order of the decomposition
    "xyz": (0, 1, 2),
    "xyt": (0, 1, 3),
    "yzt": (1, 2, 3),
    "xzt": (0, 2, 3),
}


@dataclass(frozen=True)
class LevelSpec:
    """Geometry of one resolution level of one grid (no pointers)."""
    table_size: int
    resolution: int
    dims: Tuple[int, ...]
    out_col: int
    dense: bool = False
    mask_id: int = 0


def dense_fits(resolution: int, n_dims: int, table_size: int) -> bool:
    """Dense-index rule of the decomposed mode: (R+1)^D <= T (oracle/grid4d_oracle.c header)."""
    return (int(resolution) + 1) ** int(n_dims) <= int(table_size)


def _stream_ptr(device: torch.device) -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _require_cuda(t: torch.Tensor, name: str) -> None:
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name} must be a torch.Tensor, got {type(t).__name__}")
    if not t.is_cuda:
        raise RuntimeError(f"{name} is on {t.device}: deepearth_b200 runs on CUDA devices only "
                           "(B200 / sm_100a); there is no CPU fallback")


# numpy mirror of DeLevelDesc (offsets taken from the ctypes structure): the static fields of a level list are
# packed once per list object, a call only patches the pointer columns (the per-field ctypes path cost 0.17 ms per
# 64 levels -- host time that sits exposed in front of every launch of a step that ends in a device->host read)
_LEVEL_DTYPE = np.dtype({
    "names": ["table", "grad_table", "table_size", "resolution", "out_col", "n_dims", "dim", "dense", "mask_id"],
    "formats": ["<u8", "<u8", "<u4", "<u4", "<u4", "u1", ("u1", (4,)), "u1", "u1"],
    "offsets": [DeLevelDesc.table.offset, DeLevelDesc.grad_table.offset, DeLevelDesc.table_size.offset,
                DeLevelDesc.resolution.offset, DeLevelDesc.out_col.offset, DeLevelDesc.n_dims.offset,
                DeLevelDesc.dim.offset, DeLevelDesc.dense.offset, DeLevelDesc.mask_id.offset],
    "itemsize": ctypes.sizeof(DeLevelDesc)})
_LEVEL_TEMPLATES: dict = {}


def _level_template(specs: Sequence[LevelSpec]) -> np.ndarray:
    hit = _LEVEL_TEMPLATES.get(id(specs))
    if hit is not None and hit[0] is specs and len(specs) == len(hit[1]):
        return hit[1]
    arr = np.zeros(len(specs), dtype=_LEVEL_DTYPE)
    for i, s in enumerate(specs):
        arr["table_size"][i] = s.table_size
        arr["resolution"][i] = s.resolution
        arr["out_col"][i] = s.out_col
        arr["n_dims"][i] = len(s.dims)
        arr["dim"][i, :len(s.dims)] = s.dims
        arr["dense"][i] = 1 if s.dense else 0
        arr["mask_id"][i] = s.mask_id
    if isinstance(specs, list):                  # plans keep their spec list alive and never mutate it
        if len(_LEVEL_TEMPLATES) > 256:
            _LEVEL_TEMPLATES.clear()
        _LEVEL_TEMPLATES[id(specs)] = (specs, arr)
    return arr


def pack_levels(specs: Sequence[LevelSpec], table_ptrs: Optional[Sequence[int]],
                grad_ptrs: Optional[Sequence[int]] = None):
    arr = _level_template(specs).copy()
    if table_ptrs is not None:
        arr["table"] = [p or 0 for p in table_ptrs]
    if grad_ptrs is not None:
        arr["grad_table"] = [p or 0 for p in grad_ptrs]
    return (DeLevelDesc * len(specs)).from_buffer_copy(arr.tobytes())


def _coords_view(coords: torch.Tensor) -> Tuple[torch.Tensor, int]:
    """Return (tensor whose data_ptr is row 0, row stride in floats); copies only if unavoidable."""
    if coords.dim() != 2:
        raise ValueError(f"coords must be (B, D), got shape {tuple(coords.shape)}")
    if coords.dtype != torch.float32:
        raise TypeError(f"coords must be float32 (the reference multiplies in the input dtype, "
                        f"models/hash_encoding.py:58; only fp32 is implemented), got {coords.dtype}")
    if coords.shape[0] > 1 and (coords.stride(1) != 1 or coords.stride(0) < coords.shape[1]):
        coords = coords.contiguous()
    elif coords.shape[0] <= 1 and coords.stride(1) != 1:
        coords = coords.contiguous()
    stride = coords.stride(0) if coords.shape[0] > 1 else max(coords.shape[1], 1)
    if stride > coords.shape[1]:
        # the C ABI reads rows of `stride` floats (float4 rows when stride == 4): make sure the last
        # row's padding columns exist in the underlying storage, else compact
        need = coords.storage_offset() + (coords.shape[0] - 1) * stride + stride
        if need > coords.untyped_storage().nbytes() // 4:
            coords = coords.contiguous()
            stride = coords.shape[1]
    return coords, int(stride)


def _mask_u8(mask: Optional[torch.Tensor], n: int, device) -> Optional[torch.Tensor]:
    if mask is None:
        return None
    _require_cuda(mask, "mask")
    if mask.shape != (n,):
        raise ValueError(f"mask must have shape ({n},), got {tuple(mask.shape)}")
    if mask.dtype == torch.bool:
        return mask.contiguous().view(torch.uint8)
    return (mask != 0).contiguous().view(torch.uint8)


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def n_corners(linear: bool, n_dims: int) -> int:
    return (1 << n_dims) if linear else 1


def hash_indices(specs: Sequence[LevelSpec], coords: torch.Tensor, linear: bool = False) -> torch.Tensor:
    """(n_levels, n_corners, B) int64 table indices; all specs must share the dimensionality."""
    _require_cuda(coords, "coords")
    coords, stride = _coords_view(coords)
    n = coords.shape[0]
    C = n_corners(linear, len(specs[0].dims))
    out = torch.empty((len(specs), C, n), dtype=torch.int64, device=coords.device)
    if n == 0:
        return out
    lib = _lib.load()
    with torch.cuda.device(coords.device):
        rc = lib.de_grid4d_hash_indices(pack_levels(specs, None), len(specs), _ptr(coords), n, stride,
                                        _lib.DE_INTERP_LINEAR if linear else _lib.DE_INTERP_NEAREST,
                                        _ptr(out), _stream_ptr(coords.device))
    _lib.check(rc, "de_grid4d_hash_indices")
    return out


def encode_forward(specs: Sequence[LevelSpec], coords: torch.Tensor, table_ptrs: Sequence[int],
                   table_dtype: torch.dtype, n_features: int, out_cols: int, linear: bool,
                   out_scale: float = 1.0, mask_a: Optional[torch.Tensor] = None,
                   mask_b: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
                   levels_per_group: int = 0, extra_flags: int = 0) -> torch.Tensor:
    _require_cuda(coords, "coords")
    coords, stride = _coords_view(coords)
    n = coords.shape[0]
    if out is None:
        out = torch.empty((n, out_cols), dtype=torch.float32, device=coords.device)
    elif out.shape != (n, out_cols) or out.dtype != torch.float32 or not out.is_contiguous():
        raise ValueError("out must be a contiguous float32 (B, out_cols) tensor")
    if n == 0:
        return out
    ma = _mask_u8(mask_a, n, coords.device)
    mb = _mask_u8(mask_b, n, coords.device)
    dt = {torch.float32: _lib.DE_F32, torch.float16: _lib.DE_F16}.get(table_dtype)
    if dt is None:
        raise TypeError(f"table dtype {table_dtype} not supported (float32 or float16)")
    lib = _lib.load()
    with torch.cuda.device(coords.device):
        rc = lib.de_grid4d_forward(pack_levels(specs, table_ptrs), len(specs), _ptr(coords), n, stride, dt,
                                   n_features, out_scale, _ptr(ma), _ptr(mb), _ptr(out), out.stride(0),
                                   (_lib.DE_INTERP_LINEAR if linear else _lib.DE_INTERP_NEAREST) | extra_flags,
                                   levels_per_group, _stream_ptr(coords.device))
    _lib.check(rc, "de_grid4d_forward")
    return out


def encode_backward(specs: Sequence[LevelSpec], coords: torch.Tensor, grad_out: torch.Tensor,
                    grad_ptrs: Sequence[int], n_features: int, linear: bool,
                    mask_a: Optional[torch.Tensor] = None, mask_b: Optional[torch.Tensor] = None,
                    levels_per_group: int = 0, extra_flags: int = 0, use_replicas: bool = True,
                    group_width: int = 0, workspace_fraction: float = 1.0) -> None:
    """Accumulate table gradients into the fp32 buffers at ``grad_ptrs`` (caller zeroes them).

    ``workspace_fraction`` < 1 hands the call a smaller scratch than it asked for (the C ABI accepts that and
    skips the replicas / regrouped rows that do not fit; tests exercise it).

    ``group_width`` = W > 0: ``grad_out``'s memory is already group-major, g[c // W][p][c % W] (what
    ``backward_group_width`` returned and ``mlp_head_backward(..., group_width=W)`` wrote); 0: ordinary rows."""
    _require_cuda(coords, "coords")
    _require_cuda(grad_out, "grad_out")
    coords, stride = _coords_view(coords)
    n = coords.shape[0]
    if n == 0:
        return
    if group_width:
        if grad_out.dtype != torch.float32 or not grad_out.is_contiguous():
            raise ValueError("a group-major grad_out must be a contiguous float32 buffer")
        extra_flags |= _lib.DE_FLAG_GRAD_GROUPED
    elif grad_out.dtype != torch.float32:
        grad_out = grad_out.float()
    if not group_width and (grad_out.dim() != 2 or grad_out.shape[0] != n or grad_out.stride(1) != 1):
        grad_out = grad_out.reshape(n, -1).contiguous()
    ma = _mask_u8(mask_a, n, coords.device)
    mb = _mask_u8(mask_b, n, coords.device)
    lib = _lib.load()
    flags = (_lib.DE_INTERP_LINEAR if linear else _lib.DE_INTERP_NEAREST) | extra_flags
    levels = pack_levels(specs, None, grad_ptrs)
    ws, ws_bytes = None, 0
    if use_replicas:
        ws_bytes = int(lib.de_grid4d_backward_workspace_bytes_ex(levels, len(specs), n_features, n, flags, levels_per_group))
        if workspace_fraction < 1.0:
            ws_bytes = (int(ws_bytes * workspace_fraction) // 256) * 256
        if ws_bytes:
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=coords.device)   # caching allocator: no cudaMalloc in steady state
    with torch.cuda.device(coords.device):
        rc = lib.de_grid4d_backward(levels, len(specs), _ptr(coords), n, stride,
                                    n_features, _ptr(ma), _ptr(mb), _ptr(grad_out),
                                    group_width if group_width else (grad_out.stride(0) if n > 1 else grad_out.shape[1]),
                                    flags, levels_per_group, _ptr(ws), ws_bytes, _stream_ptr(coords.device))
    _lib.check(rc, "de_grid4d_backward")


def sort_points(coords: torch.Tensor, dims: Sequence[int], bits: int = 5) -> torch.Tensor:
    """perm (int32, B): the points ordered by the cell they occupy on a 2^bits-per-axis grid over coordinate columns
    ``dims`` (x-major) -- the pre-pass of the sort-by-cell backward (csrc/grid4d_sorted_kernels.cu)."""
    _require_cuda(coords, "coords")
    coords, stride = _coords_view(coords)
    n = coords.shape[0]
    perm = torch.empty(n, dtype=torch.int32, device=coords.device)
    if n == 0:
        return perm
    lib = _lib.load()
    ws_bytes = int(lib.de_grid4d_sort_workspace_bytes(len(dims), int(bits)))
    if not ws_bytes:
        raise ValueError(f"bits * len(dims) must be in 1..16, got bits={bits}, dims={tuple(dims)}")
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=coords.device)
    d = (ctypes.c_uint8 * len(dims))(*[int(v) for v in dims])
    with torch.cuda.device(coords.device):
        rc = lib.de_grid4d_sort_points(_ptr(coords), n, stride, d, len(dims), int(bits), _ptr(perm), _ptr(ws), ws_bytes,
                                       _stream_ptr(coords.device))
    _lib.check(rc, "de_grid4d_sort_points")
    return perm


def encode_backward_sorted(specs: Sequence[LevelSpec], coords: torch.Tensor, grad_out: torch.Tensor,
                           grad_ptrs: Sequence[int], n_features: int, perm: torch.Tensor,
                           mask_a: Optional[torch.Tensor] = None, mask_b: Optional[torch.Tensor] = None) -> None:
    """Warp-aggregated scatter of 3-D multilinear levels over the points in ``perm`` order (opt-in alternative to the
    gradient replicas for densely colliding levels); accumulates into the fp32 buffers at ``grad_ptrs``."""
    _require_cuda(coords, "coords")
    _require_cuda(grad_out, "grad_out")
    coords, stride = _coords_view(coords)
    n = coords.shape[0]
    if n == 0:
        return
    if perm.dtype != torch.int32 or perm.shape != (n,) or not perm.is_contiguous():
        raise ValueError("perm must be a contiguous int32 tensor with one entry per point")
    if grad_out.dtype != torch.float32:
        grad_out = grad_out.float()
    if grad_out.dim() != 2 or grad_out.shape[0] != n or grad_out.stride(1) != 1:
        grad_out = grad_out.reshape(n, -1).contiguous()
    ma = _mask_u8(mask_a, n, coords.device)
    mb = _mask_u8(mask_b, n, coords.device)
    with torch.cuda.device(coords.device):
        rc = _lib.load().de_grid4d_backward_sorted(pack_levels(specs, None, grad_ptrs), len(specs), _ptr(coords), n, stride,
                                                   n_features, _ptr(ma), _ptr(mb), _ptr(grad_out),
                                                   grad_out.stride(0) if n > 1 else grad_out.shape[1], _lib.DE_INTERP_LINEAR,
                                                   _ptr(perm), _stream_ptr(coords.device))
    _lib.check(rc, "de_grid4d_backward_sorted")


def backward_group_width(specs: Sequence[LevelSpec], n_points: int, n_features: int, linear: bool,
                         levels_per_group: int = 0, extra_flags: int = 0) -> int:
    """W > 0 when ``encode_backward`` for these arguments would regroup the gradient rows (large tables => few levels
    per group): a producer may then write g[c // W][p][c % W] itself and pass ``group_width=W``.  0 otherwise."""
    if n_points <= 0:
        return 0
    flags = (_lib.DE_INTERP_LINEAR if linear else _lib.DE_INTERP_NEAREST) | extra_flags
    levels = pack_levels(specs, None, [16] * len(specs))
    return int(_lib.load().de_grid4d_backward_group_width(levels, len(specs), n_features, int(n_points), flags, levels_per_group))


def tables_to_f16(specs: Sequence[LevelSpec], src_ptrs: Sequence[int], dst_ptrs: Sequence[int],
                  n_features: int, scale: float, device: torch.device) -> None:
    lib = _lib.load()
    arr = (ctypes.c_void_p * len(dst_ptrs))(*dst_ptrs)
    with torch.cuda.device(device):
        rc = lib.de_tables_to_f16(pack_levels(specs, src_ptrs), len(specs), n_features, arr, scale,
                                  _stream_ptr(device))
    _lib.check(rc, "de_tables_to_f16")


def sum_f32(x: torch.Tensor) -> torch.Tensor:
    """fp64 sum of an fp32 CUDA tensor (0-d tensor on the device)."""
    _require_cuda(x, "x")
    x = x.contiguous()
    res = torch.zeros((), dtype=torch.float64, device=x.device)
    with torch.cuda.device(x.device):
        rc = _lib.load().de_sum_f32(_ptr(x), x.numel(), _ptr(res), _stream_ptr(x.device))
    _lib.check(rc, "de_sum_f32")
    return res


class EncodePlan:
    """Static description of one encode call: level geometry + numerics switches."""

    def __init__(self, specs: Sequence[LevelSpec], n_features: int, out_cols: int, linear: bool,
                 table_dtype: torch.dtype = torch.float32, table_scale: float = 1.0,
                 levels_per_group: int = 0, bwd_levels_per_group: int = 0, extra_flags: int = 0):
        self.specs = list(specs)
        self.extra_flags = int(extra_flags)   # DE_FLAG_* A/B switches (include/de_grid4d.h)
        self.grad_sync = None                 # distributed.GradSync: overlapped all-reduce inside backward
        self.grouped_token = None             # set per call by Grid4DEncoder.forward (fused head writes grads group-major)
        self.n_features = int(n_features)
        self.out_cols = int(out_cols)
        self.linear = bool(linear)
        self.table_dtype = table_dtype
        self.table_scale = float(table_scale)
        self.levels_per_group = int(levels_per_group)
        self.bwd_levels_per_group = int(bwd_levels_per_group)
        self.grad_offsets: List[int] = []
        off = 0
        for s in self.specs:
            self.grad_offsets.append(off)
            off += s.table_size * self.n_features
        self.grad_numel = off


class Grid4DFunction(torch.autograd.Function):
    """autograd bridge: forward = de_grid4d_forward, backward = de_grid4d_backward.

    Coordinates get no gradient (reference: graph cut at models/hash_encoding.py:61); every
    table gets a dense fp32 gradient of its own shape, all of them views of one flat buffer so a
    data-parallel caller can all-reduce it in a single collective.
    """

    @staticmethod
    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, coords, mask_a, mask_b, plan: EncodePlan, working_ptrs, *weights):
        # weights: the autograd leaves -- one (T, F) parameter per level, or one flat (L, T, F) parameter per encoder; in
        # plan order either way, so the flat gradient buffer splits by numel.  working_ptrs: per-LEVEL device pointers of
        # the tables the kernel reads (required with flat parameters)
        ptrs = working_ptrs if working_ptrs is not None else [w.data_ptr() for w in weights]
        if len(ptrs) != len(plan.specs):
            raise ValueError("Grid4DFunction: one table pointer per level is required")
        out = encode_forward(plan.specs, coords, ptrs, plan.table_dtype, plan.n_features, plan.out_cols,
                             plan.linear, 1.0 / plan.table_scale if plan.table_scale != 1.0 else 1.0,
                             mask_a, mask_b, levels_per_group=plan.levels_per_group, extra_flags=plan.extra_flags)
        ctx.plan = plan
        ctx.weight_shapes = [tuple(w.shape) for w in weights]
        ctx.save_for_backward(coords, mask_a, mask_b)
        ctx.set_materialize_grads(False)
        return out

    @staticmethod
    @torch.amp.custom_bwd(device_type="cuda")
    def backward(ctx, grad_out):
        plan: EncodePlan = ctx.plan
        n_w = len(ctx.weight_shapes)
        if grad_out is None or not any(ctx.needs_input_grad[5:]):
            return (None,) * (5 + n_w)
        coords, mask_a, mask_b = ctx.saved_tensors
        flat = torch.zeros(plan.grad_numel, dtype=torch.float32, device=grad_out.device)
        base = flat.data_ptr()
        W = 0
        if plan.grouped_token is not None:
            from .mlp_head import GROUPED_GRADS
            tagged = GROUPED_GRADS.pop(plan.grouped_token, None)
            if tagged is not None:
                if tagged[0] != grad_out.data_ptr() or not grad_out.is_contiguous():
                    raise RuntimeError("Grid4DFunction.backward: the fused head wrote a group-major gradient buffer, but "
                                       "autograd delivered a different tensor (features consumed more than once?)")
                W = tagged[1]
        if plan.grad_sync is not None:
            # data parallel: all-reduce slices of the flat buffer on a side stream while later level runs scatter
            plan.grad_sync.backward(plan.specs, coords, grad_out, flat, plan.grad_offsets, plan.n_features,
                                    plan.linear, mask_a, mask_b, levels_per_group=plan.bwd_levels_per_group,
                                    extra_flags=plan.extra_flags, group_width=W)
        else:
            encode_backward(plan.specs, coords, grad_out, [base + 4 * o for o in plan.grad_offsets],
                            plan.n_features, plan.linear, mask_a, mask_b,
                            levels_per_group=plan.bwd_levels_per_group, extra_flags=plan.extra_flags, group_width=W)
        # gradients as views of the flat buffer, split by numel in plan order: one view + unbind per run of equally shaped
        # (T, F) tables (64 slice + view calls were ~0.2 ms of host time per step), one view per flat (L, T, F) parameter
        grads: list = []
        shapes = ctx.weight_shapes
        i, o = 0, 0
        while i < n_w:
            j = i + 1
            while j < n_w and shapes[j] == shapes[i] and len(shapes[i]) == 2:
                j += 1
            numel = 1
            for d in shapes[i]:
                numel *= d
            if j - i > 1:
                grads.extend(flat[o:o + (j - i) * numel].view(j - i, *shapes[i]).unbind(0))
            else:
                grads.append(flat[o:o + numel].view(*shapes[i]))
            o += (j - i) * numel
            i = j
        if not all(ctx.needs_input_grad[5:]):
            grads = [g if ctx.needs_input_grad[5 + k] else None for k, g in enumerate(grads)]
        return (None, None, None, None, None, *grads)
