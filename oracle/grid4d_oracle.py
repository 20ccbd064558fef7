"""CPU ORACLE bindings (test infrastructure, NOT product code).

Three restatements of the reference algorithm live here; see grid4d_oracle.c for the
stated math and the parity-pinning status:

* ``COracle``     -- ctypes binding of libgrid4d_oracle.so (fast, OpenMP; the checker the
                     GPU parity tests use).
* ``numpy_*``     -- vectorised numpy int64 restatement of hash_coords / nearest forward
                     (models/hash_encoding.py:47-96); cross-checks the C code.
* grid4d_torch.py -- op-by-op torch port used only as the timed CPU baseline.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libgrid4d_oracle.so")

PRIMES = np.array([1, 2654435761, 805459861, 3674653429], dtype=np.int64)  # hash_encoding.py:64

GRID_DIMS = {  # SPECIFICATIONS.MD:1168-1183 column selection and concat order
    "xyz": (0, 1, 2),
    "xyt": (0, 1, 3),
    "yzt": (1, 2, 3),
    "xzt": (0, 2, 3),
}


class _OrcLevel(ctypes.Structure):
    _fields_ = [
        ("table_off", ctypes.c_int64),
        ("T", ctypes.c_int32),
        ("R", ctypes.c_int32),
        ("out_col", ctypes.c_int32),
        ("D", ctypes.c_int32),
        ("dim", ctypes.c_int32 * 4),
        ("dense", ctypes.c_int32),
        ("mask_id", ctypes.c_int32),
    ]


@dataclass
class Level:
    """One resolution level of one grid (host-side description shared by the tests)."""

    T: int
    R: int
    dims: Sequence[int]
    out_col: int
    table_off: int = 0  # in entries
    dense: bool = False
    mask_id: int = 0

    @property
    def D(self) -> int:
        return len(self.dims)


def build_oracle(force: bool = False) -> str:
    """Compile libgrid4d_oracle.so with the committed Makefile (gcc only)."""
    src = os.path.join(_HERE, "grid4d_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE, "-B"], check=True, capture_output=True)
    return _LIB_PATH


def make_levels(grids: Sequence[Sequence[int]], resolutions: Sequence[Sequence[int]],
                table_sizes: Sequence[int], F: int, dense_fallback: bool,
                mask_ids: Optional[Sequence[int]] = None) -> List[Level]:
    """Lay out levels grid-major / level-minor, tables packed back to back.

    ``dense_fallback``: a level uses the dense index iff (R+1)^D <= T (grid4d_oracle.c header).
    """
    levels: List[Level] = []
    off = 0
    col = 0
    for gi, dims in enumerate(grids):
        for R in resolutions[gi]:
            T = int(table_sizes[gi])
            dense = bool(dense_fallback and (int(R) + 1) ** len(dims) <= T)
            levels.append(Level(T=T, R=int(R), dims=tuple(dims), out_col=col, table_off=off,
                                dense=dense, mask_id=0 if mask_ids is None else int(mask_ids[gi])))
            off += T
            col += F
    return levels


class COracle:
    def __init__(self) -> None:
        build_oracle()
        self.lib = ctypes.CDLL(_LIB_PATH)
        self.lib.orc_corners.restype = ctypes.c_int

    @staticmethod
    def _pack(levels: Sequence[Level]):
        arr = (_OrcLevel * len(levels))()
        for i, lv in enumerate(levels):
            arr[i].table_off = lv.table_off
            arr[i].T = lv.T
            arr[i].R = lv.R
            arr[i].out_col = lv.out_col
            arr[i].D = lv.D
            for d in range(lv.D):
                arr[i].dim[d] = lv.dims[d]
            arr[i].dense = int(lv.dense)
            arr[i].mask_id = lv.mask_id
        return arr

    @staticmethod
    def _p(a: Optional[np.ndarray]):
        return None if a is None else a.ctypes.data_as(ctypes.c_void_p)

    def hash_indices(self, levels, coords: np.ndarray, linear: bool) -> List[np.ndarray]:
        coords = np.ascontiguousarray(coords, dtype=np.float32)
        n, stride = coords.shape
        counts = [(1 << lv.D) if linear else 1 for lv in levels]
        out = np.empty(sum(counts) * n, dtype=np.int64)
        self.lib.orc_hash_indices(self._pack(levels), len(levels), self._p(coords),
                                  ctypes.c_int64(n), stride, int(linear), self._p(out))
        res, o = [], 0
        for c in counts:
            res.append(out[o:o + c * n].reshape(c, n))
            o += c * n
        return res

    def forward(self, levels, coords: np.ndarray, tables: np.ndarray, F: int, linear: bool,
                out_scale: float = 1.0, mask_a=None, mask_b=None, out_cols: Optional[int] = None):
        coords = np.ascontiguousarray(coords, dtype=np.float32)
        n, stride = coords.shape
        assert tables.dtype in (np.float32, np.float16) and tables.flags.c_contiguous
        dt = 0 if tables.dtype == np.float32 else 1
        cols = out_cols if out_cols is not None else max(lv.out_col for lv in levels) + F
        out = np.zeros((n, cols), dtype=np.float32)
        ma = None if mask_a is None else np.ascontiguousarray(mask_a, dtype=np.uint8)
        mb = None if mask_b is None else np.ascontiguousarray(mask_b, dtype=np.uint8)
        self.lib.orc_forward(self._pack(levels), len(levels), self._p(coords), ctypes.c_int64(n),
                             stride, self._p(tables), dt, F, ctypes.c_float(out_scale),
                             self._p(ma), self._p(mb), self._p(out), ctypes.c_int64(cols),
                             int(linear))
        return out

    def backward(self, levels, coords: np.ndarray, grad_out: np.ndarray, n_entries: int, F: int,
                 linear: bool, mask_a=None, mask_b=None) -> np.ndarray:
        coords = np.ascontiguousarray(coords, dtype=np.float32)
        grad_out = np.ascontiguousarray(grad_out, dtype=np.float32)
        n, stride = coords.shape
        grads = np.zeros((n_entries, F), dtype=np.float64)
        ma = None if mask_a is None else np.ascontiguousarray(mask_a, dtype=np.uint8)
        mb = None if mask_b is None else np.ascontiguousarray(mask_b, dtype=np.uint8)
        self.lib.orc_backward(self._pack(levels), len(levels), self._p(coords), ctypes.c_int64(n),
                              stride, F, self._p(ma), self._p(mb), self._p(grad_out),
                              ctypes.c_int64(grad_out.shape[1]), self._p(grads), int(linear))
        return grads

    def geo2xyzt_norm(self, geo: np.ndarray, t: Optional[np.ndarray], bbox_min, bbox_span, t_range,
                      want_ecef: bool = False):
        geo = np.ascontiguousarray(geo, dtype=np.float64)
        n = geo.shape[0]
        tt = None if t is None else np.ascontiguousarray(t, dtype=np.float64)
        bmin = np.ascontiguousarray(bbox_min, dtype=np.float64)
        bspan = np.ascontiguousarray(bbox_span, dtype=np.float64)
        tr = np.ascontiguousarray(t_range, dtype=np.float64)
        xyzt = np.empty((n, 4), dtype=np.float32)
        ecef = np.empty((n, 3), dtype=np.float64) if want_ecef else None
        self.lib.orc_geo2xyzt_norm(self._p(geo), self._p(tt), self._p(bmin), self._p(bspan),
                                   self._p(tr), self._p(xyzt), self._p(ecef), ctypes.c_int64(n))
        return (xyzt, ecef) if want_ecef else xyzt


# --------------------------------------------------------------------------------------------
# numpy restatement (nearest mode) -- independent of the C file, used to cross-check it
# --------------------------------------------------------------------------------------------
def numpy_hash_coords(coords: np.ndarray, resolution: int, table_size: int) -> np.ndarray:
    """models/hash_encoding.py:47-74 in numpy: fp32 scale, floor, int64 prime-XOR, floor-mod."""
    coords = np.asarray(coords, dtype=np.float32)
    scaled = coords * np.float32(resolution)                       # :58
    grid = np.floor(scaled).astype(np.int64)                       # :61
    h = np.zeros(coords.shape[0], dtype=np.int64)
    with np.errstate(over="ignore"):
        for i in range(coords.shape[1]):                           # :68-69
            h ^= grid[:, i] * PRIMES[i]
    return np.mod(h, np.int64(table_size))                         # :72 (numpy % is floor-mod)


def numpy_forward_nearest(coords: np.ndarray, tables: Sequence[np.ndarray],
                          resolutions: Sequence[int]) -> np.ndarray:
    """models/hash_encoding.py:76-96: per-level gather, concatenated level-major."""
    feats = []
    for tab, R in zip(tables, resolutions):
        feats.append(tab[numpy_hash_coords(coords, R, tab.shape[0])])
    return np.concatenate(feats, axis=-1)


def numpy_backward_nearest(coords: np.ndarray, grad_out: np.ndarray, resolutions: Sequence[int],
                           T: int, F: int) -> List[np.ndarray]:
    """aten::embedding_dense_backward per level, accumulated in fp64."""
    grads = []
    for l, R in enumerate(resolutions):
        g = np.zeros((T, F), dtype=np.float64)
        np.add.at(g, numpy_hash_coords(coords, R, T), grad_out[:, l * F:(l + 1) * F].astype(np.float64))
        grads.append(g)
    return grads
