"""Geodetic prologue -- host mirror of the reference GeospatialConverter's position path.

Reference: /root/reference/geospatial/geo2xyz.py (``geodetic_to_xyz`` :128-160, ``update_bbox``
:305-325, ``xyz_to_norm`` :327-332), utils.py ``_safe_div`` :54-77, data_structures.py ``BoundingBox``
:46-100.  The reference runs ~20 fp64 ATen ops and a device->host bbox round trip per call; here one
fused kernel (csrc/geo2xyz.cu) turns (lat, lon, alt[, t]) rows into packed fp32 (x, y, z, t) rows in
[0,1], ready for the Grid4D encoder.  Orientation matrices / CSV I/O are camera-pose tooling outside
this path and are not provided.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional, Sequence, Tuple

import torch

from . import _lib
from .functional import _ptr, _require_cuda, _stream_ptr

_DT = {torch.float64: _lib.DE_F64, torch.float32: _lib.DE_F32}


@dataclass
class BoundingBox:
    """Axis-aligned ECEF box used for normalisation (data_structures.py:46-100)."""
    min_x: float
    min_y: float
    min_z: float
    max_x: float
    max_y: float
    max_z: float

    @property
    def min_point(self) -> Tuple[float, float, float]:
        return (self.min_x, self.min_y, self.min_z)

    @property
    def max_point(self) -> Tuple[float, float, float]:
        return (self.max_x, self.max_y, self.max_z)

    @property
    def span(self) -> Tuple[float, float, float]:
        return (self.max_x - self.min_x, self.max_y - self.min_y, self.max_z - self.min_z)


def _dbl3(v: Sequence[float]):
    return (ctypes.c_double * len(v))(*[float(x) for x in v])


def _geo_arg(geo: torch.Tensor) -> torch.Tensor:
    _require_cuda(geo, "geo")
    if geo.shape[-1] != 3:
        raise ValueError("last dimension must be (lat, lon, alt)")   # geo2xyz.py:146
    if geo.dtype not in _DT:
        geo = geo.to(torch.float64)
    return geo.reshape(-1, 3).contiguous()


class GeospatialConverter:
    """geodetic (lat deg, lon deg, alt m) -> normalised fp32 xyzt.

    Two deliberate differences from the reference's ``xyz_to_norm`` (geo2xyz.py:327-332), both in INTEGRATION.md:
    * the reference always grows the bounding box with the batch; here that is ``update_bbox=True`` and the default is a
      PRE-FITTED box (a training loop must not change its normalisation batch by batch);
    * the reference returns fp64 whenever the span is large (``_norm_eff``); the encoder consumes fp32 coordinates
      (models/hash_encoding.py:58 multiplies in the input dtype and every caller feeds fp32), so the normalised rows are
      computed in fp64 and rounded ONCE to fp32 -- at R = 2^19 an fp32 coordinate in [0, 1] keeps ~5 fractional bits of
      the cell position, exactly as in the reference's own fp32 pipeline."""

    def __init__(self, *, device: str = "cuda", bbox: Optional[BoundingBox] = None,
                 t_range: Optional[Tuple[float, float]] = None) -> None:
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("deepearth_b200.GeospatialConverter runs on CUDA devices only")
        self._bbox = bbox
        self.t_range = t_range
        self._affine: Optional[Tuple[Tuple[float, float, float], Tuple[float, float, float]]] = None

    # dataset-statistics normalisation (SPECIFICATIONS.MD:707-733, 843-858) -------------------------------
    @classmethod
    def from_statistics(cls, center: Sequence[float], scale: float, *, bound: Optional[float] = None,
                        t_range: Optional[Tuple[float, float]] = None, device: str = "cuda") -> "GeospatialConverter":
        """Converter for the spec's ``xyz_norm = (ecef - center) / scale`` (``center`` = mean ECEF, ``scale`` =
        2 * max per-axis std, SPECIFICATIONS.MD:729, :853-854).  With ``bound`` the result is also mapped from
        [-bound, bound] to [0, 1] the way the upstream Grid4D encoder does before hashing (``grid4d_bound = 1.6``,
        SPECIFICATIONS.MD:294): ``(v + bound) / (2 bound)``.  Both are one affine map, folded into the offset and span
        the prologue kernel already divides by in fp64 -- no extra pass, no extra rounding step."""
        scale = float(scale)
        if not scale > 0.0:
            raise ValueError("scale must be positive")
        if bound is not None and not float(bound) > 0.0:
            raise ValueError("bound must be positive")                      # SPECIFICATIONS.MD:297-298
        c = tuple(float(v) for v in center)
        if len(c) != 3:
            raise ValueError("center must have three components")
        conv = cls(device=device, t_range=t_range)
        if bound is None:
            conv._affine = (c, (scale, scale, scale))
        else:
            b = float(bound)
            conv._affine = (tuple(v - b * scale for v in c), (2.0 * b * scale,) * 3)
        return conv

    def fit_statistics(self, geo: torch.Tensor) -> dict:
        """``center`` / ``scale`` of a point set as the spec's preprocessor computes them (SPECIFICATIONS.MD:843-858):
        mean ECEF and twice the largest per-axis standard deviation.  Returns Python floats (one host read)."""
        ecef = self.geodetic_to_xyz(geo).reshape(-1, 3)
        stats = torch.cat([ecef.mean(dim=0), (ecef.std(dim=0).max() * 2.0).reshape(1)]).tolist()
        return {"center": tuple(stats[:3]), "scale": stats[3]}

    @property
    def bbox(self) -> Optional[BoundingBox]:
        return self._bbox

    def reset_bbox(self) -> None:
        self._bbox = None

    def update_bbox(self, geo: torch.Tensor) -> BoundingBox:
        """Fold the ECEF extent of ``geo`` into the running box (geo2xyz.py:305-320)."""
        g = _geo_arg(geo)
        inf = float("inf")
        b = self._bbox
        init = [inf, inf, inf, -inf, -inf, -inf] if b is None else [*b.min_point, *b.max_point]
        mm = torch.tensor(init, dtype=torch.float64, device=g.device)
        with torch.cuda.device(g.device):
            rc = _lib.load().de_geo_bbox(_ptr(g), _DT[g.dtype], _ptr(mm), g.shape[0], _stream_ptr(g.device))
        _lib.check(rc, "de_geo_bbox")
        v = mm.tolist()   # the one host read-back, as in the reference (BoundingBox holds Python floats)
        self._bbox = BoundingBox(*v)
        return self._bbox

    def geodetic_to_xyzt_norm(self, geo: torch.Tensor, t: Optional[torch.Tensor] = None, *,
                              update_bbox: bool = False, return_ecef: bool = False):
        """(…,3) geodetic [+ (…,) time] -> (N,4) fp32 normalised rows (and optionally (N,3) fp64 ECEF).

        With ``update_bbox=True`` this is the reference's ``xyz_to_norm(geodetic_to_xyz(geo))`` (the box
        grows to include the batch first); otherwise the pre-fitted box is used, which is what a training
        loop wants.
        """
        g = _geo_arg(geo)
        if self._affine is not None:
            if update_bbox:
                raise ValueError("a statistics-normalised converter has no bounding box to update")
            offset, span = self._affine
        else:
            if update_bbox or self._bbox is None:
                self.update_bbox(g)
            offset, span = self._bbox.min_point, self._bbox.span
        n = g.shape[0]
        tt = None
        t_range = None
        if t is not None:
            _require_cuda(t, "t")
            tt = t.reshape(-1)
            if tt.dtype not in _DT:
                tt = tt.to(torch.float64)
            tt = tt.contiguous()
            if tt.shape[0] != n:
                raise ValueError("t must have one value per point")
            if self.t_range is None:
                raise ValueError("t given but the converter has no t_range")
            t_range = _dbl3(self.t_range)
        xyzt = torch.empty((n, 4), dtype=torch.float32, device=g.device)
        ecef = torch.empty((n, 3), dtype=torch.float64, device=g.device) if return_ecef else None
        with torch.cuda.device(g.device):
            rc = _lib.load().de_geo2xyzt_norm(_ptr(g), _DT[g.dtype], _ptr(tt), _DT[tt.dtype] if tt is not None else 0,
                                              _dbl3(offset), _dbl3(span), t_range,
                                              _ptr(xyzt), _ptr(ecef), n, _stream_ptr(g.device))
        _lib.check(rc, "de_geo2xyzt_norm")
        return (xyzt, ecef) if return_ecef else xyzt

    # timestamps from date columns (SPECIFICATIONS.MD:740-768) ----------------------------------------------
    def timestamps_from_components(self, components: torch.Tensor, *, default_month: int = 6, default_day: int = 15,
                                   default_hour: int = 12, normalise: bool = False) -> torch.Tensor:
        """(n, k) columns [year, month, day, hour, minute, second][:k] (int32 / int64 / float64) -> (n,) fp64 UTC epoch
        seconds, or with ``normalise=True`` (n,) fp32 ``(ts - t_min) / (t_max - t_min)`` -- what the spec's
        ``_process_temporal`` computes row by row with ``datetime(...).timestamp()``.  One launch, no host loop."""
        _require_cuda(components, "components")
        if components.dim() != 2 or not 1 <= components.shape[1] <= 6:
            raise ValueError("components must be (n, 1..6): year[, month, day, hour, minute, second]")
        if components.dtype == torch.int32:
            dt = _lib.DE_I32
        else:
            components = components.to(torch.float64)
            dt = _lib.DE_F64
        components = components.contiguous()
        n, k = components.shape
        if normalise and self.t_range is None:
            raise ValueError("normalise=True but the converter has no t_range")
        out = torch.empty(n, dtype=torch.float32 if normalise else torch.float64, device=components.device)
        with torch.cuda.device(components.device):
            rc = _lib.load().de_time_from_components(
                _ptr(components), dt, n, k, k, int(default_month), int(default_day), int(default_hour),
                _dbl3(self.t_range) if self.t_range is not None else None,
                None if normalise else _ptr(out), _ptr(out) if normalise else None, _stream_ptr(components.device))
        _lib.check(rc, "de_time_from_components")
        return out

    # dense fields (SPECIFICATIONS.MD:1963-2088) -----------------------------------------------------------
    def _norm_constants(self):
        if self._affine is not None:
            return self._affine
        if self._bbox is None:
            raise ValueError("the converter has no bounding box yet (update_bbox / pass bbox=...)")
        return self._bbox.min_point, self._bbox.span

    def field_desc(self, lat_range, lon_range, n_lat: int, n_lon: int, altitude: float = 0.0, time: Optional[float] = None):
        """ctypes descriptor of the regular grid ``np.linspace(*lat_range, n_lat)`` x ``np.linspace(*lon_range, n_lon)``."""
        if time is not None and self.t_range is None:
            raise ValueError("time given but the converter has no t_range")
        return _lib.DeFieldDesc(float(lat_range[0]), float(lat_range[1]), float(lon_range[0]), float(lon_range[1]),
                                float(altitude), float(time) if time is not None else 0.0, int(n_lat), int(n_lon),
                                1 if time is not None else 0, 0)

    def field_xyzt(self, lat_range, lon_range, n_lat: int, n_lon: int, altitude: float = 0.0,
                   time: Optional[float] = None) -> torch.Tensor:
        """(n_lat * n_lon, 4) normalised fp32 rows of a regular lat/lon grid (latitude-major), generated on the device:
        bit-identical to ``geodetic_to_xyzt_norm`` over the same ``np.linspace`` points."""
        offset, span = self._norm_constants()
        desc = self.field_desc(lat_range, lon_range, n_lat, n_lon, altitude, time)
        xyzt = torch.empty((int(n_lat) * int(n_lon), 4), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            rc = _lib.load().de_field_xyzt(ctypes.byref(desc), _dbl3(offset), _dbl3(span),
                                           _dbl3(self.t_range) if self.t_range is not None else None, _ptr(xyzt),
                                           _stream_ptr(self.device))
        _lib.check(rc, "de_field_xyzt")
        return xyzt

    # reference-named conveniences --------------------------------------------------------------------
    def geodetic_to_xyz(self, geo: torch.Tensor) -> torch.Tensor:
        """(…,3) geodetic -> (…,3) fp64 ECEF metres (geo2xyz.py:146-160, position branch only)."""
        g = _geo_arg(geo)
        keep = (self._bbox, self._affine)
        self._affine = None
        if self._bbox is None:
            self._bbox = BoundingBox(0, 0, 0, 1, 1, 1)
        try:
            _, ecef = self.geodetic_to_xyzt_norm(g, None, return_ecef=True)
        finally:
            self._bbox, self._affine = keep
        return ecef.reshape(*geo.shape[:-1], 3)


def geo2xyz(lat, lon, alt, *, device="cuda"):
    """``x, y, z = geo2xyz(lat, lon, elev)`` -- the function several reference callers import but the
    reference never defines (core/plant_data_loader.py:144, florida_plants_deepearth.py:93).  Returns ECEF
    metres (fp64 tensors) for scalars or tensors."""
    lat, lon, alt = (torch.as_tensor(v, dtype=torch.float64, device=device) for v in (lat, lon, alt))
    geo = torch.stack(torch.broadcast_tensors(lat, lon, alt), -1)
    xyz = GeospatialConverter(device=device).geodetic_to_xyz(geo)
    return xyz[..., 0], xyz[..., 1], xyz[..., 2]
