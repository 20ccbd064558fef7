"""Torch restatement of the reference encoder path (ORACLE / CPU BASELINE, not product code).

Two uses, both on the checker side only:

* ``--impl reference`` / ``cpu_baseline`` in bench.py: the reference is pure PyTorch eager
  (SURVEY.md section 0), so its CPU cost *is* this op sequence -- one fp32 multiply, floor,
  int64 cast, D int64 multiplies + XORs, a remainder and an embedding gather per level, a
  concat per encoder, and autograd's embedding_dense_backward.  ``RefHashEncoding`` issues
  exactly those ATen ops (models/hash_encoding.py:47-96) so the timing is representative.
* large-size parity on the GPU box: the same functions run on CUDA tensors through ATen, giving
  a device-resident oracle for 2^22-point checks that the C oracle would take minutes for.

The decomposed/interpolated mode follows the math stated in grid4d_oracle.c ("parity unpinned").
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch
import torch.nn as nn
import torch.nn.functional as Fnn

_PRIMES = (1, 2654435761, 805459861, 3674653429)  # models/hash_encoding.py:64


def hash_cells(cells: torch.Tensor, table_size: int) -> torch.Tensor:
    """int64 prime-XOR hash of integer cells (…, D) -> (…,), floor-mod T (:64-72)."""
    primes = torch.tensor(_PRIMES[: cells.shape[-1]], device=cells.device, dtype=torch.long)
    h = torch.zeros(cells.shape[:-1], device=cells.device, dtype=torch.long)
    for i in range(cells.shape[-1]):
        h ^= cells[..., i] * primes[i]
    return h % table_size


def hash_coords(coords: torch.Tensor, resolution: int, table_size: int) -> torch.Tensor:
    """models/hash_encoding.py:47-74."""
    return hash_cells(torch.floor(coords * resolution).long(), table_size)


class RefHashEncoding(nn.Module):
    """Same parameters / op sequence as the reference MultiResolutionHashEncoding (:11-96)."""

    def __init__(self, n_levels: int, n_features_per_level: int, resolutions: Sequence[int],
                 hash_table_size: int, coords_dim: int):
        super().__init__()
        self.resolutions = list(resolutions)[:n_levels]
        self.hash_table_size = hash_table_size
        self.coords_dim = coords_dim
        self.hash_tables = nn.ModuleList(
            [nn.Embedding(hash_table_size, n_features_per_level) for _ in self.resolutions])
        for emb in self.hash_tables:
            nn.init.uniform_(emb.weight, -1e-4, 1e-4)

    def forward(self, coords: torch.Tensor) -> torch.Tensor:
        per_level = [emb(hash_coords(coords, R, self.hash_table_size))
                     for emb, R in zip(self.hash_tables, self.resolutions)]
        return torch.cat(per_level, dim=-1)


class RefGrid4DFeatures(nn.Module):
    """Feature half of the reference Grid4DEncoder (models/encoders.py:24-39, :68-82)."""

    def __init__(self, n_spatial_levels=16, n_temporal_levels=8, n_features_per_level=2,
                 spatial_resolutions=None, temporal_resolutions=None, hash_table_size=2 ** 19):
        super().__init__()
        sres = spatial_resolutions or [2 ** i for i in range(4, 4 + n_spatial_levels)]
        tres = temporal_resolutions or [2 ** i for i in range(2, 2 + n_temporal_levels)]
        self.spatial_encoder = RefHashEncoding(n_spatial_levels, n_features_per_level, sres,
                                               hash_table_size, 3)
        self.temporal_encoder = RefHashEncoding(n_temporal_levels, n_features_per_level, tres,
                                                hash_table_size // 4, 1)

    def forward(self, xyzt, spatial_mask=None, temporal_mask=None):
        s = self.spatial_encoder(xyzt[:, :3])
        if spatial_mask is not None:
            s = s * spatial_mask.unsqueeze(-1).float()
        t = self.temporal_encoder(xyzt[:, 3:4])
        if temporal_mask is not None:
            t = t * temporal_mask.unsqueeze(-1).float()
        return torch.cat([s, t], dim=-1)


# ---------------------------------------------------------------------------------------------
# decomposed / interpolated mode (stated math, grid4d_oracle.c header)
# ---------------------------------------------------------------------------------------------
def level_indices_weights(coords: torch.Tensor, dims: Sequence[int], R: int, T: int, dense: bool):
    """-> (idx (B, 2^D) int64, w (B, 2^D) fp32) for one level."""
    D = len(dims)
    c = coords[:, list(dims)]
    s = c * float(R)                     # fp32 multiply (:58)
    fl = torch.floor(s)
    g = fl.long()                        # (:61)
    w = s - fl
    idxs, ws = [], []
    for k in range(1 << D):
        off = torch.tensor([(k >> d) & 1 for d in range(D)], device=coords.device)
        cell = g + off
        if dense:
            strides = torch.tensor([(R + 1) ** d for d in range(D)], device=coords.device)
            idx = (cell * strides).sum(-1) % T
        else:
            idx = hash_cells(cell, T)
        wk = None
        for d in range(D):
            u = w[:, d] if (k >> d) & 1 else (1.0 - w[:, d])
            wk = u if wk is None else wk * u
        idxs.append(idx)
        ws.append(wk)
    return torch.stack(idxs, 1), torch.stack(ws, 1)


def grid4d_forward(coords: torch.Tensor, tables: List[torch.Tensor], level_specs, linear: bool,
                   out_scale: float = 1.0) -> torch.Tensor:
    """level_specs: list of (dims, R, T, dense); tables[i]: (T, F).  Differentiable w.r.t. tables."""
    feats = []
    for tab, (dims, R, T, dense) in zip(tables, level_specs):
        if linear:
            idx, w = level_indices_weights(coords, dims, R, T, dense)
            v = Fnn.embedding(idx, tab.float() if tab.dtype != torch.float32 else tab)  # (B,C,F)
            f = (v * w.unsqueeze(-1)).sum(1)
        else:
            c = coords[:, list(dims)]
            f = Fnn.embedding(hash_coords(c, R, T), tab)
            f = f.float() if f.dtype != torch.float32 else f
        feats.append(f * out_scale if out_scale != 1.0 else f)
    return torch.cat(feats, dim=-1)


# ---------------------------------------------------------------------------------------------
# geodetic prologue (geospatial/geo2xyz.py:146-160, :327-332; utils.py:54-77)
# ---------------------------------------------------------------------------------------------
def geo2xyzt_norm(geo: torch.Tensor, t: Optional[torch.Tensor], bbox_min, bbox_span, t_range):
    a = 6378137.0
    f = 1 / 298.257223563
    e2 = 2 * f - f * f
    geo = geo.to(torch.float64)
    lat, lon, alt = torch.deg2rad(geo[..., 0]), torch.deg2rad(geo[..., 1]), geo[..., 2]
    sl, cl, so, co = torch.sin(lat), torch.cos(lat), torch.sin(lon), torch.cos(lon)
    N = a / torch.sqrt(1 - e2 * sl ** 2)
    xyz = torch.stack(((N + alt) * cl * co, (N + alt) * cl * so, (N * (1 - e2) + alt) * sl), -1)
    bmin = torch.as_tensor(bbox_min, dtype=torch.float64, device=geo.device)
    span = torch.as_tensor(bbox_span, dtype=torch.float64, device=geo.device)
    zero = span.abs() < 1e-9
    norm = torch.where(zero, torch.full_like(xyz, 0.5), (xyz - bmin) / torch.where(zero, torch.ones_like(span), span))
    if t is None:
        tn = torch.zeros(geo.shape[0], dtype=torch.float64, device=geo.device)
    else:
        tspan = float(t_range[1]) - float(t_range[0])
        tn = torch.full_like(t, 0.5, dtype=torch.float64) if abs(tspan) < 1e-9 else \
            (t.to(torch.float64) - float(t_range[0])) / tspan
    return torch.cat([norm, tn.unsqueeze(-1)], -1).float(), xyz
