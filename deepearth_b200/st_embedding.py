"""SpatialTemporalEmbedding -- B200 drop-in for /root/reference/models/cross_modal_fusion.py:99-165.

Same constructor, parameter names / shapes / init order (state dicts interchange: ``spatial_embed_x``,
``spatial_embed_y`` (1, R, D/4), ``temporal_embed.{0,2}.{weight,bias}``, ``modality_embeds.<name>`` (1, 1, D)) and the same
``forward(tokens, spatial_positions, temporal_positions, modality_name)``.  The forward is two launches of
``de_st_embedding_forward`` (csrc/st_embed.cu): one streaming pass for the table gathers + modality + residual, one
fp32 tile GEMM for the temporal MLP whose left operand GELU(t * w1 + b1) is generated on the fly; the backward is
``de_st_embedding_backward``.  CUDA tensors only -- there is no CPU path.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from . import _lib
from .functional import _ptr, _require_cuda, _stream_ptr


def _rows(x: torch.Tensor, n: int, min_cols: int, name: str) -> torch.Tensor:
    _require_cuda(x, name)
    if x.dtype != torch.float32:
        raise TypeError(f"{name} must be float32, got {x.dtype}")
    x = x.reshape(n, -1)
    if x.shape[1] < min_cols:
        raise ValueError(f"{name} needs at least {min_cols} trailing column(s)")
    if x.stride(1) != 1 or (n > 1 and x.stride(0) < x.shape[1]):
        x = x.contiguous()
    return x


class _StEmbedFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, tokens, pos, t, ex, ey, w1, b1, w2, b2, mod, check, use_tc=True):
        D = tokens.shape[-1]
        tok2 = tokens.reshape(-1, D)
        if not tok2.is_contiguous():
            tok2 = tok2.contiguous()
        n = tok2.shape[0]
        out = torch.empty_like(tok2)
        pos2 = _rows(pos, n, 2, "spatial_positions") if pos is not None else None
        t1 = _rows(t, n, 1, "temporal_positions")[:, 0].contiguous() if t is not None else None
        R = ex.shape[1]
        err = torch.zeros(1, dtype=torch.int32, device=tokens.device) if pos is not None else None
        w1v = w1.reshape(-1).contiguous()
        lib = _lib.load() if n else None
        # temporal MLP on the tensor cores (split-bf16 products, fp32 accuracy) when the width allows it
        tc = bool(n) and use_tc and t1 is not None and bool(lib.de_st_embedding_tc_supported(D))
        if n:
            args = (_ptr(tok2), n, D, _ptr(pos2), pos2.stride(0) if pos2 is not None and n > 1 else (pos2.shape[1] if pos2 is not None else 0),
                    _ptr(t1), _ptr(ex.contiguous()), _ptr(ey.contiguous()), R, _ptr(w1v), _ptr(b1.contiguous()),
                    _ptr(w2.contiguous()), _ptr(b2.contiguous()), _ptr(mod.reshape(-1).contiguous() if mod is not None else None),
                    _ptr(out), _ptr(err))
            with torch.cuda.device(tokens.device):
                if tc:
                    ws_bytes = int(lib.de_st_embedding_tc_workspace_bytes(n, D, 0))
                    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=tokens.device)
                    rc = lib.de_st_embedding_forward_tc(*args, _ptr(ws), ws_bytes, _stream_ptr(tokens.device))
                else:
                    rc = lib.de_st_embedding_forward(*args, _stream_ptr(tokens.device))
            _lib.check(rc, "de_st_embedding_forward_tc" if tc else "de_st_embedding_forward")
            if check and err is not None and int(err.item()):
                raise IndexError(f"SpatialTemporalEmbedding: a spatial index (position * {R}).long() is out of range for "
                                 f"{R} table rows (models/cross_modal_fusion.py:145-149 raises here too)")
        ctx.save_for_backward(pos2, t1, w1v, b1, w2)
        ctx.meta = (tuple(tokens.shape), R, ex.shape, w1.shape, mod.shape if mod is not None else None)
        ctx.use_tc = use_tc
        return out.view(tokens.shape)

    @staticmethod
    def backward(ctx, g):
        pos2, t1, w1v, b1, w2 = ctx.saved_tensors
        shape, R, ex_shape, w1_shape, mod_shape = ctx.meta
        D = shape[-1]
        g2 = g.reshape(-1, D)
        if g2.dtype != torch.float32 or not g2.is_contiguous():
            g2 = g2.float().contiguous()
        n = g2.shape[0]
        need = ctx.needs_input_grad
        dev = g.device
        z = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)      # noqa: E731
        d_ex = z(*ex_shape) if pos2 is not None and (need[3] or need[4]) else None
        d_ey = z(*ex_shape) if d_ex is not None else None
        temporal = t1 is not None and any(need[5:9])
        d_w1 = z(*w1_shape) if temporal else None
        d_b1 = z(D // 2) if temporal else None
        d_w2 = z(D // 2, D // 2) if temporal else None
        d_b2 = z(D // 2) if temporal else None
        d_mod = z(*mod_shape) if mod_shape is not None and need[9] else None
        if n:
            lib = _lib.load()
            tc = temporal and ctx.use_tc and bool(lib.de_st_embedding_tc_supported(D))
            args = (_ptr(g2), n, D, _ptr(pos2 if d_ex is not None else None),
                    (pos2.stride(0) if n > 1 else pos2.shape[1]) if pos2 is not None else 0,
                    _ptr(t1 if temporal else None), R, _ptr(w1v), _ptr(b1), _ptr(w2.contiguous()), _ptr(d_ex), _ptr(d_ey), _ptr(d_w1),
                    _ptr(d_b1), _ptr(d_w2), _ptr(d_b2), _ptr(d_mod))
            with torch.cuda.device(dev):
                if tc:
                    ws_bytes = int(lib.de_st_embedding_tc_workspace_bytes(n, D, 1))
                    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
                    rc = lib.de_st_embedding_backward_tc(*args, _ptr(ws), ws_bytes, _stream_ptr(dev))
                else:
                    scratch = torch.empty((n, D // 2), dtype=torch.float32, device=dev) if temporal else None
                    rc = lib.de_st_embedding_backward(*args, _ptr(scratch), _stream_ptr(dev))
            _lib.check(rc, "de_st_embedding_backward_tc" if tc else "de_st_embedding_backward")
        return (g if need[0] else None, None, None, d_ex, d_ey, d_w1, d_b1, d_w2, d_b2, d_mod, None, None)


class SpatialTemporalEmbedding(nn.Module):
    """Learnable spatial-temporal(-modal) embeddings added to universal tokens (cross_modal_fusion.py:99-165)."""

    def __init__(self, universal_dim: int, max_spatial_resolution: int = 64, check_indices: bool = True,
                 tensor_core_mlp: bool = True):
        super().__init__()
        if universal_dim % 4:
            raise ValueError(f"universal_dim={universal_dim} must be a multiple of 4")
        self.universal_dim = universal_dim
        self.check_indices = bool(check_indices)   # True: raise IndexError like the reference (one device->host read)
        # True: the temporal MLP's second Linear runs on the tensor cores as a split-bf16 product at fp32 accuracy
        # (csrc/st_embed_tc.cu; universal_dim % 128 == 0), False: the fp32 SIMT tile GEMM of csrc/st_embed.cu
        self.tensor_core_mlp = bool(tensor_core_mlp)
        # construction order = the reference's, so a shared seed gives identical parameters (:107-123)
        self.spatial_embed_x = nn.Parameter(torch.randn(1, max_spatial_resolution, universal_dim // 4) * 0.02)
        self.spatial_embed_y = nn.Parameter(torch.randn(1, max_spatial_resolution, universal_dim // 4) * 0.02)
        self.temporal_embed = nn.Sequential(
            nn.Linear(1, universal_dim // 2),
            nn.GELU(),
            nn.Linear(universal_dim // 2, universal_dim // 2),
        )
        self.modality_embeds = nn.ParameterDict()

    def add_modality(self, name: str):
        ref = self.spatial_embed_x
        self.modality_embeds[name] = nn.Parameter(
            torch.randn(1, 1, self.universal_dim, device=ref.device, dtype=ref.dtype) * 0.02)

    def forward(self, tokens: torch.Tensor, spatial_positions: Optional[torch.Tensor] = None,
                temporal_positions: Optional[torch.Tensor] = None, modality_name: Optional[str] = None) -> torch.Tensor:
        _require_cuda(tokens, "tokens")
        if tokens.dim() != 3 or tokens.shape[-1] != self.universal_dim:
            raise ValueError(f"tokens must be (B, N, {self.universal_dim}), got {tuple(tokens.shape)}")
        if tokens.dtype != torch.float32:
            raise TypeError(f"tokens must be float32, got {tokens.dtype}")
        for q, nm in ((spatial_positions, "spatial_positions"), (temporal_positions, "temporal_positions")):
            if q is not None and q.requires_grad:
                raise RuntimeError(f"{nm} requires grad: positions are data on this path and receive no gradient")
        mod = None
        if modality_name is not None and modality_name in self.modality_embeds:
            mod = self.modality_embeds[modality_name]
        lin1, lin2 = self.temporal_embed[0], self.temporal_embed[2]
        return _StEmbedFunction.apply(tokens, spatial_positions, temporal_positions, self.spatial_embed_x,
                                      self.spatial_embed_y, lin1.weight, lin1.bias, lin2.weight, lin2.bias, mod,
                                      self.check_indices, self.tensor_core_mlp)
