"""Tensor-core MLP head: Linear -> LayerNorm -> GELU -> Linear at hidden = out = 64 in ONE kernel.

Mirrors ``Grid4DEncoder.projection`` (/root/reference/models/encoders.py:47-52).  The parameters stay in an
ordinary ``nn.Sequential`` (same state-dict keys as the reference: ``projection.0.weight`` ...); this module
only changes how the forward is computed: ``de_mlp_head_forward`` (deepearth_b200/csrc/mlp_head.cu) runs both
GEMMs as bf16 ``tcgen05.mma`` with fp32 accumulators in tensor memory, and the hidden row never leaves the SM.
The backward pass is a second tcgen05 kernel (``de_mlp_head_backward``, csrc/mlp_head_bwd2.cu: hidden-row recompute,
LayerNorm / activation backward in registers, weight gradients accumulated in tensor memory); with
``fused_backward=False`` it recomputes through the fp32 PyTorch modules (cuBLAS) instead.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch
import torch.nn as nn

from . import _lib
from .functional import _ptr, _require_cuda, _stream_ptr

_ACT = {"none": 0, "relu": 1, "gelu": 2}


def mlp_head_supported(in_dim: int, hidden: int, out_dim: int) -> bool:
    return bool(_lib.load().de_mlp_head_supported(int(in_dim), int(hidden), int(out_dim)))


def mlp_head_forward(x: torch.Tensor, w1_bf16: torch.Tensor, b1: Optional[torch.Tensor],
                     ln_gamma: Optional[torch.Tensor], ln_beta: Optional[torch.Tensor], ln_eps: float,
                     layernorm: bool, activation: str, w2_bf16: torch.Tensor, b2: Optional[torch.Tensor],
                     out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x (n, in_dim) fp32 CUDA -> (n, 64) fp32.  Weights are bf16 in nn.Linear.weight layout."""
    _require_cuda(x, "x")
    if x.dim() != 2 or x.dtype != torch.float32:
        raise TypeError("x must be a 2-D float32 tensor")
    if x.stride(1) != 1 or x.stride(0) % 4 or x.data_ptr() % 16:
        x = x.contiguous()
    n, in_dim = x.shape
    hidden, out_dim = w1_bf16.shape[0], w2_bf16.shape[0]
    if w1_bf16.dtype != torch.bfloat16 or w2_bf16.dtype != torch.bfloat16:
        raise TypeError("weights must be bfloat16")
    if w1_bf16.shape != (hidden, in_dim) or w2_bf16.shape != (out_dim, hidden):
        raise ValueError("weight shapes do not chain")
    if out is None:
        out = torch.empty((n, out_dim), dtype=torch.float32, device=x.device)
    if n == 0:
        return out
    lib = _lib.load()
    with torch.cuda.device(x.device):
        rc = lib.de_mlp_head_forward(_ptr(x), n, in_dim, x.stride(0), _ptr(w1_bf16.contiguous()), _ptr(b1),
                                     _ptr(ln_gamma), _ptr(ln_beta), float(ln_eps), int(bool(layernorm)),
                                     _ACT[activation], _ptr(w2_bf16.contiguous()), _ptr(b2), hidden, out_dim,
                                     _ptr(out), out.stride(0), _stream_ptr(x.device))
    _lib.check(rc, "de_mlp_head_forward")
    return out


def mlp_head_backward(x: torch.Tensor, grad_y: torch.Tensor, w1_bf16: torch.Tensor, w1t_bf16: torch.Tensor,
                      w2t_bf16: torch.Tensor, b1, ln_gamma, ln_beta, ln_eps: float, layernorm: bool, activation: str,
                      has_b2: bool = True, group_width: int = 0):
    """One tcgen05 kernel: (grad_x, grad_w1, grad_b1, grad_gamma, grad_beta, grad_w2, grad_b2), fp32.

    ``group_width`` = W > 0: grad_x's memory is written group-major, g[c // W][p][c % W], for
    ``functional.encode_backward(..., group_width=W)``; the returned tensor still has shape (n, in_dim)."""
    _require_cuda(x, "x")
    _require_cuda(grad_y, "grad_y")
    n, in_dim = x.shape
    hidden = w1_bf16.shape[0]
    if x.stride(1) != 1 or x.stride(0) % 4 or x.data_ptr() % 16:
        x = x.contiguous()
    grad_y = grad_y.float()
    if grad_y.stride(1) != 1 or grad_y.stride(0) % 4 or grad_y.data_ptr() % 16:
        grad_y = grad_y.contiguous()
    dev = x.device
    gx = torch.empty((n, in_dim), dtype=torch.float32, device=dev)
    # one zeroed buffer for every accumulated gradient: w1 | w2 | b1 | gamma | beta | b2
    sizes = [hidden * in_dim, hidden * hidden, hidden, hidden, hidden, hidden]
    flat = torch.zeros(sum(sizes), dtype=torch.float32, device=dev)
    gw1, gw2, gb1, gg, gbeta, gb2 = torch.split(flat, sizes)
    if n:
        lib = _lib.load()
        with torch.cuda.device(dev):
            rc = lib.de_mlp_head_backward(_ptr(x), _ptr(grad_y), n, in_dim, x.stride(0), grad_y.stride(0), _ptr(w1_bf16),
                                          _ptr(w1t_bf16), _ptr(w2t_bf16), _ptr(b1), _ptr(ln_gamma), _ptr(ln_beta),
                                          float(ln_eps), int(bool(layernorm)), _ACT[activation], hidden, hidden,
                                          _ptr(gx), -int(group_width) if group_width else gx.stride(0), _ptr(gw1), _ptr(gb1), _ptr(gg), _ptr(gbeta),
                                          _ptr(gw2), _ptr(gb2), _stream_ptr(dev))
        _lib.check(rc, "de_mlp_head_backward")
    return gx, gw1.view(hidden, in_dim), gb1, gg, gbeta, gw2.view(hidden, hidden), gb2


def projection_supported(in_dim: int, hidden: int, out_dim: int) -> bool:
    return bool(_lib.load().de_projection_supported(int(in_dim), int(hidden), int(out_dim)))


def gemm_bf16(a: torch.Tensor, b: torch.Tensor, bias: Optional[torch.Tensor] = None, out_dtype=torch.float32,
              out: Optional[torch.Tensor] = None, transposed: bool = False, accumulate: bool = False) -> torch.Tensor:
    """TMA / tcgen05 GEMM kernel (csrc/proj_kernels.cu), bf16 operands, fp32 accumulation.
    transposed=False: C (M, N) = A (M, K) @ B (N, K)^T + bias;  transposed=True: C (M, N) = A (K, M)^T @ B (K, N) (the
    contraction runs over rows).  accumulate=True adds into ``out`` (fp32), the K range split over the CTAs."""
    _require_cuda(a, "a")
    _require_cuda(b, "b")
    kd = 0 if transposed else 1
    if a.dtype != torch.bfloat16 or b.dtype != torch.bfloat16 or a.dim() != 2 or b.dim() != 2 or a.shape[kd] != b.shape[kd]:
        raise TypeError("gemm_bf16 expects bf16 matrices with a common contraction extent")
    if a.stride(1) != 1 or a.stride(0) % 8 or a.data_ptr() % 16:
        a = a.contiguous()
    if b.stride(1) != 1 or b.stride(0) % 8 or b.data_ptr() % 16:
        b = b.contiguous()
    k = a.shape[kd]
    m, n = a.shape[1 - kd], b.shape[1 - kd]
    if out is None:
        out = (torch.zeros if accumulate else torch.empty)((m, n), dtype=out_dtype, device=a.device)
    code = {torch.float32: _lib.DE_F32, torch.bfloat16: _lib.DE_BF16}[out.dtype]
    if m and k:
        with torch.cuda.device(a.device):
            rc = _lib.load().de_gemm_bf16(_ptr(a), a.stride(0), _ptr(b), b.stride(0), m, n, k,
                                          _ptr(bias.float().contiguous() if bias is not None else None), _ptr(out),
                                          out.stride(0), code, int(transposed), int(accumulate), _stream_ptr(a.device))
        _lib.check(rc, "de_gemm_bf16")
    return out


def projection_forward(x: torch.Tensor, w1_bf16: torch.Tensor, b1: torch.Tensor, ln_gamma: torch.Tensor,
                       ln_beta: torch.Tensor, ln_eps: float, w2_bf16: torch.Tensor, b2: Optional[torch.Tensor],
                       want_stats: bool = False, out_dtype: torch.dtype = torch.float32):
    """x (n, in_dim) fp32 -> (y (n, out) fp32 | bf16, hidden (n, H) bf16 = GELU(LN(W1 x + b1)), stats (n, 2) | None).
    ``out_dtype=torch.bfloat16``: y rounded once in the GEMM epilogue (what nn.Linear returns under autocast)."""
    _require_cuda(x, "x")
    if x.dim() != 2 or x.dtype != torch.float32:
        raise TypeError("x must be a 2-D float32 tensor")
    if x.stride(1) != 1 or x.stride(0) % 4 or x.data_ptr() % 16:
        x = x.contiguous()
    n, in_dim = x.shape
    hidden, out_dim = w1_bf16.shape[0], w2_bf16.shape[0]
    if out_dtype not in (torch.float32, torch.bfloat16):
        raise TypeError("out_dtype must be torch.float32 or torch.bfloat16")
    y = torch.empty((n, out_dim), dtype=out_dtype, device=x.device)
    hid = torch.empty((n, hidden), dtype=torch.bfloat16, device=x.device)
    stats = torch.empty((n, 2), dtype=torch.float32, device=x.device) if want_stats else None
    if n:
        with torch.cuda.device(x.device):
            rc = _lib.load().de_projection_forward_ex(
                _ptr(x), n, in_dim, x.stride(0), _ptr(w1_bf16), _ptr(b1), _ptr(ln_gamma), _ptr(ln_beta), float(ln_eps),
                _ptr(w2_bf16), _ptr(b2), hidden, out_dim, _ptr(hid), _ptr(stats), _ptr(y),
                _lib.DE_BF16 if out_dtype == torch.bfloat16 else _lib.DE_F32, y.stride(0), _stream_ptr(x.device))
        _lib.check(rc, "de_projection_forward_ex")
    return y, hid, stats


def projection_backward(x, grad_y, w1_bf16, w1t_bf16, w2t_bf16, b1, ln_gamma, ln_beta, hidden_bf16, stats,
                        want_grad_x: bool = True):
    """Backward of ``projection_forward``: (grad_x | None, grad_w1, grad_b1, grad_gamma, grad_beta, grad_w2, grad_b2), fp32.
    ``grad_y`` fp32, or bf16 (the gradient of a bf16 output: consumed in place, no cast pass)."""
    _require_cuda(x, "x")
    _require_cuda(grad_y, "grad_y")
    n, in_dim = x.shape
    hidden, out_dim = w1_bf16.shape[0], w2t_bf16.shape[1]
    if x.stride(1) != 1 or x.stride(0) % 4 or x.data_ptr() % 16:
        x = x.contiguous()
    if grad_y.dtype != torch.bfloat16:
        grad_y = grad_y.float()
    if grad_y.stride(1) != 1 or grad_y.stride(0) % (8 if grad_y.dtype == torch.bfloat16 else 4) or grad_y.data_ptr() % 16:
        grad_y = grad_y.contiguous()
    dev = x.device
    gx = torch.empty((n, in_dim), dtype=torch.float32, device=dev) if want_grad_x else None
    sizes = [hidden * in_dim, out_dim * hidden, hidden, hidden, hidden, out_dim]
    flat = torch.zeros(sum(sizes), dtype=torch.float32, device=dev)
    gw1, gw2, gb1, gg, gbeta, gb2 = torch.split(flat, sizes)
    if n:
        lib = _lib.load()
        ws_bytes = int(lib.de_projection_backward_workspace_bytes(n, in_dim, hidden, out_dim))
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            rc = lib.de_projection_backward_ex(
                _ptr(x), _ptr(grad_y), _lib.DE_BF16 if grad_y.dtype == torch.bfloat16 else _lib.DE_F32, n, in_dim, x.stride(0),
                grad_y.stride(0), _ptr(w1_bf16), _ptr(w1t_bf16), _ptr(w2t_bf16),
                _ptr(b1), _ptr(ln_gamma), _ptr(ln_beta), hidden, out_dim, _ptr(hidden_bf16), _ptr(stats),
                _ptr(gx), gx.stride(0) if gx is not None else in_dim, _ptr(gw1), _ptr(gb1), _ptr(gg), _ptr(gbeta), _ptr(gw2),
                _ptr(gb2), _ptr(ws), ws_bytes, _stream_ptr(dev))
        _lib.check(rc, "de_projection_backward_ex")
    return gx, gw1.view(hidden, in_dim), gb1, gg, gbeta, gw2.view(out_dim, hidden), gb2


# token -> (data_ptr, W) of a gradient buffer the head backward wrote group-major; consumed (popped) by
# functional.Grid4DFunction.backward of the same Grid4DEncoder.forward call.  The features tensor between the two
# autograd nodes is internal to that call, so the gradient reaches the consumer as the very buffer written here;
# the consumer checks the pointer and fails loudly otherwise.
GROUPED_GRADS: dict = {}


class _HeadFunction(torch.autograd.Function):
    """forward = fused tensor-core kernel; backward = recompute through the fp32 modules (cuBLAS)."""

    @staticmethod
    def forward(ctx, x, head, *params):
        ctx.head = head
        ctx.grouped = head._grouped_next     # (token, W) set by Grid4DEncoder.forward for this call, or None
        head._grouped_next = None
        if head.wide and head.fused_backward:
            y, hid, stats = head.fused_wide(x, want_stats=True, out_dtype=head._out_dtype_next)
            ctx.save_for_backward(x, hid, stats)        # the bf16 activations and the LayerNorm row statistics
            return y
        ctx.save_for_backward(x)
        return head.fused(x)

    @staticmethod
    def backward(ctx, grad_y):
        x = ctx.saved_tensors[0]
        head = ctx.head
        params = list(head.seq.parameters())
        if head.fused_backward and head.wide and len(ctx.saved_tensors) == 3:
            by_id = head.backward_wide(x, ctx.saved_tensors[1], ctx.saved_tensors[2], grad_y, ctx.needs_input_grad[0])
            gx = by_id.pop("x")
            gp = [by_id.get(id(p)) if ctx.needs_input_grad[2 + i] else None for i, p in enumerate(params)]
            return (gx, None, *gp)
        if head.fused_backward and not head.wide:
            W = ctx.grouped[1] if (ctx.grouped and ctx.needs_input_grad[0]) else 0
            by_id = head.backward_fused(x, grad_y, group_width=W)
            gx = by_id.pop("x") if ctx.needs_input_grad[0] else None
            if W:
                # tell the hash-grid backward (Grid4DFunction.backward, same token) that this buffer is group-major.
                # Entries are popped by their consumer; one whose consumer never ran (autograd.grad on the head's
                # parameters only, an exception) must not outlive the next call
                if len(GROUPED_GRADS) >= 4:
                    GROUPED_GRADS.clear()
                GROUPED_GRADS[ctx.grouped[0]] = (gx.data_ptr(), W)
            gp = [by_id.get(id(p)) if ctx.needs_input_grad[2 + i] else None for i, p in enumerate(params)]
            return (gx, None, *gp)
        with torch.enable_grad():
            xd = x.detach().requires_grad_(ctx.needs_input_grad[0])
            y = head.seq(xd)
            wanted = ([xd] if ctx.needs_input_grad[0] else []) + [p for i, p in enumerate(params) if ctx.needs_input_grad[2 + i]]
            grads = list(torch.autograd.grad(y, wanted, grad_y, allow_unused=True))
        gx = grads.pop(0) if ctx.needs_input_grad[0] else None
        gp = [grads.pop(0) if ctx.needs_input_grad[2 + i] else None for i in range(len(params))]
        return (gx, None, *gp)


class FusedProjection:
    """Computes an ``nn.Sequential(Linear, LayerNorm, GELU|ReLU, Linear)`` (or without the LayerNorm) with the
    fused kernel.  Holds no parameters of its own: bf16 copies of the Linear weights are re-derived whenever the
    fp32 parameters change."""

    def __init__(self, seq: nn.Sequential, fused_backward: bool = True):
        self.fused_backward = bool(fused_backward)   # False: gradients by fp32 recompute through the modules (cuBLAS)
        mods = list(seq)
        if not (isinstance(mods[0], nn.Linear) and isinstance(mods[-1], nn.Linear)):
            raise TypeError("expected Linear ... Linear")
        mid = mods[1:-1]
        self.seq = seq
        self.lin1, self.lin2 = mods[0], mods[-1]
        self.norm = next((m for m in mid if isinstance(m, nn.LayerNorm)), None)
        acts = [m for m in mid if not isinstance(m, nn.LayerNorm)]
        if len(acts) > 1:
            raise TypeError("at most one activation between the Linear layers")
        if not acts:
            self.activation = "none"
        elif isinstance(acts[0], nn.GELU) and getattr(acts[0], "approximate", "none") == "none":
            self.activation = "gelu"
        elif isinstance(acts[0], nn.ReLU):
            self.activation = "relu"
        else:
            raise TypeError(f"unsupported activation {acts[0]!r}")
        self._key = None
        self._w1 = self._w2 = self._w1t = self._w2t = None
        self._grouped_next = None
        self._out_dtype_next = torch.float32
        # hidden widths other than 64 (768 by default, 2048 in the integrated model) take the two-kernel path of
        # csrc/proj_kernels.cu; it needs the full Linear-LayerNorm-GELU-Linear structure with biases
        self.wide = (self.lin1.out_features != 64 and self.norm is not None and self.activation == "gelu"
                     and self.norm.elementwise_affine and self.lin1.bias is not None
                     and self.lin2.in_features == self.lin1.out_features)

    def supported(self) -> bool:
        if self.wide:
            return projection_supported(self.lin1.in_features, self.lin1.out_features, self.lin2.out_features)
        return mlp_head_supported(self.lin1.in_features, self.lin1.out_features, self.lin2.out_features)

    def _weights(self):
        key = (self.lin1.weight._version, self.lin2.weight._version, self.lin1.weight.data_ptr(), self.lin2.weight.data_ptr())
        if key != self._key:
            with torch.no_grad():
                self._w1 = self.lin1.weight.detach().to(torch.bfloat16).contiguous()
                self._w2 = self.lin2.weight.detach().to(torch.bfloat16).contiguous()
                self._w1t = self._w1.t().contiguous()
                self._w2t = self._w2.t().contiguous()
            self._key = key
        return self._w1, self._w2

    def backward_fused(self, x: torch.Tensor, grad_y: torch.Tensor, group_width: int = 0) -> dict:
        """Gradients by the tcgen05 backward kernel, keyed by id(parameter) (+ "x")."""
        w1, _ = self._weights()
        n = self.norm
        gx, gw1, gb1, gg, gbeta, gw2, gb2 = mlp_head_backward(
            x, grad_y, w1, self._w1t, self._w2t, self.lin1.bias, n.weight if n is not None else None,
            n.bias if n is not None else None, n.eps if n is not None else 1e-5, n is not None, self.activation,
            group_width=group_width)
        out = {"x": gx, id(self.lin1.weight): gw1, id(self.lin2.weight): gw2}
        if self.lin1.bias is not None:
            out[id(self.lin1.bias)] = gb1
        if self.lin2.bias is not None:
            out[id(self.lin2.bias)] = gb2
        if n is not None and n.weight is not None:
            out[id(n.weight)] = gg
            out[id(n.bias)] = gbeta
        return out

    def fused_wide(self, x: torch.Tensor, want_stats: bool = False, out_dtype: torch.dtype = torch.float32):
        w1, w2 = self._weights()
        n = self.norm
        return projection_forward(x, w1, self.lin1.bias, n.weight, n.bias, n.eps, w2, self.lin2.bias, want_stats=want_stats,
                                  out_dtype=out_dtype)

    def backward_wide(self, x, hid, stats, grad_y, want_grad_x: bool = True) -> dict:
        """Gradients of the wide projection by the tcgen05 kernels (csrc/proj_bwd_kernels.cu), keyed by id(parameter) (+ "x")."""
        w1, _ = self._weights()
        n = self.norm
        gx, gw1, gb1, gg, gbeta, gw2, gb2 = projection_backward(x, grad_y, w1, self._w1t, self._w2t, self.lin1.bias, n.weight,
                                                               n.bias, hid, stats, want_grad_x)
        out = {"x": gx, id(self.lin1.weight): gw1, id(self.lin2.weight): gw2, id(self.lin1.bias): gb1,
               id(n.weight): gg, id(n.bias): gbeta}
        if self.lin2.bias is not None:
            out[id(self.lin2.bias)] = gb2
        return out

    def fused(self, x: torch.Tensor) -> torch.Tensor:
        w1, w2 = self._weights()
        n = self.norm
        if self.wide:
            return projection_forward(x, w1, self.lin1.bias, n.weight, n.bias, n.eps, w2, self.lin2.bias, out_dtype=self._out_dtype_next)[0]
        return mlp_head_forward(x, w1, self.lin1.bias, n.weight if n is not None else None,
                                n.bias if n is not None else None, n.eps if n is not None else 1e-5,
                                n is not None, self.activation, w2, self.lin2.bias)

    def __call__(self, x: torch.Tensor, out_dtype: torch.dtype = torch.float32) -> torch.Tensor:
        """``out_dtype=torch.bfloat16`` (wide path only): y leaves the second GEMM in bf16 and its gradient is taken in bf16 --
        nn.Linear's behaviour under ``torch.autocast(bfloat16)``, without a cast pass on either side."""
        if out_dtype != torch.float32 and not self.wide:
            raise TypeError("out_dtype other than float32 needs the wide projection path")
        self._out_dtype_next = out_dtype
        try:
            return self._call(x)
        finally:
            self._out_dtype_next = torch.float32

    def _call(self, x: torch.Tensor) -> torch.Tensor:
        needs_grad = torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in self.seq.parameters()))
        if needs_grad:
            return _HeadFunction.apply(x, self, *self.seq.parameters())
        return self.fused(x)
