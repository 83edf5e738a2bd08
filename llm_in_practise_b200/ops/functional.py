"""Differentiable building blocks over the raw kernels, for models that compose their own attention
(MLA, notebook variants) instead of using the fused ``attention`` op.

    linear(x, w, b)            x @ w^T + b                          (tcgen05 GEMM fwd / dgrad / wgrad)
    matmul_nt(a, b, alpha)     alpha * a @ b^T, batched, strided     (scores = q k^T)
    matmul_nn(p, v)            p @ v, batched, strided               (context = P V)
    masked_softmax(s, causal, window)                                (no mask tensor)

All accept up to two leading batch dims with arbitrary strides (head-major views of token-major
buffers) and fall back to PyTorch on CPU.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _native as nv
from .attention import _softmax_bwd, _softmax_fwd
from .gemm import gemm
from .linear import _LinearFn


def linear(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor] = None) -> torch.Tensor:
    shape = x.shape
    x2 = x.reshape(-1, shape[-1])
    if not nv.use_native(x2):
        y = x2 @ w.t().to(x2.dtype)
        if b is not None:
            y = y + b.to(y.dtype)
        return y.view(*shape[:-1], w.shape[0])
    if not x2.is_contiguous():
        x2 = x2.contiguous()
    return _LinearFn.apply(x2, w, b, x2.dtype).view(*shape[:-1], w.shape[0])


class _MatmulNT(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b, alpha, out_dtype):
        ctx.save_for_backward(a, b)
        ctx.alpha = alpha
        return gemm(a, b, alpha=alpha, out_dtype=out_dtype)

    @staticmethod
    def backward(ctx, dy):
        a, b = ctx.saved_tensors
        dy = dy.to(a.dtype) if dy.dtype != a.dtype else dy
        if dy.stride(-1) != 1:
            dy = dy.contiguous()
        da = db = None
        if ctx.needs_input_grad[0]:
            da = gemm(dy, b, b_mn=True, alpha=ctx.alpha)            # dy @ b
        if ctx.needs_input_grad[1]:
            db = gemm(dy, a, a_mn=True, b_mn=True, alpha=ctx.alpha)  # dy^T @ a
        return da, db, None, None


class _MatmulNN(torch.autograd.Function):
    @staticmethod
    def forward(ctx, p, v):
        ctx.save_for_backward(p, v)
        return gemm(p, v, b_mn=True)

    @staticmethod
    def backward(ctx, dy):
        p, v = ctx.saved_tensors
        if dy.stride(-1) != 1:
            dy = dy.contiguous()
        dp = dv = None
        if ctx.needs_input_grad[0]:
            dp = gemm(dy, v, out_dtype=p.dtype)                      # dy @ v^T
        if ctx.needs_input_grad[1]:
            dv = gemm(p, dy, a_mn=True, b_mn=True)                   # p^T @ dy
        return dp, dv


def matmul_nt(a: torch.Tensor, b: torch.Tensor, alpha: float = 1.0, out_dtype=None) -> torch.Tensor:
    """alpha * a @ b^T for a [.., M, K], b [.., N, K]."""
    if not nv.use_native(a):
        return (alpha * (a.float() @ b.float().transpose(-1, -2))).to(out_dtype or a.dtype)
    return _MatmulNT.apply(a, b, alpha, out_dtype or a.dtype)


def matmul_nn(p: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
    """p @ v for p [.., M, K], v [.., K, N]."""
    if not nv.use_native(p):
        return (p.float() @ v.float()).to(p.dtype)
    return _MatmulNN.apply(p, v)


class _MaskedSoftmax(torch.autograd.Function):
    @staticmethod
    def forward(ctx, s, causal, window, out_dtype):
        p = _softmax_fwd(s.contiguous(), causal, window, out_dtype)
        ctx.save_for_backward(p)
        ctx.in_dtype = s.dtype
        return p

    @staticmethod
    def backward(ctx, dp):
        (p,) = ctx.saved_tensors
        dpf = dp.contiguous().float() if p.dtype != torch.float32 else dp.contiguous()
        return _softmax_bwd(p, dpf, 1.0, ctx.in_dtype if ctx.in_dtype == p.dtype else p.dtype).to(ctx.in_dtype), None, None, None


def masked_softmax(s: torch.Tensor, causal: bool = True, window: int = 0, out_dtype=None) -> torch.Tensor:
    """Row softmax of scores [.., rows, cols] with an implicit causal (+ sliding window) mask."""
    out_dtype = out_dtype or s.dtype
    ok = s.dtype == torch.float32 and out_dtype in (torch.float32, torch.bfloat16, torch.float16)
    if not (nv.use_native(s) and ok):
        rows, cols = s.shape[-2:]
        sf = s.float()
        if causal:
            i = torch.arange(rows, device=s.device)[:, None] + (cols - rows)
            j = torch.arange(cols, device=s.device)[None, :]
            mask = j > i
            if window > 0:
                mask = mask | (j < i - window + 1)
            sf = sf.masked_fill(mask, float("-inf"))
        return torch.softmax(sf, dim=-1).to(out_dtype)
    return _MaskedSoftmax.apply(s, causal, window, out_dtype)
