"""K5 — RMSNorm / LayerNorm with the residual add fused in (``csrc/norm.cu``).

``rms_norm(x, w, eps, residual=None)`` and ``layer_norm(x, w, b, eps, residual=None)`` return
``(y, h)`` where ``h = x + residual`` is the updated residual stream (``h is x`` when no residual is
given). Both are autograd Functions over the CUDA kernels; on CPU they evaluate the same maths
with PyTorch, which is also the oracle the GPU tests compare against.

Reference call sites: nn.LayerNorm in GPTLike (LLM_Distributed_Trainning/PyTorch/ddp_basics/
ddp_gpt_wikitext2.py:99-116), HF Qwen3RMSNorm in every Fine-Tuning/*.py script.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Tuple

import torch

from . import _native as nv

vp, i32, i64, f32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float


def _ref_norm(x, w, b, eps, residual, layer_norm):
    h = x if residual is None else (x.float() + residual.float()).to(x.dtype)
    hf = h.float()
    if layer_norm:
        mean = hf.mean(-1, keepdim=True)
        var = hf.var(-1, unbiased=False, keepdim=True)
        y = (hf - mean) * torch.rsqrt(var + eps)
    else:
        y = hf * torch.rsqrt(hf.pow(2).mean(-1, keepdim=True) + eps)
    y = y * w.float()
    if b is not None:
        y = y + b.float()
    return y.to(x.dtype), h


def _fwd_native(x2, res2, w, b, eps, layer_norm, want_h):
    rows, cols = x2.shape
    y = torch.empty_like(x2)
    h = torch.empty_like(x2) if (res2 is not None and want_h) else None
    stats = torch.empty(rows, 2, dtype=torch.float32, device=x2.device)
    fn = nv.proto("b200_norm_fwd", [i32, vp, vp, vp, vp, vp, vp, vp, i32, i32, f32, i32, i32, vp])
    nv.check(fn(int(layer_norm), nv.ptr(x2), nv.ptr(res2), nv.ptr(w), nv.ptr(b), nv.ptr(y), nv.ptr(h), nv.ptr(stats),
                rows, cols, eps, nv.dtype_code(x2.dtype), nv.dtype_code(w.dtype), nv.stream()), "norm_fwd")
    nv.count_launch()
    return y, (h if h is not None else x2), stats


_scratch = {}


def _bwd_scratch(device, floats):
    key = (device.index if device.index is not None else torch.cuda.current_device())
    buf = _scratch.get(key)
    if buf is None or buf.numel() < floats:
        buf = torch.empty(floats, dtype=torch.float32, device=device)
        _scratch[key] = buf
    return buf


def _bwd_native(dy, h, w, stats, dres, layer_norm, need_dw, need_db):
    rows, cols = h.shape
    dx = torch.empty_like(h)
    dw = torch.zeros_like(w) if need_dw else None
    db = torch.zeros_like(w) if need_db else None
    scratch = None
    if need_dw or need_db:
        parts = nv.lib().b200_norm_bwd_parts(rows)
        scratch = _bwd_scratch(h.device, 2 * parts * cols)
    fn = nv.proto("b200_norm_bwd", [i32, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp])
    nv.check(fn(int(layer_norm), nv.ptr(dy), nv.ptr(h), nv.ptr(w), nv.ptr(stats), nv.ptr(dres), nv.ptr(dx),
                nv.ptr(dw), nv.ptr(db), nv.ptr(scratch), 0, rows, cols, nv.dtype_code(h.dtype),
                nv.dtype_code(w.dtype), nv.stream()), "norm_bwd")
    nv.count_launch(1 + int(need_dw) + int(need_db))
    return dx, dw, db


class _NormFn(torch.autograd.Function):
    """y, h = norm(x (+ residual)). Gradients: the residual stream receives dh + dx."""

    @staticmethod
    def forward(ctx, x, residual, w, b, eps, layer_norm):
        shape = x.shape
        x2 = x.reshape(-1, shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        r2 = None
        if residual is not None:
            r2 = residual.reshape(-1, shape[-1])
            if not r2.is_contiguous():
                r2 = r2.contiguous()
        y, h, stats = _fwd_native(x2, r2, w, b, eps, layer_norm, True)
        ctx.save_for_backward(h, w, stats)
        ctx.layer_norm = layer_norm
        ctx.has_res = residual is not None
        ctx.has_b = b is not None
        ctx.shape = shape
        return y.view(shape), h.view(shape)

    @staticmethod
    def backward(ctx, dy, dh):
        h, w, stats = ctx.saved_tensors
        cols = ctx.shape[-1]
        dy2 = dy.reshape(-1, cols)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dres = None
        if dh is not None:
            dres = dh.reshape(-1, cols)
            if not dres.is_contiguous():
                dres = dres.contiguous()
        need_dw = ctx.needs_input_grad[2]
        need_db = ctx.has_b and ctx.needs_input_grad[3]
        dx, dw, db = _bwd_native(dy2, h, w, stats, dres, ctx.layer_norm, need_dw, need_db)
        dx = dx.view(ctx.shape)
        return dx, (dx if ctx.has_res else None), dw, db, None, None


def rms_norm(x: torch.Tensor, w: torch.Tensor, eps: float = 1e-6,
             residual: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    if nv.use_native(x):
        return _NormFn.apply(x, residual, w, None, eps, False)
    return _ref_norm(x, w, None, eps, residual, False)


def layer_norm(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor], eps: float = 1e-5,
               residual: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    if nv.use_native(x):
        return _NormFn.apply(x, residual, w, b, eps, True)
    return _ref_norm(x, w, b, eps, residual, True)


class RMSNorm(torch.nn.Module):
    def __init__(self, dim: int, eps: float = 1e-6, dtype=None, device=None):
        super().__init__()
        self.weight = torch.nn.Parameter(torch.ones(dim, dtype=dtype, device=device))
        self.eps = eps

    def forward(self, x, residual=None):
        return rms_norm(x, self.weight, self.eps, residual)


class LayerNorm(torch.nn.Module):
    def __init__(self, dim: int, eps: float = 1e-5, dtype=None, device=None):
        super().__init__()
        self.weight = torch.nn.Parameter(torch.ones(dim, dtype=dtype, device=device))
        self.bias = torch.nn.Parameter(torch.zeros(dim, dtype=dtype, device=device))
        self.eps = eps

    def forward(self, x, residual=None):
        return layer_norm(x, self.weight, self.bias, self.eps, residual)
