"""K8 — cross-entropy with the gradient produced in the forward launch (``csrc/loss.cu``), and the
lm_head + loss fusion built on it.

``cross_entropy(logits, targets)``        mean CE over non-ignored rows (nn.CrossEntropyLoss semantics).
``linear_cross_entropy(x, w, targets)``   ``CE(x @ w^T, targets)`` computed in token chunks: each
    chunk's logits exist only between its GEMM and its CE kernel, which overwrites them in place
    with d loss / d logits; the chunk's dx (and dw for a trainable / tied head) are produced right
    away, so the backward pass of the head is a multiply by the upstream scalar.

Reference: nn.CrossEntropyLoss in LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py:282,304;
HF ``ForCausalLM`` loss (shifted labels, ignore_index = -100) used by Fine-Tuning/*.py.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch
import torch.nn.functional as F

from . import _native as nv
from .gemm import gemm

vp, i32, i64, f32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float


def _ce_native(logits2, targets, grad_scale, write_grad, ignore_index):
    rows, vocab = logits2.shape
    loss_rows = torch.empty(rows, dtype=torch.float32, device=logits2.device)
    fn = nv.proto("b200_cross_entropy", [vp, vp, vp, vp, i64, i32, i64, i64, f32, i32, i32, vp])
    nv.check(fn(nv.ptr(logits2), nv.ptr(targets), nv.ptr(loss_rows), None, rows, vocab, logits2.stride(0),
                ignore_index, grad_scale, int(write_grad), nv.dtype_code(logits2.dtype), nv.stream()), "cross_entropy")
    nv.count_launch()
    return loss_rows


class _CrossEntropyFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, targets, ignore_index, n_valid):
        vocab = logits.shape[-1]
        l2 = logits.reshape(-1, vocab)
        t = targets.reshape(-1).contiguous()
        need_grad = logits.requires_grad
        work = l2.clone() if need_grad else l2  # the kernel overwrites its input with the gradient
        if not work.is_contiguous():
            work = work.contiguous()
        inv = 1.0 / max(n_valid, 1)
        loss_rows = _ce_native(work, t, inv, need_grad, ignore_index)
        if need_grad:
            ctx.save_for_backward(work)
        ctx.shape = logits.shape
        return loss_rows.sum() * inv

    @staticmethod
    def backward(ctx, g):
        (dlogits,) = ctx.saved_tensors
        return (dlogits * g.to(dlogits.dtype)).view(ctx.shape), None, None, None


def cross_entropy(logits: torch.Tensor, targets: torch.Tensor, ignore_index: int = -100,
                  n_valid: Optional[int] = None) -> torch.Tensor:
    """Mean cross-entropy over rows whose target != ignore_index. ``n_valid`` (the number of such
    rows) can be passed to avoid a device->host sync; it is computed otherwise."""
    vocab = logits.shape[-1]
    es = 4 if logits.dtype == torch.float32 else 2
    if nv.use_native(logits) and (vocab * es) % 16 == 0:
        if n_valid is None:
            n_valid = int((targets != ignore_index).sum().item())
        return _CrossEntropyFn.apply(logits, targets, ignore_index, n_valid)
    return F.cross_entropy(logits.reshape(-1, vocab).float(), targets.reshape(-1), ignore_index=ignore_index)


class _LinearCEFn(torch.autograd.Function):
    """loss = mean CE(x @ w^T). Gradients are formed in forward (scaled by 1/n_valid) and multiplied
    by the upstream scalar in backward."""

    @staticmethod
    def forward(ctx, x, w, targets, ignore_index, n_valid, chunk_tokens, w_main_grad):
        M, K = x.shape
        V = w.shape[0]
        ld = (V + 7) // 8 * 8  # padded leading dimension keeps every logits row 16-byte aligned
        need_dx = x.requires_grad
        need_dw = w.requires_grad
        inv = 1.0 / max(n_valid, 1)
        dx = torch.empty_like(x) if need_dx else None
        dw = None
        if need_dw:
            dw = w_main_grad if w_main_grad is not None else torch.zeros(V, K, dtype=torch.float32, device=x.device)
        loss_sum = torch.zeros((), dtype=torch.float32, device=x.device)
        t = targets.reshape(-1).contiguous()
        chunk = min(chunk_tokens, M)
        buf = torch.empty(chunk, ld, dtype=x.dtype, device=x.device)
        for s in range(0, M, chunk):
            e = min(M, s + chunk)
            lg = buf[: e - s, :V]
            gemm(x[s:e], w, out=lg)
            rows = _ce_native(lg, t[s:e], inv, need_dx or need_dw, ignore_index)
            loss_sum += rows.sum()
            if need_dx:
                gemm(lg, w, b_mn=True, out=dx[s:e])
            if need_dw:
                gemm(lg, x[s:e], a_mn=True, b_mn=True, out=dw, beta=1.0)
        ctx.save_for_backward(dx, dw if (need_dw and w_main_grad is None) else None)
        ctx.dw_in_place = need_dw and w_main_grad is not None
        ctx.w_dtype = w.dtype
        return loss_sum * inv

    @staticmethod
    def backward(ctx, g):
        dx, dw = ctx.saved_tensors
        gdx = None if dx is None else dx * g.to(dx.dtype)
        gdw = None if dw is None else (dw * g).to(ctx.w_dtype)
        return gdx, gdw, None, None, None, None, None


def linear_cross_entropy(x: torch.Tensor, w: torch.Tensor, targets: torch.Tensor, ignore_index: int = -100,
                         n_valid: Optional[int] = None, chunk_tokens: int = 4096) -> torch.Tensor:
    """Fused lm_head + cross-entropy. x: [tokens, K], w: [vocab, K], targets: [tokens].

    If ``w`` carries a ``main_grad`` fp32 buffer (flat-arena parameters) the weight gradient is
    accumulated there directly and autograd sees no gradient for ``w``; this assumes the upstream
    gradient of the loss is 1 (the usual ``loss.backward()``), which the training engines guarantee
    by folding loss scaling into ``n_valid``."""
    if nv.use_native(x) and x.dtype in (torch.bfloat16, torch.float16):
        if n_valid is None:
            n_valid = int((targets != ignore_index).sum().item())
        return _LinearCEFn.apply(x.contiguous(), w, targets, ignore_index, n_valid, chunk_tokens,
                                 getattr(w, "main_grad", None))
    logits = x.float() @ w.float().t()
    return F.cross_entropy(logits, targets.reshape(-1), ignore_index=ignore_index)
