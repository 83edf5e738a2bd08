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


def _ce_native(logits2, targets, grad_scale, write_grad, ignore_index, row_scale=None):
    rows, vocab = logits2.shape
    loss_rows = torch.empty(rows, dtype=torch.float32, device=logits2.device)
    fn = nv.proto("b200_cross_entropy_rows", [vp, vp, vp, vp, i64, i32, i64, i64, f32, i32, vp, i32, vp])
    nv.check(fn(nv.ptr(logits2), nv.ptr(targets), nv.ptr(loss_rows), None, rows, vocab, logits2.stride(0),
                ignore_index, grad_scale, int(write_grad), None if row_scale is None else nv.ptr(row_scale),
                nv.dtype_code(logits2.dtype), nv.stream()), "cross_entropy")
    nv.count_launch()
    return loss_rows


class _CrossEntropyFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, targets, ignore_index, row_scale):
        vocab = logits.shape[-1]
        l2 = logits.reshape(-1, vocab)
        t = targets.reshape(-1).contiguous()
        need_grad = logits.requires_grad
        work = l2.clone() if need_grad else l2  # the kernel overwrites its input with the gradient
        if not work.is_contiguous():
            work = work.contiguous()
        loss_rows = _ce_native(work, t, 1.0, need_grad, ignore_index, row_scale)
        if need_grad:
            ctx.save_for_backward(work)
        ctx.shape = logits.shape
        return torch.dot(loss_rows, row_scale)

    @staticmethod
    def backward(ctx, g):
        (dlogits,) = ctx.saved_tensors
        return (dlogits * g.to(dlogits.dtype)).view(ctx.shape), None, None, None


def _row_scale(targets: torch.Tensor, ignore_index: int, n_valid: Optional[int]) -> torch.Tensor:
    """Per-row weights of the mean over non-ignored rows, formed on the device: without ``n_valid`` the count never
    travels to the host, so the loss does not stall the launch queue (small models are launch-bound)."""
    valid = (targets.reshape(-1) != ignore_index).to(torch.float32)
    if n_valid is not None:
        return valid * (1.0 / max(n_valid, 1))
    return valid / valid.sum().clamp_min_(1.0)


def cross_entropy(logits: torch.Tensor, targets: torch.Tensor, ignore_index: int = -100,
                  n_valid: Optional[int] = None) -> torch.Tensor:
    """Mean cross-entropy over rows whose target != ignore_index (nn.CrossEntropyLoss semantics). The normalisation is
    computed on the device; ``n_valid`` (the number of such rows) may be passed when the caller knows it."""
    vocab = logits.shape[-1]
    es = 4 if logits.dtype == torch.float32 else 2
    if nv.use_native(logits) and (vocab * es) % 16 == 0:
        return _CrossEntropyFn.apply(logits, targets, ignore_index, _row_scale(targets, ignore_index, n_valid))
    return F.cross_entropy(logits.reshape(-1, vocab).float(), targets.reshape(-1), ignore_index=ignore_index)


# Upstream gradient the training loop will send into the loss (1 / grad_accum, times the fp16 loss scale), declared
# BEFORE the forward pass. With it declared, a head weight that lives in a flat arena gets its gradient accumulated
# straight into ``main_grad`` during the forward launch sequence, already scaled; without it (None) the weight gradient
# is kept in a temporary and added to ``main_grad`` in backward, multiplied by whatever gradient autograd delivers.
_DECLARED_GRAD_SCALE: Optional[float] = None


class grad_scale_context:
    """``with grad_scale_context(1 / grad_accum * loss_scale): loss = model(x, y)`` — the engines and training loops
    of this package wrap their forward in it; ``loss.backward(gradient=that scale)`` (or ``(loss * scale).backward()``)
    must follow."""

    def __init__(self, scale: Optional[float]):
        self.scale, self.prev = scale, None

    def __enter__(self):
        global _DECLARED_GRAD_SCALE
        self.prev, _DECLARED_GRAD_SCALE = _DECLARED_GRAD_SCALE, self.scale
        return self

    def __exit__(self, *exc):
        global _DECLARED_GRAD_SCALE
        _DECLARED_GRAD_SCALE = self.prev
        return False


_VOCAB_CHUNK = int(__import__("os").environ.get("B200_CE_VOCAB_CHUNK", "16384"))
_VOCAB_CHUNK_MIN_BYTES = int(__import__("os").environ.get("B200_CE_VOCAB_MIN_MB", "64")) << 20


def _vocab_chunked(x, w, t, ignore_index, row_scale, gscale, need_dx, dw):
    """lm_head + CE over vocabulary chunks of ``_VOCAB_CHUNK`` columns (csrc/loss.cu, K8). The [M, C] chunk of logits
    (32 MB for 2048 tokens x 8192 columns) lives in L2 between its GEMM and the kernels that consume it.

    Frozen head (``dw is None``): ONE pass, flash-attention style — per chunk the kernel turns the logits into
    p = exp(l - running max) and returns the row rescale alpha; ``acc = alpha * acc + p @ W_chunk`` runs as a GEMM with
    a per-row beta; at the end ``dx = (acc / sum - W[target]) * scale``. Same two GEMMs as the materialising version.
    Trainable / tied head: pass 1 collects the row statistics only, pass 2 recomputes each chunk, turns it into
    d loss / d logits and feeds ``dx += dl @ W_chunk`` and ``dW_chunk += dl^T @ x`` (one extra lm_head GEMM)."""
    M, K = x.shape
    V = w.shape[0]
    C = min(_VOCAB_CHUNK, (V + 7) // 8 * 8)
    dev = x.device
    dt = nv.dtype_code(x.dtype)
    buf = torch.empty(M, C, dtype=x.dtype, device=dev)
    m_run = torch.full((M,), float("-inf"), dtype=torch.float32, device=dev)
    s_run = torch.zeros(M, dtype=torch.float32, device=dev)
    tgt = torch.zeros(M, dtype=torch.float32, device=dev)
    loss_rows = torch.empty(M, dtype=torch.float32, device=dev)
    f_on = nv.proto("b200_ce_chunk_online", [vp, i64, i64, i32, i32, vp, i64, vp, vp, vp, vp, i32, i32, vp])
    f_gr = nv.proto("b200_ce_chunk_grad", [vp, i64, i64, i32, i32, vp, i64, vp, vp, vp, f32, i32, vp])
    f_fin = nv.proto("b200_ce_online_finish", [vp, i64, vp, i64, i32, vp, i64, vp, vp, vp, vp, f32, vp, i64, vp, i64, i32, vp])
    chunks = [(c0, min(V, c0 + C)) for c0 in range(0, V, C)]

    def logits(c0, c1):
        cw = c1 - c0
        if cw < C:
            buf[:, cw:].fill_(float("-inf"))             # padded columns contribute exp() = 0
        gemm(x, w[c0:c1], out=buf[:, :cw])
        return cw

    def online(c0, alpha, write_p):
        nv.check(f_on(nv.ptr(buf), M, buf.stride(0), C, c0, nv.ptr(t), ignore_index, nv.ptr(m_run), nv.ptr(s_run), nv.ptr(tgt),
                      None if alpha is None else nv.ptr(alpha), int(write_p), dt, nv.stream()), "ce_chunk_online")
        nv.count_launch()

    dx = None
    if dw is None:
        acc = torch.zeros(M, K, dtype=torch.float32, device=dev) if need_dx else None
        alpha = torch.empty(M, dtype=torch.float32, device=dev)
        for c0, c1 in chunks:
            cw = logits(c0, c1)
            online(c0, alpha, need_dx)
            if need_dx:
                gemm(buf[:, :cw], w[c0:c1], b_mn=True, out=acc, beta_rows=alpha)
        if need_dx:
            dx = torch.empty_like(x)
        nv.check(f_fin(None if acc is None else nv.ptr(acc), K, nv.ptr(w), w.stride(0), K, nv.ptr(t), ignore_index, nv.ptr(m_run),
                       nv.ptr(s_run), nv.ptr(tgt), nv.ptr(row_scale), gscale, None if dx is None else nv.ptr(dx), K,
                       nv.ptr(loss_rows), M, dt, nv.stream()), "ce_online_finish")
        nv.count_launch()
    else:
        for c0, c1 in chunks:
            logits(c0, c1)
            online(c0, None, False)
        loss_rows = torch.where(t != ignore_index, m_run + torch.log(s_run) - tgt, torch.zeros_like(tgt))
        acc = torch.zeros(M, K, dtype=torch.float32, device=dev) if need_dx else None
        for c0, c1 in chunks:
            cw = logits(c0, c1)
            nv.check(f_gr(nv.ptr(buf), M, buf.stride(0), C, c0, nv.ptr(t), ignore_index, nv.ptr(m_run), nv.ptr(s_run),
                          nv.ptr(row_scale), gscale, dt, nv.stream()), "ce_chunk_grad")
            nv.count_launch()
            if need_dx:
                gemm(buf[:, :cw], w[c0:c1], b_mn=True, out=acc, beta=1.0)
            gemm(buf[:, :cw], x, a_mn=True, b_mn=True, out=dw[c0:c1], beta=1.0)
        if need_dx:
            dx = acc.to(x.dtype)
    return torch.dot(loss_rows, row_scale), dx


class _LinearCEFn(torch.autograd.Function):
    """loss = mean CE(x @ w^T). Gradients are formed in forward (scaled by 1/n_valid, and by the declared upstream
    scale when there is one) and multiplied by the (remaining) upstream factor in backward."""

    @staticmethod
    def forward(ctx, x, w, targets, ignore_index, n_valid, chunk_tokens, w_main_grad, need_dx, need_dw, declared,
                row_scale=None):
        M, K = x.shape
        V = w.shape[0]
        ld = (V + 7) // 8 * 8  # padded leading dimension keeps every logits row 16-byte aligned
        pre = 1.0 if declared is None else float(declared)     # folded into the gradients formed here
        inv = 1.0 if row_scale is not None else 1.0 / max(n_valid, 1)   # row_scale already holds the normalisation
        dw = None
        in_place = need_dw and w_main_grad is not None and declared is not None
        if need_dw:
            dw = w_main_grad if in_place else torch.zeros(V, K, dtype=torch.float32, device=x.device)
        t = targets.reshape(-1).contiguous()
        if row_scale is not None and M * V * x.element_size() > _VOCAB_CHUNK_MIN_BYTES:
            # vocabulary-chunked: no [tokens, vocab] tensor; one pass for a frozen head, two for a trainable one
            loss, dx = _vocab_chunked(x, w, t, ignore_index, row_scale, inv * pre, need_dx, dw)
            ctx.save_for_backward(dx, None if (in_place or not need_dw) else dw)
            ctx.pre = pre
            ctx.main_grad = w_main_grad if (need_dw and not in_place) else None
            ctx.hook = None
            if ctx.main_grad is not None:
                from ..utils.arena import note_use
                note_use(w, True)
                ctx.hook = getattr(w, "_grad_ready", None)
            ctx.w_dtype = w.dtype
            return loss
        dx = torch.empty_like(x) if need_dx else None
        loss_sum = torch.zeros((), dtype=torch.float32, device=x.device)
        chunk = min(chunk_tokens, M)
        buf = torch.empty(chunk, ld, dtype=x.dtype, device=x.device)
        for s in range(0, M, chunk):
            e = min(M, s + chunk)
            lg = buf[: e - s, :V]
            gemm(x[s:e], w, out=lg)
            rs = None if row_scale is None else row_scale[s:e]
            rows = _ce_native(lg, t[s:e], inv * pre, need_dx or need_dw, ignore_index, rs)
            loss_sum += rows.sum() if rs is None else torch.dot(rows, rs)
            if need_dx:
                gemm(lg, w, b_mn=True, out=dx[s:e])
            if need_dw:
                gemm(lg, x[s:e], a_mn=True, b_mn=True, out=dw, beta=1.0)
        ctx.save_for_backward(dx, None if (in_place or not need_dw) else dw)
        ctx.pre = pre
        ctx.main_grad = w_main_grad if (need_dw and not in_place) else None
        ctx.hook = None
        if ctx.main_grad is not None:           # this op's backward adds into main_grad: one outstanding contribution
            from ..utils.arena import note_use
            note_use(w, True)
            ctx.hook = getattr(w, "_grad_ready", None)
        ctx.w_dtype = w.dtype
        return loss_sum * inv

    @staticmethod
    def backward(ctx, g):
        dx, dw = ctx.saved_tensors
        f = g if ctx.pre == 1.0 else g / ctx.pre            # what is left of the upstream gradient (1 when it was declared right)
        gdx = None if dx is None else dx * f.to(dx.dtype)
        gdw = None
        if dw is not None:
            if ctx.main_grad is not None:
                ctx.main_grad.add_(dw * f)
            else:
                gdw = (dw * f).to(ctx.w_dtype)
        # the in-place path put its contribution into main_grad during forward (nothing outstanding); the other one
        # has just landed: tell the data-parallel wrapper (it waits for the remaining uses of a tied weight)
        if ctx.hook is not None:
            ctx.hook()
        return (gdx, gdw) + (None,) * 9


def micro_batch_row_scale(targets: torch.Tensor, micro_batches: int, ignore_index: int = -100) -> torch.Tensor:
    """float[tokens]: 1 / (micro_batches * n_valid(m)) for the rows of micro-batch m, 0 for ignored rows — the weights
    with which ``sum(loss_rows * scale)`` equals the mean over micro-batches of each micro-batch's mean loss (what
    ``gradient_accumulation_steps`` separate passes with ``loss / steps`` produce). Computed on the device."""
    valid = (targets.reshape(micro_batches, -1) != ignore_index).to(torch.float32)
    n = valid.sum(1, keepdim=True).clamp_min_(1.0)
    return (valid / (n * micro_batches)).reshape(-1)


def linear_cross_entropy(x: torch.Tensor, w: torch.Tensor, targets: torch.Tensor, ignore_index: int = -100,
                         n_valid: Optional[int] = None, chunk_tokens: int = 4096,
                         row_scale: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Fused lm_head + cross-entropy. x: [tokens, K], w: [vocab, K], targets: [tokens].
    ``row_scale`` (float[tokens], device) replaces the 1 / n_valid normalisation by per-row weights
    (``micro_batch_row_scale``); with it no host read-back of the valid-token count happens.

    Under ``torch.no_grad()`` (evaluation) no gradient of any kind is formed. If ``w`` carries a ``main_grad`` fp32
    buffer (flat-arena parameters) the weight gradient ends up there and autograd sees none for ``w``: accumulated
    during forward when the upstream gradient was declared with ``grad_scale_context`` (the engines do), otherwise
    held in a temporary and added in backward times the gradient autograd delivers — correct for any
    ``(loss * c).backward()``."""
    if nv.use_native(x) and x.dtype in (torch.bfloat16, torch.float16):
        if row_scale is None:
            row_scale = _row_scale(targets, ignore_index, n_valid)      # no host read-back of the valid count
        ge = torch.is_grad_enabled()
        return _LinearCEFn.apply(x.contiguous(), w, targets, ignore_index, n_valid or 1, chunk_tokens,
                                 getattr(w, "main_grad", None), ge and x.requires_grad, ge and w.requires_grad,
                                 _DECLARED_GRAD_SCALE if ge else None, row_scale)
    logits = x.float() @ w.float().t()
    if row_scale is not None:
        rows = F.cross_entropy(logits, targets.reshape(-1), ignore_index=ignore_index, reduction="none")
        return torch.dot(rows, row_scale.to(rows.dtype))
    return F.cross_entropy(logits, targets.reshape(-1), ignore_index=ignore_index)
