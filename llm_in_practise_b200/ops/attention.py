"""K4 — causal / sliding-window / bidirectional multi-head attention with grouped KV heads.

Two native paths share this front end:

* ``flash`` (``csrc/flash_attn.cu``; 16-bit dtypes, head_dim 32 / 64 / 96 / 128, no window; probability dropout in-kernel): 64 x 64
  tiles, online softmax forward, recomputing backward, no L x L tensor in memory in either direction.
* ``tiled`` (always available): batched tcgen05 GEMMs (``csrc/gemm_tcgen05.cu``) around fused
  masked-softmax kernels (``csrc/loss.cu``); scores exist per (batch, head) as a transient.

Both read q, k and v *in place* from the fused qkv projection ``[tokens, (H + 2 Hkv) * D]`` through
strided TMA descriptors — no head transposes, no ``repeat_kv`` — and write the context as
``[tokens, H * D]`` ready for the output projection. The causal mask is implicit; the reference
builds an L x L boolean mask on every forward (ddp_gpt_wikitext2.py:93-95) and HF Qwen3 relies on
SDPA / flash-attn behind ``AutoModelForCausalLM`` (Fine-Tuning/*.py).
"""
from __future__ import annotations

import ctypes
import math
from typing import Optional

import torch

from . import _native as nv
from .gemm import gemm

vp, i32, i64, f32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float


_BLOCK_Q = int(__import__("os").environ.get("B200_ATTN_BLOCK_Q", "512"))   # query rows per block of the blocked causal path


def _softmax_fwd(scores, causal, window, out_dtype):
    *lead, rows, cols = scores.shape
    batch = 1
    for d in lead:
        batch *= d
    P = torch.empty(scores.shape, dtype=out_dtype, device=scores.device)
    fn = nv.proto("b200_softmax_fwd", [vp, vp, i64, i32, i32, i32, i32, f32, i32, i32, vp])
    nv.check(fn(nv.ptr(scores), nv.ptr(P), batch, rows, cols, int(causal), int(window), 1.0,
                nv.dtype_code(scores.dtype), nv.dtype_code(out_dtype), nv.stream()), "softmax_fwd")
    nv.count_launch()
    return P


def _softmax_bwd(P, dP, scale, out_dtype):
    cols = P.shape[-1]
    dS = torch.empty(P.shape, dtype=out_dtype, device=P.device)
    fn = nv.proto("b200_softmax_bwd", [vp, vp, vp, i64, i32, f32, i32, i32, i32, vp])
    nv.check(fn(nv.ptr(P), nv.ptr(dP), nv.ptr(dS), P.numel() // cols, cols, scale, nv.dtype_code(P.dtype),
                nv.dtype_code(dP.dtype), nv.dtype_code(out_dtype), nv.stream()), "softmax_bwd")
    nv.count_launch()
    return dS


def _split_heads(qkv, B, S, H, Hkv, D):
    """Strided [B, heads, S, D] views of q, k, v inside the fused projection."""
    t = qkv.view(B, S, H + 2 * Hkv, D)
    q = t[:, :, :H].permute(0, 2, 1, 3)
    k = t[:, :, H:H + Hkv].permute(0, 2, 1, 3)
    v = t[:, :, H + Hkv:].permute(0, 2, 1, 3)
    return q, k, v


class _BlockedCausalAttentionFn(torch.autograd.Function):
    """Causal attention in query blocks of ``bs`` rows: block i only multiplies against keys [0, end of block i), so the
    score / probability matrices hold ~S^2/2 entries instead of S^2 and the work above the diagonal (half of it at long
    sequence length) is never done. Same kernels as the one-shot path: strided tcgen05 GEMMs straight out of the fused
    qkv projection, the row softmax with its implicit causal offset (last query row <-> last key column), accumulation
    of dK / dV across blocks in the GEMM epilogue (beta = 1). Used for S > 512 (the 4k-token LoRA configuration)."""

    @staticmethod
    def forward(ctx, qkv, B, S, H, Hkv, D, scale, bs):
        q, k, v = _split_heads(qkv, B, S, H, Hkv, D)
        out = torch.empty(B, S, H, D, dtype=qkv.dtype, device=qkv.device)
        o4 = out.permute(0, 2, 1, 3)
        Ps = []
        for q0 in range(0, S, bs):
            q1 = min(S, q0 + bs)
            sc = gemm(q[:, :, q0:q1], k[:, :, :q1], alpha=scale, out_dtype=torch.float32)      # [B, H, bs, q1]
            P = _softmax_fwd(sc, True, 0, qkv.dtype)
            del sc
            gemm(P, v[:, :, :q1], b_mn=True, out=o4[:, :, q0:q1])
            Ps.append(P)
        ctx.save_for_backward(qkv, *Ps)
        ctx.cfg = (B, S, H, Hkv, D, scale, bs)
        return out.view(B * S, H * D)

    @staticmethod
    def backward(ctx, dout):
        qkv, *Ps = ctx.saved_tensors
        B, S, H, Hkv, D, scale, bs = ctx.cfg
        g = H // Hkv
        q, k, v = _split_heads(qkv, B, S, H, Hkv, D)
        dO = dout.contiguous().view(B, S, H, D).permute(0, 2, 1, 3)
        dqkv = torch.empty_like(qkv)
        dq, dk, dv = _split_heads(dqkv, B, S, H, Hkv, D)
        dk_h = torch.zeros(B, H, S, D, dtype=torch.float32, device=qkv.device)
        dv_h = torch.zeros(B, H, S, D, dtype=torch.float32, device=qkv.device)
        for i, q0 in enumerate(range(0, S, bs)):
            q1 = min(S, q0 + bs)
            P = Ps[i]
            dP = gemm(dO[:, :, q0:q1], v[:, :, :q1], out_dtype=torch.float32)
            dS = _softmax_bwd(P, dP, scale, qkv.dtype)
            del dP
            gemm(dS, k[:, :, :q1], b_mn=True, out=dq[:, :, q0:q1])
            gemm(dS, q[:, :, q0:q1], a_mn=True, b_mn=True, out=dk_h[:, :, :q1], beta=1.0)
            gemm(P, dO[:, :, q0:q1], a_mn=True, b_mn=True, out=dv_h[:, :, :q1], beta=1.0)
        if g == 1:
            dk.copy_(dk_h)
            dv.copy_(dv_h)
        else:
            dk.copy_(dk_h.view(B, Hkv, g, S, D).sum(2))
            dv.copy_(dv_h.view(B, Hkv, g, S, D).sum(2))
        return dqkv, None, None, None, None, None, None, None


class _TiledAttentionFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, B, S, H, Hkv, D, causal, window, scale, dropout_p):
        q, k, v = _split_heads(qkv, B, S, H, Hkv, D)
        scores = gemm(q, k, alpha=scale, out_dtype=torch.float32)            # [B, H, S, S]
        P = _softmax_fwd(scores, causal, window, qkv.dtype)
        del scores
        seed = 0
        Pd = P
        if dropout_p > 0.0:
            # attention-probability dropout (nn.MultiheadAttention(dropout=p)): Philox mask, regenerated in backward
            from .elementwise import _DropoutFn, next_seed
            seed = next_seed()
            Pd = _DropoutFn.apply(P, dropout_p, seed)
        out = torch.empty(B, S, H, D, dtype=qkv.dtype, device=qkv.device)
        gemm(Pd, v, b_mn=True, out=out.permute(0, 2, 1, 3))                  # context, written as [B, S, H, D]
        ctx.save_for_backward(qkv, P)
        ctx.cfg = (B, S, H, Hkv, D, scale, dropout_p, seed)
        return out.view(B * S, H * D)

    @staticmethod
    def backward(ctx, dout):
        qkv, P = ctx.saved_tensors
        B, S, H, Hkv, D, scale, dropout_p, seed = ctx.cfg
        g = H // Hkv
        q, k, v = _split_heads(qkv, B, S, H, Hkv, D)
        dO = dout.contiguous().view(B, S, H, D).permute(0, 2, 1, 3)         # [B, H, S, D] strided
        dqkv = torch.empty_like(qkv)
        dq, dk, dv = _split_heads(dqkv, B, S, H, Hkv, D)
        dP = gemm(dO, v, out_dtype=torch.float32)                            # [B, H, S, S]
        Pv = P
        if dropout_p > 0.0:
            from .elementwise import _DropoutFn
            dP = _DropoutFn.apply(dP, dropout_p, seed)                       # same mask, same 1/(1-p) scale
            Pv = _DropoutFn.apply(P, dropout_p, seed)                        # the dropped P that multiplied V
        dS = _softmax_bwd(P, dP, scale, qkv.dtype)
        del dP
        gemm(dS, k, b_mn=True, out=dq)                                       # dQ = dS K
        if g == 1:
            gemm(dS, q, a_mn=True, b_mn=True, out=dk)                        # dK = dS^T Q
            gemm(Pv, dO, a_mn=True, b_mn=True, out=dv)                       # dV = P^T dO
        else:
            dk_h = gemm(dS, q, a_mn=True, b_mn=True, out_dtype=torch.float32)
            dv_h = gemm(Pv, dO, a_mn=True, b_mn=True, out_dtype=torch.float32)
            dk.copy_(dk_h.view(B, Hkv, g, S, D).sum(2))
            dv.copy_(dv_h.view(B, Hkv, g, S, D).sum(2))
        return dqkv, None, None, None, None, None, None, None, None, None


def _ref_attention(qkv, B, S, H, Hkv, D, causal, window, scale, dropout_p=0.0):
    q, k, v = _split_heads(qkv, B, S, H, Hkv, D)
    g = H // Hkv
    kf = k.float().repeat_interleave(g, dim=1)
    vf = v.float().repeat_interleave(g, dim=1)
    s = torch.matmul(q.float(), kf.transpose(-1, -2)) * scale
    if causal:
        i = torch.arange(S, device=qkv.device)
        mask = i[None, :] > i[:, None]
        if window and window > 0:
            mask = mask | (i[None, :] < i[:, None] - window + 1)
        s = s.masked_fill(mask, float("-inf"))
    p = torch.softmax(s, dim=-1)
    if dropout_p > 0.0:
        p = torch.nn.functional.dropout(p, dropout_p, True)
    o = torch.matmul(p, vf)
    return o.permute(0, 2, 1, 3).reshape(B * S, H * D).to(qkv.dtype)


_flash_ok = None


def _flash_available() -> bool:
    global _flash_ok
    if _flash_ok is None:
        try:
            _flash_ok = hasattr(nv.lib(), "b200_flash_attn_fwd") and __import__("os").environ.get("B200_FLASH", "1") != "0"
        except Exception:
            _flash_ok = False
    return _flash_ok


def attention(qkv: torch.Tensor, B: int, S: int, n_heads: int, n_kv_heads: int, head_dim: int,
              causal: bool = True, window: int = 0, scale: Optional[float] = None,
              impl: str = "auto", dropout_p: float = 0.0) -> torch.Tensor:
    """softmax(Q K^T * scale + mask) V over the fused projection ``qkv`` [B*S, (H + 2 Hkv) * D].
    Returns the context as [B*S, H * D]."""
    if scale is None:
        scale = 1.0 / math.sqrt(head_dim)
    if not nv.use_native(qkv) or (S % 8 and not causal) or qkv.dtype == torch.float32 and False:
        return _ref_attention(qkv, B, S, n_heads, n_kv_heads, head_dim, causal, window, scale, dropout_p)
    if impl in ("auto", "flash") and _flash_available() and qkv.dtype in (torch.bfloat16, torch.float16) \
            and head_dim in (32, 64, 96, 128) and window == 0:
        from .flash import flash_attention
        seed = 0
        if dropout_p > 0.0:
            from .elementwise import next_seed
            seed = next_seed()
        return flash_attention(qkv.contiguous(), B, S, n_heads, n_kv_heads, head_dim, causal, scale, dropout_p, seed)
    if S % 8:
        # the score / probability matrices need 16-byte row strides: pad the sequence to a multiple of 8; under the
        # causal mask the real rows never see the padded keys, so the result is exact
        Sp = (S + 7) // 8 * 8
        W = qkv.shape[1]
        qp = torch.nn.functional.pad(qkv.view(B, S, W), (0, 0, 0, Sp - S)).reshape(B * Sp, W)
        out = attention(qp, B, Sp, n_heads, n_kv_heads, head_dim, causal, window, scale, impl, dropout_p)
        return out.view(B, Sp, -1)[:, :S].reshape(B * S, -1)
    qkv = qkv.contiguous()
    if causal and window == 0 and dropout_p == 0.0 and S > _BLOCK_Q and S % 8 == 0:
        return _BlockedCausalAttentionFn.apply(qkv, B, S, n_heads, n_kv_heads, head_dim, scale, _BLOCK_Q)
    return _TiledAttentionFn.apply(qkv, B, S, n_heads, n_kv_heads, head_dim, causal, window, scale, dropout_p)


def attention_with_past(qkv: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, pos0: int, n_heads: int,
                        n_kv_heads: int, head_dim: int, scale: Optional[float] = None) -> torch.Tensor:
    """Attention of T new tokens of ONE sequence over its KV-cache row: chunked prefill and prefix-cache hits
    (vLLM's ``enable_chunked_prefill`` / ``enable_prefix_caching``, Deployment/Ray/serve_run_examples/deepseek.py:29-34).

    ``qkv`` [T, (H + 2 Hkv) * D] holds the new tokens (positions pos0 .. pos0 + T - 1, already rotated); ``k_cache`` /
    ``v_cache`` [Hkv, L, D] already contain their keys / values at those positions and the past before them. Query row r
    sees cache columns 0 .. pos0 + r (the causal offset of the row-softmax kernel). Query blocks of ``_BLOCK_Q`` rows keep
    the score transient small. Returns the context [T, H * D]."""
    T = qkv.shape[0]
    H, Hkv, D = n_heads, n_kv_heads, head_dim
    if scale is None:
        scale = 1.0 / math.sqrt(D)
    q = qkv.view(1, T, H + 2 * Hkv, D)[:, :, :H].permute(0, 2, 1, 3)           # [1, H, T, D] strided view
    out = torch.empty(1, T, H, D, dtype=qkv.dtype, device=qkv.device)
    o4 = out.permute(0, 2, 1, 3)
    native = nv.use_native(qkv)
    for q0 in range(0, T, _BLOCK_Q):
        q1 = min(T, q0 + _BLOCK_Q)
        kv_len = pos0 + q1
        k = k_cache[:, :kv_len].unsqueeze(0)                                   # [1, Hkv, kv_len, D]
        v = v_cache[:, :kv_len].unsqueeze(0)
        if native and kv_len % 8 == 0:
            sc = gemm(q[:, :, q0:q1], k, alpha=scale, out_dtype=torch.float32)                 # [1, H, rows, kv_len]
            P = _softmax_fwd(sc, True, 0, qkv.dtype)
            gemm(P, v, b_mn=True, out=o4[:, :, q0:q1])
        else:
            g = H // Hkv
            kf = k.float().repeat_interleave(g, 1)
            vf = v.float().repeat_interleave(g, 1)
            sc = torch.matmul(q[:, :, q0:q1].float(), kf.transpose(-1, -2)) * scale
            rows = torch.arange(q0, q1, device=qkv.device).view(-1, 1) + pos0
            cols = torch.arange(kv_len, device=qkv.device).view(1, -1)
            sc = sc.masked_fill(cols > rows, float("-inf"))
            o4[:, :, q0:q1] = torch.matmul(sc.softmax(-1), vf).to(qkv.dtype)
    return out.view(T, H * D)
