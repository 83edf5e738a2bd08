"""Flash attention (csrc/flash_attn.cu): forward keeps only the context and the row log-sum-exp, backward recomputes
the probabilities tile by tile — no [B, H, S, S] tensor in either direction. Replaces the SDPA / flash-attn call behind
HF Qwen3 (Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:101-112) and the masked ``nn.MultiheadAttention`` of the GPT
scripts (LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py:92-96).
"""
from __future__ import annotations

import ctypes

import torch

from . import _native as nv

vp, i32, f32, u64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_ulonglong


def flash_supported(qkv: torch.Tensor, head_dim: int) -> bool:
    return qkv.dtype in (torch.bfloat16, torch.float16) and head_dim in (32, 64, 96, 128)


_TC = __import__("os").environ.get("B200_FLASH_TC", "1") != "0"
_TC_BWD = __import__("os").environ.get("B200_FLASH_TC_BWD", "1") != "0"


def _fwd(qkv, B, S, H, Hkv, D, causal, scale, want_lse=True, dropout_p=0.0, seed=0):
    out = torch.empty(B * S, H * D, dtype=qkv.dtype, device=qkv.device)
    lse = torch.empty(B, H, S, dtype=torch.float32, device=qkv.device) if want_lse else None
    if _TC and D == 128 and dropout_p == 0.0:
        # tcgen05 / TMEM / TMA forward (csrc/flash_tc.cu); the mma.sync kernel below keeps the other head sizes and dropout
        ft = nv.proto("b200_flash_attn_fwd_tc", [vp, vp, vp, i32, i32, i32, i32, i32, i32, f32, i32, vp])
        rc = ft(nv.ptr(qkv), nv.ptr(out), nv.ptr(lse), B, S, H, Hkv, D, int(causal), float(scale), nv.dtype_code(qkv.dtype),
                nv.stream())
        if rc == 0:
            nv.count_launch()
            return out, lse
        if rc != 10006:
            nv.check(rc, "flash_attn_fwd_tc")
    fn = nv.proto("b200_flash_attn_fwd", [vp, vp, vp, i32, i32, i32, i32, i32, i32, f32, f32, u64, i32, vp])
    nv.check(fn(nv.ptr(qkv), nv.ptr(out), nv.ptr(lse), B, S, H, Hkv, D, int(causal), float(scale), float(dropout_p), int(seed),
                nv.dtype_code(qkv.dtype), nv.stream()), "flash_attn_fwd")
    nv.count_launch()
    return out, lse


class _FlashAttentionFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, B, S, H, Hkv, D, causal, scale, dropout_p, seed):
        out, lse = _fwd(qkv, B, S, H, Hkv, D, causal, scale, True, dropout_p, seed)
        ctx.save_for_backward(qkv, out, lse)
        ctx.cfg = (B, S, H, Hkv, D, causal, scale, dropout_p, seed)
        return out

    @staticmethod
    def backward(ctx, dout):
        qkv, out, lse = ctx.saved_tensors
        B, S, H, Hkv, D, causal, scale, dropout_p, seed = ctx.cfg
        dout = dout.contiguous()
        dqkv = torch.empty_like(qkv)
        delta = torch.empty(B, H, S, dtype=torch.float32, device=qkv.device)
        if _TC_BWD and D == 128 and dropout_p == 0.0:
            # tcgen05 backward (csrc/flash_bwd_tc.cu): dK/dV pass + dQ pass, no atomics, no fp32 staging buffer
            ft = nv.proto("b200_flash_attn_bwd_tc", [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, f32, i32, vp])
            rc = ft(nv.ptr(qkv), nv.ptr(out), nv.ptr(dout), nv.ptr(lse), nv.ptr(delta), nv.ptr(dqkv), B, S, H, Hkv, D, int(causal),
                    float(scale), nv.dtype_code(qkv.dtype), nv.stream())
            if rc != 10006:
                nv.check(rc, "flash_attn_bwd_tc")
                nv.count_launch(3)   # delta, dK/dV pass, dQ pass
                return dqkv, None, None, None, None, None, None, None, None, None
        dqkv32 = torch.empty(qkv.shape, dtype=torch.float32, device=qkv.device)
        fn = nv.proto("b200_flash_attn_bwd", [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, f32, f32, u64, i32, vp])
        nv.check(fn(nv.ptr(qkv), nv.ptr(out), nv.ptr(dout), nv.ptr(lse), nv.ptr(delta), nv.ptr(dqkv32), nv.ptr(dqkv),
                    B, S, H, Hkv, D, int(causal), float(scale), float(dropout_p), int(seed), nv.dtype_code(qkv.dtype),
                    nv.stream()), "flash_attn_bwd")
        nv.count_launch(4)   # memset, delta, backward, cast
        return dqkv, None, None, None, None, None, None, None, None, None


def flash_attention(qkv: torch.Tensor, B: int, S: int, H: int, Hkv: int, D: int, causal: bool, scale: float,
                    dropout_p: float = 0.0, seed: int = 0) -> torch.Tensor:
    """Context [B*S, H*D] of dropout(softmax(Q K^T * scale + mask)) V over the fused projection ``qkv``
    [B*S, (H + 2 Hkv) * D]. The dropout mask is a pure function of (seed, batch, head, row, key)."""
    if not (qkv.requires_grad and torch.is_grad_enabled()):
        return _fwd(qkv, B, S, H, Hkv, D, causal, scale, False, dropout_p, seed)[0]
    return _FlashAttentionFn.apply(qkv, B, S, H, Hkv, D, causal, scale, dropout_p, seed)


def flash_attention_varlen(qkv: torch.Tensor, cu_seqlens: torch.Tensor, max_seqlen: int, H: int, Hkv: int, D: int,
                           scale: float, causal: bool = True) -> torch.Tensor:
    """Forward-only attention over sequences packed back to back in ``qkv`` [T, (H + 2 Hkv) * D]; sequence b is rows
    [cu_seqlens[b], cu_seqlens[b+1]) (int32 device tensor). The prefill of a serving iteration: several prompts of
    different lengths in one pass (what vLLM's scheduler does for the reference's `vllm serve` deployments,
    LLM_on_Kubernetes/Inference_Platfrom/README.md:1504-1509)."""
    n = cu_seqlens.numel() - 1
    if not nv.use_native(qkv) or not flash_supported(qkv, D):
        from .attention import _ref_attention
        cu = cu_seqlens.tolist()
        outs = [_ref_attention(qkv[cu[i]:cu[i + 1]], 1, cu[i + 1] - cu[i], H, Hkv, D, causal, 0, scale) for i in range(n)]
        return torch.cat(outs, 0)
    qkv = qkv.contiguous()
    out = torch.empty(qkv.shape[0], H * D, dtype=qkv.dtype, device=qkv.device)
    fn = nv.proto("b200_flash_attn_varlen_fwd", [vp, vp, vp, i32, i32, i32, i32, i32, i32, f32, i32, vp])
    nv.check(fn(nv.ptr(qkv), nv.ptr(out), nv.ptr(cu_seqlens), n, int(max_seqlen), H, Hkv, D, int(causal), float(scale),
                nv.dtype_code(qkv.dtype), nv.stream()), "flash_attn_varlen_fwd")
    nv.count_launch()
    return out
