"""K3 (small side) — wrappers for the skinny LoRA products in ``csrc/lora.cu`` (bf16), each with a
PyTorch oracle. See ``ops/linear.py`` for how they combine with the tcgen05 GEMM into a LoRA layer.

Reference: peft ``lora.Linear`` configured by ``LoraConfig(r, lora_alpha, lora_dropout, target_modules)``
in Fine-Tuning/qwen3-8b-lora-dist.py:118-125 and Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:113-120.
"""
from __future__ import annotations

import ctypes

import torch

from . import _native as nv

vp, i32, i64, u64, f32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_ulonglong, ctypes.c_float


def keep_mask(shape, seed: int, p: float, device) -> torch.Tensor:
    """The Philox keep-mask the kernels regenerate (materialised via the dropout kernel; oracle use)."""
    from .elementwise import _DropoutFn
    ones = torch.ones(shape, dtype=torch.float32, device=device)
    return _DropoutFn.apply(ones, p, seed) > 0


def lora_down(x: torch.Tensor, a: torch.Tensor, alpha: float = 1.0, seed: int = 0, p: float = 0.0) -> torch.Tensor:
    """T[M, r] = alpha * dropout(x)[M, K] @ a[r, K]^T."""
    M, K = x.shape
    r = a.shape[0]
    if not (nv.use_native(x) and x.dtype == torch.bfloat16):
        xd = x.float()
        if p > 0:
            xd = torch.nn.functional.dropout(xd, p, True)
        return (alpha * xd @ a.float().t()).to(x.dtype)
    if x.stride(1) != 1:
        x = x.contiguous()
    out = torch.empty(M, r, dtype=x.dtype, device=x.device)
    fn = nv.proto("b200_lora_down", [vp, vp, vp, i32, i32, i32, i64, i64, i64, f32, u64, f32, vp])
    nv.check(fn(nv.ptr(x), nv.ptr(a), nv.ptr(out), M, K, r, x.stride(0), a.stride(0), out.stride(0), alpha, seed, p,
                nv.stream()), "lora_down")
    nv.count_launch()
    return out


def lora_wgrad(pm: torch.Tensor, x: torch.Tensor, out: torch.Tensor, alpha: float = 1.0, seed: int = 0,
               p: float = 0.0) -> torch.Tensor:
    """out[r, K] += alpha * pm[M, r]^T @ dropout(x)[M, K]   (out: fp32, accumulated in place)."""
    M, K = x.shape
    r = pm.shape[1]
    if not (nv.use_native(x) and x.dtype == torch.bfloat16):
        xd = x.float()
        if p > 0:
            xd = torch.nn.functional.dropout(xd, p, True)
        out += alpha * pm.float().t() @ xd
        return out
    assert out.dtype == torch.float32 and out.stride(1) == 1
    fn = nv.proto("b200_lora_wgrad", [vp, vp, vp, i32, i32, i32, i64, i64, i64, f32, u64, f32, vp])
    nv.check(fn(nv.ptr(pm), nv.ptr(x), nv.ptr(out), M, K, r, pm.stride(0), x.stride(0), out.stride(0), alpha, seed, p,
                nv.stream()), "lora_wgrad")
    nv.count_launch()
    return out


def lora_dx(dt: torch.Tensor, a: torch.Tensor, dx: torch.Tensor, alpha: float = 1.0, seed: int = 0,
            p: float = 0.0) -> torch.Tensor:
    """dx[M, K] += mask/(1-p) * alpha * dt[M, r] @ a[r, K]   (in place)."""
    M, K = dx.shape
    r = a.shape[0]
    if not (nv.use_native(dx) and dx.dtype == torch.bfloat16):
        dx += (alpha * dt.float() @ a.float()).to(dx.dtype)
        return dx
    fn = nv.proto("b200_lora_dx", [vp, vp, vp, i32, i32, i32, i64, i64, i64, f32, u64, f32, vp])
    nv.check(fn(nv.ptr(dt), nv.ptr(a), nv.ptr(dx), M, K, r, dt.stride(0), a.stride(0), dx.stride(0), alpha, seed, p,
                nv.stream()), "lora_dx")
    nv.count_launch()
    return dx
