"""Decode-step projection with a fused tail (``b200_gemv_fused``, csrc/gemv.cu): one launch computes

    y   = x W^T                       (dense 16-bit, NF4 or GPTQ/AWQ 4-bit weights, <= 64 tokens)
    y   = sum over TP ranks of y      (row-parallel layer: partials staged in symmetric memory, epoch flags over NVLink)
    res = res + y                     (residual stream)
    out = RMSNorm(res) * norm_weight  (the input of the NEXT projection)

which is what a decoder layer does after ``o_proj`` and after ``down_proj``. Stand-alone it would be a GEMV, an
all-reduce kernel and a residual/norm kernel (vLLM: row-parallel linear + custom all-reduce + fused_add_rms_norm — the
stack behind the reference's ``--tensor-parallel-size`` knob, Quantization/README.md:175). The epoch and the barrier tickets
live on the device, so the launch is captured in the decode CUDA graph at every TP degree.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Tuple

import torch

from . import _native as nv
from .w4 import NF4Weight, linear_lut

vp, i32, i64, f32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float
MAX_TOKENS = 64
CHANNEL = 5


class _Args(ctypes.Structure):
    _fields_ = [("x", vp), ("w", vp), ("scale", vp), ("zs", vp), ("lut", vp), ("bias", vp),
                ("kind", i32), ("M", i32), ("N", i64), ("K", i32), ("group", i32), ("x_rs", i64), ("w_rs", i64),
                ("lin_a", f32), ("lin_off", f32), ("dtype", i32),
                ("world", i32), ("rank", i32), ("channel", i32), ("stage_ptrs", vp), ("flag_ptrs", vp), ("half_elems", i64),
                ("state", vp), ("res_in", vp), ("res_out", vp), ("norm_w", vp), ("x_out", vp), ("eps", f32)]


class FusedTail:
    """Staging buffer + device state of the fused tail for one model replica (one per TP group, or local at TP 1)."""

    def __init__(self, hidden: int, device, tp=None, dtype=torch.bfloat16):
        self.hidden = hidden
        self.world = tp.world if tp is not None else 1
        self.rank = tp.rank if tp is not None else 0
        self.half = MAX_TOKENS * hidden
        self.state = torch.zeros(4, dtype=torch.int32, device=device)
        if self.world > 1:
            from ..parallel.symm import SymmBuffer
            self.stage = SymmBuffer(2 * self.half, dtype, device, tp.group)
            self.flags = tp.flags.flags
            self._stage_c = ctypes.cast(self.stage.c_ptrs(), vp)
            self._flags_c = ctypes.cast(self.flags.c_ptrs(), vp)
        else:
            self.local = torch.empty(self.half, dtype=dtype, device=device)
            self._arr = (ctypes.c_void_p * 8)(self.local.data_ptr(), *([0] * 7))
            self._stage_c = ctypes.cast(self._arr, vp)
            self._flags_c = None


def _weight_of(lin):
    """(kind, w, scale, zs, lut, group, lin_a, lin_off, N, K, bias) of a FusedLinear (or the inner one of a row-parallel layer)."""
    q = lin.nf4
    if q is None:
        w = lin.weight
        return 0, w, None, None, None, 128, 0.0, 0.0, w.shape[0], w.shape[1], lin.bias
    if isinstance(q, NF4Weight):
        return 1, q.packed, q.scales(), None, None, q.group, 0.0, 0.0, q.N, q.K, lin.bias
    if hasattr(q, "shards"):
        return None
    lut = q._lut()
    lin_p = linear_lut(lut) if q.K % 128 == 0 else None
    if lin_p is not None:
        return 2, q.packed, q.scale, q.zs, None, q.group, lin_p[0], lin_p[1], q.N, q.K, lin.bias
    return 1, q.packed, q.scale, q.zs, lut, q.group, 0.0, 0.0, q.N, q.K, lin.bias


def supported(lin, x: torch.Tensor, hidden: int) -> bool:
    inner = getattr(lin, "inner", lin)
    if not (x.is_cuda and nv.use_native(x)) or x.dtype not in (torch.bfloat16, torch.float16):
        return False
    if x.shape[0] > MAX_TOKENS or hidden % 8 or hidden > 8192 or getattr(inner, "has_lora", False):
        return False
    w = _weight_of(inner)
    if w is None:
        return False
    kind, wt, _, _, _, group, _, _, N, K, _ = w
    if kind == 0 and wt.dtype != x.dtype:
        return False
    return K % 128 == 0 and N == hidden and (kind == 0 or (group % 32 == 0 and K % group == 0))


def gemv_fused(lin, x: torch.Tensor, res: Optional[torch.Tensor], norm_w: Optional[torch.Tensor], eps: float,
               tail: FusedTail) -> Tuple[torch.Tensor, torch.Tensor]:
    """-> (normed, new_residual), both [M, hidden] in x's dtype. ``lin``: FusedLinear or RowParallelLinear."""
    inner = getattr(lin, "inner", lin)
    kind, w, scale, zs, lut, group, la, lo, N, K, bias = _weight_of(inner)
    M = x.shape[0]
    x = x.contiguous()
    out = torch.empty(M, N, dtype=x.dtype, device=x.device)
    res_out = torch.empty(M, N, dtype=x.dtype, device=x.device)
    if tail.world > 1:
        bias = getattr(lin, "bias", None) if tail.rank == 0 else None      # the row-parallel bias is added once
    a = _Args()
    a.x, a.w = x.data_ptr(), w.data_ptr()
    a.scale = None if scale is None else scale.data_ptr()
    a.zs = None if zs is None else zs.data_ptr()
    a.lut = None if lut is None else lut.data_ptr()
    a.bias = None if bias is None else bias.data_ptr()
    a.kind, a.M, a.N, a.K, a.group = kind, M, N, K, group
    a.x_rs, a.w_rs = x.stride(0), (w.stride(0) if kind == 0 else 0)
    a.lin_a, a.lin_off, a.dtype = la, lo, nv.dtype_code(x.dtype)
    a.world, a.rank, a.channel = tail.world, tail.rank, CHANNEL
    a.stage_ptrs, a.flag_ptrs, a.half_elems = tail._stage_c, tail._flags_c, tail.half
    a.state = tail.state.data_ptr()
    a.res_in = None if res is None else res.data_ptr()
    a.res_out = res_out.data_ptr()
    a.norm_w = None if norm_w is None else norm_w.data_ptr()
    a.x_out = out.data_ptr()
    a.eps = eps
    fn = nv.proto("b200_gemv_fused", [ctypes.POINTER(_Args), vp])
    nv.check(fn(ctypes.byref(a), nv.stream()), "gemv_fused")
    nv.count_launch()
    return out, res_out
