"""K2/K14 — 4-bit weights: NF4 (QLoRA) and GPTQ/AWQ-style W4A16 (``csrc/w4.cu``).

Storage convention (shared by both families, see the kernel file): ``packed[N, K/2]`` uint8 with K
contiguous, element 2j in the high nibble; ``scale[N, K/group]`` fp32; optional ``zs[N, K/group]``
fp32 (= zero_point * scale); a 16-entry code book. ``w = lut[q] * scale - zs``.

``NF4Weight``  reproduces bitsandbytes' ``Params4bit`` semantics used by the reference's
``BitsAndBytesConfig(load_in_4bit, nf4, double_quant, compute bf16)``
(Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:101-106): block size 64, absmax per block, and the
nested 8-bit quantisation of absmax (block 256, dynamic code book, mean offset).
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional, Sequence

import torch

from . import _native as nv

vp, i32, i64, f32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float

NF4_CODE = torch.tensor([
    -1.0, -0.6961928009986877, -0.5250730514526367, -0.39491748809814453, -0.28444138169288635,
    -0.18477343022823334, -0.09105003625154495, 0.0, 0.07958029955625534, 0.16093020141124725,
    0.24611230194568634, 0.33791524171829224, 0.44070982933044434, 0.5626170039176941,
    0.7229568362236023, 1.0], dtype=torch.float32)


def create_dynamic_map(signed: bool = True, max_exponent_bits: int = 7, total_bits: int = 8) -> torch.Tensor:
    """The 8-bit "dynamic tree" code book of bitsandbytes (used for nested absmax quantisation and
    for the 8-bit optimizer states): values in [-1, 1] (signed) or [0, 1], denser near zero."""
    data = []
    non_sign_bits = total_bits - 1
    additional_items = 2 ** (non_sign_bits - max_exponent_bits) - 1
    for i in range(max_exponent_bits):
        fraction_items = int(2 ** (i + non_sign_bits - max_exponent_bits) + 1 if signed
                             else 2 ** (i + non_sign_bits - max_exponent_bits + 1) + 1)
        boundaries = torch.linspace(0.1, 1, fraction_items)
        means = (boundaries[:-1] + boundaries[1:]) / 2.0
        data += ((10 ** (-(max_exponent_bits - 1) + i)) * means).tolist()
        if signed:
            data += (-(10 ** (-(max_exponent_bits - 1) + i)) * means).tolist()
    if additional_items > 0:
        boundaries = torch.linspace(0.1, 1, additional_items + 1)
        means = (boundaries[:-1] + boundaries[1:]) / 2.0
        data += ((10 ** (-(max_exponent_bits - 1) + i)) * means).tolist()
        if signed:
            data += (-(10 ** (-(max_exponent_bits - 1) + i)) * means).tolist()
    data.append(0)
    data.append(1.0)
    data += [0] * (2 ** total_bits - len(data))
    data.sort()
    return torch.tensor(data, dtype=torch.float32)


def quantize_blockwise_8bit(x: torch.Tensor, code: torch.Tensor, blocksize: int = 256):
    """Nearest-code 8-bit block-wise quantisation (PyTorch; cold path: load time / checkpointing)."""
    n = x.numel()
    pad = (-n) % blocksize
    xf = torch.cat([x.reshape(-1).float(), x.new_zeros(pad, dtype=torch.float32)]) if pad else x.reshape(-1).float()
    blocks = xf.view(-1, blocksize)
    absmax = blocks.abs().amax(dim=1)
    normed = blocks / absmax.clamp_min(1e-38).unsqueeze(1)
    code = code.to(x.device)
    mid = (code[1:] + code[:-1]) / 2
    q = torch.bucketize(normed.reshape(-1), mid).to(torch.uint8)
    return q[:n], absmax


def dequantize_blockwise_8bit(q: torch.Tensor, absmax: torch.Tensor, code: torch.Tensor, blocksize: int = 256):
    code = code.to(q.device)
    idx = torch.arange(q.numel(), device=q.device) // blocksize
    return code[q.long()] * absmax[idx]


# ------------------------------------------------------------------------------------------------
# NF4
# ------------------------------------------------------------------------------------------------
def _nf4_quantize_ref(w: torch.Tensor):
    flat = w.reshape(-1, 64).float()
    absmax = flat.abs().amax(dim=1)
    normed = flat / absmax.clamp_min(1e-38).unsqueeze(1)
    code = NF4_CODE.to(w.device)
    q = (normed.unsqueeze(-1) - code).abs().argmin(dim=-1).to(torch.uint8)
    packed = (q[:, 0::2] << 4) | q[:, 1::2]
    return packed.reshape(-1), absmax


def _w4_dequant_ref(packed, scale, zs, N, K, group, lut, dtype):
    lut = (NF4_CODE if lut is None else lut).to(packed.device)
    p = packed.view(N, K // 2)
    q = torch.stack([p >> 4, p & 15], dim=-1).view(N, K).long()
    w = lut[q] * scale.view(N, K // group).repeat_interleave(group, dim=1)
    if zs is not None:
        w = w - zs.view(N, K // group).repeat_interleave(group, dim=1)
    return w.to(dtype)


def nf4_quantize(w: torch.Tensor):
    """-> (packed uint8 [n/2], absmax fp32 [n/64])."""
    if w.numel() % 64:
        raise ValueError("NF4 needs a multiple of 64 values")
    if not nv.use_native(w):
        return _nf4_quantize_ref(w)
    w = w.contiguous()
    packed = torch.empty(w.numel() // 2, dtype=torch.uint8, device=w.device)
    absmax = torch.empty(w.numel() // 64, dtype=torch.float32, device=w.device)
    fn = nv.proto("b200_nf4_quantize", [vp, vp, vp, i64, i32, vp])
    nv.check(fn(nv.ptr(w), nv.ptr(packed), nv.ptr(absmax), w.numel(), nv.dtype_code(w.dtype), nv.stream()), "nf4_quantize")
    nv.count_launch()
    return packed, absmax


def w4_dequant(packed, scale, zs, N: int, K: int, group: int, lut: Optional[torch.Tensor] = None,
               dtype=torch.bfloat16, out: Optional[torch.Tensor] = None,
               shards: Optional[Sequence[Sequence[int]]] = None, rows_per_shard: int = 0,
               transpose: bool = False) -> torch.Tensor:
    """Expand 4-bit weights to ``dtype`` [N, K] (``transpose``: [K, N], the layout the dgrad GEMM wants). With
    ``shards`` = [(packed_ptr, scale_ptr, zs_ptr), ...] the rows are read from several (possibly peer-GPU) buffers
    of ``rows_per_shard`` rows each."""
    want = out.dtype if out is not None else dtype
    if not (packed.is_cuda and nv.use_native(packed)) or want == torch.float32:   # fp32 only in calibration searches
        res = _w4_dequant_ref(packed, scale, zs, N, K, group, lut, dtype)
        if transpose:
            res = res.t().contiguous()
        if out is not None:
            out.copy_(res)
            return out
        return res
    if out is None:
        out = torch.empty((K, N) if transpose else (N, K), dtype=dtype, device=packed.device)
    if shards is None:
        shards = [(packed.data_ptr(), scale.data_ptr(), 0 if zs is None else zs.data_ptr())]
        rows_per_shard = N
    n = len(shards)
    PA = (ctypes.c_void_p * n)(*[s[0] for s in shards])
    SA = (ctypes.c_void_p * n)(*[s[1] for s in shards])
    ZA = (ctypes.c_void_p * n)(*[s[2] or None for s in shards])
    fn = nv.proto("b200_w4_dequant_t" if transpose else "b200_w4_dequant",
                  [vp, vp, vp, i32, i64, vp, i64, i32, i32, vp, i64, i32, vp])
    nv.check(fn(ctypes.cast(PA, vp), ctypes.cast(SA, vp), ctypes.cast(ZA, vp), n, rows_per_shard, nv.ptr(out), N, K,
                group, nv.ptr(lut), out.stride(0), nv.dtype_code(out.dtype), nv.stream()), "w4_dequant")
    nv.count_launch()
    return out


_LIN_ON = __import__("os").environ.get("B200_W4_LINEAR", "1") != "0"


def linear_lut(lut) -> Optional[tuple]:
    """(a, off) when ``lut[c] == a * (c - off)`` with an integer ``off`` (RTN / GPTQ / AWQ integer codes), else None.
    Decided once per table tensor (one device read), never inside a graph capture."""
    if lut is None or not _LIN_ON:
        return None
    hit = getattr(lut, "_b200_lin", 0)      # cached on the tensor object (an address can be reused by another table)
    if hit != 0:
        return hit
    if lut.is_cuda and torch.cuda.is_current_stream_capturing():
        return None
    v = lut.detach().float().cpu().tolist()
    a = v[1] - v[0]
    res = None
    if len(v) == 16 and a != 0.0:
        off = -v[0] / a
        if abs(off - round(off)) < 1e-6 and abs(off) <= 64 and all(abs(v[c] - a * (c - round(off))) <= 1e-6 * max(1.0, abs(v[c])) for c in range(16)):
            res = (float(a), float(round(off)))
    lut._b200_lin = res
    return res


def w4_gemv(x, packed, scale, zs, N: int, K: int, group: int, lut=None, bias=None) -> torch.Tensor:
    """y[m, :] = x[m, :] @ dequant(W)^T for 1 <= m <= 64 rows (decode / serving steps)."""
    M = x.shape[0]
    if not nv.use_native(x):
        w = _w4_dequant_ref(packed, scale, zs, N, K, group, lut, torch.float32)
        y = x.float() @ w.t()
        if bias is not None:
            y = y + bias.float()
        return y.to(x.dtype)
    x = x.contiguous()
    y = torch.empty(M, N, dtype=x.dtype, device=x.device)
    lin = linear_lut(lut) if K % 128 == 0 else None
    if lin is not None:   # integer codes: table-free conversion (csrc/gemv.cu, LIN)
        f32 = ctypes.c_float
        fl = nv.proto("b200_w4_gemv_linear", [vp, vp, vp, vp, vp, vp, i32, i64, i32, i32, i64, i64, f32, f32, i32, vp])
        nv.check(fl(nv.ptr(x), nv.ptr(packed), nv.ptr(scale), nv.ptr(zs), nv.ptr(bias), nv.ptr(y), M, N, K, group,
                    x.stride(0), y.stride(0), lin[0], lin[1], nv.dtype_code(x.dtype), nv.stream()), "w4_gemv_linear")
        nv.count_launch()
        return y
    if M > 8:     # 9..64 tokens: token groups of 8 share one pass over the codes (csrc/gemv.cu)
        fb = nv.proto("b200_w4_gemv_batched", [vp, vp, vp, vp, vp, vp, vp, i32, i64, i32, i32, i64, i64, i32, vp])
        nv.check(fb(nv.ptr(x), nv.ptr(packed), nv.ptr(scale), nv.ptr(zs), nv.ptr(lut), nv.ptr(bias), nv.ptr(y), M, N, K,
                    group, x.stride(0), y.stride(0), nv.dtype_code(x.dtype), nv.stream()), "w4_gemv_batched")
        nv.count_launch()
        return y
    fn = nv.proto("b200_w4_gemv", [vp, vp, vp, vp, vp, vp, vp, i32, i64, i32, i32, i64, i64, i32, vp])
    nv.check(fn(nv.ptr(x), nv.ptr(packed), nv.ptr(scale), nv.ptr(zs), nv.ptr(lut), nv.ptr(bias), nv.ptr(y), M, N, K,
                group, x.stride(0), y.stride(0), nv.dtype_code(x.dtype), nv.stream()), "w4_gemv")
    nv.count_launch()
    return y


@dataclass
class NF4Weight:
    """A frozen NF4-quantised [N, K] matrix (bitsandbytes Params4bit equivalent).

    ``absmax`` is kept either in fp32 or — with ``double_quant`` — as uint8 codes + per-256 second-level
    absmax + mean offset (4.127 bits/param in total); ``scales()`` returns the fp32 view either way."""
    packed: torch.Tensor
    N: int
    K: int
    absmax: Optional[torch.Tensor] = None         # fp32 [N*K/64]           (single quant, or decoded cache)
    absmax_q: Optional[torch.Tensor] = None       # uint8 [N*K/64]          (double quant)
    absmax2: Optional[torch.Tensor] = None        # fp32 [ceil(N*K/64/256)]
    code2: Optional[torch.Tensor] = None          # fp32 [256]
    offset: float = 0.0
    group: int = 64

    @classmethod
    def quantize(cls, w: torch.Tensor, double_quant: bool = True, keep_decoded: bool = True) -> "NF4Weight":
        N, K = w.shape
        packed, absmax = nf4_quantize(w)
        if not double_quant:
            return cls(packed=packed, N=N, K=K, absmax=absmax)
        offset = float(absmax.mean().item())
        code2 = create_dynamic_map(signed=True).to(w.device)
        q, absmax2 = quantize_blockwise_8bit(absmax - offset, code2, 256)
        self = cls(packed=packed, N=N, K=K, absmax_q=q, absmax2=absmax2, code2=code2, offset=offset)
        if keep_decoded:
            self.absmax = self.scales()
        return self

    def scales(self) -> torch.Tensor:
        if self.absmax is not None:
            return self.absmax
        n = self.absmax_q.numel()
        if nv.use_native(self.absmax_q):
            out = torch.empty(n, dtype=torch.float32, device=self.absmax_q.device)
            fn = nv.proto("b200_nf4_absmax_decode", [vp, vp, vp, f32, vp, i64, vp])
            nv.check(fn(nv.ptr(self.absmax_q), nv.ptr(self.code2), nv.ptr(self.absmax2), self.offset, nv.ptr(out), n,
                        nv.stream()), "nf4_absmax_decode")
            nv.count_launch()
            return out
        return dequantize_blockwise_8bit(self.absmax_q, self.absmax2, self.code2, 256) + self.offset

    def dequantize(self, dtype=torch.bfloat16, out=None, transpose: bool = False) -> torch.Tensor:
        return w4_dequant(self.packed, self.scales(), None, self.N, self.K, self.group, None, dtype, out,
                          transpose=transpose)

    def can_transpose(self) -> bool:
        return self.N % 64 == 0 and self.K % 64 == 0

    def nbytes(self) -> int:
        n = self.packed.numel()
        if self.absmax_q is not None:
            n += self.absmax_q.numel() + self.absmax2.numel() * 4 + 1024
        else:
            n += self.absmax.numel() * 4
        return n

    def to(self, device) -> "NF4Weight":
        mv = lambda t: None if t is None else t.to(device)
        return NF4Weight(mv(self.packed), self.N, self.K, mv(self.absmax), mv(self.absmax_q), mv(self.absmax2),
                         mv(self.code2), self.offset, self.group)
