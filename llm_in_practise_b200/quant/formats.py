"""W4A16 weight container and the on-disk formats of the reference's quantisation lesson.

``W4Weight``  group-wise 4-bit weight in the kernel layout of ``csrc/w4.cu`` (packed [N, K/2], K contiguous, even
              element in the high nibble; scale [N, K/g] fp32; zs = zero * scale or None for a fixed zero of 8).
``quantize_rtn``  round-to-nearest group quantiser (symmetric or asymmetric) — the arithmetic every other method ends in.

Readers / writers (bit layouts reconstructed from the formats' public descriptions):
* GPTQ (gptqmodel ``model.save``; Quantization/GPTQModel/quantize_qwen3_4b_gptq.py:16-46): ``qweight`` int32 [K/8, N]
  (8 nibbles per word along K, least-significant first), ``qzeros`` int32 [K/g, N/8] (nibbles along N, stored as
  zero - 1 in the classic v1 checkpoints), ``scales`` fp16 [K/g, N], ``g_idx`` int32 [K]; ``quantize_config.json``
  {bits 4, group_size 128, desc_act false}.
* compressed-tensors ``pack-quantized`` (llm-compressor ``oneshot``; Quantization/LLM-Compressor/*/quantize_*.py):
  ``weight_packed`` int32 [N, K/8] (nibbles along K, LSB first, values offset by 8 when symmetric),
  ``weight_scale`` [N, K/g], ``weight_zero_point`` int32 [N/8, K/g] (asymmetric only, nibbles along N),
  ``weight_shape`` [2]; ``quantization_config`` inside ``config.json``.
* AutoAWQ ``gemm`` (the published ``Qwen3-*-AWQ`` checkpoints the serving recipes load with ``quantization: awq``,
  Deployment/Ray/serve_deploy_examples/qwen3_app_minimal.yaml:9-13): ``qweight`` int32 [K, N/8] and ``qzeros`` int32
  [K/g, N/8] — eight OUTPUT columns per word, nibble i holding column ``8*b + (0,2,4,6,1,3,5,7)[i]`` — ``scales`` fp16
  [K/g, N]; ``w = (q - zero) * scale`` with no zero offset; ``quantization_config`` {quant_method awq, bits 4,
  group_size 128, zero_point true, version gemm}.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional, Tuple

import torch

from ..ops.w4 import w4_dequant, w4_gemv

INT4_LUT = torch.arange(16, dtype=torch.float32)


@dataclass
class W4Weight:
    packed: torch.Tensor            # uint8 [N * K / 2]
    scale: torch.Tensor             # fp32 [N * K / group]
    zs: Optional[torch.Tensor]      # fp32 [N * K / group] = zero_point * scale
    N: int
    K: int
    group: int = 128
    lut: Optional[torch.Tensor] = None   # 16 floats; None => NF4 table

    def _lut(self):
        if self.lut is None:
            self.lut = INT4_LUT.to(self.packed.device)
        elif self.lut.device != self.packed.device:
            self.lut = self.lut.to(self.packed.device)
        return self.lut

    def dequantize(self, dtype=torch.bfloat16, out=None, transpose: bool = False) -> torch.Tensor:
        return w4_dequant(self.packed, self.scale, self.zs, self.N, self.K, self.group, self._lut(), dtype, out,
                          transpose=transpose)

    def can_transpose(self) -> bool:
        return self.N % 64 == 0 and self.K % 64 == 0

    def gemv(self, x: torch.Tensor, bias=None) -> torch.Tensor:
        return w4_gemv(x, self.packed, self.scale, self.zs, self.N, self.K, self.group, self._lut(), bias)

    def to(self, device) -> "W4Weight":
        mv = lambda t: None if t is None else t.to(device)  # noqa: E731
        return W4Weight(mv(self.packed), mv(self.scale), mv(self.zs), self.N, self.K, self.group, mv(self.lut))

    def codes(self) -> torch.Tensor:
        """uint8 [N, K] with values 0..15."""
        p = self.packed.view(self.N, self.K // 2)
        return torch.stack([p >> 4, p & 15], dim=-1).view(self.N, self.K)

    def zero_points(self) -> torch.Tensor:
        if self.zs is None:
            return torch.full((self.N, self.K // self.group), 8, dtype=torch.int32, device=self.packed.device)
        return torch.round(self.zs.view(self.N, -1) / self.scale.view(self.N, -1)).to(torch.int32)


def pack_codes(q: torch.Tensor) -> torch.Tensor:
    """uint8 codes [N, K] -> packed bytes [N*K/2] (even element high nibble)."""
    q = q.to(torch.uint8)
    return ((q[:, 0::2] << 4) | q[:, 1::2]).reshape(-1).contiguous()


def quantize_rtn(w: torch.Tensor, group: int = 128, symmetric: bool = True) -> W4Weight:
    N, K = w.shape
    g = w.float().view(N, K // group, group)
    if symmetric:
        amax = g.abs().amax(-1).clamp_min(1e-10)
        scale = amax / 7.0                      # codes 1..15 <-> [-7, 7] * scale, zero point 8
        q = torch.clamp(torch.round(g / scale.unsqueeze(-1)) + 8, 0, 15)
        zs = scale * 8.0
    else:
        lo, hi = g.amin(-1), g.amax(-1)
        scale = ((hi - lo) / 15.0).clamp_min(1e-10)
        zero = torch.clamp(torch.round(-lo / scale), 0, 15)
        q = torch.clamp(torch.round(g / scale.unsqueeze(-1)) + zero.unsqueeze(-1), 0, 15)
        zs = scale * zero
    return W4Weight(pack_codes(q.view(N, K)), scale.reshape(-1).contiguous(), zs.reshape(-1).contiguous(), N, K, group,
                    INT4_LUT.to(w.device))


# ------------------------------------------------------------------------------------------------
# int32 nibble packing helpers
# ------------------------------------------------------------------------------------------------
def _pack_int32(x: torch.Tensor, dim: int) -> torch.Tensor:
    """Pack 8 consecutive 4-bit values along ``dim`` into one int32, least-significant nibble first."""
    x = x.to(torch.int64).movedim(dim, -1)
    x = x.reshape(*x.shape[:-1], x.shape[-1] // 8, 8)
    shifts = torch.arange(0, 32, 4, device=x.device, dtype=torch.int64)
    v = (x << shifts).sum(-1)
    v = torch.where(v >= 2 ** 31, v - 2 ** 32, v).to(torch.int32)
    return v.movedim(-1, dim)


def _unpack_int32(v: torch.Tensor, dim: int) -> torch.Tensor:
    v = v.to(torch.int64) & 0xFFFFFFFF
    v = v.movedim(dim, -1)
    shifts = torch.arange(0, 32, 4, device=v.device, dtype=torch.int64)
    x = ((v.unsqueeze(-1) >> shifts) & 15).reshape(*v.shape[:-1], v.shape[-1] * 8)
    return x.movedim(-1, dim).to(torch.uint8)


# ------------------------------------------------------------------------------------------------
# GPTQ
# ------------------------------------------------------------------------------------------------
def to_gptq(w: W4Weight, zero_offset: int = 1) -> Dict[str, torch.Tensor]:
    q = w.codes()                                            # [N, K]
    z = w.zero_points()                                      # [N, K/g]
    return {"qweight": _pack_int32(q.t().contiguous(), 0),                       # [K/8, N]
            "qzeros": _pack_int32((z.t().contiguous() - zero_offset) & 15, 1),   # [K/g, N/8]
            "scales": w.scale.view(w.N, -1).t().contiguous().to(torch.float16),
            "g_idx": (torch.arange(w.K, device=q.device) // w.group).to(torch.int32)}


def from_gptq(t: Dict[str, torch.Tensor], group: int = 128, zero_offset: int = 1) -> W4Weight:
    qw, qz, sc = t["qweight"], t["qzeros"], t["scales"]
    K, N = qw.shape[0] * 8, qw.shape[1]
    g_idx = t.get("g_idx")
    if g_idx is not None and not torch.equal(g_idx.to(torch.int64), torch.arange(K, device=g_idx.device) // group):
        raise NotImplementedError("desc_act=True (permuted g_idx) checkpoints are not supported; the reference uses desc_act=False")
    q = _unpack_int32(qw, 0).t().contiguous()                # [N, K]
    z = ((_unpack_int32(qz, 1).to(torch.int32) + zero_offset) & 15).t().contiguous().float()   # [N, K/g]
    s = sc.t().contiguous().float()
    return W4Weight(pack_codes(q), s.reshape(-1), (z * s).reshape(-1), N, K, group, INT4_LUT.to(qw.device))


# ------------------------------------------------------------------------------------------------
# AutoAWQ (gemm)
# ------------------------------------------------------------------------------------------------
_AWQ_ORDER = (0, 2, 4, 6, 1, 3, 5, 7)        # nibble i of a word holds column 8 * block + _AWQ_ORDER[i]


def _awq_unpack(v: torch.Tensor) -> torch.Tensor:
    """int32 [R, N/8] -> uint8 [R, N] in natural column order."""
    x = _unpack_int32(v, 1)                                  # column 8 * b + i = nibble i of word b
    R = x.shape[0]
    out = torch.empty_like(x).view(R, -1, 8)
    out[:, :, list(_AWQ_ORDER)] = x.view(R, -1, 8)
    return out.view(R, -1)


def _awq_pack(x: torch.Tensor) -> torch.Tensor:
    R = x.shape[0]
    return _pack_int32(x.view(R, -1, 8)[:, :, list(_AWQ_ORDER)].reshape(R, -1), 1)


def to_awq(w: W4Weight) -> Dict[str, torch.Tensor]:
    return {"qweight": _awq_pack(w.codes().t().contiguous()),                    # [K, N/8]
            "qzeros": _awq_pack(w.zero_points().t().contiguous()),               # [K/g, N/8]
            "scales": w.scale.view(w.N, -1).t().contiguous().to(torch.float16)}  # [K/g, N]


def from_awq(t: Dict[str, torch.Tensor], group: int = 128) -> W4Weight:
    qw, sc = t["qweight"], t["scales"]
    K, N = qw.shape[0], qw.shape[1] * 8
    q = _awq_unpack(qw).t().contiguous()                     # [N, K]
    s = sc.t().contiguous().float()                          # [N, K/g]
    z = _awq_unpack(t["qzeros"]).t().contiguous().float() if "qzeros" in t else torch.full_like(s, 8.0)
    return W4Weight(pack_codes(q), s.reshape(-1), (z * s).reshape(-1), N, K, group, INT4_LUT.to(qw.device))


# ------------------------------------------------------------------------------------------------
# compressed-tensors pack-quantized
# ------------------------------------------------------------------------------------------------
def to_compressed_tensors(w: W4Weight, symmetric: bool) -> Dict[str, torch.Tensor]:
    q = w.codes()
    out = {"weight_packed": _pack_int32(q, 1),                                   # [N, K/8]
           "weight_scale": w.scale.view(w.N, -1).to(torch.bfloat16),
           "weight_shape": torch.tensor([w.N, w.K], dtype=torch.int64)}
    if not symmetric:
        out["weight_zero_point"] = _pack_int32(w.zero_points(), 0)                # [N/8, K/g]
    return out


def from_compressed_tensors(t: Dict[str, torch.Tensor], group: int = 128) -> W4Weight:
    wp = t["weight_packed"]
    N, K = (int(x) for x in t["weight_shape"]) if "weight_shape" in t else (wp.shape[0], wp.shape[1] * 8)
    q = _unpack_int32(wp, 1)[:, :K]
    s = t["weight_scale"].float().view(N, -1)
    if "weight_zero_point" in t:
        z = _unpack_int32(t["weight_zero_point"], 0)[:N].float()
    else:
        z = torch.full_like(s, 8.0)
    return W4Weight(pack_codes(q), s.reshape(-1).contiguous(), (z * s).reshape(-1).contiguous(), N, K, group,
                    INT4_LUT.to(wp.device))
