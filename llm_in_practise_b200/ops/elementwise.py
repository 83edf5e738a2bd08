"""K6/K7/K9/K10 — fused element-wise ops (``csrc/elementwise.cu``) with PyTorch oracles.

* ``qknorm_rope``  — per-head RMSNorm of q and k (HF Qwen3 ``q_norm``/``k_norm``) + rotary embedding
  on a slice of the fused qkv projection, rotate-half (Qwen3) or interleaved (DeepSeekLike) layout.
* ``swiglu``       — ``silu(gate) * up`` on the fused gate|up projection (HF ``Qwen3MLP``).
* ``gelu``         — exact erf GELU (``nn.GELU`` in GPTLike).
* ``embedding``    — token gather (+ learned / sinusoidal positions), scatter-add backward.
* ``dropout``      — Philox mask regenerated in backward from (seed), never stored.

Reference: Fine-Tuning/*.py (HF Qwen3 modules), LLM_Distributed_Trainning/PyTorch/ddp_basics/
ddp_gpt_wikitext2.py:86-165, …/transformer_basics/DeepSeekLike_spare_MoE_wikitext2.py:131-174.
"""
from __future__ import annotations

import ctypes
import math
from typing import Optional

import torch
import torch.nn.functional as F

from . import _native as nv

vp, i32, i64, u64, f32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_ulonglong, ctypes.c_float


# ------------------------------------------------------------------------------------------------
# rotary tables
# ------------------------------------------------------------------------------------------------
def _yarn_correction_dim(num_rotations: float, dim: int, base: float, max_pos: int) -> float:
    return dim * math.log(max_pos / (num_rotations * 2 * math.pi)) / (2 * math.log(base))


def rope_inv_freq(head_dim: int, theta: float, scaling: Optional[dict] = None, max_position_embeddings: Optional[int] = None,
                  device=None):
    """(inverse frequencies [head_dim / 2] fp32, attention factor) for plain RoPE or the ``rope_scaling`` of an HF config
    (``rope_type`` / ``type``: ``linear``, ``yarn``, ``llama3``) — e.g. DeepSeek-R1-0528-Qwen3-8B ships YaRN with factor 4,
    which the reference's fine-tune script switches off with ``rope_scaling=None``
    (Fine-Tuning/deepseek-r1-0528-qwen3-8b-qlora.dist.py:98-112) and its vLLM servers leave on."""
    idx = torch.arange(0, head_dim, 2, dtype=torch.float32, device=device)
    inv = 1.0 / (theta ** (idx / head_dim))
    kind = (scaling or {}).get("rope_type", (scaling or {}).get("type", "default"))
    if not scaling or kind in ("default", None):
        return inv, 1.0
    factor = float(scaling.get("factor", 1.0))
    if kind == "linear":
        return inv / factor, 1.0
    if kind == "yarn":
        orig = int(scaling.get("original_max_position_embeddings") or max_position_embeddings or 0)
        if orig <= 0:
            raise ValueError("yarn rope scaling needs original_max_position_embeddings")
        beta_fast, beta_slow = float(scaling.get("beta_fast") or 32), float(scaling.get("beta_slow") or 1)
        att = scaling.get("attention_factor", scaling.get("attn_factor"))
        msc = lambda s, m=1.0: 1.0 if s <= 1 else 0.1 * m * math.log(s) + 1.0  # noqa: E731
        if att is None:
            m, ma = scaling.get("mscale"), scaling.get("mscale_all_dim")
            att = msc(factor, m) / msc(factor, ma) if (m and ma) else msc(factor)
        low = _yarn_correction_dim(beta_fast, head_dim, theta, orig)
        high = _yarn_correction_dim(beta_slow, head_dim, theta, orig)
        if scaling.get("truncate", True):
            low, high = math.floor(low), math.ceil(high)
        low, high = max(low, 0), min(high, head_dim - 1)
        if low == high:
            high += 0.001
        ramp = ((torch.arange(head_dim // 2, dtype=torch.float32, device=device) - low) / (high - low)).clamp(0, 1)
        keep = 1.0 - ramp                                       # 1: high-frequency dims keep their frequency, 0: interpolated
        return (inv / factor) * (1 - keep) + inv * keep, float(att)
    if kind == "llama3":
        orig = int(scaling["original_max_position_embeddings"])
        lo_f, hi_f = float(scaling.get("low_freq_factor", 1.0)), float(scaling.get("high_freq_factor", 4.0))
        wavelen = 2 * math.pi / inv
        smooth = ((orig / wavelen - lo_f) / (hi_f - lo_f)).clamp(0, 1)
        mixed = (1 - smooth) * inv / factor + smooth * inv
        out = torch.where(wavelen > orig / lo_f, inv / factor, torch.where(wavelen < orig / hi_f, inv, mixed))
        return out, 1.0
    raise NotImplementedError(f"rope scaling type {kind!r} is not implemented (linear, yarn and llama3 are)")


def rope_tables(max_pos: int, head_dim: int, theta: float = 10000.0, device=None, scaling: Optional[dict] = None,
                max_position_embeddings: Optional[int] = None):
    """cos/sin tables [max_pos, head_dim/2] in fp32 (angles = pos * theta^(-2i/d)); with ``scaling`` the frequencies follow the
    HF ``rope_scaling`` rule and YaRN's attention factor is folded into the tables (as transformers does)."""
    inv_freq, att = rope_inv_freq(head_dim, theta, scaling, max_position_embeddings, device)
    ang = torch.outer(torch.arange(max_pos, dtype=torch.float32, device=device), inv_freq)
    if att != 1.0:
        return (ang.cos() * att).contiguous(), (ang.sin() * att).contiguous()
    return ang.cos().contiguous(), ang.sin().contiguous()


def _ref_qknorm_rope(qk, qw, kw, cos, sin, pos, n_q, n_k, D, interleaved, eps):
    # qk: [tokens, (n_q + n_k) * D]
    T = qk.shape[0]
    x = qk.view(T, n_q + n_k, D)
    if qw is not None:
        xf = x.float()
        xn = (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)).to(qk.dtype)
        w = torch.cat([qw.view(1, 1, D).expand(1, n_q, D), kw.view(1, 1, D).expand(1, n_k, D)], dim=1)
        x = (xn * w).to(qk.dtype)
    c = cos[pos.long()].unsqueeze(1)  # [T, 1, D/2]
    s = sin[pos.long()].unsqueeze(1)
    xf = x.float()
    if interleaved:
        a, b = xf[..., 0::2], xf[..., 1::2]
        out = torch.stack([a * c - b * s, b * c + a * s], dim=-1).flatten(-2)
    else:
        a, b = xf[..., : D // 2], xf[..., D // 2:]
        out = torch.cat([a * c - b * s, b * c + a * s], dim=-1)
    return out.to(qk.dtype).reshape(T, (n_q + n_k) * D)


class _QKNormRopeFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, qw, kw, cos, sin, pos, n_q, n_k, n_v, D, interleaved, eps):
        T = qkv.shape[0]
        out = torch.empty(T, (n_q + n_k + n_v) * D, dtype=qkv.dtype, device=qkv.device)
        fn = nv.proto("b200_qknorm_rope_fwd", [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i64, i64, i32, f32, vp, i32, vp])
        nv.check(fn(nv.ptr(qkv), nv.ptr(out), nv.ptr(qw), nv.ptr(kw), nv.ptr(cos), nv.ptr(sin), nv.ptr(pos), T, n_q, n_k,
                    n_v, D, qkv.stride(0), out.stride(0), int(interleaved), eps, None, nv.dtype_code(qkv.dtype),
                    nv.stream()), "qknorm_rope_fwd")
        nv.count_launch()
        ctx.save_for_backward(qkv, qw, kw, cos, sin, pos)
        ctx.cfg = (n_q, n_k, n_v, D, interleaved, eps)
        return out

    @staticmethod
    def backward(ctx, g):
        qkv, qw, kw, cos, sin, pos = ctx.saved_tensors
        n_q, n_k, n_v, D, interleaved, eps = ctx.cfg
        T = qkv.shape[0]
        g = g.contiguous()
        dqkv = torch.empty_like(qkv)
        need_w = qw is not None and (ctx.needs_input_grad[1] or ctx.needs_input_grad[2])
        dwq = torch.zeros(D, dtype=torch.float32, device=qkv.device) if need_w else None
        dwk = torch.zeros(D, dtype=torch.float32, device=qkv.device) if need_w else None
        fn = nv.proto("b200_qknorm_rope_bwd", [vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i64, i64, i64, i32, f32, vp, vp, i32, vp])
        nv.check(fn(nv.ptr(g), nv.ptr(qkv), nv.ptr(dqkv), nv.ptr(qw), nv.ptr(kw), nv.ptr(cos), nv.ptr(sin), nv.ptr(pos),
                    T, n_q, n_k, n_v, D, g.stride(0), qkv.stride(0), dqkv.stride(0), int(interleaved), eps, nv.ptr(dwq),
                    nv.ptr(dwk), nv.dtype_code(qkv.dtype), nv.stream()), "qknorm_rope_bwd")
        nv.count_launch()
        return (dqkv, dwq.to(qw.dtype) if need_w else None, dwk.to(kw.dtype) if need_w else None,
                None, None, None, None, None, None, None, None, None)


def qknorm_rope(qkv: torch.Tensor, qw: Optional[torch.Tensor], kw: Optional[torch.Tensor], cos: torch.Tensor,
                sin: torch.Tensor, pos: torch.Tensor, n_q: int, n_k: int, head_dim: int,
                interleaved: bool = False, eps: float = 1e-6) -> torch.Tensor:
    """Normalise (optional) and rotate the q and k heads of the fused projection ``qkv``
    ([tokens, (n_q + n_k + n_v) * D], q heads first, then k, then v). Returns a tensor of the same
    shape; v heads are copied through so the attention kernel reads a single tensor."""
    n_v = qkv.shape[1] // head_dim - n_q - n_k
    if nv.use_native(qkv):
        return _QKNormRopeFn.apply(qkv.contiguous(), qw, kw, cos, sin, pos, n_q, n_k, n_v, head_dim, interleaved, eps)
    qk = _ref_qknorm_rope(qkv[:, : (n_q + n_k) * head_dim].contiguous(), qw, kw, cos, sin, pos, n_q, n_k,
                          head_dim, interleaved, eps)
    return torch.cat([qk, qkv[:, (n_q + n_k) * head_dim:]], dim=1) if n_v else qk


# ------------------------------------------------------------------------------------------------
# SwiGLU / GELU
# ------------------------------------------------------------------------------------------------
class _SwiGLUFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, gu):
        rows, F2 = gu.shape
        Fh = F2 // 2
        h = torch.empty(rows, Fh, dtype=gu.dtype, device=gu.device)
        fn = nv.proto("b200_swiglu_fwd", [vp, vp, i64, i32, i32, vp])
        nv.check(fn(nv.ptr(gu), nv.ptr(h), rows, Fh, nv.dtype_code(gu.dtype), nv.stream()), "swiglu_fwd")
        nv.count_launch()
        ctx.save_for_backward(gu)
        return h

    @staticmethod
    def backward(ctx, dh):
        (gu,) = ctx.saved_tensors
        rows, F2 = gu.shape
        dgu = torch.empty_like(gu)
        fn = nv.proto("b200_swiglu_bwd", [vp, vp, vp, i64, i32, i32, vp])
        nv.check(fn(nv.ptr(dh.contiguous()), nv.ptr(gu), nv.ptr(dgu), rows, F2 // 2, nv.dtype_code(gu.dtype), nv.stream()),
                 "swiglu_bwd")
        nv.count_launch()
        return dgu


def swiglu(gate_up: torch.Tensor) -> torch.Tensor:
    """``silu(gate) * up`` for ``gate_up = [..., 2F]`` (gate in the first F columns)."""
    shape = gate_up.shape
    gu = gate_up.reshape(-1, shape[-1])
    if nv.use_native(gu):
        return _SwiGLUFn.apply(gu.contiguous()).view(*shape[:-1], shape[-1] // 2)
    Fh = shape[-1] // 2
    g, u = gu[:, :Fh], gu[:, Fh:]
    return (F.silu(g.float()).to(gu.dtype) * u).view(*shape[:-1], Fh)


class _GeluFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        y = torch.empty_like(x)
        fn = nv.proto("b200_gelu_fwd", [vp, vp, i64, i32, vp])
        nv.check(fn(nv.ptr(x), nv.ptr(y), x.numel(), nv.dtype_code(x.dtype), nv.stream()), "gelu_fwd")
        nv.count_launch()
        ctx.save_for_backward(x)
        return y

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        dx = torch.empty_like(x)
        fn = nv.proto("b200_gelu_bwd", [vp, vp, vp, i64, i32, vp])
        nv.check(fn(nv.ptr(dy.contiguous()), nv.ptr(x), nv.ptr(dx), x.numel(), nv.dtype_code(x.dtype), nv.stream()), "gelu_bwd")
        nv.count_launch()
        return dx


def gelu(x: torch.Tensor) -> torch.Tensor:
    vec = 4 if x.dtype == torch.float32 else 8
    if nv.use_native(x) and x.numel() % vec == 0:
        return _GeluFn.apply(x.contiguous())
    return F.gelu(x)


# ------------------------------------------------------------------------------------------------
# dropout
# ------------------------------------------------------------------------------------------------
class _DropoutFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, p, seed):
        y = torch.empty_like(x)
        fn = nv.proto("b200_dropout", [vp, vp, i64, u64, f32, i32, vp])
        nv.check(fn(nv.ptr(x), nv.ptr(y), x.numel(), seed, p, nv.dtype_code(x.dtype), nv.stream()), "dropout")
        nv.count_launch()
        ctx.p, ctx.seed = p, seed
        return y

    @staticmethod
    def backward(ctx, dy):
        dy = dy.contiguous()
        dx = torch.empty_like(dy)
        fn = nv.proto("b200_dropout", [vp, vp, i64, u64, f32, i32, vp])
        nv.check(fn(nv.ptr(dy), nv.ptr(dx), dy.numel(), ctx.seed, ctx.p, nv.dtype_code(dy.dtype), nv.stream()), "dropout")
        nv.count_launch()
        return dx, None, None


_seed_state = {"base": 0x5EED, "ctr": 0}


def manual_seed(seed: int) -> None:
    """Seed of the Philox streams used by dropout / LoRA dropout (per-rank offsets are the caller's job)."""
    _seed_state["base"] = int(seed) & 0xFFFFFFFFFFFF
    _seed_state["ctr"] = 0


def next_seed() -> int:
    _seed_state["ctr"] += 1
    return (_seed_state["base"] * 0x9E3779B97F4A7C15 + _seed_state["ctr"] * 0xD1342543DE82EF95) & 0xFFFFFFFFFFFFFFFF


def set_step(step: int, micro: int = 0) -> None:
    """Training engines call this once per micro-batch: dropout seeds become a pure function of
    (base seed, step, micro-batch, layer id), so an activation-checkpoint recompute regenerates
    exactly the masks of the original forward."""
    _seed_state["step"] = int(step)
    _seed_state["micro"] = int(micro)


def layer_seed(uid: int) -> int:
    s = _seed_state
    x = (s["base"] * 0x9E3779B97F4A7C15) ^ ((s.get("step", 0) + 1) * 0xBF58476D1CE4E5B9) \
        ^ ((s.get("micro", 0) + 1) * 0x94D049BB133111EB) ^ ((uid + 1) * 0xD6E8FEB86659FD93)
    x &= 0xFFFFFFFFFFFFFFFF
    x ^= x >> 31
    x = (x * 0x7FB5D329728EA185) & 0xFFFFFFFFFFFFFFFF
    x ^= x >> 27
    return x


def dropout(x: torch.Tensor, p: float, training: bool = True, seed: Optional[int] = None) -> torch.Tensor:
    if not training or p <= 0.0:
        return x
    if nv.use_native(x):
        return _DropoutFn.apply(x.contiguous(), float(p), next_seed() if seed is None else seed)
    return F.dropout(x, p, True)


# ------------------------------------------------------------------------------------------------
# embedding
# ------------------------------------------------------------------------------------------------
class _EmbeddingFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ids, table, pos_table, period, scale):
        flat = ids.reshape(-1).contiguous()
        out = torch.empty(flat.numel(), table.shape[1], dtype=table.dtype, device=table.device)
        fn = nv.proto("b200_embedding_fwd", [vp, vp, vp, vp, i64, i32, i32, f32, i32, vp])
        nv.check(fn(nv.ptr(flat), nv.ptr(table), nv.ptr(pos_table), nv.ptr(out), flat.numel(), table.shape[1], period,
                    scale, nv.dtype_code(table.dtype), nv.stream()), "embedding_fwd")
        nv.count_launch()
        if table.requires_grad:
            from ..utils.arena import note_use
            note_use(table, False)              # its gradient arrives through autograd (possibly after a tied head's)
        ctx.save_for_backward(flat)
        ctx.meta = (table.shape, table.dtype, pos_table is not None, period, scale,
                    None if pos_table is None else pos_table.shape)
        return out.view(*ids.shape, table.shape[1])

    @staticmethod
    def backward(ctx, dout):
        (flat,) = ctx.saved_tensors
        tshape, tdtype, has_pos, period, scale, pshape = ctx.meta
        dout2 = dout.reshape(-1, tshape[1]).contiguous()
        dtable = None
        if ctx.needs_input_grad[1]:
            acc = torch.zeros(tshape, dtype=torch.float32, device=dout.device)
            fn = nv.proto("b200_embedding_bwd", [vp, vp, vp, i64, i32, f32, i32, vp])
            nv.check(fn(nv.ptr(flat), nv.ptr(dout2), nv.ptr(acc), flat.numel(), tshape[1], scale,
                        nv.dtype_code(dout2.dtype), nv.stream()), "embedding_bwd")
            nv.count_launch()
            dtable = acc.to(tdtype)
        dpos = None
        if has_pos and ctx.needs_input_grad[2]:
            g = dout2.float().view(-1, period, tshape[1]).sum(0)
            dpos = torch.zeros(pshape, dtype=torch.float32, device=dout.device)
            dpos[:period] = g
            dpos = dpos.to(tdtype)
        return None, dtable, dpos, None, None


def embedding(ids: torch.Tensor, table: torch.Tensor, pos_table: Optional[torch.Tensor] = None,
              scale: float = 1.0) -> torch.Tensor:
    """``table[ids] * scale (+ pos_table[position])``; ids [..., L], position = index along the last dim."""
    period = ids.shape[-1]
    vec = 4 if table.dtype == torch.float32 else 8
    if nv.use_native(table) and table.shape[1] % vec == 0:
        return _EmbeddingFn.apply(ids, table, pos_table, period, float(scale))
    out = F.embedding(ids, table) * scale if scale != 1.0 else F.embedding(ids, table)
    if pos_table is not None:
        out = out + pos_table[:period]
    return out
