"""Attention and block variants of ``Transformer/Transformer_Advanced.ipynb`` on this framework's kernels.

* ``GroupedQueryAttention`` (cell 6) / ``MultiQueryAttention`` (cell 8): fewer K/V heads than Q heads; the attention
  kernels read the shared K/V head through a batch-index divisor, nothing is repeated in memory.
* ``MultiHeadLatentAttention`` (cell 10): K and V are up-projected from one shared latent ``C_kv`` (queries from their own
  latent ``C_q``), rotary embedding is applied to the last ``d_rope`` channels of the latents only ("decoupled RoPE",
  first/second-half pairing) and the rotated ``C_kv`` is returned as the KV cache.
* ``LocalAttention`` (cell 12): every token attends to the tokens within ±``window_size`` (the notebook loops over
  positions in Python; here it is one banded-mask softmax).
* ``ResiDualBlock`` (cell 17), ``ParallelBlock`` (cell 19), ``StochasticDepth`` + ``StochasticDepthBlock`` (cell 21).
* ``DenseMoEDecoder`` (cell 24): decoder-only stack whose FFN is a mixture of experts evaluated densely under a top-k mask.

All linears are ``FusedLinear`` (tcgen05 GEMMs on the GPU, plain matmul on CPU); every module is differentiable.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch
from torch import nn

from ..ops import elementwise as ew
from ..ops.functional import masked_softmax, matmul_nn, matmul_nt
from ..ops.linear import FusedLinear
from ..ops.norm import LayerNorm as _FusedLayerNorm
from .transformer import MultiHeadSelfAttention


class LayerNorm(_FusedLayerNorm):
    """The fused norm returns (normalised, residual stream); these blocks only want the first."""

    def forward(self, x):  # noqa: D102
        shape = x.shape
        return super().forward(x.reshape(-1, shape[-1]))[0].view(shape)


class GroupedQueryAttention(MultiHeadSelfAttention):
    def __init__(self, d_model: int, num_heads: int, num_kv_heads: int, causal: bool = False, dtype=torch.float32, device=None):
        assert num_heads % num_kv_heads == 0, "num_heads must be a multiple of num_kv_heads"
        super().__init__(d_model, num_heads, 0.0, causal=causal, n_kv_head=num_kv_heads, dtype=dtype, device=device)


class MultiQueryAttention(GroupedQueryAttention):
    def __init__(self, d_model: int, num_heads: int, causal: bool = False, dtype=torch.float32, device=None):
        super().__init__(d_model, num_heads, 1, causal, dtype, device)


def _lin(i, o, dtype, device, bias=True):
    return FusedLinear(i, [("w", o)], bias=bias, dtype=dtype, device=device, init="xavier")


def _heads(x, B, S, H):
    return x.view(B, S, H, -1).transpose(1, 2)            # [B, H, S, D] (a view; the GEMM takes the strides)


def _sdpa(q, k, v, causal: bool, window: int = 0, band: int = 0):
    """softmax(q k^T / sqrt(D)) v for [B, H, S, D] operands; ``band`` > 0 keeps |i - j| <= band."""
    D = q.shape[-1]
    s = matmul_nt(q, k, 1.0 / math.sqrt(D), out_dtype=torch.float32)
    if band > 0:
        S = s.shape[-1]
        i = torch.arange(S, device=s.device)
        s = s.masked_fill((i[:, None] - i[None, :]).abs() > band, float("-inf"))
        p = torch.softmax(s, dim=-1).to(q.dtype)
    else:
        p = masked_softmax(s, causal=causal, window=window, out_dtype=q.dtype)
    return matmul_nn(p, v)


class MultiHeadLatentAttention(nn.Module):
    def __init__(self, d_model: int, num_heads: int, d_latent: int, d_rope: int = 64, causal: bool = False,
                 dtype=torch.float32, device=None):
        super().__init__()
        assert d_model % num_heads == 0 and d_latent >= d_rope and d_rope % 2 == 0
        self.h, self.d_latent, self.d_rope, self.causal = num_heads, d_latent, d_rope, causal
        self.W_dkv, self.W_dq = _lin(d_model, d_latent, dtype, device), _lin(d_model, d_latent, dtype, device)
        self.W_uq, self.W_uk, self.W_uv = (_lin(d_latent, d_model, dtype, device) for _ in range(3))
        self.W_o = _lin(d_model, d_model, dtype, device)
        half = d_rope // 2
        self.register_buffer("inv_freq", 1.0 / (10000 ** (torch.arange(half, dtype=torch.float32, device=device) / half)), persistent=False)

    def apply_decoupled_rope(self, x: torch.Tensor, positions: torch.Tensor) -> torch.Tensor:
        nope, rope = x[..., : self.d_latent - self.d_rope], x[..., self.d_latent - self.d_rope:]
        ang = positions[:, None].float() * self.inv_freq[None, :]
        cos, sin = ang.cos().to(x.dtype), ang.sin().to(x.dtype)
        x1, x2 = rope[..., : self.d_rope // 2], rope[..., self.d_rope // 2:]
        return torch.cat([nope, x1 * cos - x2 * sin, x1 * sin + x2 * cos], dim=-1)

    def forward(self, x: torch.Tensor, past_c_kv: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        B, S, d = x.shape
        start = 0 if past_c_kv is None else past_c_kv.shape[1]
        pos = torch.arange(start, start + S, device=x.device)
        c_kv = self.apply_decoupled_rope(self.W_dkv(x), pos)
        c_q = self.apply_decoupled_rope(self.W_dq(x), pos)
        if past_c_kv is not None:
            c_kv = torch.cat([past_c_kv, c_kv], dim=1)         # the latent IS the cache: d_latent numbers per token
        T = c_kv.shape[1]
        q = _heads(self.W_uq(c_q), B, S, self.h)
        k, v = _heads(self.W_uk(c_kv), B, T, self.h), _heads(self.W_uv(c_kv), B, T, self.h)
        if past_c_kv is None:
            ctx = _sdpa(q, k, v, self.causal)
        else:                                                   # decode: new queries see every cached position
            s = matmul_nt(q, k, 1.0 / math.sqrt(q.shape[-1]), out_dtype=torch.float32)
            ctx = matmul_nn(torch.softmax(s, -1).to(q.dtype), v)
        return self.W_o(ctx.transpose(1, 2).reshape(B, S, d)), c_kv


class LocalAttention(nn.Module):
    def __init__(self, d_model: int, num_heads: int, window_size: int, dtype=torch.float32, device=None):
        super().__init__()
        assert d_model % num_heads == 0
        self.h, self.w = num_heads, window_size
        self.W_q, self.W_k, self.W_v, self.W_o = (_lin(d_model, d_model, dtype, device) for _ in range(4))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        B, S, d = x.shape
        q, k, v = (_heads(f(x), B, S, self.h) for f in (self.W_q, self.W_k, self.W_v))
        return self.W_o(_sdpa(q, k, v, False, band=self.w).transpose(1, 2).reshape(B, S, d))


class _FFN(nn.Module):
    def __init__(self, d_model, d_ff, dropout, act="gelu", dtype=torch.float32, device=None):
        super().__init__()
        self.fc1, self.fc2, self.p, self.act = _lin(d_model, d_ff, dtype, device), _lin(d_ff, d_model, dtype, device), dropout, act

    def forward(self, x):
        h = self.fc1(x)
        h = ew.gelu(h) if self.act == "gelu" else torch.relu(h)
        return ew.dropout(self.fc2(h), self.p, self.training)


class ResiDualBlock(nn.Module):
    """y = x + LN3(y2), y2 = y1 + FFN(LN2(y1)), y1 = LN1(x) + Attn(LN1(x))."""

    def __init__(self, d_model: int, n_heads: int, d_ff: int, dropout: float = 0.1, causal: bool = True, dtype=torch.float32, device=None):
        super().__init__()
        self.attn = MultiHeadSelfAttention(d_model, n_heads, dropout, causal=causal, dtype=dtype, device=device)
        self.ffn = _FFN(d_model, d_ff, dropout, dtype=dtype, device=device)
        self.ln1, self.ln2, self.ln3 = (LayerNorm(d_model, dtype=dtype, device=device) for _ in range(3))
        self.p = dropout

    def forward(self, x):
        x_ln = self.ln1(x)
        y1 = x_ln + ew.dropout(self.attn(x_ln), self.p, self.training)
        y2 = y1 + self.ffn(self.ln2(y1))
        return x + self.ln3(y2)


class ParallelBlock(nn.Module):
    """y = x + drop(LN(x) + Attn(LN(x)) + FFN(LN(x))): attention and FFN read the same normalised input."""

    def __init__(self, d_model: int, n_heads: int, d_ff: int, dropout: float = 0.1, causal: bool = True, dtype=torch.float32, device=None):
        super().__init__()
        self.ln = LayerNorm(d_model, dtype=dtype, device=device)
        self.attn = MultiHeadSelfAttention(d_model, n_heads, dropout, causal=causal, dtype=dtype, device=device)
        self.ffn = _FFN(d_model, d_ff, dropout, dtype=dtype, device=device)
        self.p = dropout

    def forward(self, x):
        x_ln = self.ln(x)
        out = x_ln + ew.dropout(self.attn(x_ln), self.p, self.training) + ew.dropout(self.ffn(x_ln), self.p, self.training)
        return x + ew.dropout(out, self.p, self.training)


class StochasticDepth(nn.Module):
    """x + residual * Bernoulli(keep) / keep, one draw per sample; identity-plus-residual at eval time."""

    def __init__(self, drop_prob: float):
        super().__init__()
        self.drop_prob = drop_prob

    def forward(self, x, residual):
        if not self.training or self.drop_prob == 0.0:
            return x + residual
        keep = 1.0 - self.drop_prob
        mask = torch.empty(x.shape[0], *([1] * (x.dim() - 1)), device=x.device, dtype=x.dtype).bernoulli_(keep)
        return x + residual * (mask / keep)


class StochasticDepthBlock(nn.Module):
    def __init__(self, d_model: int, nhead: int, dim_ff: int, drop_prob: float = 0.1, causal: bool = False, dtype=torch.float32, device=None):
        super().__init__()
        self.ln1, self.ln2 = LayerNorm(d_model, dtype=dtype, device=device), LayerNorm(d_model, dtype=dtype, device=device)
        self.attn = MultiHeadSelfAttention(d_model, nhead, 0.0, causal=causal, dtype=dtype, device=device)
        self.ffn = _FFN(d_model, dim_ff, 0.0, act="relu", dtype=dtype, device=device)
        self.sd1, self.sd2 = StochasticDepth(drop_prob), StochasticDepth(drop_prob)

    def forward(self, x):
        x = self.sd1(x, self.attn(self.ln1(x)))
        return self.sd2(x, self.ffn(self.ln2(x)))


class DenseMoEFeedForward(nn.Module):
    """Top-k gate, every expert evaluated on every token, outputs combined under the (renormalised) top-k mask."""

    def __init__(self, d_model: int, d_ff: int, num_experts: int = 4, top_k: int = 2, dtype=torch.float32, device=None):
        super().__init__()
        self.k = top_k
        self.gate = _lin(d_model, num_experts, dtype, device, bias=False)
        self.experts = nn.ModuleList([_FFN(d_model, d_ff, 0.0, dtype=dtype, device=device) for _ in range(num_experts)])

    def forward(self, x):
        probs = torch.softmax(self.gate(x).float(), dim=-1)
        topv, topi = probs.topk(self.k, dim=-1)
        w = torch.zeros_like(probs).scatter(-1, topi, topv / topv.sum(-1, keepdim=True)).to(x.dtype)
        return sum(w[..., e:e + 1] * expert(x) for e, expert in enumerate(self.experts))


class DenseMoEDecoder(nn.Module):
    def __init__(self, vocab_size: int, d_model: int = 128, n_heads: int = 4, d_ff: int = 256, n_layers: int = 2, num_experts: int = 4,
                 top_k: int = 2, max_len: int = 512, dtype=torch.float32, device=None):
        super().__init__()
        self.tok = nn.Parameter(torch.randn(vocab_size, d_model, dtype=dtype, device=device) * 0.02)
        self.pos = nn.Parameter(torch.randn(max_len, d_model, dtype=dtype, device=device) * 0.02)
        self.layers = nn.ModuleList()
        for _ in range(n_layers):
            self.layers.append(nn.ModuleDict({
                "ln1": LayerNorm(d_model, dtype=dtype, device=device), "ln2": LayerNorm(d_model, dtype=dtype, device=device),
                "attn": MultiHeadSelfAttention(d_model, n_heads, 0.0, causal=True, dtype=dtype, device=device),
                "moe": DenseMoEFeedForward(d_model, d_ff, num_experts, top_k, dtype, device)}))
        self.ln_f = LayerNorm(d_model, dtype=dtype, device=device)
        self.head = _lin(d_model, vocab_size, dtype, device, bias=False)

    def forward(self, idx: torch.Tensor, targets: Optional[torch.Tensor] = None):
        B, S = idx.shape
        x = ew.embedding(idx, self.tok, self.pos).view(B, S, -1)
        for l in self.layers:
            x = x + l["attn"](l["ln1"](x))
            x = x + l["moe"](l["ln2"](x))
        logits = self.head(self.ln_f(x))
        if targets is None:
            return logits
        from ..ops.loss import cross_entropy
        return logits, cross_entropy(logits.view(B * S, -1), targets.reshape(-1))
