"""Generic transformer building blocks on the sm_100a kernels.

``TransformerBlock`` covers every hand-written block shape of the reference with a few switches:

* pre-LN decoder block of GPTLike (LayerNorm -> MHA(causal) -> Dropout -> +res; LayerNorm -> Linear(d,4d) ->
  GELU -> Linear(4d,d) -> Dropout -> +res) — LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py:86-125;
* post-LN ``nn.TransformerEncoderLayer(activation="gelu", batch_first=True)`` of minigpt2 and SimpleTransformer
  (llm-demo/minigpt2/model.py:44-54, …/transformer_basics/GPTLike_wikitext2.py:95-138), with or without a mask.

Differences from the torch modules it replaces: q|k|v is one fused projection (one tcgen05 GEMM, bias in the
epilogue); attention reads q, k, v in place (no head transposes, no L x L mask tensor); the residual add is fused
into the following LayerNorm; dropout masks are Philox streams regenerated in backward, never stored.
``load_torch_mha_state`` / ``export_torch_mha_state`` convert to and from the ``nn.MultiheadAttention`` layout
(``in_proj_weight``, ``in_proj_bias``, ``out_proj.*``) so reference ``.pth`` checkpoints load.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
from torch import nn

from ..ops import elementwise as ew
from ..ops.attention import attention
from ..ops.linear import FusedLinear
from ..ops.norm import layer_norm


def sinusoidal_table(n_pos: int, dim: int, device=None, dtype=torch.float32) -> torch.Tensor:
    pe = torch.zeros(n_pos, dim, device=device)
    position = torch.arange(0, n_pos, dtype=torch.float32, device=device).unsqueeze(1)
    div = torch.exp(torch.arange(0, dim, 2, device=device).float() * (-math.log(10000.0) / dim))
    pe[:, 0::2] = torch.sin(position * div)
    pe[:, 1::2] = torch.cos(position * div)[:, : dim // 2]
    return pe.to(dtype)


class MultiHeadSelfAttention(nn.Module):
    def __init__(self, d_model: int, n_head: int, dropout: float = 0.0, causal: bool = True, bias: bool = True,
                 n_kv_head: Optional[int] = None, window: int = 0, dtype=torch.float32, device=None, init: str = "xavier"):
        super().__init__()
        assert d_model % n_head == 0
        self.n_head, self.n_kv = n_head, n_kv_head or n_head
        self.head_dim = d_model // n_head
        self.causal, self.window, self.p = causal, window, dropout
        D = self.head_dim
        self.qkv = FusedLinear(d_model, [("q_proj", n_head * D), ("k_proj", self.n_kv * D), ("v_proj", self.n_kv * D)],
                               bias=bias, dtype=dtype, device=device, init=init)
        self.out_proj = FusedLinear(n_head * D, [("o_proj", d_model)], bias=bias, dtype=dtype, device=device, init=init)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        B, S, _ = x.shape
        qkv = self.qkv(x.reshape(B * S, -1))
        Sp = S
        if S % 8 and qkv.is_cuda:
            # the tiled kernels want 16-byte aligned score rows: pad the sequence, masked out by causality
            Sp = (S + 7) // 8 * 8
            qkv = torch.nn.functional.pad(qkv.view(B, S, -1), (0, 0, 0, Sp - S)).reshape(B * Sp, -1)
        ctx = attention(qkv, B, Sp, self.n_head, self.n_kv, self.head_dim, causal=self.causal, window=self.window,
                        dropout_p=self.p if self.training else 0.0)
        if Sp != S:
            ctx = ctx.view(B, Sp, -1)[:, :S].reshape(B * S, -1)
        return self.out_proj(ctx).view(B, S, -1)

    # ---- nn.MultiheadAttention checkpoint layout -------------------------------------------------
    def load_torch_mha_state(self, sd: dict, prefix: str) -> None:
        with torch.no_grad():
            self.qkv.weight.copy_(sd[prefix + "in_proj_weight"])
            self.qkv.bias.copy_(sd[prefix + "in_proj_bias"])
            self.out_proj.weight.copy_(sd[prefix + "out_proj.weight"])
            self.out_proj.bias.copy_(sd[prefix + "out_proj.bias"])

    def export_torch_mha_state(self, prefix: str) -> dict:
        return {prefix + "in_proj_weight": self.qkv.weight.detach().clone(),
                prefix + "in_proj_bias": self.qkv.bias.detach().clone(),
                prefix + "out_proj.weight": self.out_proj.weight.detach().clone(),
                prefix + "out_proj.bias": self.out_proj.bias.detach().clone()}


class FeedForward(nn.Module):
    def __init__(self, d_model: int, hidden: int, dropout: float = 0.0, dtype=torch.float32, device=None, init="xavier"):
        super().__init__()
        self.fc1 = FusedLinear(d_model, [("fc1", hidden)], bias=True, dtype=dtype, device=device, init=init)
        self.fc2 = FusedLinear(hidden, [("fc2", d_model)], bias=True, dtype=dtype, device=device, init=init)
        self.p = dropout

    def forward(self, x):
        h = ew.gelu(self.fc1(x))
        return ew.dropout(self.fc2(h), self.p, self.training)


class TransformerBlock(nn.Module):
    def __init__(self, d_model: int, n_head: int, dropout: float = 0.1, causal: bool = True, norm_first: bool = True,
                 ff_mult: float = 4.0, eps: float = 1e-5, dtype=torch.float32, device=None, init: str = "xavier",
                 attn_dropout: Optional[float] = None, window: int = 0):
        super().__init__()
        self.norm_first, self.eps, self.p = norm_first, eps, dropout
        self.attn = MultiHeadSelfAttention(d_model, n_head, dropout if attn_dropout is None else attn_dropout, causal,
                                           dtype=dtype, device=device, init=init, window=window)
        self.mlp = FeedForward(d_model, int(d_model * ff_mult), dropout, dtype=dtype, device=device, init=init)
        self.ln1_w = nn.Parameter(torch.ones(d_model, dtype=dtype, device=device))
        self.ln1_b = nn.Parameter(torch.zeros(d_model, dtype=dtype, device=device))
        self.ln2_w = nn.Parameter(torch.ones(d_model, dtype=dtype, device=device))
        self.ln2_b = nn.Parameter(torch.zeros(d_model, dtype=dtype, device=device))

    def forward(self, delta: torch.Tensor, residual: Optional[torch.Tensor]):
        """Streams (delta, residual): the hidden state is ``delta + residual``; the add happens inside the
        next LayerNorm kernel. Returns the new pair."""
        if self.norm_first:
            x, residual = layer_norm(delta, self.ln1_w, self.ln1_b, self.eps, residual)
            a = ew.dropout(self.attn(x), self.p, self.training)
            x, residual = layer_norm(a, self.ln2_w, self.ln2_b, self.eps, residual)
            return self.mlp(x), residual
        # post-LN (nn.TransformerEncoderLayer default): x = LN(x + SA(x)); x = LN(x + FF(x))
        x = delta if residual is None else delta + residual
        a = ew.dropout(self.attn(x), self.p, self.training)
        x, _ = layer_norm(a, self.ln1_w, self.ln1_b, self.eps, x)
        f = self.mlp(x)
        x, _ = layer_norm(f, self.ln2_w, self.ln2_b, self.eps, x)
        return x, None
