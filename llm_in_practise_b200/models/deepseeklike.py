"""DeepSeekLike: causal multi-head *latent* attention + RoPE + mixture-of-experts FFN, after
LLM_Distributed_Trainning/PyTorch/transformer_basics/DeepSeekLike_wikitext2.py (complex-number RoPE, dense
expert loop) and …/DeepSeekLike_spare_MoE_wikitext2.py (real cos/sin RoPE, gather/scatter experts).

Semantics kept from the reference:
* full q/k/v projections -> RoPE on q,k (pairs (x[2i], x[2i+1]); the complex and the real formulation are the same
  rotation) -> one shared ``Linear(head_dim, latent)`` each for q, k, v applied per head (latent defaults to
  head_dim // 4) -> scores in latent space scaled by 1/sqrt(latent) -> causal softmax -> ``Linear(latent, head_dim)``
  -> out_proj (…_spare_MoE…:180-233).
* router ``Linear(d, E)``; **top-k on the raw logits, softmax over the k selected values**; ``num_shared`` always-on
  experts averaged; every expert is Linear(d,4d) -> GELU -> Linear(4d,d) (…:255-329).

B200 mapping (K16/K17): q|k|v is one fused tcgen05 GEMM; RoPE is one kernel over the fused projection; latent attention
runs on the flash kernels with the latent vectors zero-padded to 32 columns (no L x L score or mask tensor; fp32 /
non-native runs keep the batched-GEMM + masked-softmax path); tokens are sorted by expert on the device so each expert runs two dense GEMMs over a
contiguous row segment (``moe_impl="sparse"``), or every expert runs on all tokens and is masked
(``moe_impl="dense"``, the first script's behaviour — same function, E/k times the FLOPs).
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn.functional as F
from torch import nn

from ..ops import elementwise as ew
from ..ops import functional as Fn
from ..ops.flash import flash_attention
from ..ops.gemm import gemm
from ..ops.linear import FusedLinear
from ..ops.loss import cross_entropy, linear_cross_entropy
from ..ops.norm import layer_norm


class CausalMLA(nn.Module):
    def __init__(self, d_model: int, n_head: int, latent_dim: Optional[int] = None, resid_dropout: float = 0.0,
                 dtype=torch.float32, device=None):
        super().__init__()
        assert d_model % n_head == 0
        self.h, self.hd = n_head, d_model // n_head
        self.latent = latent_dim if latent_dim is not None else max(1, self.hd // 4)
        kw = dict(dtype=dtype, device=device, init="normal")
        self.qkv = FusedLinear(d_model, [("q_proj", d_model), ("k_proj", d_model), ("v_proj", d_model)], bias=True, **kw)
        self.q_compress = FusedLinear(self.hd, [("c", self.latent)], bias=True, **kw)
        self.k_compress = FusedLinear(self.hd, [("c", self.latent)], bias=True, **kw)
        self.v_compress = FusedLinear(self.hd, [("c", self.latent)], bias=True, **kw)
        self.decompress = FusedLinear(self.latent, [("d", self.hd)], bias=True, **kw)
        self.out_proj = FusedLinear(d_model, [("o_proj", d_model)], bias=True, **kw)
        self.p = resid_dropout

    FLASH_LATENT = 32      # smallest head size of the flash kernels (csrc/flash_attn.cu)

    def _flash_ok(self, x, lat) -> bool:
        from ..ops import _native as nv
        return (nv.use_native(x) and x.dtype in (torch.bfloat16, torch.float16) and lat <= self.FLASH_LATENT
                and self.q_compress.weight is not None and self.q_compress.weight.dtype == x.dtype)

    def forward(self, x, cos, sin, pos):
        B, L, D = x.shape
        H, hd, lat = self.h, self.hd, self.latent
        qkv = self.qkv(x.reshape(B * L, D))
        qkv = ew.qknorm_rope(qkv, None, None, cos, sin, pos, H, H, hd, interleaved=True)
        q, k, v = (qkv[:, i * D:(i + 1) * D].contiguous().view(B * L * H, hd) for i in range(3))
        if self._flash_ok(x, lat):
            # K17: latent attention on the flash kernels. The three per-head compressions write latent vectors zero-padded
            # to the kernel's smallest head size (32): zero q / k columns leave the scores unchanged, the extra v columns
            # come out zero and are dropped before `decompress`. No L x L score or mask tensor in either direction.
            pad = self.FLASH_LATENT - lat
            def comp(lin, t):
                return Fn.linear(t, F.pad(lin.weight, (0, 0, 0, pad)), F.pad(lin.bias, (0, pad))).view(B * L, H * self.FLASH_LATENT)
            lat_qkv = torch.cat([comp(self.q_compress, q), comp(self.k_compress, k), comp(self.v_compress, v)], dim=1)
            o = flash_attention(lat_qkv, B, L, H, H, self.FLASH_LATENT, True, 1.0 / math.sqrt(lat))
            o = o.view(B * L * H, self.FLASH_LATENT)[:, :lat]
            o = self.decompress(o.contiguous()).view(B * L, D)
            return ew.dropout(self.out_proj(o), self.p, self.training).view(B, L, D)
        # per-head compression into the latent space; pad the latent dim to a multiple of 8 for 16-byte rows
        qc = self.q_compress(q).view(B, L, H, lat).permute(0, 2, 1, 3)
        kc = self.k_compress(k).view(B, L, H, lat).permute(0, 2, 1, 3)
        vc = self.v_compress(v).view(B, L, H, lat).permute(0, 2, 1, 3)
        if lat % 8 or L % 8:
            # shapes the TMA path cannot address directly (rows not 16-byte multiples): plain batched matmul
            s = torch.matmul(qc.float(), kc.float().transpose(-1, -2)) / math.sqrt(lat)
            p = Fn.masked_softmax(s, True, 0, torch.float32)
            o = torch.matmul(p, vc.float()).to(x.dtype)
        else:
            s = Fn.matmul_nt(qc, kc, 1.0 / math.sqrt(lat), torch.float32)
            p = Fn.masked_softmax(s, True, 0, x.dtype)
            o = Fn.matmul_nn(p, vc)                                     # [B, H, L, lat]
        o = self.decompress(o.permute(0, 2, 1, 3).reshape(B * L * H, lat)).view(B * L, D)
        return ew.dropout(self.out_proj(o), self.p, self.training).view(B, L, D)


class Expert(nn.Module):
    def __init__(self, d_model, hidden, dropout=0.0, dtype=torch.float32, device=None):
        super().__init__()
        self.fc1 = FusedLinear(d_model, [("fc1", hidden)], bias=True, dtype=dtype, device=device, init="normal")
        self.fc2 = FusedLinear(hidden, [("fc2", d_model)], bias=True, dtype=dtype, device=device, init="normal")
        self.p = dropout

    def forward(self, x):
        return ew.dropout(self.fc2(ew.gelu(self.fc1(x))), self.p, self.training)


class MoEFeedForward(nn.Module):
    def __init__(self, d_model, hidden, num_experts=8, top_k=2, num_shared=2, dropout=0.0, moe_impl="sparse",
                 dtype=torch.float32, device=None):
        super().__init__()
        assert 1 <= top_k <= num_experts
        self.E, self.k, self.ns, self.impl, self.p = num_experts, top_k, max(0, num_shared), moe_impl, dropout
        self.router = FusedLinear(d_model, [("router", num_experts)], bias=True, dtype=dtype, device=device, init="normal")
        self.experts = nn.ModuleList([Expert(d_model, hidden, dropout, dtype, device) for _ in range(num_experts)])
        self.shared_experts = nn.ModuleList([Expert(d_model, hidden, dropout, dtype, device) for _ in range(self.ns)])
        self.last_load = None  # tokens routed to each expert in the last forward (observability)

    def _grouped_ok(self, xf) -> bool:
        from ..ops import _native as nv
        ex = self.experts[0]
        return (nv.use_native(xf) and xf.dtype in (torch.bfloat16, torch.float16) and ex.fc1.weight is not None
                and ex.fc1.weight.dtype == xf.dtype and xf.shape[1] % 8 == 0 and ex.fc1.weight.shape[0] % 8 == 0
                and not getattr(ex.fc1, "has_lora", False) and ex.fc1.nf4 is None)

    def forward(self, x):
        B, L, D = x.shape
        xf = x.reshape(B * L, D)
        T = xf.shape[0]
        shared = sum(ex(xf) for ex in self.shared_experts) / max(1, self.ns) if self.ns else torch.zeros_like(xf)
        gates = self.router(xf).float()                                  # [T, E]
        top_v, top_i = torch.topk(gates, self.k, dim=-1)
        top_p = torch.softmax(top_v, dim=-1).to(xf.dtype)                # softmax over the k selected logits
        if self.impl == "dense":
            moe = torch.zeros_like(xf)
            for e, ex in enumerate(self.experts):
                w = (top_p * (top_i == e).to(top_p.dtype)).sum(-1, keepdim=True)   # gate weight or 0
                moe = moe + ex(xf) * w
        elif self.impl == "sparse" and self._grouped_ok(xf):
            # K16: tokens counting-sorted by expert on the device, every expert's rows padded to whole 128-row tiles, ONE
            # grouped tcgen05 GEMM per projection (ops/moe.py); no host read of the per-expert counts on the forward path
            from ..ops.moe import moe_ffn
            w1 = torch.stack([ex.fc1.weight for ex in self.experts])
            b1 = torch.stack([ex.fc1.bias for ex in self.experts])
            w2 = torch.stack([ex.fc2.weight for ex in self.experts])
            b2 = torch.stack([ex.fc2.bias for ex in self.experts])
            ex_p = self.experts[0].p                            # the reference's Expert ends in its own dropout
            moe, d = moe_ffn(xf, top_i, top_p, w1, b1, w2, b2, ew.gelu,
                             post=(lambda y: ew.dropout(y, ex_p, True)) if (ex_p and self.training) else None)
            self.last_load = d.counts
        else:
            # sort the (token, slot) assignments by expert: each expert sees one contiguous row segment
            flat_e = top_i.reshape(-1)
            order = torch.argsort(flat_e, stable=True)
            tok = order // self.k
            counts = torch.bincount(flat_e, minlength=self.E)
            self.last_load = counts
            xs = xf.index_select(0, tok)                                  # [T*k, D] permuted inputs
            ws = top_p.reshape(-1).index_select(0, order).unsqueeze(1)
            outs = []
            start = 0
            for e, n in enumerate(counts.tolist()):
                if n:
                    outs.append(self.experts[e](xs[start:start + n]))
                start += n
            ys = torch.cat(outs, 0) * ws
            moe = torch.zeros_like(xf).index_add_(0, tok, ys)
        return ew.dropout((shared + moe).view(B, L, D), self.p, self.training)


class DeepSeekBlock(nn.Module):
    def __init__(self, d_model, n_head, mlp_ratio=4.0, dropout=0.1, latent_dim=None, num_experts=8, top_k=2,
                 num_shared=2, moe_impl="sparse", dtype=torch.float32, device=None):
        super().__init__()
        self.ln1_w = nn.Parameter(torch.ones(d_model, dtype=dtype, device=device))
        self.ln1_b = nn.Parameter(torch.zeros(d_model, dtype=dtype, device=device))
        self.ln2_w = nn.Parameter(torch.ones(d_model, dtype=dtype, device=device))
        self.ln2_b = nn.Parameter(torch.zeros(d_model, dtype=dtype, device=device))
        self.attn = CausalMLA(d_model, n_head, latent_dim, dropout, dtype, device)
        self.mlp = MoEFeedForward(d_model, int(d_model * mlp_ratio), num_experts, top_k, num_shared, dropout, moe_impl,
                                  dtype, device)

    def forward(self, delta, residual, cos, sin, pos):
        x, residual = layer_norm(delta, self.ln1_w, self.ln1_b, 1e-5, residual)
        a = self.attn(x, cos, sin, pos)
        x, residual = layer_norm(a, self.ln2_w, self.ln2_b, 1e-5, residual)
        return self.mlp(x), residual


class DeepSeekLike(nn.Module):
    def __init__(self, vocab_size=30000, block_size=256, n_layer=6, n_head=8, d_model=768, dropout=0.1,
                 latent_dim=None, num_experts=8, top_k=2, num_shared=2, rope_theta=10000.0, moe_impl="sparse",
                 dtype=torch.float32, device=None):
        super().__init__()
        assert d_model % n_head == 0
        self.block_size, self.d_model, self.n_head, self.p = block_size, d_model, n_head, dropout
        self.rope_theta = rope_theta
        self.tok_emb = nn.Parameter(torch.empty(vocab_size, d_model, dtype=dtype, device=device))
        nn.init.normal_(self.tok_emb, std=0.02)
        self.blocks = nn.ModuleList([DeepSeekBlock(d_model, n_head, 4.0, dropout, latent_dim, num_experts, top_k,
                                                   num_shared, moe_impl, dtype, device) for _ in range(n_layer)])
        self.ln_f_w = nn.Parameter(torch.ones(d_model, dtype=dtype, device=device))
        self.ln_f_b = nn.Parameter(torch.zeros(d_model, dtype=dtype, device=device))
        self._rope = None

    def forward(self, idx, targets=None, n_valid=None):
        B, L = idx.shape
        if L > self.block_size:
            idx, L = idx[:, : self.block_size], self.block_size
        dev = idx.device
        if self._rope is None or self._rope[0].device != dev:
            self._rope = ew.rope_tables(self.block_size, self.d_model // self.n_head, self.rope_theta, dev)
        cos, sin = self._rope
        pos = torch.arange(L, device=dev, dtype=torch.int32).repeat(B)
        delta = ew.dropout(ew.embedding(idx, self.tok_emb), self.p, self.training)
        residual = None
        for blk in self.blocks:
            delta, residual = blk(delta, residual, cos, sin, pos)
        x, _ = layer_norm(delta, self.ln_f_w, self.ln_f_b, 1e-5, residual)
        x2 = x.reshape(B * L, -1)
        if targets is None:
            return gemm(x2, self.tok_emb).view(B, L, -1)
        t = targets[:, :L].reshape(-1)
        if x2.dtype == torch.float32 and x2.is_cuda:
            return cross_entropy(Fn.linear(x2, self.tok_emb), t, -100, n_valid)
        return linear_cross_entropy(x2, self.tok_emb, t, -100, n_valid)
