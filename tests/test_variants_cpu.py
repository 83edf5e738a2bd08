"""Transformer_Advanced notebook variants: shapes, semantics against plain PyTorch formulas, gradients."""
import math

import torch

from llm_in_practise_b200.models.attention_variants import (DenseMoEDecoder, GroupedQueryAttention, LocalAttention,
                                                            MultiHeadLatentAttention, MultiQueryAttention, ParallelBlock,
                                                            ResiDualBlock, StochasticDepth, StochasticDepthBlock)


def test_gqa_mqa_shapes_and_param_savings():
    x = torch.rand(2, 16, 64)
    g, m = GroupedQueryAttention(64, 8, 2), MultiQueryAttention(64, 8)
    assert g(x).shape == x.shape and m(x).shape == x.shape
    assert g.qkv.out_features == 64 + 2 * 2 * 8 and m.qkv.out_features == 64 + 2 * 8


def test_mla_cache_and_incremental_decode():
    torch.manual_seed(0)
    mla = MultiHeadLatentAttention(64, 4, 32, 16, causal=True)
    x = torch.rand(2, 10, 64)
    out, c = mla(x)
    assert out.shape == (2, 10, 64) and c.shape == (2, 10, 32)
    # decoding token 9 from the cached latents of tokens 0..8 gives the same output as the full causal pass
    _, c9 = mla(x[:, :9])
    o_last, c10 = mla(x[:, 9:], past_c_kv=c9)
    assert torch.allclose(c10, c, atol=1e-5) and torch.allclose(o_last[:, 0], out[:, 9], atol=1e-4)
    # the no-position part of the latent is untouched by RoPE
    raw = mla.W_dkv(x)
    assert torch.allclose(c[..., :16], raw[..., :16], atol=1e-6)


def test_local_attention_matches_position_loop():
    torch.manual_seed(1)
    la = LocalAttention(32, 4, 3)
    x = torch.rand(1, 12, 32)
    out = la(x)
    q, k, v = (f(x).view(1, 12, 4, 8).transpose(1, 2) for f in (la.W_q, la.W_k, la.W_v))
    ctx = torch.zeros_like(q)
    for i in range(12):
        lo, hi = max(0, i - 3), min(12, i + 4)
        a = torch.softmax(q[:, :, i:i + 1] @ k[:, :, lo:hi].transpose(-1, -2) / math.sqrt(8), -1)
        ctx[:, :, i:i + 1] = a @ v[:, :, lo:hi]
    ref = la.W_o(ctx.transpose(1, 2).reshape(1, 12, 32))
    assert torch.allclose(out, ref, atol=1e-5)


def test_blocks_and_stochastic_depth():
    torch.manual_seed(2)
    x = torch.randn(4, 6, 32, requires_grad=True)
    for blk in (ResiDualBlock(32, 4, 64, 0.0), ParallelBlock(32, 4, 64, 0.0), StochasticDepthBlock(32, 4, 16, 0.5)):
        y = blk(x)
        assert y.shape == x.shape
        y.sum().backward()
        assert all(p.grad is not None for p in blk.parameters())
    sd = StochasticDepth(0.5)
    sd.train()
    r = torch.ones(1000, 1, 1)
    y = sd(torch.zeros(1000, 1, 1), r)
    assert set(y.unique().tolist()) <= {0.0, 2.0} and 0.35 < (y > 0).float().mean() < 0.65
    sd.eval()
    assert torch.equal(sd(torch.zeros(3, 1, 1), torch.ones(3, 1, 1)), torch.ones(3, 1, 1))


def test_dense_moe_decoder_trains():
    torch.manual_seed(3)
    m = DenseMoEDecoder(50, 32, 4, 64, 2, 4, 2, 16)
    idx = torch.randint(0, 50, (4, 16))
    opt = torch.optim.Adam(m.parameters(), 1e-2)
    first = None
    for _ in range(15):
        opt.zero_grad()
        _, loss = m(idx, idx.roll(-1, 1))
        loss.backward()
        opt.step()
        first = first or loss.item()
    assert loss.item() < first
