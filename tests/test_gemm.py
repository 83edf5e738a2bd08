"""K1/K3 GEMM: tcgen05 kernel vs a plain fp32 PyTorch reference of the same product."""
import pytest
import torch

from llm_in_practise_b200.ops import gemm as G


def _ref(a, b, a_mn=False, b_mn=False):
    A = a.transpose(-1, -2) if a_mn else a
    B = b.transpose(-1, -2) if b_mn else b
    return A.float() @ B.float().transpose(-1, -2)


def test_cpu_fallback_matches_reference():
    x = torch.randn(37, 48)
    w = torch.randn(29, 48)
    y = G.gemm(x, w)
    assert torch.allclose(y, x @ w.t(), atol=1e-5)
    dy = torch.randn(37, 29)
    assert torch.allclose(G.linear_dgrad(dy, w), dy @ w, atol=1e-5)
    assert torch.allclose(G.linear_wgrad(dy, x), dy.t() @ x, atol=1e-4)


def _check(out, ref, tol):
    err = (out.float() - ref).abs().max().item()
    scale = ref.abs().max().item() + 1e-6
    assert err / scale < tol, f"rel err {err / scale:.3e} (abs {err:.3e})"


def _check_elem(out, ref):
    """Per-element bound for a 16-bit-input, fp32-accumulate product: |a - b| <= 4e-3 |b| + 1e-3 max|b| (the half-ulp of a
    bf16 output is 3.9e-3 |b|; a dropped k-block of a K = 17 408 product moves single elements by far more than that,
    while it would slip under a 2e-2 max-norm bound)."""
    d = (out.float() - ref).abs()
    bound = 4e-3 * ref.abs() + 1e-3 * ref.abs().max()
    bad = d > bound
    assert not bool(bad.any()), f"{int(bad.sum())} elements out of bound, worst excess {(d - bound).max().item():.3e}"


@pytest.mark.gpu
@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (256, 512, 128), (1024, 5120, 5120), (1000, 1000, 520),
                                   (77, 40, 72), (4096, 30528, 768), (128, 32, 4096)])
def test_gemm_bf16_kmajor(M, N, K):
    x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    w = torch.randn(N, K, device="cuda", dtype=torch.bfloat16) / K ** 0.5
    y = G.gemm(x, w)
    _check_elem(y, _ref(x, w))


@pytest.mark.gpu
def test_gemm_bf16_long_k_per_element():
    """K = 17 408 (down_proj of Qwen3-14B): every element within the bf16 rounding bound, stream-K and 2-CTA tiles included."""
    for M, N, K in ((1024, 5120, 17408), (300, 5120, 17408)):
        x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
        w = torch.randn(N, K, device="cuda", dtype=torch.bfloat16) / K ** 0.5
        _check_elem(G.gemm(x, w), _ref(x, w))


@pytest.mark.gpu
@pytest.mark.parametrize("M,N,K", [(256, 512, 384), (1024, 5120, 1024), (515, 200, 136)])
def test_gemm_bf16_dgrad_wgrad(M, N, K):
    x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    w = torch.randn(N, K, device="cuda", dtype=torch.bfloat16) / K ** 0.5
    dy = torch.randn(M, N, device="cuda", dtype=torch.bfloat16)
    dx = G.linear_dgrad(dy, w)
    _check_elem(dx, dy.float() @ w.float())
    dw = G.linear_wgrad(dy, x, out_dtype=torch.float32)
    _check_elem(dw, dy.float().t() @ x.float())
    # accumulate into an existing gradient (beta = 1)
    dw2 = dw.clone()
    G.linear_wgrad(dy, x, out=dw2, beta=1.0)
    _check(dw2, 2 * (dy.float().t() @ x.float()), 2e-2)


@pytest.mark.gpu
def test_gemm_fp32_tf32():
    M, N, K = 512, 768, 264
    x = torch.randn(M, K, device="cuda")
    w = torch.randn(N, K, device="cuda") / K ** 0.5
    dy = torch.randn(M, N, device="cuda")
    _check(G.gemm(x, w), x @ w.t(), 5e-3)
    _check(G.linear_dgrad(dy, w), dy @ w, 5e-3)
    _check(G.linear_wgrad(dy, x), dy.t() @ x, 5e-3)


@pytest.mark.gpu
def test_gemm_bias_alpha_fp16():
    M, N, K = 300, 260, 192
    x = torch.randn(M, K, device="cuda", dtype=torch.float16)
    w = torch.randn(N, K, device="cuda", dtype=torch.float16) / K ** 0.5
    bias = torch.randn(N, device="cuda", dtype=torch.float16)
    y = G.gemm(x, w, bias=bias, alpha=0.5)
    _check(y, 0.5 * (x.float() @ w.float().t()) + bias.float(), 1e-2)


@pytest.mark.gpu
@pytest.mark.parametrize("r", [8, 16, 64])
def test_gemm_lora_extra_kblock(r):
    M, N, K = 1024, 1536, 1024
    x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    w = torch.randn(N, K, device="cuda", dtype=torch.bfloat16) / K ** 0.5
    t = torch.randn(M, r, device="cuda", dtype=torch.bfloat16)
    lb = torch.randn(N, r, device="cuda", dtype=torch.bfloat16)
    y = G.gemm(x, w, lora=(t, lb))
    _check(y, x.float() @ w.float().t() + t.float() @ lb.float().t(), 2e-2)


@pytest.mark.gpu
def test_gemm_batched_gqa_strided():
    # attention-shaped: Q [B,S,H,D] view, K [B,S,Hkv,D] view, scores[b,h] = Q K^T; then P V (b_mn)
    B, S, H, Hkv, D = 2, 256, 8, 2, 128
    qkv = torch.randn(B, S, (H + 2 * Hkv) * D, device="cuda", dtype=torch.bfloat16)
    q = qkv[..., : H * D].view(B, S, H, D).permute(0, 2, 1, 3)
    k = qkv[..., H * D:(H + Hkv) * D].view(B, S, Hkv, D).permute(0, 2, 1, 3)
    v = qkv[..., (H + Hkv) * D:].view(B, S, Hkv, D).permute(0, 2, 1, 3)
    s = G.gemm(q, k, alpha=D ** -0.5, out_dtype=torch.float32)
    kr = k.repeat_interleave(H // Hkv, dim=1)
    vr = v.repeat_interleave(H // Hkv, dim=1)
    _check(s, (q.float() @ kr.float().transpose(-1, -2)) * D ** -0.5, 2e-2)
    p = torch.softmax(s, dim=-1).to(torch.bfloat16)
    o = G.gemm(p, v, b_mn=True)
    _check(o, p.float() @ vr.float(), 2e-2)


# ------------------------------------------------------------------------------------------------
# 4-bit B operand expanded inside the GEMM
# ------------------------------------------------------------------------------------------------
class _FakeSharded:
    """Two row shards in separate buffers on one GPU: same descriptor layout as ZeRO-3's ShardedNF4."""

    def __init__(self, q, world=2):
        self.N, self.K, self.group = q.N, q.K, q.group
        self.rows = q.N // world
        sc = q.scales()
        self._bufs = []
        self.shards = []
        for r in range(world):
            p = q.packed.view(q.N, q.K // 2)[r * self.rows:(r + 1) * self.rows].contiguous().clone()
            s = sc.view(q.N, -1)[r * self.rows:(r + 1) * self.rows].contiguous().clone()
            self._bufs.append((p, s))
            self.shards.append((p.data_ptr(), s.data_ptr(), 0))
        self.packed = self._bufs[0][0]


@pytest.mark.gpu
@pytest.mark.parametrize("M,N,K,b_mn", [(1024, 1024, 512, False), (300, 256, 64, False), (1024, 5120, 17408, False),
                                        (512, 7168, 5120, False), (1024, 512, 1024, True), (1024, 5120, 7168, True),
                                        (260, 256, 128, True)])
def test_gemm_w4_nf4_fused(M, N, K, b_mn):
    from llm_in_practise_b200.ops.w4 import NF4Weight, _w4_dequant_ref
    torch.manual_seed(0)
    rows, kw = (K, N) if b_mn else (N, K)            # stored weight W [rows, kw]
    w = torch.randn(rows, kw, device="cuda", dtype=torch.bfloat16) / kw ** 0.5
    q = NF4Weight.quantize(w, double_quant=True)
    # oracle: the plain torch expansion of the codes (not the sibling dequant kernel), rounded to the GEMM's input dtype
    wd = _w4_dequant_ref(q.packed, q.scales(), None, rows, kw, q.group, None, torch.float32).bfloat16().float()
    x = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    ref = x.float() @ (wd if b_mn else wd.t())
    y = G.gemm_w4(x, q, b_mn=b_mn, force=True)
    assert y is not None, "fused kernel refused a supported shape"
    _check(y, ref, 2e-2)
    # low-rank term + bias ride along as in the dense kernel
    a2 = torch.randn(M, 16, device="cuda", dtype=torch.bfloat16)
    b2 = torch.randn(N, 16, device="cuda", dtype=torch.bfloat16) / 4
    bias = torch.randn(N, device="cuda", dtype=torch.bfloat16)
    y2 = G.gemm_w4(x, q, b_mn=b_mn, lora=(a2, b2), bias=bias, force=True)
    _check(y2, ref + a2.float() @ b2.float().t() + bias.float(), 2e-2)
    # row shards in separate buffers (the ZeRO-3 layout): identical result
    if rows % 256 == 0:
        y3 = G.gemm_w4(x, _FakeSharded(q), b_mn=b_mn, force=True)
        assert y3 is not None and torch.equal(y3, y)


@pytest.mark.gpu
@pytest.mark.parametrize("sym", [True, False])
def test_gemm_w4_int4_fp16(sym):
    from llm_in_practise_b200.quant.formats import quantize_rtn
    torch.manual_seed(1)
    w = torch.randn(1024, 2048, device="cuda") / 45
    q = quantize_rtn(w, 128, sym)
    x = torch.randn(512, 2048, device="cuda", dtype=torch.float16)
    wd = q.dequantize(torch.float16).float()
    y = G.gemm_w4(x, q, force=True)
    assert y is not None
    _check(y, x.float() @ wd.t(), 2e-2)
    dy = torch.randn(512, 1024, device="cuda", dtype=torch.float16)
    dx = G.gemm_w4(dy, q, b_mn=True, force=True)
    _check(dx, dy.float() @ wd, 2e-2)
