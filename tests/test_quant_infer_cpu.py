"""W4A16 formats / quantisers / quantised checkpoints and KV-cache generation — CPU reference paths."""
import math

import pytest
import torch

from llm_in_practise_b200.inference import GenerationConfig, Generator, TextIteratorStreamer, sample_next
from llm_in_practise_b200.models.qwen3 import QWEN3_PRESETS, Qwen3Config, Qwen3ForCausalLM
from llm_in_practise_b200.quant import (AWQModifier, GPTQModel, GPTQModifier, QuantizeConfig, from_compressed_tensors,
                                        from_gptq, gptq_quantize, load_quantized, oneshot, perplexity, quantize_rtn,
                                        save_quantized, to_compressed_tensors, to_gptq)


def tiny(dtype=torch.float32, **kw):
    cfg = Qwen3Config(vocab_size=320, hidden_size=128, intermediate_size=256, num_hidden_layers=2, num_attention_heads=4,
                      num_key_value_heads=2, head_dim=32, max_position_embeddings=256, **kw)
    torch.manual_seed(0)
    return Qwen3ForCausalLM(cfg, dtype=dtype, device="cpu", init_std=0.08)


@pytest.mark.parametrize("sym", [True, False])
def test_rtn_roundtrip_and_formats(sym):
    torch.manual_seed(1)
    w = torch.randn(64, 256)
    q = quantize_rtn(w, 128, sym)
    d = q.dequantize(torch.float32)
    step = (w.view(64, 2, 128).abs().amax(-1) / 7.0) if sym else ((w.view(64, 2, 128).amax(-1) - w.view(64, 2, 128).amin(-1)) / 15)
    assert ((d - w).view(64, 2, 128).abs().amax(-1) <= step * 0.5 + 1e-5).all()
    g = from_gptq(to_gptq(q), 128)
    assert torch.equal(g.packed, q.packed)
    # GPTQ stores fp16 scales
    torch.testing.assert_close(g.dequantize(torch.float32), d, rtol=2e-3, atol=2e-3)
    c = from_compressed_tensors(to_compressed_tensors(q, sym), 128)
    assert torch.equal(c.packed, q.packed)
    torch.testing.assert_close(c.dequantize(torch.float32), d, rtol=1e-2, atol=1e-2)


def test_gptq_beats_rtn_on_correlated_inputs():
    torch.manual_seed(2)
    K, N, T = 256, 96, 2048
    mix = torch.randn(K, K) * 0.3 + torch.eye(K)
    x = torch.randn(T, K) @ mix
    w = torch.randn(N, K) * 0.1
    H = 2 * x.t() @ x
    ref = x @ w.t()
    e_rtn = ((x @ quantize_rtn(w, 128, True).dequantize(torch.float32).t()) - ref).pow(2).mean()
    e_gptq = ((x @ gptq_quantize(w, H, 128, True).dequantize(torch.float32).t()) - ref).pow(2).mean()
    assert e_gptq < 0.8 * e_rtn, (float(e_gptq), float(e_rtn))


@pytest.mark.parametrize("method,fmt", [("gptq", "gptq"), ("awq", "compressed-tensors"), ("rtn", "compressed-tensors")])
def test_quantize_model_save_load(tmp_path, method, fmt):
    m = tiny()
    torch.manual_seed(3)
    calib = [torch.randint(0, 320, (1, 48)) for _ in range(6)]
    test = [torch.randint(0, 320, (2, 32))]
    with torch.no_grad():
        ref_logits = m(test[0])["logits"]
    if method == "gptq":
        gm = GPTQModel(m, QuantizeConfig(bits=4, group_size=128, desc_act=False))
        gm.quantize([c[0].tolist() for c in calib], batch_size=1)
        gm.save(str(tmp_path), fmt)
    else:
        mod = AWQModifier(targets="Linear", scheme="W4A16", ignore=["lm_head"]) if method == "awq" else \
            __import__("llm_in_practise_b200.quant", fromlist=["x"]).QuantizationModifier(scheme="W4A16_SYM")
        oneshot(m, [mod], dataset=[{"input_ids": c[0].tolist()} for c in calib], output_dir=str(tmp_path), save_format=fmt)
    with torch.no_grad():
        q_logits = m(test[0])["logits"]
    rel = (q_logits - ref_logits).norm() / ref_logits.norm()
    assert rel < 0.25, float(rel)
    m2 = load_quantized(str(tmp_path), dtype=torch.float32, device="cpu")
    with torch.no_grad():
        l2 = m2(test[0])["logits"]
    torch.testing.assert_close(l2, q_logits, rtol=3e-2, atol=3e-2)
    assert math.isfinite(perplexity(m2, test))


def test_awq_fold_preserves_function_before_rounding():
    """Folding s into the producer and 1/s into the consumer must leave the dense model unchanged."""
    from llm_in_practise_b200.quant.algorithms import ActivationStats
    from llm_in_practise_b200.quant.api import _awq_fold_layer
    m = tiny(attention_bias=True)
    ids = torch.randint(0, 320, (2, 24))
    with torch.no_grad():
        ref = m(ids)["logits"]
        layer = m.layers[0]
        stats = {k: ActivationStats(l.in_features, "cpu", False) for k, l in layer.fused_linears().items()}
        hooks = [l.register_forward_pre_hook(lambda mod, a, s=stats[k]: s.add(a[0])) for k, l in layer.fused_linears().items()]
        m(ids)
        for h in hooks:
            h.remove()
        _awq_fold_layer(layer, stats, QuantizeConfig(sym=False, method="awq"))
        out = m(ids)["logits"]
    torch.testing.assert_close(out, ref, rtol=1e-3, atol=1e-3)


def test_kv_cache_decode_matches_full_forward():
    m = tiny()
    g = Generator(m, max_len=64)
    ids = torch.randint(0, 320, (2, 9))
    out = g.generate(ids, max_new_tokens=6, do_sample=False)
    assert out.shape == (2, 15)
    with torch.no_grad():
        for t in range(9, 15):
            full = m(out[:, :t])["logits"][:, -1]
            assert torch.equal(full.argmax(-1), out[:, t]), t
    out2, lp = g.generate(ids, max_new_tokens=6, do_sample=False, return_logprobs=True)
    assert torch.equal(out, out2) and lp.shape == (2, 6) and (lp <= 0).all()


def test_sampling_processors():
    torch.manual_seed(0)
    lg = torch.tensor([[2.0, 1.0, 0.5, -1.0, -3.0]])
    assert sample_next(lg, GenerationConfig(do_sample=False)).item() == 0
    prev = torch.tensor([[0]])
    assert sample_next(lg, GenerationConfig(do_sample=False, repetition_penalty=4.0), prev).item() == 1
    cfg = GenerationConfig(do_sample=True, top_k=2, temperature=1.0)
    assert {sample_next(lg, cfg).item() for _ in range(50)} <= {0, 1}
    cfg = GenerationConfig(do_sample=True, top_p=0.5)
    assert {sample_next(lg, cfg).item() for _ in range(50)} == {0}
    cfg = GenerationConfig(do_sample=True, top_p=0.8)
    assert {sample_next(lg, cfg).item() for _ in range(200)} == {0, 1}


def test_eos_stops_and_streamer():
    m = tiny()
    g = Generator(m, max_len=64)
    ids = torch.randint(0, 320, (1, 5))
    first = g.generate(ids, max_new_tokens=4, do_sample=False)[0, 5].item()
    out = g.generate(ids, max_new_tokens=8, do_sample=False, eos_token_id=[first])
    assert out.shape[1] == 6

    class Tok:
        def decode(self, ids, **_):
            return "".join(chr(65 + i % 26) for i in ids)
    st = TextIteratorStreamer(Tok())
    g.generate(ids, max_new_tokens=5, do_sample=False, streamer=st)
    assert len("".join(st)) == 5


def test_packed_prefill_matches_per_prompt_prefill():
    """Several prompts of different lengths in one packed pass == one prefill each (logits and cached K/V)."""
    from llm_in_practise_b200.inference.generate import KVCache
    m = tiny()
    g = Generator(m, max_len=64)
    c = m.config
    prompts = [[5, 9, 200, 3, 17], [11, 12], [7] * 13]
    mk = lambda: KVCache(c.num_hidden_layers, 4, c.num_key_value_heads, 64, c.head_dim, torch.float32, "cpu")
    ca, cb = mk(), mk()
    slots = [2, 0, 3]
    packed = g._prefill_packed(prompts, ca, slots)
    for i, (p, s) in enumerate(zip(prompts, slots)):
        one = g._prefill(torch.tensor([p]), cb, torch.tensor([s]))
        torch.testing.assert_close(packed[i:i + 1], one, rtol=1e-4, atol=1e-4)
        for li in range(c.num_hidden_layers):
            torch.testing.assert_close(ca.k[li][s, :, :len(p)], cb.k[li][s, :, :len(p)], rtol=1e-5, atol=1e-5)
            torch.testing.assert_close(ca.v[li][s, :, :len(p)], cb.v[li][s, :, :len(p)], rtol=1e-5, atol=1e-5)


def test_linear_code_table_detection():
    """Integer code books (RTN / GPTQ / AWQ) are recognised as a * (code - off); NF4 and arbitrary tables are not."""
    from llm_in_practise_b200.ops.w4 import NF4_CODE, linear_lut
    lut = torch.arange(16, dtype=torch.float32)
    assert linear_lut(lut) == (1.0, 0.0)
    assert linear_lut(lut - 8) == (1.0, 8.0)
    assert linear_lut((lut - 8) * 0.25) == (0.25, 8.0)
    assert linear_lut(NF4_CODE.clone().float()) is None
    assert linear_lut(lut * lut) is None
    assert linear_lut(lut - 7.5) is None          # offset must be an integer
