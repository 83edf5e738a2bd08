"""GPU numerics: decode attention, KV-cache generation, 4-bit decode path, GPTQ quantisation on device."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _tiny(dtype=torch.bfloat16):
    from llm_in_practise_b200.models.qwen3 import Qwen3Config, Qwen3ForCausalLM
    cfg = Qwen3Config(vocab_size=512, hidden_size=256, intermediate_size=512, num_hidden_layers=2, num_attention_heads=4,
                      num_key_value_heads=2, head_dim=64, max_position_embeddings=512)
    torch.manual_seed(0)
    return Qwen3ForCausalLM(cfg, dtype=dtype, device="cuda", init_std=0.05)


@pytest.mark.parametrize("D,dtype", [(128, torch.bfloat16), (64, torch.float16), (32, torch.float32)])
def test_decode_attn_matches_reference(D, dtype):
    from llm_in_practise_b200.inference.generate import _decode_attn
    torch.manual_seed(0)
    B, H, Hkv, L = 3, 8, 2, 96
    qkv = torch.randn(B, (H + 2 * Hkv) * D, device="cuda", dtype=dtype)
    ck = torch.randn(B, Hkv, L, D, device="cuda", dtype=dtype)
    cv = torch.randn(B, Hkv, L, D, device="cuda", dtype=dtype)
    t = 70
    pos = torch.full((B,), t, device="cuda", dtype=torch.int32)
    ck0, cv0 = ck.clone(), cv.clone()
    out = _decode_attn(qkv, ck, cv, pos, H, Hkv, D, D ** -0.5)
    x = qkv.view(B, H + 2 * Hkv, D).float()
    ck0[:, :, t], cv0[:, :, t] = x[:, H:H + Hkv].to(dtype), x[:, H + Hkv:].to(dtype)
    assert torch.equal(ck[:, :, t], ck0[:, :, t]) and torch.equal(cv[:, :, t], cv0[:, :, t])
    k = ck0[:, :, :t + 1].float().repeat_interleave(H // Hkv, 1)
    v = cv0[:, :, :t + 1].float().repeat_interleave(H // Hkv, 1)
    s = torch.einsum("bhd,bhtd->bht", x[:, :H], k) * D ** -0.5
    ref = torch.einsum("bht,bhtd->bhd", s.softmax(-1), v).reshape(B, H * D)
    tol = 2e-2 if dtype != torch.float32 else 1e-4
    torch.testing.assert_close(out.float(), ref, rtol=tol, atol=tol)


def test_generate_matches_full_forward():
    from llm_in_practise_b200.inference import Generator
    m = _tiny()
    g = Generator(m, max_len=128)
    ids = torch.randint(0, 512, (2, 11), device="cuda")
    out, lp = g.generate(ids, max_new_tokens=8, do_sample=False, return_logprobs=True)
    assert out.shape == (2, 19)
    # teacher-forced logits of the generated sequence must rank the chosen token (near-)top at every step
    with torch.no_grad():
        full = m(out[:, :-1].contiguous())["logits"].float()
    for t in range(11, 19):
        row = full[:, t - 1]
        chosen = row.gather(1, out[:, t:t + 1]).squeeze(1)
        assert (row.max(-1).values - chosen < 0.15).all(), t


def test_w4_decode_path_and_quantize_on_device():
    from llm_in_practise_b200.inference import Generator
    from llm_in_practise_b200.quant import QuantizeConfig, quantize_model
    m = _tiny()
    calib = [torch.randint(0, 512, (1, 64), device="cuda") for _ in range(4)]
    ids = torch.randint(0, 512, (1, 16), device="cuda")
    with torch.no_grad():
        ref = m(ids)["logits"].float()
    quantize_model(m, calib, QuantizeConfig(method="gptq", group_size=128, sym=True))
    with torch.no_grad():
        q = m(ids)["logits"].float()
    assert (q - ref).norm() / ref.norm() < 0.3
    g = Generator(m, max_len=64)
    out = g.generate(ids, max_new_tokens=6, do_sample=False)       # M=1 decode -> fused w4 GEMV kernels
    with torch.no_grad():
        full = m(out[:, :-1].contiguous())["logits"].float()
    for t in range(16, 22):
        row = full[:, t - 1]
        assert (row.max(-1).values - row.gather(1, out[:, t:t + 1]).squeeze(1) < 0.2).all(), t


def test_cuda_graph_decode_matches_eager():
    from llm_in_practise_b200.inference import Generator
    from llm_in_practise_b200.quant.formats import quantize_rtn
    m = _tiny()
    for layer in m.layers:
        for lin in layer.fused_linears().values():
            lin.set_nf4(quantize_rtn(lin.weight.data, 128, True))
    ids = torch.randint(0, 512, (3, 9), device="cuda")
    g = Generator(m, max_len=64)
    a = g.generate(ids, max_new_tokens=12, do_sample=False, cuda_graph=True)
    assert len(g._graphs) == 1
    b = g.generate(ids, max_new_tokens=12, do_sample=False, cuda_graph=False)
    assert torch.equal(a, b)
    # a second generation reuses... a new cache means a new graph; the engine keeps one cache and one graph per batch size
    from llm_in_practise_b200.serve import EngineArgs, LLMEngine
    eng = LLMEngine(m, None, EngineArgs(max_num_seqs=4, max_model_len=64))
    from llm_in_practise_b200.inference import GenerationConfig
    reqs = eng.generate([ids[i].tolist() for i in range(3)], GenerationConfig(max_new_tokens=12, do_sample=False))
    assert [r.output_ids for r in reqs] == [a[i, 9:].tolist() for i in range(3)]
    assert 1 <= len(eng.gen._graphs) <= 3


def test_packed_prefill_matches_per_prompt_prefill_gpu():
    """Varlen flash kernel + scattered cache writes: one packed pass over prompts of different lengths (including ones
    that cross 64-row tile boundaries) equals a prefill per prompt."""
    from llm_in_practise_b200.inference import Generator
    from llm_in_practise_b200.inference.generate import KVCache
    m = _tiny()
    g = Generator(m, max_len=256)
    c = m.config
    gen = torch.Generator().manual_seed(1)
    prompts = [torch.randint(0, 512, (n,), generator=gen).tolist() for n in (5, 130, 64, 77)]
    mk = lambda: KVCache(c.num_hidden_layers, 5, c.num_key_value_heads, 256, c.head_dim, torch.bfloat16, "cuda")
    ca, cb = mk(), mk()
    slots = [4, 0, 2, 1]
    packed = g._prefill_packed(prompts, ca, slots)
    for i, (p, s) in enumerate(zip(prompts, slots)):
        one = g._prefill(torch.tensor([p], device="cuda"), cb, torch.tensor([s], device="cuda"))
        assert (packed[i] - one[0]).abs().max() < 0.05 * one.abs().max()
        assert packed[i].argmax() == one[0].argmax()
        assert torch.equal(ca.k[0][s, :, :len(p)], cb.k[0][s, :, :len(p)])
    # CUDA-graph path: bucketed token count, padded sequence count, padding tokens land in the scratch cache row
    cc = mk()
    for _ in range(2):                       # capture, then replay
        graphed = g._prefill_packed(prompts, cc, slots, scratch_slot=3)
        assert (graphed - packed).abs().max() < 0.02 * packed.abs().max()
    other = [torch.randint(0, 512, (n,), generator=gen).tolist() for n in (31, 200)]   # another bucket, other lengths
    g2 = g._prefill_packed(other, cc, [0, 1], scratch_slot=3)
    e2 = g._prefill_packed(other, cb, [0, 1])
    assert (g2 - e2).abs().max() < 0.02 * e2.abs().max()
    assert torch.equal(cc.k[1][1, :, :200], cb.k[1][1, :, :200])


@pytest.mark.parametrize("D,dtype,G", [(128, torch.bfloat16, 4), (64, torch.float16, 2), (32, torch.float32, 4)])
def test_paged_decode_attention_matches_slot_pool(D, dtype, G):
    """b200_decode_attn{,_split}_paged: the same K/V scattered over a shuffled pool of 16-token blocks (block table per
    sequence) gives the same output as the contiguous [slot, head, position] cache, and the new token's K/V lands in the
    right physical block; sequences of different lengths, slots out of order."""
    from llm_in_practise_b200.inference.generate import _decode_attn
    torch.manual_seed(1)
    B, Hkv, L, BS = 3, 2, 96, 16
    H = Hkv * G
    qkv = torch.randn(B, (H + 2 * Hkv) * D, device="cuda", dtype=dtype)
    ck = torch.randn(4, Hkv, L, D, device="cuda", dtype=dtype)
    cv = torch.randn(4, Hkv, L, D, device="cuda", dtype=dtype)
    slots = torch.tensor([2, 0, 3], device="cuda", dtype=torch.int32)
    pos = torch.tensor([70, 5, 47], device="cuda", dtype=torch.int32)
    nbs = L // BS
    perm = torch.randperm(4 * nbs, device="cuda") + 1                            # physical block 0 stays unused
    table = perm.view(4, nbs).to(torch.int32).contiguous()
    pk = torch.zeros(4 * nbs + 1, Hkv, BS, D, device="cuda", dtype=dtype)
    pv = torch.zeros_like(pk)
    for s in range(4):
        for j in range(nbs):
            pk[table[s, j]] = ck[s, :, j * BS:(j + 1) * BS]
            pv[table[s, j]] = cv[s, :, j * BS:(j + 1) * BS]
    want = _decode_attn(qkv, ck, cv, pos, H, Hkv, D, D ** -0.5, slots)
    got = _decode_attn(qkv, pk, pv, pos, H, Hkv, D, D ** -0.5, slots, table)
    torch.testing.assert_close(got.float(), want.float(), rtol=1e-3, atol=1e-3)
    for b in range(B):
        s, t = int(slots[b]), int(pos[b])
        blk = int(table[s, t // BS])
        assert torch.equal(pk[blk, :, t % BS], ck[s, :, t]) and torch.equal(pv[blk, :, t % BS], cv[s, :, t])


def test_paged_engine_matches_slot_pool_engine_gpu():
    """LLMEngine with block_size 16 (paged cache, CUDA-graph decode buckets, packed prefill, shared prefix blocks) emits
    the same greedy tokens as the slot-pool engine."""
    from llm_in_practise_b200.inference import GenerationConfig
    from llm_in_practise_b200.serve import EngineArgs, LLMEngine
    model = _tiny()
    torch.manual_seed(2)
    system = torch.randint(0, 512, (48,)).tolist()
    prompts = [system + torch.randint(0, 512, (n,)).tolist() for n in (9, 21, 4)] + [torch.randint(0, 512, (70,)).tolist()]
    outs = []
    for bs in (0, 16):
        eng = LLMEngine(model, None, EngineArgs(max_num_seqs=4, max_model_len=256, max_prefill_tokens=64, warmup=False,
                                                block_size=bs))
        reqs = [eng.submit(p, GenerationConfig(max_new_tokens=12, do_sample=False)) for p in prompts[:2]]
        for _ in range(3):
            eng.step()
        reqs += [eng.submit(p, GenerationConfig(max_new_tokens=12, do_sample=False)) for p in prompts[2:]]
        n = 0
        while not all(r.done.is_set() for r in reqs):
            eng.step()
            n += 1
            assert n < 500
        outs.append([r.output_ids for r in reqs])
        if bs:
            assert eng._reserved == 0 and sum(r.cached_tokens for r in reqs) >= 48
    assert outs[0] == outs[1]
