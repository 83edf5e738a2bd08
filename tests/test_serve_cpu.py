"""Serving stack on CPU: continuous batching vs. single-sequence generation, OpenAI API schema, moderation wrapper."""
import asyncio
import json

import httpx
import pytest
import torch
from fastapi.testclient import TestClient

from llm_in_practise_b200.data.tokenizers import ByteTokenizer
from llm_in_practise_b200.inference import GenerationConfig, Generator
from llm_in_practise_b200.models.qwen3 import Qwen3Config, Qwen3ForCausalLM
from llm_in_practise_b200.serve import EngineArgs, LLMEngine
from llm_in_practise_b200.serve.moderation import (build_moderation_app, http_backend, interpret_guard_output,
                                                   make_openai_moderation_response, parse_native_scodes)
from llm_in_practise_b200.serve.openai_server import build_app


@pytest.fixture(scope="module")
def model():
    cfg = Qwen3Config(vocab_size=300, hidden_size=64, intermediate_size=128, num_hidden_layers=2, num_attention_heads=4,
                      num_key_value_heads=2, head_dim=16, max_position_embeddings=256)
    torch.manual_seed(0)
    return Qwen3ForCausalLM(cfg, dtype=torch.float32, device="cpu", init_std=0.1)


def test_continuous_batching_matches_sequential(model):
    torch.manual_seed(1)
    prompts = [torch.randint(0, 300, (n,)).tolist() for n in (5, 9, 3, 12, 7)]
    g = Generator(model, max_len=64)
    want = [g.generate(torch.tensor([p]), max_new_tokens=6 + i, do_sample=False)[0, len(p):].tolist() for i, p in enumerate(prompts)]
    eng = LLMEngine(model, None, EngineArgs(max_num_seqs=2, max_model_len=64, max_prefill_per_step=1,
                                              enable_prefix_caching=False))   # forces queueing + slot reuse
    reqs = [eng.submit(p, GenerationConfig(max_new_tokens=6 + i, do_sample=False)) for i, p in enumerate(prompts)]
    assert eng.num_waiting == 5
    while not all(r.done.is_set() for r in reqs):
        eng.step()
        assert eng.num_running <= 2
    assert [r.output_ids for r in reqs] == want
    assert all(r.finish_reason == "length" and r.ttft is not None for r in reqs)
    assert sorted(eng.free) == [0, 1] and eng.metrics.generation_tokens == sum(len(w) for w in want)


def test_chunked_prefill_and_prefix_cache_match_plain_generation(model):
    """enable_chunked_prefill: a prompt longer than the token budget is prefilled in pieces between decode steps of the
    running sequences; enable_prefix_caching: a prompt that starts like an earlier one reuses that row's K/V (in place
    when the earlier sequence has finished, by copy while it is still running). Outputs equal one-by-one generation."""
    torch.manual_seed(3)
    system = torch.randint(0, 300, (40,)).tolist()                    # a shared "system prompt" of 2.5 blocks
    prompts = [system + torch.randint(0, 300, (n,)).tolist() for n in (7, 11, 3)] + [torch.randint(0, 300, (50,)).tolist()]
    g = Generator(model, max_len=96)
    want = [g.generate(torch.tensor([p]), max_new_tokens=5, do_sample=False)[0, len(p):].tolist() for p in prompts]
    eng = LLMEngine(model, None, EngineArgs(max_num_seqs=3, max_model_len=96, max_prefill_per_step=4, max_prefill_tokens=24,
                                            prefix_block=16))
    r0 = eng.submit(prompts[0], GenerationConfig(max_new_tokens=5, do_sample=False))
    eng.step()                                                          # 47 tokens > 24: first piece only
    assert r0 in eng.prefilling and 0 < r0.prefilled < len(prompts[0]) and eng.num_running == 0
    eng.step()
    assert eng.num_running == 1 and r0.first_token_time is not None
    r1 = eng.submit(prompts[1], GenerationConfig(max_new_tokens=5, do_sample=False))   # r0 still running: prefix by copy
    eng.step()
    assert r1.cached_tokens == 32 and r1.slot != r0.slot
    while not (r0.done.is_set() and r1.done.is_set()):
        eng.step()
    r2 = eng.submit(prompts[2], GenerationConfig(max_new_tokens=5, do_sample=False))   # both finished: continues in a cached row
    r3 = eng.submit(prompts[3], GenerationConfig(max_new_tokens=5, do_sample=False))   # no shared prefix, chunked
    while not (r2.done.is_set() and r3.done.is_set()):
        eng.step()
    assert r2.cached_tokens == 32 and r3.cached_tokens == 0
    assert [r.output_ids for r in (r0, r1, r2, r3)] == want
    m = eng.metrics
    assert m.prefix_hits == 64 and m.prefix_queries == sum(len(p) for p in prompts)
    assert len(eng.free) + len(eng._cached) == 3                        # every row is free or a cached prefix


def test_packed_admission_respects_token_budget(model):
    """Several prompts admitted in one iteration run as one packed prefill; the token budget splits a burst over
    iterations (a single over-long prompt is still admitted alone); results equal one-by-one generation."""
    torch.manual_seed(2)
    prompts = [torch.randint(0, 300, (n,)).tolist() for n in (6, 6, 6, 20, 4)]
    g = Generator(model, max_len=64)
    want = [g.generate(torch.tensor([p]), max_new_tokens=5, do_sample=False)[0, len(p):].tolist() for p in prompts]
    eng = LLMEngine(model, None, EngineArgs(max_num_seqs=8, max_model_len=64, max_prefill_per_step=8, max_prefill_tokens=13,
                                              enable_chunked_prefill=False, enable_prefix_caching=False))
    reqs = [eng.submit(p, GenerationConfig(max_new_tokens=5, do_sample=False)) for p in prompts]
    eng.step()
    assert eng.num_running == 2 and eng.num_waiting == 3          # 6 + 6 fit the budget of 13, the third prompt does not
    eng.step()
    assert eng.num_waiting == 2                                    # third admitted alone (20 does not fit behind it)
    eng.step()
    assert eng.num_waiting == 1                                    # the 20-token prompt exceeds the budget but goes alone
    while not all(r.done.is_set() for r in reqs):
        eng.step()
    assert [r.output_ids for r in reqs] == want


def test_openai_endpoints(model):
    tok = ByteTokenizer()
    eng = LLMEngine(model, tok, EngineArgs(max_num_seqs=4, max_model_len=128), model_name="tiny").start()
    try:
        c = TestClient(build_app(eng, prompt_style="deepseek"))
        body = {"model": "tiny", "messages": [{"role": "system", "content": "s"}, {"role": "user", "content": "hi"}],
                "max_tokens": 5, "temperature": 0.0}
        r = c.post("/v1/chat/completions", json=body)
        assert r.status_code == 200, r.text
        j = r.json()
        assert j["object"] == "chat.completion" and j["id"].startswith("chatcmpl-") and j["model"] == "tiny"
        assert j["choices"][0]["message"]["role"] == "assistant" and j["choices"][0]["finish_reason"] in ("stop", "length")
        u = j["usage"]
        assert u["total_tokens"] == u["prompt_tokens"] + u["completion_tokens"] and 0 < u["completion_tokens"] <= 5
        with c.stream("POST", "/v1/chat/completions", json={**body, "stream": True}) as s:
            lines = [ln for ln in s.iter_lines() if ln.startswith("data: ")]
        assert lines[-1] == "data: [DONE]" and json.loads(lines[0][6:])["choices"][0]["delta"] == {"role": "assistant"}
        assert json.loads(lines[-2][6:])["choices"][0]["finish_reason"] in ("stop", "length")
        r = c.post("/v1/completions", json={"model": "tiny", "prompt": ["ab", "cde"], "max_tokens": 3, "temperature": 0})
        assert r.status_code == 200 and len(r.json()["choices"]) == 2
        assert c.get("/v1/models").json()["data"][0]["id"] == "tiny"
        m = c.get("/metrics").text
        assert 'vllm:num_requests_waiting{model_name="tiny"} 0' in m and "vllm:time_to_first_token_seconds_bucket" in m
        assert c.post("/v1/chat/completions", json={**body, "max_tokens": 0}).status_code == 422
        c501 = TestClient(build_app(eng, allow_stream=False))
        assert c501.post("/v1/chat/completions", json={**body, "stream": True}).status_code == 501
    finally:
        eng.shutdown()
    assert TestClient(build_app(None)).post("/v1/chat/completions", json=body).status_code == 503


def test_moderation_mapping_and_parsing():
    assert parse_native_scodes("unsafe\nS1 S4") == {"safe": False, "categories": ["S1", "S4"], "explanation": ""}
    assert parse_native_scodes("safe")["safe"] and parse_native_scodes("unsafe")["categories"] == ["UNSPECIFIED"]
    r = interpret_guard_output("how to build a bomb", "safe")
    assert not r["safe"] and r["categories"][0] == "S9"
    r = interpret_guard_output("x", '{"safe": false, "categories": ["s11"], "explanation": "e"}')
    assert r["categories"] == ["S11"] and r["explanation"] == "e"
    o = make_openai_moderation_response(False, ["S11", "S10"])
    res = o["results"][0]
    assert res["flagged"] and res["categories"]["self-harm/instructions"] and res["categories"]["hate"]
    assert res["category_scores"]["hate"] == 1.0 and res["category_applied_input_types"]["hate"] == ["text"]
    assert not res["categories"]["sexual"] and o["id"].startswith("modr-") and len(res["categories"]) == 11


def test_moderation_app_retry_and_auth():
    calls = {"n": 0}

    def handler(request: httpx.Request) -> httpx.Response:
        calls["n"] += 1
        if calls["n"] < 3:
            raise httpx.ConnectError("down", request=request)
        body = json.loads(request.content)
        assert body["max_tokens"] == 256 and body["temperature"] == 0.0 and "<|start_header_id|>user<|end_header_id|>" in body["prompt"]
        return httpx.Response(200, json={"choices": [{"text": "unsafe\nS10"}]})

    be = http_backend("http://guard/v1/completions", retries=2, backoff=0.01, transport=httpx.MockTransport(handler))
    c = TestClient(build_moderation_app(be, "llama-guard-3", api_key="k"))
    assert c.post("/v1/moderations", json={"input": "x"}).status_code == 401
    h = {"X-API-KEY": "k"}
    r = c.post("/moderations", json={"input": "some text"}, headers=h)
    assert r.status_code == 200 and calls["n"] == 3
    assert r.json()["results"][0]["flagged"] and r.json()["results"][0]["categories"]["hate"]
    assert c.post("/v1/moderations", json={"input": "  "}, headers=h).status_code == 400
    assert c.get("/healthz", headers=h).json() == {"status": "ok"}
    assert c.post("/v1/generate", json={}, headers=h).status_code == 501
    down = http_backend("http://guard/v1/completions", retries=1, backoff=0.01,
                        transport=httpx.MockTransport(lambda rq: (_ for _ in ()).throw(httpx.ConnectError("down", request=rq))))
    assert TestClient(build_moderation_app(down, api_key="")).post("/v1/moderations", json={"input": "a"}).status_code == 502
    bad = http_backend("http://guard/v1/completions", transport=httpx.MockTransport(lambda rq: httpx.Response(500, text="boom")))
    assert TestClient(build_moderation_app(bad, api_key="")).post("/v1/moderations", json={"input": "a"}).status_code == 502


def test_loadtest_against_app(model):
    from llm_in_practise_b200.serve.loadtest import PROFILES, run_load
    tok = ByteTokenizer()
    eng = LLMEngine(model, tok, EngineArgs(max_num_seqs=4, max_model_len=256), model_name="qwen3").start()
    PROFILES["mini"] = {"payload": {"model": "qwen3", "messages": [{"role": "user", "content": "hi"}], "max_tokens": 4,
                                    "temperature": 0.0}, "wait": (0.0, 0.01)}
    try:
        out = asyncio.run(run_load("http://test", "mini", users=3, duration=30, max_requests_per_user=2,
                                   transport=httpx.ASGITransport(app=build_app(eng))))
    finally:
        eng.shutdown()
    assert out["requests"] == 6 and out["failures"] == 0 and out["completion_tokens"] > 0
    assert out["aggregate_tokens_per_second"] > 0


def test_cache_gateway_levels_and_metrics():
    """Exact (L2) and semantic (L3) response caches in front of a chat endpoint: miss -> upstream, identical request -> L2,
    re-worded request -> L3 (then promoted to L2), sampled / streaming requests bypass, upstream failures become 502."""
    from fastapi.testclient import TestClient
    from llm_in_practise_b200.serve.cache_gateway import build_cache_gateway, hashed_ngram_embedding
    calls = []

    async def upstream(payload):
        if payload["messages"][-1]["content"] == "boom":
            raise RuntimeError("backend down")
        calls.append(payload)
        return {"id": f"cmpl-{len(calls)}", "choices": [{"message": {"role": "assistant", "content": f"answer {len(calls)}"}}],
                "usage": {"completion_tokens": 2}}

    c = TestClient(build_cache_gateway(upstream, ttl=60, semantic_threshold=0.9))
    q1 = {"model": "m", "messages": [{"role": "user", "content": "How do I restart the nginx service on the server?"}]}
    r = c.post("/v1/chat/completions", json=q1)
    assert r.status_code == 200 and r.headers["X-Cache"] == "MISS" and len(calls) == 1
    r2 = c.post("/v1/chat/completions", json=q1)
    assert r2.headers["X-Cache"] == "HIT-L2" and r2.json() == r.json() and len(calls) == 1
    q2 = {"model": "m", "messages": [{"role": "user", "content": "how do I restart the nginx service on the server ?"}]}
    r3 = c.post("/v1/chat/completions", json=q2)
    assert r3.headers["X-Cache"] == "HIT-L3" and r3.json() == r.json() and len(calls) == 1
    assert c.post("/v1/chat/completions", json=q2).headers["X-Cache"] == "HIT-L2"          # promoted
    q3 = {"model": "m", "messages": [{"role": "user", "content": "Write a haiku about tensor cores."}]}
    assert c.post("/v1/chat/completions", json=q3).headers["X-Cache"] == "MISS" and len(calls) == 2
    assert c.post("/v1/chat/completions", json=dict(q1, temperature=0.8)).headers["X-Cache"] == "BYPASS" and len(calls) == 3
    assert c.post("/v1/chat/completions", json={"model": "m", "messages": []}).status_code == 400
    assert c.post("/v1/chat/completions", json={"model": "m", "messages": [{"role": "user", "content": "boom"}]}).status_code == 502
    m = c.get("/metrics").text
    assert 'cache_gateway_requests_total{result="l2_hit"} 2' in m and 'result="l3_hit"} 1' in m and 'result="miss"} 3' in m
    assert 'result="bypass"} 1' in m and 'result="upstream_error"} 1' in m and "cache_gateway_latency_seconds_bucket" in m
    assert c.get("/healthz").json() == {"status": "ok"}
    a, b = hashed_ngram_embedding("restart nginx now"), hashed_ngram_embedding("completely different text about GPUs")
    assert abs(float(a.norm()) - 1) < 1e-5 and float(a @ b) < 0.5


def test_stop_strings_and_abort(model):
    """`stop` truncates the completion at the first match (finish_reason "stop"), also while streaming, where the rest of the
    generation is aborted; LLMEngine.abort frees the KV slot of a running request and drops a waiting one."""
    import json
    tok = ByteTokenizer()
    eng = LLMEngine(model, tok, EngineArgs(max_num_seqs=2, max_model_len=128), model_name="tiny").start()
    try:
        c = TestClient(build_app(eng, prompt_style="deepseek"))
        body = {"model": "tiny", "messages": [{"role": "user", "content": "hello there"}], "max_tokens": 24, "temperature": 0.0}
        full = c.post("/v1/chat/completions", json=body).json()["choices"][0]["message"]["content"]
        assert len(full) >= 6
        stop = full[3:5]
        want = full[:full.find(stop)]
        j = c.post("/v1/chat/completions", json=dict(body, stop=[stop, "\x00never"])).json()
        assert j["choices"][0]["message"]["content"] == want.strip() or j["choices"][0]["message"]["content"] == want
        assert j["choices"][0]["finish_reason"] == "stop"
        with c.stream("POST", "/v1/chat/completions", json=dict(body, stream=True, stop=stop)) as r:
            chunks = [json.loads(ln[6:]) for ln in r.iter_lines() if ln.startswith("data: {")]
        text = "".join(ch["choices"][0]["delta"].get("content", "") for ch in chunks)
        assert stop not in text and want.startswith(text.rstrip()) or text == want
        assert chunks[-1]["choices"][0]["finish_reason"] == "stop"
        cj = c.post("/v1/completions", json={"model": "tiny", "prompt": "abc", "max_tokens": 12, "temperature": 0.0, "stop": "\x00never"}).json()
        assert cj["choices"][0]["finish_reason"] == "length"
    finally:
        eng.shutdown()
    eng2 = LLMEngine(model, None, EngineArgs(max_num_seqs=1, max_model_len=64))
    a = eng2.submit([1, 2, 3], GenerationConfig(max_new_tokens=40, do_sample=False))
    b = eng2.submit([4, 5], GenerationConfig(max_new_tokens=3, do_sample=False))
    eng2.step(); eng2.step()
    assert eng2.num_running == 1 and eng2.num_waiting == 1
    eng2.abort(a)
    while not b.done.is_set():
        eng2.step()
    assert a.finish_reason == "abort" and a.done.is_set() and 0 < len(a.output_ids) < 40
    assert b.finish_reason == "length" and len(b.output_ids) == 3 and sorted(eng2.free) == [0]
    w = eng2.submit([7], GenerationConfig(max_new_tokens=2, do_sample=False))
    w2 = eng2.submit([8], GenerationConfig(max_new_tokens=2, do_sample=False))
    eng2.abort(w2)
    assert w2.finish_reason == "abort" and eng2.num_waiting == 1
    while not w.done.is_set():
        eng2.step()
    assert len(w.output_ids) == 2


def _drain(eng, reqs):
    n = 0
    while not all(r.done.is_set() for r in reqs):
        eng.step()
        n += 1
        assert n < 2000
    return n


def test_paged_kv_cache_matches_slot_pool(model):
    """block_size > 0: the KV cache is a pool of 16-token blocks with a block table (vLLM's PagedAttention layout). Same
    tokens as one-by-one generation through chunked prefill, a prefix shared with a RUNNING sequence (both table rows list
    the same physical blocks — nothing is copied), a prefix continued in a finished sequence's row, and queueing when the
    pool cannot promise a request its blocks; every block returns to the pool."""
    torch.manual_seed(3)
    system = torch.randint(0, 300, (40,)).tolist()
    prompts = [system + torch.randint(0, 300, (n,)).tolist() for n in (7, 11, 3)] + [torch.randint(0, 300, (50,)).tolist()]
    g = Generator(model, max_len=96)
    want = [g.generate(torch.tensor([p]), max_new_tokens=5, do_sample=False)[0, len(p):].tolist() for p in prompts]
    eng = LLMEngine(model, None, EngineArgs(max_num_seqs=3, max_model_len=96, max_prefill_per_step=4, max_prefill_tokens=24,
                                            prefix_block=16, block_size=16))
    assert eng.paged and eng.cache.k[0].shape == (3 * 6 + 1, 2, 16, 16)
    total = len(eng._block_ref) - 1
    r0 = eng.submit(prompts[0], GenerationConfig(max_new_tokens=5, do_sample=False))
    eng.step()
    assert r0 in eng.prefilling and 0 < r0.prefilled < len(prompts[0])
    eng.step()
    r1 = eng.submit(prompts[1], GenerationConfig(max_new_tokens=5, do_sample=False))   # r0 still running: shared blocks
    eng.step()
    assert r1.cached_tokens == 32 and r1.slot != r0.slot
    assert eng._slot_blocks[r1.slot][:2] == eng._slot_blocks[r0.slot][:2]
    assert all(eng._block_ref[b] == 2 for b in eng._slot_blocks[r0.slot][:2])
    _drain(eng, [r0, r1])
    r2 = eng.submit(prompts[2], GenerationConfig(max_new_tokens=5, do_sample=False))
    r3 = eng.submit(prompts[3], GenerationConfig(max_new_tokens=5, do_sample=False))
    _drain(eng, [r2, r3])
    assert r2.cached_tokens == 32 and r3.cached_tokens == 0
    assert [r.output_ids for r in (r0, r1, r2, r3)] == want
    # accounting: nothing reserved, every block is free or held by a cached row, reference counts match the table rows
    assert eng._reserved == 0
    held = {}
    for slot, blocks in eng._slot_blocks.items():
        for b in blocks:
            held[b] = held.get(b, 0) + 1
    assert all(eng._block_ref[b] == n for b, n in held.items())
    assert len(eng._free_blocks) + len(held) == total and 0 not in held


def test_paged_kv_cache_small_pool_queues_and_recovers(model):
    """A pool smaller than max_num_seqs x max_model_len: admissions wait until finished sequences give their blocks back
    (cached prefix rows are evicted first); outputs still equal sequential generation."""
    torch.manual_seed(5)
    prompts = [torch.randint(0, 300, (n,)).tolist() for n in (20, 33, 18, 40, 25, 9)]
    g = Generator(model, max_len=96)
    want = [g.generate(torch.tensor([p]), max_new_tokens=8, do_sample=False)[0, len(p):].tolist() for p in prompts]
    eng = LLMEngine(model, None, EngineArgs(max_num_seqs=4, max_model_len=96, max_prefill_per_step=4, max_prefill_tokens=64,
                                            block_size=16, num_kv_blocks=9))        # 144 tokens of KV for 4 rows x 96
    reqs = [eng.submit(p, GenerationConfig(max_new_tokens=8, do_sample=False)) for p in prompts]
    peak = 0
    n = 0
    while not all(r.done.is_set() for r in reqs):
        eng.step()
        peak = max(peak, eng.num_running + len(eng.prefilling))
        assert len(eng._free_blocks) - eng._reserved >= 0
        n += 1
        assert n < 2000
    assert [r.output_ids for r in reqs] == want
    assert peak < 4                                                     # the block pool, not the row count, was the limit
    assert 0.0 <= eng.kv_cache_usage() <= 1.0


def test_paged_pool_rejects_or_caps_requests_it_can_never_hold(model):
    import pytest as _pt
    eng = LLMEngine(model, None, EngineArgs(max_num_seqs=2, max_model_len=96, block_size=16, num_kv_blocks=3))   # 48 tokens of KV
    with _pt.raises(ValueError):
        eng.submit(list(range(60)), GenerationConfig(max_new_tokens=4, do_sample=False))
    r = eng.submit(list(range(20)), GenerationConfig(max_new_tokens=64, do_sample=False))                       # worst case 84 tokens
    assert r.cfg.max_new_tokens == 48 - 8 - 20
    _drain(eng, [r])
    assert len(r.output_ids) == 20 and r.finish_reason == "length" and eng._reserved == 0


@pytest.mark.parametrize("block_size", [0, 16])
def test_host_tier_restores_an_evicted_prefix(model, block_size):
    """kv_offload_gb: a cached prefix that loses its HBM row goes to host memory; a later prompt with that prefix gets it copied
    back (cached_tokens counts it, only the suffix is computed) and generates exactly what a cold engine generates."""
    torch.manual_seed(11)
    system = torch.randint(0, 300, (48,)).tolist()
    others = [torch.randint(0, 300, (40,)).tolist() for _ in range(3)]
    first, again = system + [7, 8, 9], system + [21, 22]
    g = Generator(model, max_len=96)
    cfg = GenerationConfig(max_new_tokens=4, do_sample=False)
    want = {tuple(p): g.generate(torch.tensor([p]), max_new_tokens=4, do_sample=False)[0, len(p):].tolist() for p in [first, again] + others}
    eng = LLMEngine(model, None, EngineArgs(max_num_seqs=2, max_model_len=96, block_size=block_size, kv_offload_gb=0.01, warmup=False))

    def run(p):
        r = eng.submit(p, cfg)
        while not r.done.is_set():
            eng.step()
        assert r.output_ids == want[tuple(p)]
        return r
    run(first)
    for p in others:                                            # two rows only: the system prompt's row is recycled
        run(p)
    assert eng.metrics.host_offloads >= 1 and eng._lookup_prefix(again) == (-1, 0)
    key, n = eng._lookup_host(again)
    assert key >= 0 and n == 48
    r = run(again)
    assert r.cached_tokens == 48 and eng.metrics.host_hits == 48
    # budget: the tier holds at most kv_offload_gb of rows and forgets the least recently used ones
    assert eng._host_bytes <= eng._host_budget and eng._host_bytes == sum(b for _, b in eng._host_rows.values())
    tiny = LLMEngine(model, None, EngineArgs(max_num_seqs=1, max_model_len=96, block_size=block_size, warmup=False,
                                             kv_offload_gb=(48 * eng._kv_token_bytes + 64) / 2 ** 30))       # room for one 48-token row
    for p in [first] + others:
        r = tiny.submit(p, cfg)
        while not r.done.is_set():
            tiny.step()
    assert len(tiny._host_rows) == 1 and tiny._lookup_host(again) == (-1, 0)      # the system prompt was the oldest entry
    assert set(tiny.gen._host_kv) == set(tiny._host_rows)                         # dropped entries free their host memory
    off = LLMEngine(model, None, EngineArgs(max_num_seqs=1, max_model_len=96, warmup=False))
    for p in (first, others[0]):
        r = off.submit(p, cfg)
        while not r.done.is_set():
            off.step()
    assert off.metrics.host_offloads == 0 and not off._host_rows                  # off by default


def test_completions_stream_ignore_eos_and_usage_chunk(model):
    """What ``vllm bench serve --backend vllm --ignore-eos`` sends: a streaming /v1/completions request with ignore_eos and
    stream_options.include_usage; the answer is text deltas, a finish chunk and a usage chunk."""
    tok = ByteTokenizer()
    eng = LLMEngine(model, tok, EngineArgs(max_num_seqs=2, max_model_len=128, warmup=False), model_name="qwen3").start()
    try:
        c = TestClient(build_app(eng))
        body = {"model": "qwen3", "prompt": "hello", "max_tokens": 6, "temperature": 0.0}
        plain = c.post("/v1/completions", json=body).json()
        first = plain["choices"][0]
        # make the engine stop early: the second generated token becomes EOS
        ids = eng.generate([tok.encode("hello")], GenerationConfig(max_new_tokens=6, do_sample=False))[0].output_ids
        tok.eos_token_id = ids[1]
        assert c.post("/v1/completions", json=body).json()["usage"]["completion_tokens"] == ids.index(ids[1]) < 6     # the EOS token itself is not output
        r = c.post("/v1/completions", json=dict(body, stream=True, ignore_eos=True, stream_options={"include_usage": True}))
        assert r.status_code == 200 and r.headers["content-type"].startswith("text/event-stream")
        events = [json.loads(l[6:]) for l in r.text.splitlines() if l.startswith("data: ") and l != "data: [DONE]"]
        assert r.text.rstrip().endswith("data: [DONE]")
        usage = [e for e in events if e.get("usage")]
        assert len(usage) == 1 and usage[0]["usage"]["completion_tokens"] == 6 and usage[0]["choices"] == []      # EOS ignored
        assert usage[0]["usage"]["prompt_tokens"] == len(tok.encode("hello"))
        text = "".join(e["choices"][0]["text"] for e in events if e["choices"])
        assert text == first["text"] and [e["choices"][0]["finish_reason"] for e in events if e["choices"]][-1] == "length"
        chat = c.post("/v1/chat/completions", json={"model": "qwen3", "messages": [{"role": "user", "content": "hi"}], "max_tokens": 4,
                                                    "temperature": 0.0, "stream": True, "ignore_eos": True,
                                                    "stream_options": {"include_usage": True}})
        ev = [json.loads(l[6:]) for l in chat.text.splitlines() if l.startswith("data: ") and l != "data: [DONE]"]
        assert ev[-1]["usage"]["completion_tokens"] == 4
    finally:
        del tok.eos_token_id
        eng.shutdown()


def test_bench_serve_client_reports_vllm_result_keys(model, tmp_path):
    from llm_in_practise_b200.serve.bench_client import format_report, load_sharegpt, random_requests, run_benchmark
    tok = ByteTokenizer()
    eng = LLMEngine(model, tok, EngineArgs(max_num_seqs=4, max_model_len=256, warmup=False), model_name="qwen3").start()
    data = [{"conversations": [{"from": "human", "value": f"question number {i}"}, {"from": "gpt", "value": "an answer of five words"}]}
            for i in range(5)] + [{"conversations": [{"from": "human", "value": "only one turn"}]}]
    path = tmp_path / "sharegpt.json"
    path.write_text(json.dumps(data))
    reqs = load_sharegpt(str(path), 8, 5, seed=1)
    assert len(reqs) == 8 and all(r.output_len == 5 for r in reqs) and len({r.prompt for r in reqs}) == 5
    assert load_sharegpt(str(path), 2, None)[0].output_len == 5                       # from the reference answer's length
    try:
        for backend in ("vllm", "openai-chat"):
            out = asyncio.run(run_benchmark("http://test", "qwen3", reqs, backend=backend, max_concurrency=3, ignore_eos=True,
                                            transport=httpx.ASGITransport(app=build_app(eng))))
            assert out["completed"] == 8 and out["failed"] == 0 and out["total_output_tokens"] == 8 * 5, out
            assert out["total_input_tokens"] > 0 and out["output_throughput"] > 0 and out["request_throughput"] > 0
            for k in ("mean_ttft_ms", "median_ttft_ms", "p99_ttft_ms", "mean_tpot_ms", "p99_tpot_ms", "mean_itl_ms", "mean_e2el_ms", "p99_e2el_ms"):
                assert k in out and out[k] >= 0
            assert out["mean_e2el_ms"] >= out["mean_ttft_ms"] > 0
            assert "Output token throughput" in format_report(out)
        paced = asyncio.run(run_benchmark("http://test", "qwen3", random_requests(4, 6, 3), request_rate=200.0,
                                          transport=httpx.ASGITransport(app=build_app(eng))))
        assert paced["completed"] == 4 and paced["request_rate"] == 200.0
        bad = asyncio.run(run_benchmark("http://test", "nope", random_requests(2, 300, 3), backend="vllm",
                                        transport=httpx.ASGITransport(app=build_app(eng))))     # prompts longer than max_model_len
        assert bad["completed"] == 0 and bad["failed"] == 2 and "HTTP 400" in bad["errors"][0]
    finally:
        eng.shutdown()
