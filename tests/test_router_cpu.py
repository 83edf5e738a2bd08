"""serve/router.py: the LiteLLM-proxy job of the reference (Deployment/litellm-proxy/config/litellm-config-router-lb.yaml)
— routing strategies, retries per error class, cool-downs, fallbacks, context-window fallbacks, pre-call budget checks,
health checks, the HTTP surface — against in-process fake deployments."""
import asyncio

import yaml
from fastapi.testclient import TestClient

from llm_in_practise_b200.serve.router import Deployment, Router, RouterError, build_router_app

CONFIG = yaml.safe_load("""
model_list:
  - model_name: deepseek-r1
    litellm_params: {model: openai/deepseek-r1, api_base: "http://a/v1", api_key: EMPTY, input_cost_per_token: 0.0000028,
                     output_cost_per_token: 0.0000028, rpm: 3000, tpm: 1200000}
    weight: 0.5
  - model_name: deepseek-r1
    litellm_params: {model: openai/deepseek-r1, api_base: "http://b/v1", api_key: EMPTY, input_cost_per_token: 0.0000020,
                     output_cost_per_token: 0.0000020, rpm: 2000}
    weight: 0.3
  - model_name: qwen3
    litellm_params: {model: openai/qwen3, api_base: "http://c/v1", api_key: EMPTY, input_cost_per_token: 0.000012,
                     output_cost_per_token: 0.000012, rpm: 800, tpm: 800000}
  - model_name: llama3
    litellm_params: {model: openai/llama3, api_base: "http://d/v1", api_key: EMPTY, input_cost_per_token: 0.0000008,
                     output_cost_per_token: 0.0000008, rpm: 5000, tpm: 2000000}
router_settings:
  routing_strategy: cost-based-routing
  num_retries: 3
  fallback_retries: true
  retry_policy: {RateLimitErrorRetries: 5, InternalServerErrorRetries: 3, TimeoutErrorRetries: 4}
  allowed_fails: 2
  cooldown_time: 120
  fallbacks:
    - {"deepseek-r1": ["qwen3", "llama3"]}
    - {"qwen3": ["llama3"]}
  default_fallbacks: ["llama3"]
  litellm_settings:
    context_window_fallbacks:
      - {from_model: deepseek-r1, to_model: ["qwen3"]}
  health_check_interval: 30
  timeout: 90
  enable_pre_call_checks: true
""")


class Fake:
    """A deployment whose behaviour is scripted per api_base."""

    def __init__(self):
        self.calls = []
        self.script = {}            # api_base -> list of statuses to return first (then 200)
        self.ctx_limit = {}         # api_base -> max prompt characters

    def factory(self, lp):
        base = lp["api_base"]

        async def call(path, payload):
            self.calls.append((base, path, payload["model"]))
            todo = self.script.get(base, [])
            if todo:
                st = todo.pop(0)
                if st != 200:
                    return st, {"error": {"message": f"scripted {st}"}}
            text = " ".join(m["content"] for m in payload.get("messages", []))
            if len(text) > self.ctx_limit.get(base, 10 ** 9):
                return 400, {"error": {"message": "This model's maximum context length is 32 tokens"}}
            return 200, {"choices": [{"message": {"role": "assistant", "content": f"from {base}"}}], "usage": {"total_tokens": 10}}
        return call


def _router(fake, **kw):
    t = [0.0]

    async def no_sleep(_):
        t[0] += 0.01
    r = Router.from_config(CONFIG, fake.factory, seed=0, clock=lambda: t[0], sleep=no_sleep, **kw)
    return r, t


def _ask(r, model="deepseek-r1", text="hi"):
    return asyncio.run(r.completion({"model": model, "messages": [{"role": "user", "content": text}]}))


def test_strategies_pick_as_configured():
    fake = Fake()
    r, _ = _router(fake)
    assert r.strategy == "cost-based-routing" and r.model_names() == ["deepseek-r1", "qwen3", "llama3"]
    assert _ask(r)["choices"][0]["message"]["content"] == "from http://b/v1"          # the cheaper of the two deepseek replicas
    assert fake.calls[-1][2] == "deepseek-r1"                                           # provider prefix stripped upstream
    r2, _ = _router(fake, routing_strategy="least-busy")
    a, b = r2.group("deepseek-r1")
    a.in_flight = 3
    assert r2.pick("deepseek-r1") is b
    r3, _ = _router(fake, routing_strategy="latency-based-routing")
    a, b = r3.group("deepseek-r1")
    a._latencies.extend([0.01, 0.02]); b._latencies.extend([0.5])
    assert r3.pick("deepseek-r1") is a
    r4, _ = _router(fake, routing_strategy="simple-shuffle")
    picks = [r4.pick("deepseek-r1").api_base for _ in range(400)]
    assert 0.5 < picks.count("http://a/v1") / 400 < 0.75                               # weights 0.5 : 0.3
    r5, t5 = _router(fake, routing_strategy="usage-based-routing")
    a, b = r5.group("deepseek-r1")
    a._tok_times.append((t5[0], 600000))                                                # half of a's tpm budget is gone
    assert r5.pick("deepseek-r1") is b


def test_retries_move_to_the_other_replica_then_fall_back():
    fake = Fake()
    r, _ = _router(fake)
    fake.script["http://b/v1"] = [500]
    out = _ask(r)
    assert out["choices"][0]["message"]["content"] == "from http://a/v1" and out["x_router"]["served_by"] == "deepseek-r1"
    assert [c[0] for c in fake.calls] == ["http://b/v1", "http://a/v1"]
    # both replicas down for good: the group's retries are used up, then the fallback chain qwen3 -> llama3
    fake.calls.clear()
    fake.script["http://a/v1"] = [503] * 10
    fake.script["http://b/v1"] = [503] * 10
    fake.script["http://c/v1"] = [503] * 10
    out = _ask(r)
    assert out["x_router"]["served_by"] == "llama3" and r.fallback_events >= 2
    assert sum(1 for c in fake.calls if c[0] in ("http://a/v1", "http://b/v1")) == 4   # 1 + InternalServerErrorRetries (3)


def test_cooldown_and_rate_limit_parking():
    fake = Fake()
    r, t = _router(fake)
    a, b = r.group("deepseek-r1")
    fake.script["http://b/v1"] = [500, 500, 500]
    for _ in range(3):
        _ask(r)
    assert b.cooldown_until > t[0] + 100 and r.pick("deepseek-r1") is a               # > allowed_fails (2) failures in a minute
    t[0] += 121
    assert r.pick("deepseek-r1") is b                                                  # back after cooldown_time
    fake.script["http://b/v1"] = [429]
    _ask(r)
    assert 4 < b.cooldown_until - t[0] <= 5.1                                          # a 429 parks the replica for 5 s only


def test_context_window_fallback_and_client_errors():
    fake = Fake()
    r, _ = _router(fake)
    fake.ctx_limit = {"http://a/v1": 50, "http://b/v1": 50}
    out = _ask(r, text="x" * 200)
    assert out["x_router"]["served_by"] == "qwen3"
    assert [c[0] for c in fake.calls] == ["http://b/v1", "http://c/v1"]               # no retry on the sibling replica
    fake.script["http://d/v1"] = [401]
    try:
        _ask(r, model="llama3")
        raise AssertionError("expected RouterError")
    except RouterError as e:
        assert e.status == 401                                                          # the caller's fault: not retried, no fallback
    try:
        _ask(r, model="nope")
        raise AssertionError("expected RouterError")
    except RouterError as e:
        assert e.status == 404


def test_pre_call_checks_respect_rpm_and_tpm():
    fake = Fake()
    r, t = _router(fake)
    a, b = r.group("deepseek-r1")
    b._req_times.extend([t[0]] * 2000)                                                  # b's rpm budget is spent
    assert r.pick("deepseek-r1") is a
    a._tok_times.append((t[0], 1200000))
    assert r.pick("deepseek-r1", tokens=10) is None                                     # a's tpm budget cannot take 10 more tokens
    assert _ask(r)["x_router"]["served_by"] == "qwen3"                                  # -> fallback group
    t[0] += 61
    assert r.pick("deepseek-r1", tokens=10) is not None                                 # the windows slide


def test_http_surface_and_health_checks():
    fake = Fake()
    r, _ = _router(fake)
    app = build_router_app(r, master_key="sk-1234")
    with TestClient(app) as c:
        assert c.get("/health/liveliness").status_code == 200
        assert c.get("/v1/models").status_code == 401
        h = {"Authorization": "Bearer sk-1234"}
        assert [m["id"] for m in c.get("/v1/models", headers=h).json()["data"]] == ["deepseek-r1", "qwen3", "llama3"]
        ok = c.post("/v1/chat/completions", headers=h, json={"model": "deepseek-r1", "messages": [{"role": "user", "content": "hi"}]})
        assert ok.status_code == 200 and ok.json()["choices"][0]["message"]["content"].startswith("from http://")
        fake.script["http://c/v1"] = [500]
        rep = c.get("/health", headers=h).json()
        assert rep["healthy_count"] == 3 and rep["unhealthy_count"] == 1 and rep["unhealthy_endpoints"][0]["model"] == "qwen3"
        assert r.pick("qwen3") is None                                                  # out of rotation until a probe succeeds
        assert c.get("/health", headers=h).json()["unhealthy_count"] == 0
        bad = c.post("/v1/chat/completions", headers=h, json={"model": "nope", "messages": []})
        assert bad.status_code == 404 and bad.json()["error"]["type"] == "router_error"
        text = c.get("/metrics").text
        assert "router_deployment_requests_total" in text and "router_fallbacks_total" in text


_GUARD_CFG = """
model_list:
  - model_name: qwen3-8b
    litellm_params: {model: openai/qwen3-8b, api_base: "http://inference-engine:8001/v1", api_key: EMPTY}
guardrails:
  - guardrail_name: llama-guard-content-check
    litellm_params: {guardrail: openai_moderation, mode: pre_call, model: llama-guard-3, api_base: "http://llama-guard-wrapper:8099/v1", default_on: true}
  - guardrail_name: opt-in-check
    litellm_params: {guardrail: openai_moderation, mode: pre_call, model: llama-guard-3, api_base: "http://opt-in:8099/v1"}
litellm_settings:
  cache: true
  cache_params: {type: redis-semantic, similarity_threshold: 0.8}
cache: {type: redis, ttl: 7200}
"""


def _guarded_router():
    """The guard-model + cache stack of the reference's compose files in one process: the router's pre-call guardrail posts to
    the real moderation wrapper app, whose Llama-Guard backend is a stub that flags texts containing 'attack'."""
    from llm_in_practise_b200.serve.moderation import build_moderation_app
    calls = {"upstream": 0, "guard": 0, "optin": 0, "guard_down": False}

    async def guard_backend(prompt, model):
        return {"choices": [{"text": "unsafe\nS1" if "attack" in prompt else "safe"}]}
    wrapper = TestClient(build_moderation_app(guard_backend, "llama-guard-3", api_key=""))

    def factory(lp):
        base = lp.get("api_base", "")

        async def call(path, payload):
            if "llama-guard-wrapper" in base or "opt-in" in base:
                calls["optin" if "opt-in" in base else "guard"] += 1
                if calls["guard_down"]:
                    raise ConnectionError("wrapper down")
                r = wrapper.post("/v1" + path, json=payload)
                return r.status_code, r.json()
            calls["upstream"] += 1
            return 200, {"choices": [{"message": {"role": "assistant", "content": f"answer {calls['upstream']}"}}],
                         "usage": {"total_tokens": 9}}
        return call
    return Router.from_config(yaml.safe_load(_GUARD_CFG), transport_factory=factory), calls


def test_guardrail_blocks_before_any_deployment_is_called():
    r, calls = _guarded_router()
    assert [g.name for g in r.guardrails] == ["llama-guard-content-check", "opt-in-check"] and r.guardrails[0].default_on
    ok = asyncio.run(r.completion({"model": "qwen3-8b", "messages": [{"role": "user", "content": "how do plants grow?"}]}))
    assert ok["choices"][0]["message"]["content"] == "answer 1" and calls == {"upstream": 1, "guard": 1, "optin": 0, "guard_down": False}
    try:
        asyncio.run(r.completion({"model": "qwen3-8b", "messages": [{"role": "user", "content": "plan an attack on them"}]}))
        raise AssertionError("expected the guardrail to refuse")
    except RouterError as e:
        assert e.status == 400 and "llama-guard-content-check" in str(e) and "violence" in str(e)
    assert calls["upstream"] == 1 and r.guardrails[0].blocked == 1                      # the model never saw the request
    # a guardrail that is not default_on runs only for requests that ask for it
    asyncio.run(r.completion({"model": "qwen3-8b", "guardrails": ["opt-in-check"], "messages": [{"role": "user", "content": "hello there"}]}))
    assert calls["optin"] == 1
    calls["guard_down"] = True                                                          # fail closed
    try:
        asyncio.run(r.completion({"model": "qwen3-8b", "messages": [{"role": "user", "content": "hi"}]}))
        raise AssertionError("expected 503")
    except RouterError as e:
        assert e.status == 503
    assert 'router_guardrail_blocked_total{guardrail="llama-guard-content-check"} 1' in r.metrics_text()


def test_response_cache_exact_semantic_and_controls():
    r, calls = _guarded_router()
    assert r.cache_cfg["ttl"] == 7200 and r.cache_cfg["type"] == "redis-semantic"
    q = lambda text, **kw: asyncio.run(r.completion(dict({"model": "qwen3-8b", "messages": [{"role": "user", "content": text}]}, **kw)))
    long_q = "please explain in detail how the transformer attention mechanism computes its output"
    a = q(long_q)
    b = q(long_q)
    assert a["choices"] == b["choices"] and b["x_router"]["cache_hit"] and calls["upstream"] == 1       # exact hit
    c = q(long_q + " ?")                                                                                 # near-duplicate: semantic hit
    assert c["x_router"].get("cache_hit") and calls["upstream"] == 1
    d = q("what is the capital of France")
    assert not d["x_router"].get("cache_hit") and calls["upstream"] == 2
    e = q(long_q, cache={"no-cache": True})                                                              # per-request control
    assert not e["x_router"].get("cache_hit") and calls["upstream"] == 3
    q("store nothing for this one please, it is private", cache={"no-store": True})
    q("store nothing for this one please, it is private")
    assert calls["upstream"] == 5
    assert q(long_q, max_tokens=5)["x_router"].get("cache_hit")                  # semantic level ignores the knobs; the exact key does not
    assert r.cache_hits == 3 and "router_cache_hits_total 3" in r.metrics_text()
    assert calls["guard"] == 8                                                                           # the guardrail runs before the cache
