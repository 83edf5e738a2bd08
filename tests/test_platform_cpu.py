"""The platform rules of LLM_on_Kubernetes/Inference_Platfrom (KEDA scaling, canary gate, preloader, cache-aware routing)
driven by the metrics text the serving engine exports."""
import asyncio
import math
import os
import random

import pytest

from llm_in_practise_b200.serve import platform as P


class Clock:
    def __init__(self):
        self.t = 1000.0

    def __call__(self):
        return self.t


def _metrics(waiting=0, ttft=(), ok=0, bad=0, mem=None):
    """A replica's /metrics: ``ttft`` = observed first-token times (s), cumulative like the real exporter."""
    bounds = (0.1, 0.25, 0.5, 0.75, 1.0, 2.5, 5.0, 7.5, 10.0)
    out = [f'vllm:num_requests_waiting{{model_name="m"}} {waiting}']
    for b in bounds:
        out.append(f'vllm:time_to_first_token_seconds_bucket{{model_name="m",le="{b}"}} {sum(1 for x in ttft if x <= b)}')
    out.append(f'vllm:time_to_first_token_seconds_bucket{{model_name="m",le="+Inf"}} {len(ttft)}')
    out.append(f'vllm:request_success_total{{model_name="m",finished_reason="stop"}} {ok}')
    out.append(f'vllm:request_failure_total{{model_name="m"}} {bad}')
    if mem is not None:
        out += [f'nvidia_gpu_memory_used_bytes{{gpu="0"}} {mem * 100}', 'nvidia_gpu_memory_total_bytes{gpu="0"} 100']
    return "\n".join(out) + "\n"


def test_prometheus_text_and_quantile():
    s = P.parse_prometheus_text('# HELP x y\nfoo{a="1",b="q\\"z"} 3\nfoo{a="2"} 4.5\nbar 7 1700000000\nbad line here\n')
    assert P.metric_sum(s, "foo") == 7.5 and P.metric_sum(s, "foo", a="2") == 4.5 and P.metric_sum(s, "bar") == 7
    assert P.metric_sum(s, "nope") == 0
    # PromQL's interpolation: 100 observations, 90 at <= 0.5 s, 10 between 0.5 and 1.0 -> the 99th percentile is 0.95 s
    q = P.histogram_quantile(0.99, [(0.5, 90), (1.0, 100), (float("inf"), 100)])
    assert abs(q - 0.95) < 1e-9
    assert P.histogram_quantile(0.5, [(0.5, 90), (1.0, 100), (float("inf"), 100)]) == pytest.approx(0.5 * 50 / 90)
    assert P.histogram_quantile(0.99, [(1.0, 0), (float("inf"), 5)]) == 1.0           # everything beyond the last finite bound
    assert math.isnan(P.histogram_quantile(0.99, [(1.0, 0), (float("inf"), 0)]))


def test_metric_history_rates_only_the_window():
    clk = Clock()
    h = P.MetricHistory(120.0, clk)
    slow = [3.0] * 50
    h.add([_metrics(ttft=slow, ok=50)])
    assert math.isnan(h.quantile(0.99, "vllm:time_to_first_token_seconds_bucket"))     # one scrape: no rate yet
    clk.t += 60
    h.add([_metrics(ttft=slow + [0.2] * 100, ok=150)])
    # only the 100 fast requests fall inside the window: the old slow ones do not count
    assert h.quantile(0.99, "vllm:time_to_first_token_seconds_bucket") < 0.25
    assert h.rate("vllm:request_success_total") == pytest.approx(100 / 60)
    clk.t += 60
    h.add([_metrics(ttft=[0.2] * 10, ok=10)])                   # the replica restarted: counters start again from zero
    assert h.rate("vllm:request_success_total") == pytest.approx(10 / 120)
    for _ in range(5):
        clk.t += 60
        h.add([_metrics(ttft=[0.2] * 10, ok=10)])
    assert len(h._scrapes) <= 4                                 # old scrapes are dropped, one stays at the far edge


def test_keda_scaler_follows_the_scaled_object():
    import yaml
    clk = Clock()
    spec = yaml.safe_load(open(os.path.join(os.path.dirname(__file__), "..", "configs", "keda_scaledobject.yaml")))
    k = P.KedaScaler.from_scaled_object(spec, clock=clk)
    assert (k.min_replicas, k.max_replicas, k.failure_threshold, k.fallback_replicas) == (1, 2, 3, 1)
    assert (k.up.stabilization_window_s, k.up.pods, k.up.period_s) == (30.0, 1, 10.0)
    assert (k.down.stabilization_window_s, k.down.pods, k.down.period_s) == (120.0, 1, 30.0)
    assert [t.threshold for t in k.triggers] == [2.0, 600.0]
    n = 1
    state = {"waiting": 0, "ttft": []}
    scrape = lambda: [_metrics(waiting=state["waiting"], ttft=state["ttft"])]
    n = k.tick(n, scrape)
    assert n == 1
    # queue depth 5 against a threshold of 2 -> ceil(5 / 2) = 3 wanted, capped at maxReplicaCount 2; the 30 s scale-up
    # window still remembers the quiet poll, so nothing moves until the spike has lasted that long
    state["waiting"] = 5
    for _ in range(2):
        clk.t += 15
        n = k.tick(n, scrape)
        assert n == 1
    clk.t += 15
    n = k.tick(n, scrape)
    assert n == 2 and k.last_values["vllm-num-requests-waiting"] == 5
    # the queue drains: scale-down waits for its 120 s window
    state["waiting"] = 0
    for _ in range(8):
        clk.t += 15
        n = k.tick(n, scrape)
        assert n == 2
    clk.t += 15
    n = k.tick(n, scrape)
    assert n == 1
    # the second trigger alone: P99 TTFT of 900 ms against 600 -> two replicas
    state["ttft"] = [0.9] * 200
    for _ in range(4):
        clk.t += 15
        n = k.tick(n, scrape)
    assert n == 2 and 600 < k.last_values["vllm-p99-ttft-ms"] <= 1000
    # Prometheus goes away: the third failed poll in a row falls back to one replica, a good poll ends the fallback
    def broken():
        raise ConnectionError("prometheus unreachable")
    assert k.tick(n, broken) == 2 and k.tick(n, broken) == 2 and not k.in_fallback
    assert k.tick(n, broken) == 1 and k.in_fallback
    clk.t += 15
    k.tick(1, scrape)
    assert not k.in_fallback and k.failures == 0


def test_keda_policy_limits_pods_per_period():
    clk = Clock()
    k = P.KedaScaler([P.queue_depth_trigger(2.0)], 1, 8, P.ScalePolicy(0.0, 1, 60.0), P.ScalePolicy(0.0, 2, 60.0), clock=clk)
    scrape = lambda: [_metrics(waiting=20)]
    n = k.tick(1, scrape)
    assert n == 2                                      # 10 wanted, one pod per minute allowed
    clk.t += 15
    assert k.tick(n, scrape) == 2
    clk.t += 50
    assert k.tick(n, scrape) == 3
    idle = lambda: [_metrics(waiting=0)]
    clk.t += 61
    assert k.tick(8, idle) == 6                        # two pods per minute on the way down
    clk.t += 10
    assert k.tick(6, idle) == 6


def _drive(rollout, clk, scrape, step_s=60, limit=200):
    phases = []
    for _ in range(limit):
        ph = rollout.advance(scrape)
        phases.append((rollout.index, rollout.weight, ph))
        if ph != "Progressing":
            break
        clk.t += step_s
    return phases


def test_canary_rollout_promotes_a_healthy_version():
    clk = Clock()
    ro = P.CanaryRollout(clock=clk)
    state = {"n": 0}

    def scrape():                                       # every scrape: 30 more requests, all fast, none failed
        state["n"] += 30
        return [_metrics(ttft=[0.3] * state["n"], ok=state["n"], bad=0, mem=0.7)]
    phases = _drive(ro, clk, scrape)
    assert phases[-1][1:] == (100, "Healthy")
    weights = [w for _, w, _ in phases]
    assert 25 in weights and 50 in weights and weights == sorted(weights)
    assert len(ro.runs) == 2 and all(r.phase == "Successful" for r in ro.runs)
    assert all(len(v) == 5 for v in ro.runs[0].results.values())          # count: 5 measurements of each metric
    # the first pause lasts 3 minutes: weight stays 25 for at least that long before the analysis starts
    t25 = sum(1 for i, w, _ in phases if w == 25 and i == 1)
    assert t25 >= 3


@pytest.mark.parametrize("fault,reason", [("slow", "p99-ttft"), ("errors", "error-rate"), ("memory", "gpu-memory-saturation")])
def test_canary_rollout_aborts_on_each_gate(fault, reason):
    clk = Clock()
    ro = P.CanaryRollout(clock=clk)
    state = {"n": 0}

    def scrape():
        state["n"] += 100
        n = state["n"]
        ttft = [6.0 if fault == "slow" else 0.3] * n
        return [_metrics(ttft=ttft, ok=n, bad=(n // 20 if fault == "errors" else 0), mem=(0.95 if fault == "memory" else 0.5))]
    phases = _drive(ro, clk, scrape)
    assert phases[-1][1:] == (0, "Degraded")                              # traffic is back on the stable version
    run = ro.runs[-1]
    assert run.phase == "Failed" and any(r == "Failed" for _, r in run.results[reason])
    rng = random.Random(0)
    assert all(ro.pick(["s1", "s2"], ["c1"], rng) != "c1" for _ in range(50))


def test_canary_inconclusive_pauses_and_traffic_split():
    clk = Clock()
    ro = P.CanaryRollout(clock=clk)
    state = {"n": 0}

    def scrape():                                       # P99 of 3 s: neither < 2 s nor >= 5 s
        state["n"] += 100
        return [_metrics(ttft=[3.0] * state["n"], ok=state["n"], mem=0.5)]
    phases = _drive(ro, clk, scrape)
    assert phases[-1][1:] == (25, "Paused") and ro.runs[-1].phase == "Inconclusive"
    rng = random.Random(1)
    share = sum(ro.pick(["s"], ["c"], rng) == "c" for _ in range(4000)) / 4000
    assert 0.22 < share < 0.28                          # canary-weight 25
    ro.promote()
    assert ro.advance(scrape) == "Progressing" and ro.weight == 50


def test_preloader_marker_mirror_and_required_files(tmp_path):
    src, dst = tmp_path / "store" / "Qwen3-8B", tmp_path / "cache" / "qwen3-8b"
    (src / "sub").mkdir(parents=True)
    (src / "config.json").write_text("{}")
    (src / "tokenizer.json").write_text("{}")
    (src / "model-00001.safetensors").write_bytes(b"x" * 1000)
    (src / "sub" / "extra.txt").write_text("hi")
    dst.mkdir(parents=True)
    (dst / "stale.bin").write_bytes(b"old")                                # --remove: not in the source any more
    assert not P.wait_until_ready(str(dst), timeout_s=0.0, sleep=lambda s: None)
    rep = P.preload_model(str(src), str(dst))
    assert rep["copied"] == 4 and rep["removed"] == 1 and not rep["skipped"] and rep["bytes"] >= 1006
    assert (dst / ".ready").exists() and (dst / "sub" / "extra.txt").read_text() == "hi" and not (dst / "stale.bin").exists()
    assert P.wait_until_ready(str(dst), timeout_s=0.0)
    (src / "config.json").write_text('{"changed": true}')
    assert P.preload_model(str(src), str(dst))["skipped"]                   # the marker short-circuits the second run
    assert (dst / "config.json").read_text() == "{}"
    os.remove(dst / "tokenizer.json")
    with pytest.raises(FileNotFoundError):
        P.preload_model(str(src), str(dst))                                # marker present but a key file is gone
    with pytest.raises(FileNotFoundError):
        P.preload_model(str(tmp_path / "nowhere"), str(tmp_path / "cache2"))


def test_prefix_index_longest_prefix_and_lru():
    ix = P.PrefixIndex(chunk=4, capacity=8)
    ix.insert("a", "system prompt one. question A")
    ix.insert("b", "system prompt two. question B")
    assert ix.match("a", "system prompt one. question C") == 28            # every whole 4-character chunk before the last letter
    assert ix.match("b", "system prompt one. question C") == 12            # "system prompt" only
    assert ix.match("c", "system prompt one.") == 0
    ix.insert("a", "z" * 32)                                               # 8 new chunks push the old ones out
    assert ix.match("a", "system prompt one. question A") == 0
    ix.forget("b")
    assert ix.match("b", "system prompt two.") == 0
    assert P.prompt_text({"messages": [{"role": "system", "content": "S"}, {"role": "user", "content": "U"}]}) == "system:S\nuser:U"
    assert P.prompt_text({"prompt": ["a", "b"]}) == "a\nb"


def _router(strategy, **kw):
    from llm_in_practise_b200.serve.router import Deployment, Router
    seen = {"r0": [], "r1": [], "r2": []}

    def transport(tag):
        async def call(path, payload):
            seen[tag].append(payload["messages"][0]["content"][:8])
            return 200, {"choices": [{"message": {"content": tag}}], "usage": {"total_tokens": 5}}
        return call
    deps = [Deployment(model_name="qwen", upstream_model="qwen", transport=transport(t), api_base=f"http://{t}/v1") for t in seen]
    return Router(deps, routing_strategy=strategy, seed=0, **kw), seen, deps


def test_router_prefix_aware_keeps_a_conversation_on_one_replica():
    r, seen, deps = _router("prefix-aware", prefix_chunk=16)
    sys_a, sys_b = "A" * 200, "B" * 200
    body = lambda s, q: {"model": "qwen", "messages": [{"role": "system", "content": s}, {"role": "user", "content": q}]}

    async def go():
        for i in range(6):
            await r.completion(body(sys_a, f"question {i}"))
            await r.completion(body(sys_b, f"question {i}"))
    asyncio.run(go())
    homes_a = [t for t, v in seen.items() if "AAAAAAAA" in v]
    homes_b = [t for t, v in seen.items() if "BBBBBBBB" in v]
    assert len(homes_a) == 1 and len(homes_b) == 1                         # each system prompt found its replica and stayed
    assert r.prefix_routed >= 10 and "router_prefix_routed_total" in r.metrics_text()


def test_router_load_aware_prefix_weighs_load_against_the_cache():
    r, seen, deps = _router("load-aware-prefix", prefix_chunk=16, prefix_weight=0.5)
    text = {"model": "qwen", "messages": [{"role": "user", "content": "C" * 300}]}

    async def metrics(waiting, usage):
        return (f'vllm:num_requests_waiting{{m="x"}} {waiting}\nvllm:gpu_cache_usage_perc{{m="x"}} {usage}\n'
                'vllm:time_to_first_token_seconds_sum 2.0\nvllm:time_to_first_token_seconds_count 10\n')
    deps[0].metrics_fn = lambda: metrics(0, 0.1)
    deps[1].metrics_fn = lambda: metrics(0, 0.1)
    deps[2].metrics_fn = lambda: metrics(0, 0.1)

    async def go():
        await r.refresh_load()
        assert deps[0].load["ttft_ms"] == pytest.approx(200.0)
        await r.completion(text)
        home = next(t for t, v in seen.items() if v)
        await r.completion(text)
        assert len(seen[home]) == 2                                        # equal load: the cached prefix decides
        # the home replica becomes the busiest on every axis: 0.4 + 0.3 (+ 0.2 ttft tie) outweighs the 0.5 prefix bonus
        for d in deps:
            busy = d.api_base == f"http://{home}/v1"
            d.metrics_fn = (lambda b=busy: metrics(9 if b else 0, 0.95 if b else 0.05))
        await r.refresh_load()
        await r.completion(text)
        assert len(seen[home]) == 2
    asyncio.run(go())


def test_rules_read_the_engines_own_metrics():
    """The exporter of serve/openai_server.py and the parser here agree: queue depth, TTFT histogram, success / failure
    counters of a live (CPU) engine drive the scaler and the canary gate."""
    import torch
    from llm_in_practise_b200.data.tokenizers import ByteTokenizer
    from llm_in_practise_b200.inference import GenerationConfig
    from llm_in_practise_b200.models.qwen3 import QWEN3_PRESETS, Qwen3ForCausalLM
    from llm_in_practise_b200.serve import EngineArgs, LLMEngine
    from llm_in_practise_b200.serve.openai_server import render_metrics
    torch.manual_seed(0)
    m = Qwen3ForCausalLM(QWEN3_PRESETS["qwen3-tiny"], dtype=torch.float32, device="cpu")
    eng = LLMEngine(m, ByteTokenizer(), EngineArgs(max_num_seqs=2, max_model_len=64, warmup=False), model_name="tiny")
    clk = Clock()
    k = P.KedaScaler(max_replicas=4, scale_up=P.ScalePolicy(0.0, 4, 15.0), clock=clk)
    hist = P.MetricHistory(120.0, clk)
    hist.add([render_metrics(eng)])
    for _ in range(7):
        eng.submit([1, 2, 3, 4], GenerationConfig(max_new_tokens=2, do_sample=False))
    text = render_metrics(eng)
    assert P.metric_sum(P.parse_prometheus_text(text), "vllm:num_requests_waiting") == 7 == eng.num_waiting
    assert k.tick(1, lambda: [text]) == 4                                  # ceil(7 / 2) = 4 replicas wanted
    while eng.num_waiting or eng.num_running:
        eng.step()
    clk.t += 60
    hist.add([render_metrics(eng)])
    assert hist.rate("vllm:request_success_total") == pytest.approx(7 / 60)
    assert hist.rate("vllm:request_failure_total") == 0
    p99 = hist.quantile(0.99, "vllm:time_to_first_token_seconds_bucket")
    assert 0 < p99 <= 10.0
    run = P.CanaryAnalysis([mt for mt in P.smoke_template() if mt.name != "gpu-memory-saturation"])
    assert run.measure(hist) == "Running" and run.results["error-rate"][0][1] == "Successful"


def test_keda_replica_set_starts_and_drains_engines():
    import torch
    from llm_in_practise_b200.data.tokenizers import ByteTokenizer
    from llm_in_practise_b200.inference import GenerationConfig
    from llm_in_practise_b200.models.qwen3 import QWEN3_PRESETS, Qwen3ForCausalLM
    from llm_in_practise_b200.serve import EngineArgs, LLMEngine
    torch.manual_seed(0)
    m = Qwen3ForCausalLM(QWEN3_PRESETS["qwen3-tiny"], dtype=torch.float32, device="cpu")
    made = []

    def factory():
        made.append(LLMEngine(m, ByteTokenizer(), EngineArgs(max_num_seqs=2, max_model_len=64, warmup=False), model_name=f"r{len(made)}"))
        return made[-1]
    clk = Clock()
    scaler = P.KedaScaler([P.queue_depth_trigger(2.0)], 1, 3, P.ScalePolicy(0.0, 4, 15.0), P.ScalePolicy(30.0, 1, 15.0), clock=clk)
    rs = P.keda_replica_set(factory, scaler, clk)
    assert len(rs.engines) == 1
    cfg = GenerationConfig(max_new_tokens=2, do_sample=False)
    reqs = [rs.engines[0].submit([1, 2, 3], cfg) for _ in range(6)]          # six waiting -> ceil(6 / 2) = 3 replicas
    assert rs.tick() == 3 and len(made) == 3
    assert rs.tick() == 3                                                    # inside the polling interval: no new decision
    reqs.append(rs.submit([4, 5, 6], cfg))                                   # routed to one of the new, idle replicas
    assert made[0].num_waiting == 6 and made[1].num_waiting + made[2].num_waiting == 1
    for e in made:
        while e.num_waiting or e.num_running:
            e.step()
    assert all(r.done.is_set() for r in reqs)
    for _ in range(6):                                                       # the lull outlasts the 30 s scale-down window
        clk.t += 15
        rs.tick()
    assert len(rs.engines) == 1 and [ev[1] for ev in rs.events].count("stopped") == 2


def test_queue_time_metric_feeds_the_load_score():
    """``vllm:request_queue_time_seconds`` (the fourth figure llm-d weighs): exported by the engine, read by ``refresh_load``."""
    import time
    import torch
    from llm_in_practise_b200.data.tokenizers import ByteTokenizer
    from llm_in_practise_b200.inference import GenerationConfig
    from llm_in_practise_b200.models.qwen3 import QWEN3_PRESETS, Qwen3ForCausalLM
    from llm_in_practise_b200.serve import EngineArgs, LLMEngine
    from llm_in_practise_b200.serve.openai_server import render_metrics
    from llm_in_practise_b200.serve.router import Deployment, Router
    torch.manual_seed(0)
    m = Qwen3ForCausalLM(QWEN3_PRESETS["qwen3-tiny"], dtype=torch.float32, device="cpu")
    eng = LLMEngine(m, ByteTokenizer(), EngineArgs(max_num_seqs=1, max_model_len=64, warmup=False), model_name="tiny")
    reqs = [eng.submit([1, 2, 3], GenerationConfig(max_new_tokens=2, do_sample=False)) for _ in range(3)]
    time.sleep(0.05)                                              # everybody waits at least this long
    while not all(r.done.is_set() for r in reqs):
        eng.step()
    s = P.parse_prometheus_text(render_metrics(eng))
    assert P.metric_sum(s, "vllm:request_queue_time_seconds_count") == 3
    assert P.metric_sum(s, "vllm:request_queue_time_seconds_sum") >= 3 * 0.05

    async def fetch():
        return render_metrics(eng)

    async def noop(path, payload):
        return 200, {}
    d = Deployment(model_name="tiny", upstream_model="tiny", transport=noop, metrics_fn=fetch)
    r = Router([d], routing_strategy="load-aware-prefix")
    asyncio.run(r.refresh_load())
    assert d.load["queue_time_ms"] >= 50.0 and d.load["ttft_ms"] > 0


def test_canary_replica_set_promotes_and_rolls_back_real_engines():
    import torch
    from llm_in_practise_b200.data.tokenizers import ByteTokenizer
    from llm_in_practise_b200.inference import GenerationConfig
    from llm_in_practise_b200.models.qwen3 import QWEN3_PRESETS, Qwen3ForCausalLM
    from llm_in_practise_b200.serve import EngineArgs, LLMEngine
    torch.manual_seed(0)
    m = Qwen3ForCausalLM(QWEN3_PRESETS["qwen3-tiny"], dtype=torch.float32, device="cpu")
    mk = lambda n: LLMEngine(m, ByteTokenizer(), EngineArgs(max_num_seqs=2, max_model_len=64, warmup=False), model_name=n)
    cfg = GenerationConfig(max_new_tokens=2, do_sample=False)
    gate = [mt for mt in P.smoke_template() if mt.name != "gpu-memory-saturation"]          # no GPU exporter on this machine
    for t in gate:
        t.count = 2
    steps = ({"setWeight": 50}, {"pause": {"duration": "1m"}}, {"analysis": {}}, {"setWeight": 100})

    def run(fail: bool):
        clk = Clock()
        stable, canary = mk("v1"), mk("v2")
        rs = P.CanaryReplicaSet([stable], [canary], P.CanaryRollout(steps, template=lambda: gate, clock=clk), seed=0)
        phase = "Progressing"
        for _ in range(12):
            reqs = [rs.submit([1, 2, 3], cfg) for _ in range(8)]
            for e in (stable, canary):
                while e.num_waiting or e.num_running:
                    e.step()
            if fail:                                            # the new version errors on what it is given
                with canary.metrics.lock:
                    canary.metrics.requests_finished["error"] = canary.metrics.requests_finished.get("error", 0) + 4
            phase = rs.tick()
            if phase != "Progressing":
                break
            clk.t += 60
        return rs, stable, canary, phase
    rs, stable, canary, phase = run(fail=False)
    assert phase == "Healthy" and rs.stable == [canary] and rs.canary == []
    assert rs.sent["canary"] > 0 and rs.sent["stable"] > 0                    # traffic was split while the rollout ran
    rs.tick()
    assert rs.draining == []                                                  # the old version drained and was shut down
    rs, stable, canary, phase = run(fail=True)
    assert phase == "Degraded" and rs.stable == [stable] and rs.canary == []
    n0 = rs.sent["stable"]
    for _ in range(5):
        rs.submit([1, 2, 3], cfg)
    assert rs.sent["stable"] == n0 + 5                                         # everything is back on the stable version
