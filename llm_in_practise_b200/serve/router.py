"""Model router / load balancer in front of several OpenAI-compatible deployments.

The reference runs a LiteLLM proxy for this and configures it in YAML (Deployment/litellm-proxy/config/
litellm-config-router-lb.yaml:1-93, described in Deployment/litellm-proxy/README.md:300-340): a ``model_list`` whose
entries share a ``model_name`` form a load-balanced group (``api_base``, ``rpm`` / ``tpm`` budgets, per-token cost,
``weight``), and ``router_settings`` choose the strategy (``cost-based-routing``, ``least-busy``, ``simple-shuffle``,
``latency-based-routing``, ``usage-based-routing``), retries per error class, cool-downs (``allowed_fails`` failures
inside a minute park a deployment for ``cooldown_time`` seconds; a 429 parks it for 5 s), ``fallbacks`` between groups,
``default_fallbacks``, ``context_window_fallbacks``, ``enable_pre_call_checks`` (rpm / tpm budgets) and periodic health
checks. This module reads the same YAML shape and does the same job in process:

* ``Router`` holds the deployments and their sliding-window counters and decides, ``await router.completion(body)``
  forwards one request with the retry / fallback policy;
* ``build_router_app`` wraps it in the HTTP surface the compose files probe (``/health/liveliness``, ``/health``,
  ``/v1/models``, ``/v1/chat/completions``, ``/v1/completions``, ``/metrics``).

``guardrails:`` entries (litellm-config-with-guard-model.yaml:18-26) run a pre-call moderation check against the
Llama-Guard wrapper, and ``litellm_settings.cache`` (litellm-config-with-cache.yaml:34-45, the ``redis-semantic`` variant
of litellm-config-redis-semantic-cache.yaml:55-60) answers repeated or near-duplicate requests from a response cache.
Two more strategies come from the Kubernetes routers of the reference: ``prefix-aware`` and ``load-aware-prefix``
(``serve/platform.py``).

A deployment's transport is any ``async (path, payload) -> (status, json)``: ``http_transport(api_base)`` for a real
endpoint, or an in-process engine / app in tests. No Redis: counters are per process (the reference shares them through
Redis only for multi-pod proxies).
"""
from __future__ import annotations

import asyncio
import random
import time
from collections import deque
from dataclasses import dataclass, field
from typing import Any, Awaitable, Callable, Deque, Dict, List, Optional, Sequence, Tuple

from fastapi import FastAPI, HTTPException, Request
from fastapi.responses import JSONResponse, PlainTextResponse

Transport = Callable[[str, Dict[str, Any]], Awaitable[Tuple[int, Dict[str, Any]]]]

STRATEGIES = ("simple-shuffle", "least-busy", "cost-based-routing", "latency-based-routing", "usage-based-routing",
              "prefix-aware", "load-aware-prefix")

# llm-d's load score (LLM_on_Kubernetes/Inference_Platfrom/08-LLM-Router/llm-d/llm-d-config.yaml:19-23)
LOAD_WEIGHTS = {"pending_requests": 0.4, "gpu_cache_usage": 0.3, "ttft_ms": 0.2, "queue_time_ms": 0.1}


def http_transport(api_base: str, api_key: Optional[str] = None, timeout: float = 90.0) -> Transport:
    import httpx

    async def call(path: str, payload: Dict[str, Any]) -> Tuple[int, Dict[str, Any]]:
        headers = {"Authorization": f"Bearer {api_key}"} if api_key and api_key != "EMPTY" else {}
        async with httpx.AsyncClient(timeout=timeout) as c:
            r = await c.post(api_base.rstrip("/") + path, json=payload, headers=headers)
        try:
            return r.status_code, r.json()
        except Exception:
            return r.status_code, {"error": {"message": r.text[:500]}}
    return call


def http_metrics(api_base: str, timeout: float = 3.0) -> Callable[[], Awaitable[str]]:
    """``GET <server root>/metrics`` of an OpenAI-compatible endpoint whose ``api_base`` ends in ``/v1``."""
    root = api_base.rstrip("/")
    root = root[:-3] if root.endswith("/v1") else root

    async def fetch() -> str:
        import urllib.request
        loop = asyncio.get_running_loop()
        return await loop.run_in_executor(None, lambda: urllib.request.urlopen(root + "/metrics", timeout=timeout).read().decode())
    return fetch


@dataclass
class Deployment:
    model_name: str                      # the group ("model" a client asks for)
    upstream_model: str                  # the name sent upstream (litellm_params.model without its provider prefix)
    transport: Transport
    api_base: str = ""
    rpm: Optional[int] = None
    tpm: Optional[int] = None
    input_cost: float = 0.0
    output_cost: float = 0.0
    weight: float = 1.0
    max_input_tokens: Optional[int] = None
    # runtime state
    in_flight: int = 0
    cooldown_until: float = 0.0
    healthy: bool = True
    _req_times: Deque[float] = field(default_factory=deque)
    _tok_times: Deque[Tuple[float, int]] = field(default_factory=deque)
    _fail_times: Deque[float] = field(default_factory=deque)
    _latencies: Deque[float] = field(default_factory=lambda: deque(maxlen=32))
    requests: int = 0
    failures: int = 0
    metrics_fn: Optional[Callable[[], Awaitable[str]]] = None      # GET of the replica's /metrics (load-aware routing)
    load: Dict[str, float] = field(default_factory=dict)           # last scraped load figures, keys of LOAD_WEIGHTS

    def _trim(self, now: float) -> None:
        while self._req_times and now - self._req_times[0] > 60.0:
            self._req_times.popleft()
        while self._tok_times and now - self._tok_times[0][0] > 60.0:
            self._tok_times.popleft()
        while self._fail_times and now - self._fail_times[0] > 60.0:
            self._fail_times.popleft()

    def rpm_used(self, now: float) -> int:
        self._trim(now)
        return len(self._req_times)

    def tpm_used(self, now: float) -> int:
        self._trim(now)
        return sum(n for _, n in self._tok_times)

    def avg_latency(self) -> float:
        return sum(self._latencies) / len(self._latencies) if self._latencies else 0.0

    def cost_per_token(self) -> float:
        return self.input_cost + self.output_cost


class RouterError(Exception):
    def __init__(self, status: int, message: str):
        super().__init__(message)
        self.status, self.message = status, message


@dataclass
class Guardrail:
    """One ``guardrails:`` entry (Deployment/litellm-proxy/config/litellm-config-with-guard-model.yaml:18-26):
    ``guardrail: openai_moderation`` in ``mode: pre_call`` posts the request's text to ``<api_base>/moderations`` (the
    Llama-Guard wrapper, ``serve/moderation.py``) before any deployment is called; a flagged text is refused with HTTP 400.
    ``default_on: false`` guardrails only run for requests that name them in their ``guardrails`` list."""
    name: str
    transport: Transport
    model: str = "omni-moderation-latest"
    mode: str = "pre_call"
    default_on: bool = False
    checks: int = 0
    blocked: int = 0


def _request_text(body: Dict[str, Any]) -> str:
    if "messages" in body:
        parts = []
        for m in body.get("messages") or []:
            c = m.get("content", "") if isinstance(m, dict) else ""
            if isinstance(c, list):                                      # OpenAI content parts
                c = " ".join(str(p.get("text", "")) for p in c if isinstance(p, dict))
            parts.append(str(c))
        return "\n".join(parts)
    p = body.get("prompt", "")
    return p if isinstance(p, str) else "\n".join(str(x) for x in p)


def _estimate_tokens(body: Dict[str, Any]) -> int:
    """Cheap prompt-size estimate for the pre-call checks (4 characters per token, what LiteLLM falls back to without a
    tokenizer), plus the requested completion budget."""
    text = body.get("prompt")
    if text is None:
        text = " ".join(str(m.get("content", "")) for m in body.get("messages", []) if isinstance(m, dict))
    if isinstance(text, list):
        text = " ".join(map(str, text))
    return len(str(text)) // 4 + 1


class Router:
    def __init__(self, deployments: Sequence[Deployment], *, routing_strategy: str = "simple-shuffle", num_retries: int = 2,
                 retry_policy: Optional[Dict[str, int]] = None, allowed_fails: int = 3, cooldown_time: float = 60.0,
                 fallbacks: Optional[Sequence[Dict[str, List[str]]]] = None, default_fallbacks: Optional[List[str]] = None,
                 context_window_fallbacks: Optional[Sequence[Dict[str, Any]]] = None, enable_pre_call_checks: bool = False,
                 fallback_retries: bool = True, rate_limit_cooldown: float = 5.0, seed: Optional[int] = None,
                 load_weights: Optional[Dict[str, float]] = None, prefix_weight: float = 0.5, prefix_chunk: int = 128,
                 guardrails: Optional[Sequence[Guardrail]] = None, cache: Optional[Dict[str, Any]] = None,
                 clock: Callable[[], float] = time.monotonic, sleep: Callable[[float], Awaitable[None]] = asyncio.sleep):
        if routing_strategy not in STRATEGIES:
            raise ValueError(f"routing_strategy must be one of {STRATEGIES}")
        self.deployments = list(deployments)
        self.strategy = routing_strategy
        self.num_retries = num_retries
        self.retry_policy = dict(retry_policy or {})
        self.allowed_fails, self.cooldown_time, self.rate_limit_cooldown = allowed_fails, cooldown_time, rate_limit_cooldown
        self.fallbacks: Dict[str, List[str]] = {}
        for item in fallbacks or []:
            self.fallbacks.update({k: list(v) for k, v in item.items()})
        self.default_fallbacks = list(default_fallbacks or [])
        self.context_fallbacks: Dict[str, List[str]] = {}
        for item in context_window_fallbacks or []:
            if "from_model" in item:
                self.context_fallbacks[item["from_model"]] = list(item.get("to_model", []))
            else:
                self.context_fallbacks.update({k: list(v) for k, v in item.items()})
        self.pre_call_checks = enable_pre_call_checks
        self.fallback_retries = fallback_retries
        self._rng = random.Random(seed)
        self._clock, self._sleep = clock, sleep
        self.fallback_events = 0
        # the two cache-aware strategies of the reference's Kubernetes routers (vLLM production-stack ``prefixaware``,
        # llm-d ``load_aware_prefix``): remember which deployment saw which prompt prefix
        from .platform import PrefixIndex
        self.prefix_index = PrefixIndex(prefix_chunk)
        self.load_weights = dict(load_weights or LOAD_WEIGHTS)
        self.prefix_weight = prefix_weight
        self.prefix_routed = 0
        self.guardrails = list(guardrails or [])
        # response cache (``litellm_settings.cache: true``): ``type: redis`` -> exact match on the request, ``redis-semantic``
        # -> nearest stored prompt above ``similarity_threshold`` as well; in process, no Redis
        self.cache_cfg = dict(cache) if cache else None
        self.cache_hits = self.cache_misses = 0
        if self.cache_cfg is not None:
            from .cache_gateway import ExactCache, SemanticCache
            self._exact = ExactCache(int(self.cache_cfg.get("max_entries", 4096)))
            self._semantic = None
            if str(self.cache_cfg.get("type", "redis")).endswith("semantic"):
                self._embed_dim = int(self.cache_cfg.get("embedding_dim", 1024))
                self._semantic = SemanticCache(self._embed_dim, int(self.cache_cfg.get("max_entries", 4096)))

    # ------------------------------------------------------------------ construction from the reference's YAML shape
    @classmethod
    def from_config(cls, cfg: Dict[str, Any], transport_factory: Optional[Callable[[Dict[str, Any]], Transport]] = None,
                    **overrides) -> "Router":
        deps = []
        for entry in cfg.get("model_list", []):
            lp = entry.get("litellm_params", {})
            factory = transport_factory or (lambda p: http_transport(p.get("api_base", ""), p.get("api_key"),
                                                                     float(cfg.get("router_settings", {}).get("timeout", 90))))
            up = str(lp.get("model", entry["model_name"]))
            deps.append(Deployment(model_name=entry["model_name"], upstream_model=up.split("/", 1)[-1], transport=factory(lp),
                                   api_base=lp.get("api_base", ""), rpm=lp.get("rpm"), tpm=lp.get("tpm"),
                                   input_cost=float(lp.get("input_cost_per_token", 0.0)),
                                   output_cost=float(lp.get("output_cost_per_token", 0.0)),
                                   weight=float(entry.get("weight", lp.get("weight", 1.0))),
                                   max_input_tokens=(entry.get("model_info") or {}).get("max_input_tokens"),
                                   metrics_fn=(http_metrics(lp["api_base"]) if (transport_factory is None and lp.get("api_base")) else None)))
        rs = dict(cfg.get("router_settings", {}))
        ls = rs.pop("litellm_settings", None) or cfg.get("litellm_settings", {}) or {}
        mk = transport_factory or (lambda p: http_transport(p.get("api_base", ""), p.get("api_key")))
        guards = []
        for g in cfg.get("guardrails") or []:
            gp = g.get("litellm_params", {})
            if gp.get("guardrail") != "openai_moderation":
                raise ValueError(f"guardrail type {gp.get('guardrail')!r} is not supported (openai_moderation is)")
            guards.append(Guardrail(g.get("guardrail_name", "guardrail"), mk(gp), gp.get("model", "omni-moderation-latest"),
                                    gp.get("mode", "pre_call"), bool(gp.get("default_on", False))))
        cache = None
        if ls.get("cache"):
            cache = dict(ls.get("cache_params") or {})
            cache.setdefault("ttl", (cfg.get("cache") or {}).get("ttl", cache.get("ttl", 3600)))
        kw = dict(routing_strategy=rs.get("routing_strategy", "simple-shuffle"), num_retries=int(rs.get("num_retries", 2)),
                  retry_policy=rs.get("retry_policy"), allowed_fails=int(rs.get("allowed_fails", 3)),
                  cooldown_time=float(rs.get("cooldown_time", 60)), fallbacks=rs.get("fallbacks"),
                  default_fallbacks=rs.get("default_fallbacks"),
                  context_window_fallbacks=rs.get("context_window_fallbacks") or ls.get("context_window_fallbacks"),
                  enable_pre_call_checks=bool(rs.get("enable_pre_call_checks", False)),
                  fallback_retries=bool(rs.get("fallback_retries", True)), load_weights=cfg.get("load_weights"),
                  guardrails=guards, cache=cache)
        kw.update(overrides)
        return cls(deps, **kw)

    # ------------------------------------------------------------------ selection
    def group(self, name: str) -> List[Deployment]:
        return [d for d in self.deployments if d.model_name == name]

    def model_names(self) -> List[str]:
        seen: List[str] = []
        for d in self.deployments:
            if d.model_name not in seen:
                seen.append(d.model_name)
        return seen

    def _available(self, name: str, tokens: int, exclude: Sequence[Deployment] = ()) -> List[Deployment]:
        now = self._clock()
        out = []
        for d in self.group(name):
            if d in exclude or not d.healthy or now < d.cooldown_until:
                continue
            if self.pre_call_checks:
                if d.rpm is not None and d.rpm_used(now) >= d.rpm:
                    continue
                if d.tpm is not None and d.tpm_used(now) + tokens > d.tpm:
                    continue
                if d.max_input_tokens is not None and tokens > d.max_input_tokens:
                    continue
            out.append(d)
        return out

    def _load_score(self, d: Deployment, cands: Sequence[Deployment]) -> float:
        """Weighted sum of the load figures, each scaled by the largest value among the candidates (0 = idle, 1 = the
        busiest on every axis). Without a scrape, the router's own counters stand in for queue depth and latency."""
        def fig(x: Deployment, k: str) -> float:
            if k in x.load:
                return float(x.load[k])
            return {"pending_requests": float(x.in_flight), "ttft_ms": x.avg_latency() * 1000.0}.get(k, 0.0)
        score = 0.0
        for k, w in self.load_weights.items():
            top = max(fig(x, k) for x in cands)
            score += w * (fig(d, k) / top if top > 0 else 0.0)
        return score

    def pick(self, name: str, tokens: int = 0, exclude: Sequence[Deployment] = (), prompt: Optional[str] = None) -> Optional[Deployment]:
        """One deployment of group ``name`` under the configured strategy, or None when none is available."""
        cands = self._available(name, tokens, exclude)
        if not cands:
            return None
        now = self._clock()
        if self.strategy in ("prefix-aware", "load-aware-prefix"):
            text = prompt or ""
            hit = {id(d): self.prefix_index.match(d.api_base or id(d), text) for d in cands}
            if self.strategy == "prefix-aware":
                # the replica holding the longest cached prefix of this prompt; nobody has one -> the least busy replica
                key = lambda d: (-hit[id(d)], d.in_flight)
            else:
                frac = {k: (v / len(text) if text else 0.0) for k, v in hit.items()}
                key = lambda d: (round(self._load_score(d, cands) - self.prefix_weight * frac[id(d)], 9), d.in_flight)
            best = min(key(d) for d in cands)
            ties = [d for d in cands if key(d) == best]
            d = ties[0] if len(ties) == 1 else self._rng.choice(ties)
            if hit[id(d)] > 0:
                self.prefix_routed += 1
            if text:
                self.prefix_index.insert(d.api_base or id(d), text)
            return d
        if self.strategy == "simple-shuffle":
            return self._rng.choices(cands, weights=[max(d.weight, 1e-9) for d in cands], k=1)[0]
        if self.strategy == "least-busy":
            key = lambda d: (d.in_flight, -d.weight)
        elif self.strategy == "cost-based-routing":
            key = lambda d: (d.cost_per_token(), d.in_flight)            # cheapest first, the less busy one on a tie
        elif self.strategy == "latency-based-routing":
            key = lambda d: (d.avg_latency(), d.in_flight)
        else:                                                            # usage-based: most head-room in its tpm / rpm budget
            def key(d):
                t = d.tpm_used(now) / d.tpm if d.tpm else 0.0
                r = d.rpm_used(now) / d.rpm if d.rpm else 0.0
                return (max(t, r), d.in_flight)
        best = min(key(d) for d in cands)
        ties = [d for d in cands if key(d) == best]
        return ties[0] if len(ties) == 1 else self._rng.choice(ties)

    # ------------------------------------------------------------------ bookkeeping
    def _note_failure(self, d: Deployment, status: int) -> None:
        now = self._clock()
        d.failures += 1
        if status == 429:
            d.cooldown_until = max(d.cooldown_until, now + self.rate_limit_cooldown)
            return
        d._fail_times.append(now)
        d._trim(now)
        if len(d._fail_times) > self.allowed_fails:
            d.cooldown_until = max(d.cooldown_until, now + self.cooldown_time)
            d._fail_times.clear()

    def _retries_for(self, status: int) -> int:
        name = {429: "RateLimitErrorRetries", 408: "TimeoutErrorRetries", 504: "TimeoutErrorRetries",
                401: "AuthenticationErrorRetries", 400: "BadRequestErrorRetries"}.get(status)
        if name is None and status >= 500:
            name = "InternalServerErrorRetries"
        return int(self.retry_policy.get(name, self.num_retries)) if name else self.num_retries

    async def _try_group(self, name: str, path: str, body: Dict[str, Any], tokens: int,
                         retries: Optional[int] = None) -> Tuple[int, Dict[str, Any], bool]:
        """-> (status, json, context_window_error). Retries inside the group, moving to another deployment after a failure."""
        tried: List[Deployment] = []
        fixed = retries is not None
        prompt = None
        if self.strategy in ("prefix-aware", "load-aware-prefix"):
            from .platform import prompt_text
            prompt = prompt_text(body)
        attempt, budget = 0, (retries if fixed else self.num_retries)
        last: Tuple[int, Dict[str, Any]] = (503, {"error": {"message": f"no healthy deployment for model '{name}'"}})
        while attempt <= budget:
            d = self.pick(name, tokens, exclude=tried, prompt=prompt) or self.pick(name, tokens, prompt=prompt)
            if d is None:
                return last[0], last[1], False
            payload = dict(body, model=d.upstream_model)
            now = self._clock()
            d._req_times.append(now)
            d.in_flight += 1
            d.requests += 1
            t0 = time.perf_counter()
            try:
                status, js = await d.transport(path, payload)
            except asyncio.TimeoutError:
                status, js = 408, {"error": {"message": "upstream timeout"}}
            except Exception as e:                                       # connection errors look like a 503 to the policy
                status, js = 503, {"error": {"message": f"{type(e).__name__}: {e}"}}
            finally:
                d.in_flight -= 1
            if status < 400:
                d._latencies.append(time.perf_counter() - t0)
                used = int((js.get("usage") or {}).get("total_tokens", tokens))
                d._tok_times.append((now, used))
                return status, js, False
            last = (status, js)
            msg = str((js.get("error") or {}).get("message", "")).lower() if isinstance(js, dict) else ""
            if status == 400 and ("context" in msg or "maximum" in msg and "token" in msg):
                return status, js, True
            self._note_failure(d, status)
            if status in (400, 401, 403, 404, 422):                      # the request itself is at fault: no retry
                return status, js, False
            if attempt == 0 and not fixed:
                budget = self._retries_for(status)                       # the first error's class sets the retry budget
            tried.append(d)
            attempt += 1
            if attempt <= budget:
                await self._sleep(min(0.05 * attempt, 0.5))
        return last[0], last[1], False

    async def completion(self, body: Dict[str, Any], path: str = "/chat/completions") -> Dict[str, Any]:
        name = body.get("model")
        if not name or not self.group(name):
            raise RouterError(404, f"unknown model '{name}'")
        await self._run_guardrails(body)
        ctl = body.get("cache") if isinstance(body.get("cache"), dict) else {}
        body = {k: v for k, v in body.items() if k not in ("cache", "guardrails")}      # proxy-only fields stay here
        ckey = self._cache_key(body) if (self.cache_cfg is not None and not body.get("stream")) else None
        if ckey is not None and not ctl.get("no-cache"):
            hit = self._cache_get(ckey, body)
            if hit is not None:
                self.cache_hits += 1
                hit.setdefault("x_router", {})["cache_hit"] = True
                return hit
            self.cache_misses += 1
        tokens = _estimate_tokens(body) + int(body.get("max_tokens") or 0)
        chain = [name] + [m for m in self.fallbacks.get(name, []) if m != name]
        chain += [m for m in self.default_fallbacks if m not in chain]
        status, js = 503, {"error": {"message": "no deployment"}}
        i = 0
        while i < len(chain):
            model = chain[i]
            if not self.group(model):
                i += 1
                continue
            if i > 0:
                self.fallback_events += 1
            status, js, ctx = await self._try_group(model, path, body, tokens,
                                                    retries=0 if (i > 0 and not self.fallback_retries) else None)
            if status < 400:
                if isinstance(js, dict):
                    js.setdefault("x_router", {})["served_by"] = model
                    if ckey is not None and not ctl.get("no-store"):
                        self._cache_put(ckey, body, js, int(ctl.get("ttl", self.cache_cfg.get("ttl", 3600))))
                return js
            if ctx:                                                      # prompt too long for this group: jump to its bigger siblings
                extra = [m for m in self.context_fallbacks.get(model, []) if m not in chain[:i + 1]]
                chain = chain[:i + 1] + extra + [m for m in chain[i + 1:] if m not in extra]
            elif status in (400, 401, 403, 404, 422):
                break
            i += 1
        raise RouterError(status, str((js.get("error") or {}).get("message", js)) if isinstance(js, dict) else str(js))

    # ------------------------------------------------------------------ guardrails and response cache
    async def _run_guardrails(self, body: Dict[str, Any]) -> None:
        asked = set(body.get("guardrails") or [])
        for g in self.guardrails:
            if g.mode != "pre_call" or not (g.default_on or g.name in asked):
                continue
            g.checks += 1
            try:
                status, js = await g.transport("/moderations", {"model": g.model, "input": _request_text(body)})
            except Exception as e:
                raise RouterError(503, f"guardrail '{g.name}' unavailable: {type(e).__name__}: {e}")
            if status >= 400 or not isinstance(js, dict):
                raise RouterError(503, f"guardrail '{g.name}' unavailable (HTTP {status})")
            res = (js.get("results") or [{}])[0]
            if res.get("flagged"):
                g.blocked += 1
                cats = sorted(k for k, v in (res.get("categories") or {}).items() if v)
                raise RouterError(400, f"Violated guardrail policy '{g.name}': " + (", ".join(cats) or "flagged"))

    @staticmethod
    def _cache_key(body: Dict[str, Any]) -> str:
        import hashlib
        import json
        knobs = {k: body.get(k) for k in ("model", "messages", "prompt", "max_tokens", "temperature", "top_p", "stop", "n", "tools")}
        return hashlib.sha256(json.dumps(knobs, sort_keys=True, ensure_ascii=False, default=str).encode("utf-8")).hexdigest()

    def _cache_get(self, key: str, body: Dict[str, Any]) -> Optional[Dict[str, Any]]:
        import json
        raw = self._exact.get(key)
        if raw is None and self._semantic is not None:
            from .cache_gateway import hashed_ngram_embedding
            q = hashed_ngram_embedding(str(body.get("model")) + "\n" + _request_text(body), self._embed_dim)
            raw, _ = self._semantic.lookup(q, float(self.cache_cfg.get("similarity_threshold", 0.8)))
        return json.loads(raw) if raw is not None else None

    def _cache_put(self, key: str, body: Dict[str, Any], js: Dict[str, Any], ttl: int) -> None:
        import json
        raw = json.dumps(js, ensure_ascii=False)
        self._exact.setex(key, ttl, raw)
        if self._semantic is not None:
            from .cache_gateway import hashed_ngram_embedding
            self._semantic.add(hashed_ngram_embedding(str(body.get("model")) + "\n" + _request_text(body), self._embed_dim), ttl, raw)

    async def health_check(self) -> Dict[str, Any]:
        """Probe every deployment with a one-token request (LiteLLM's /health): unhealthy ones leave the rotation until a
        later probe succeeds."""
        healthy, unhealthy = [], []
        for d in self.deployments:
            try:
                status, _ = await d.transport("/chat/completions", {"model": d.upstream_model, "max_tokens": 1,
                                                                    "messages": [{"role": "user", "content": "ping"}]})
            except Exception:
                status = 503
            d.healthy = status < 400
            (healthy if d.healthy else unhealthy).append({"model": d.model_name, "api_base": d.api_base})
        return {"healthy_endpoints": healthy, "unhealthy_endpoints": unhealthy, "healthy_count": len(healthy),
                "unhealthy_count": len(unhealthy)}

    async def refresh_load(self) -> None:
        """Scrape every deployment that has a ``metrics_fn`` and keep the four figures llm-d weighs: queue depth, KV-cache
        usage, mean time to first token and (its stand-in here) mean queueing share of the end-to-end latency."""
        from .platform import metric_sum, parse_prometheus_text
        for d in self.deployments:
            if d.metrics_fn is None:
                continue
            try:
                s = parse_prometheus_text(await d.metrics_fn())
            except Exception:
                continue
            n = metric_sum(s, "vllm:time_to_first_token_seconds_count")
            ttft = metric_sum(s, "vllm:time_to_first_token_seconds_sum") / n * 1000.0 if n else 0.0
            d.load = {"pending_requests": metric_sum(s, "vllm:num_requests_waiting"),
                      "gpu_cache_usage": metric_sum(s, "vllm:gpu_cache_usage_perc"), "ttft_ms": ttft,
                      "queue_time_ms": (metric_sum(s, "vllm:request_queue_time_seconds_sum") * 1000.0 /
                                        (metric_sum(s, "vllm:request_queue_time_seconds_count") or n)) if n else 0.0}

    def metrics_text(self) -> str:
        lines = ["# TYPE router_deployment_requests_total counter", "# TYPE router_deployment_failures_total counter",
                 "# TYPE router_deployment_in_flight gauge", "# TYPE router_deployment_cooldown gauge"]
        now = self._clock()
        for d in self.deployments:
            lab = f'{{model="{d.model_name}",api_base="{d.api_base}"}}'
            lines += [f"router_deployment_requests_total{lab} {d.requests}", f"router_deployment_failures_total{lab} {d.failures}",
                      f"router_deployment_in_flight{lab} {d.in_flight}",
                      f"router_deployment_cooldown{lab} {1 if now < d.cooldown_until or not d.healthy else 0}"]
        lines += ["# TYPE router_fallbacks_total counter", f"router_fallbacks_total {self.fallback_events}",
                  "# TYPE router_prefix_routed_total counter", f"router_prefix_routed_total {self.prefix_routed}",
                  "# TYPE router_cache_hits_total counter", f"router_cache_hits_total {self.cache_hits}",
                  "# TYPE router_cache_misses_total counter", f"router_cache_misses_total {self.cache_misses}",
                  "# TYPE router_guardrail_blocked_total counter"]
        lines += [f'router_guardrail_blocked_total{{guardrail="{g.name}"}} {g.blocked}' for g in self.guardrails]
        return "\n".join(lines) + "\n"


def build_router_app(router: Router, health_check_interval: float = 0.0, master_key: Optional[str] = None) -> FastAPI:
    """FastAPI surface of the proxy. ``health_check_interval`` > 0 starts the background probe loop (the reference sets 30)."""
    from contextlib import asynccontextmanager

    @asynccontextmanager
    async def lifespan(app_: FastAPI):
        task = None
        if health_check_interval > 0:
            async def loop():
                while True:
                    await router.health_check()
                    await router.refresh_load()
                    await asyncio.sleep(health_check_interval)
            task = asyncio.create_task(loop())
        yield
        if task is not None:
            task.cancel()

    app = FastAPI(title="llm_in_practise_b200 model router", lifespan=lifespan)
    app.state.router = router

    def _auth(request: Request) -> None:
        if master_key and request.headers.get("authorization", "") != f"Bearer {master_key}":
            raise HTTPException(status_code=401, detail="invalid proxy key")

    @app.get("/health/liveliness")
    async def live():
        return "I'm alive!"

    @app.get("/health")
    async def health(request: Request):
        _auth(request)
        return await router.health_check()

    @app.get("/v1/models")
    async def models(request: Request):
        _auth(request)
        return {"object": "list", "data": [{"id": m, "object": "model", "owned_by": "router"} for m in router.model_names()]}

    @app.get("/metrics")
    async def metrics():
        return PlainTextResponse(router.metrics_text(), media_type="text/plain; version=0.0.4")

    async def _forward(request: Request, path: str):
        _auth(request)
        body = await request.json()
        if body.get("stream"):
            raise HTTPException(status_code=400, detail="the router forwards non-streaming requests; stream from the deployment directly")
        try:
            return JSONResponse(await router.completion(body, path))
        except RouterError as e:
            return JSONResponse({"error": {"message": e.message, "type": "router_error", "code": e.status}}, status_code=e.status)

    @app.post("/v1/chat/completions")
    async def chat(request: Request):
        return await _forward(request, "/chat/completions")

    @app.post("/v1/completions")
    async def completions(request: Request):
        return await _forward(request, "/completions")

    return app
