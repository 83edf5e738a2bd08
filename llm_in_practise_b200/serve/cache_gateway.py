"""Response-cache gateway in front of an OpenAI-compatible chat endpoint.

The reference describes this component inline in its Kubernetes guide (LLM_on_Kubernetes/Inference_Platfrom/README.md:3021-3230:
a small web service between the ingress and vLLM): ``POST /v1/chat/completions`` is answered from

* **L2** — an exact-match cache keyed by the SHA-256 of the conversation text (+ model and sampling knobs), with a TTL;
* **L3** — a semantic cache: the conversation is embedded and the nearest stored conversation is reused when its cosine
  similarity exceeds a threshold (0.95 in the reference);

and otherwise forwarded upstream, the answer being written to both levels. Streaming requests and sampled requests
(``temperature > 0`` unless ``cache_sampled``) bypass the caches. ``GET /metrics`` exposes hit / miss counters and latency
histograms in the Prometheus text format, ``GET /healthz`` the liveness probe the manifests expect.

Design here: the gateway is storage- and transport-agnostic. ``upstream`` is any ``async (payload) -> dict`` (the in-process
``serve.openai_server`` app, an ``LLMEngine``, or an HTTP forwarder built by ``http_upstream``); the exact store is a bounded
LRU dict (a Redis-like object with ``get`` / ``setex`` can be passed instead); ``embed`` is any ``(text) -> 1-D tensor`` — the
default is a hashed character n-gram vector, good enough to catch re-worded duplicates without a second model on the GPU;
pass the serving model's own mean-pooled embeddings for a learned one.
"""
from __future__ import annotations

import asyncio
import hashlib
import json
import time
from collections import OrderedDict
from typing import Any, Awaitable, Callable, Dict, List, Optional, Tuple

import torch
from fastapi import FastAPI, HTTPException, Request
from fastapi.responses import JSONResponse, PlainTextResponse

Upstream = Callable[[Dict[str, Any]], Awaitable[Dict[str, Any]]]


def conversation_text(messages: List[Dict[str, Any]]) -> str:
    return "\n".join(f"{m.get('role', '')}: {m.get('content', '')}" for m in messages)


def hashed_ngram_embedding(text: str, dim: int = 1024, n: int = 3) -> torch.Tensor:
    """L2-normalised bag of hashed character n-grams (no model, deterministic across processes)."""
    v = torch.zeros(dim)
    t = " " + " ".join(text.lower().split()) + " "
    for i in range(max(len(t) - n + 1, 1)):
        h = int.from_bytes(hashlib.blake2b(t[i:i + n].encode("utf-8"), digest_size=8).digest(), "little")
        v[h % dim] += 1.0 if (h >> 63) == 0 else -1.0
    return v / v.norm().clamp_min(1e-12)


class ExactCache:
    """Bounded LRU with per-entry expiry (the in-process stand-in for the Redis the reference deploys)."""

    def __init__(self, max_entries: int = 4096):
        self.max_entries = max_entries
        self._d: "OrderedDict[str, Tuple[float, str]]" = OrderedDict()

    def get(self, key: str) -> Optional[str]:
        item = self._d.get(key)
        if item is None:
            return None
        if item[0] < time.time():
            del self._d[key]
            return None
        self._d.move_to_end(key)
        return item[1]

    def setex(self, key: str, ttl: int, value: str) -> None:
        self._d[key] = (time.time() + ttl, value)
        self._d.move_to_end(key)
        while len(self._d) > self.max_entries:
            self._d.popitem(last=False)


class SemanticCache:
    """Dense matrix of unit vectors; lookup is one matrix-vector product (runs on the GPU when ``device`` is one)."""

    def __init__(self, dim: int, max_entries: int = 4096, device: str = "cpu"):
        self.keys = torch.zeros(max_entries, dim, device=device)
        self.values: List[Optional[Tuple[float, str]]] = [None] * max_entries
        self.n = 0
        self.cursor = 0

    def lookup(self, q: torch.Tensor, threshold: float) -> Tuple[Optional[str], float]:
        if self.n == 0:
            return None, 0.0
        sims = self.keys[: self.n] @ q.to(self.keys.device)
        best = int(sims.argmax())
        score = float(sims[best])
        item = self.values[best]
        if score < threshold or item is None or item[0] < time.time():
            return None, score
        return item[1], score

    def add(self, q: torch.Tensor, ttl: int, value: str) -> None:
        i = self.cursor
        self.keys[i] = q.to(self.keys.device)
        self.values[i] = (time.time() + ttl, value)
        self.cursor = (i + 1) % self.keys.shape[0]          # ring buffer: the oldest entry is overwritten
        self.n = min(self.n + 1, self.keys.shape[0])


class _Metrics:
    BUCKETS = (0.005, 0.01, 0.025, 0.05, 0.1, 0.25, 0.5, 1.0, 2.5, 5.0, 10.0)

    def __init__(self):
        self.count: Dict[str, int] = {"l2_hit": 0, "l3_hit": 0, "miss": 0, "bypass": 0, "upstream_error": 0}
        self.hist: Dict[str, List[int]] = {}
        self.sum: Dict[str, float] = {}

    def observe(self, source: str, seconds: float) -> None:
        h = self.hist.setdefault(source, [0] * (len(self.BUCKETS) + 1))
        for i, b in enumerate(self.BUCKETS):
            if seconds <= b:
                h[i] += 1
        h[-1] += 1
        self.sum[source] = self.sum.get(source, 0.0) + seconds

    def render(self) -> str:
        out = ["# TYPE cache_gateway_requests_total counter"]
        out += [f'cache_gateway_requests_total{{result="{k}"}} {v}' for k, v in self.count.items()]
        out.append("# TYPE cache_gateway_latency_seconds histogram")
        for src, h in self.hist.items():
            out += [f'cache_gateway_latency_seconds_bucket{{source="{src}",le="{b}"}} {h[i]}' for i, b in enumerate(self.BUCKETS)]
            out.append(f'cache_gateway_latency_seconds_bucket{{source="{src}",le="+Inf"}} {h[-1]}')
            out.append(f'cache_gateway_latency_seconds_sum{{source="{src}"}} {self.sum[src]:.6f}')
            out.append(f'cache_gateway_latency_seconds_count{{source="{src}"}} {h[-1]}')
        return "\n".join(out) + "\n"


def http_upstream(base_url: str, timeout: float = 120.0) -> Upstream:
    """Forwarder to a running OpenAI-compatible server (``requests`` in a worker thread)."""
    import requests

    async def call(payload: Dict[str, Any]) -> Dict[str, Any]:
        def post():
            r = requests.post(base_url.rstrip("/") + "/v1/chat/completions", json=payload, timeout=timeout)
            r.raise_for_status()
            return r.json()
        return await asyncio.get_running_loop().run_in_executor(None, post)
    return call


def build_cache_gateway(upstream: Upstream, *, ttl: int = 3600, semantic_threshold: float = 0.95, semantic: bool = True,
                        embed: Optional[Callable[[str], torch.Tensor]] = None, embed_dim: int = 1024, exact_store=None,
                        max_entries: int = 4096, cache_sampled: bool = False, device: str = "cpu") -> FastAPI:
    app = FastAPI(title="llm_in_practise_b200 cache gateway")
    exact = exact_store if exact_store is not None else ExactCache(max_entries)
    embed_fn = embed or (lambda t: hashed_ngram_embedding(t, embed_dim))
    sem = SemanticCache(embed_dim, max_entries, device) if semantic else None
    metrics = _Metrics()
    app.state.metrics, app.state.exact, app.state.semantic = metrics, exact, sem

    def key_of(body: Dict[str, Any], text: str) -> str:
        knobs = {k: body.get(k) for k in ("model", "max_tokens", "temperature", "top_p", "stop")}
        return "chat:" + hashlib.sha256((text + "\x00" + json.dumps(knobs, sort_keys=True)).encode("utf-8")).hexdigest()

    @app.get("/healthz")
    async def healthz():
        return {"status": "ok"}

    @app.get("/metrics")
    async def prom():
        return PlainTextResponse(metrics.render(), media_type="text/plain; version=0.0.4")

    @app.post("/v1/chat/completions")
    async def chat(request: Request):
        t0 = time.perf_counter()
        body = await request.json()
        messages = body.get("messages")
        if not isinstance(messages, list) or not messages:
            raise HTTPException(status_code=400, detail="messages must be a non-empty list")
        cacheable = not body.get("stream") and (cache_sampled or not float(body.get("temperature", 0) or 0) > 0)

        async def forward() -> Dict[str, Any]:
            try:
                return await upstream(body)
            except HTTPException:
                raise
            except Exception as e:   # noqa: BLE001 - anything upstream becomes a 502 for the client
                metrics.count["upstream_error"] += 1
                raise HTTPException(status_code=502, detail=f"upstream failed: {e}") from e

        if not cacheable:
            metrics.count["bypass"] += 1
            ans = await forward()
            metrics.observe("upstream", time.perf_counter() - t0)
            return JSONResponse(ans, headers={"X-Cache": "BYPASS"})
        text = conversation_text(messages)
        key = key_of(body, text)
        hit = exact.get(key)
        if hit is not None:
            metrics.count["l2_hit"] += 1
            metrics.observe("l2", time.perf_counter() - t0)
            return JSONResponse(json.loads(hit), headers={"X-Cache": "HIT-L2"})
        q = None
        if sem is not None:
            q = embed_fn(text)
            val, score = sem.lookup(q, semantic_threshold)
            if val is not None:
                metrics.count["l3_hit"] += 1
                exact.setex(key, ttl, val)                       # promote: the next identical request is an L2 hit
                metrics.observe("l3", time.perf_counter() - t0)
                return JSONResponse(json.loads(val), headers={"X-Cache": "HIT-L3", "X-Cache-Similarity": f"{score:.4f}"})
        metrics.count["miss"] += 1
        ans = await forward()
        blob = json.dumps(ans)
        exact.setex(key, ttl, blob)
        if sem is not None and q is not None:
            sem.add(q, ttl, blob)
        metrics.observe("upstream", time.perf_counter() - t0)
        return JSONResponse(ans, headers={"X-Cache": "MISS"})

    return app
