"""Load generator for an OpenAI-compatible endpoint (Locust is not in the image; this is an asyncio equivalent of
the reference's two Locust files).

* ``qps`` profile  — a fixed chat payload (system + long user prompt, ``max_tokens 128, temperature 0``) fired by
  ``users`` simulated users that wait U(0.05, 0.15) s between requests, i.e. ~10 QPS per user
  (Deployment/Ray/scripts/locustfile.py:4-24).
* ``tps`` profile  — ``max_tokens 256, temperature 0.7``, think time U(1, 3) s, and per request
  ``completion_tokens / wall`` reported as tokens-per-second (Deployment/Ray/scripts/locustfile-TPS.py:21-46).

Reports request count, failures, latency percentiles and aggregate / per-request tokens per second.
``python -m llm_in_practise_b200.serve.loadtest --host http://127.0.0.1:8000 --profile tps --users 8 --duration 60``
"""
from __future__ import annotations

import argparse
import asyncio
import json
import random
import statistics
import time
from dataclasses import dataclass, field
from typing import Dict, List, Optional

PROFILES: Dict[str, dict] = {
    "qps": {"payload": {"model": "qwen3", "messages": [
        {"role": "system", "content": "You are a helpful assistant."},
        {"role": "user", "content": "请写一个关于大型语言模型自动扩缩容的详细技术报告，长度约为1024个 token 的提示词。"}],
        "max_tokens": 128, "temperature": 0.0, "stream": False}, "wait": (0.05, 0.15)},
    "tps": {"payload": {"model": "qwen3", "messages": [
        {"role": "system", "content": "You are a helpful assistant."},
        {"role": "user", "content": "请写一篇约200字的短文，介绍中国的长城。"}],
        "max_tokens": 256, "temperature": 0.7}, "wait": (1.0, 3.0)},
}


@dataclass
class LoadStats:
    latencies: List[float] = field(default_factory=list)
    tps: List[float] = field(default_factory=list)
    completion_tokens: int = 0
    failures: int = 0
    wall: float = 0.0

    def summary(self) -> dict:
        lat = sorted(self.latencies)
        pct = lambda p: lat[min(len(lat) - 1, int(p * len(lat)))] if lat else float("nan")  # noqa: E731
        return {"requests": len(lat) + self.failures, "failures": self.failures,
                "rps": len(lat) / self.wall if self.wall else 0.0,
                "latency_ms": {"p50": pct(0.5) * 1e3, "p90": pct(0.9) * 1e3, "p99": pct(0.99) * 1e3,
                               "mean": statistics.fmean(lat) * 1e3 if lat else float("nan")},
                "completion_tokens": self.completion_tokens,
                "aggregate_tokens_per_second": self.completion_tokens / self.wall if self.wall else 0.0,
                "per_request_tokens_per_second_mean": statistics.fmean(self.tps) if self.tps else 0.0}


async def _user(client, url: str, payload: dict, wait, deadline: float, stats: LoadStats, max_requests: Optional[int]):
    n = 0
    while time.perf_counter() < deadline and (max_requests is None or n < max_requests):
        t0 = time.perf_counter()
        try:
            resp = await client.post(url, json=payload)
            dt = time.perf_counter() - t0
            if resp.status_code != 200:
                stats.failures += 1
            else:
                toks = int(resp.json().get("usage", {}).get("completion_tokens", 0))
                stats.latencies.append(dt)
                stats.completion_tokens += toks
                if dt > 0:
                    stats.tps.append(toks / dt)
        except Exception:
            stats.failures += 1
        n += 1
        await asyncio.sleep(random.uniform(*wait))


async def run_load(host: str, profile: str = "qps", users: int = 4, duration: float = 30.0, model: Optional[str] = None,
                   max_requests_per_user: Optional[int] = None, transport=None, timeout: float = 300.0) -> dict:
    import httpx
    prof = PROFILES[profile]
    payload = dict(prof["payload"])
    if model:
        payload["model"] = model
    stats = LoadStats()
    t0 = time.perf_counter()
    async with httpx.AsyncClient(base_url=host, timeout=timeout, transport=transport) as client:
        await asyncio.gather(*[_user(client, "/v1/chat/completions", payload, prof["wait"], t0 + duration, stats,
                                     max_requests_per_user) for _ in range(users)])
    stats.wall = time.perf_counter() - t0
    return stats.summary()


def main(argv=None) -> None:
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--host", default="http://127.0.0.1:8000")
    ap.add_argument("--profile", choices=sorted(PROFILES), default="qps")
    ap.add_argument("--users", type=int, default=4)
    ap.add_argument("--duration", type=float, default=30.0)
    ap.add_argument("--model", default=None)
    a = ap.parse_args(argv)
    print(json.dumps(asyncio.run(run_load(a.host, a.profile, a.users, a.duration, a.model)), ensure_ascii=False, indent=2))


if __name__ == "__main__":
    main()
