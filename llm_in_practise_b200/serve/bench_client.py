"""HTTP benchmark client for an OpenAI-compatible endpoint, with the flags and the result file of ``vllm bench serve``.

The reference measures every stage of its inference platform from a client pod with
``vllm bench serve --backend vllm --base-url … --model … --dataset-name sharegpt --dataset-path … --num-prompts 128
--request-rate inf --max-concurrency C --ignore-eos --sharegpt-output-len 256 --save-result --result-dir … --result-filename …``
(LLM_on_Kubernetes/Inference_Platfrom/README.md:1361-1375, 1535-1549, 1614-1628; the pod is 04-BenchMark/benchmark-client.yaml).
This module sends the same traffic — streaming requests, bounded concurrency, optional Poisson arrivals — and writes the
same result keys (``request_throughput``, ``output_throughput``, ``total_token_throughput``, mean / median / std / p99 of
``ttft_ms``, ``tpot_ms``, ``itl_ms``, ``e2el_ms``), so the reference's result tables can be filled from either tool. The
serving side accepts what either client sends (``ignore_eos``, ``stream_options.include_usage``, streaming
``/v1/completions``).

    python -m llm_in_practise_b200.recipes.tools bench-serve --base-url http://127.0.0.1:8000 --model qwen3-8b \\
        --dataset-name sharegpt --dataset-path sharegpt.json --num-prompts 128 --max-concurrency 16 --ignore-eos \\
        --sharegpt-output-len 256 --save-result --result-dir results --result-filename c16.json
"""
from __future__ import annotations

import argparse
import asyncio
import json
import math
import os
import random
import statistics
import time
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional


@dataclass
class SampleRequest:
    prompt: str
    output_len: int
    prompt_len: int = 0                 # tokens, when a tokenizer was given (else filled from the server's usage chunk)


@dataclass
class RequestResult:
    ok: bool = False
    error: str = ""
    start: float = 0.0
    ttft: float = 0.0
    latency: float = 0.0
    itl: List[float] = field(default_factory=list)
    prompt_tokens: int = 0
    output_tokens: int = 0
    text: str = ""


def load_sharegpt(path: str, num_prompts: int, output_len: Optional[int], seed: int = 0, tokenizer=None) -> List[SampleRequest]:
    """ShareGPT JSON (a list of {"conversations": [{"from", "value"}, ...]}): the first turn is the prompt, the second one's
    length the wanted output length unless ``output_len`` fixes it; conversations with fewer than two turns are skipped."""
    with open(path, encoding="utf-8") as f:
        data = json.load(f)
    rows = []
    for d in data:
        conv = d.get("conversations") or d.get("conversation") or []
        if len(conv) < 2:
            continue
        prompt, answer = str(conv[0].get("value", "")), str(conv[1].get("value", ""))
        if not prompt.strip():
            continue
        n_out = output_len or (len(tokenizer.encode(answer)) if tokenizer is not None else max(1, len(answer.split())))
        rows.append(SampleRequest(prompt, int(n_out), len(tokenizer.encode(prompt)) if tokenizer is not None else 0))
    random.Random(seed).shuffle(rows)
    if not rows:
        raise ValueError(f"{path} holds no two-turn conversation")
    return [rows[i % len(rows)] for i in range(num_prompts)]


def random_requests(num_prompts: int, input_len: int, output_len: int, seed: int = 0) -> List[SampleRequest]:
    """``--dataset-name random``: ``input_len`` pseudo-words per prompt (distinct prompts, so prefix caching does not help)."""
    rng = random.Random(seed)
    words = ["alpha", "beta", "gamma", "delta", "omega", "sigma", "kappa", "theta", "lambda", "zeta", "node", "cache", "token", "model"]
    return [SampleRequest(" ".join(rng.choice(words) + str(rng.randrange(1000)) for _ in range(input_len)), output_len) for _ in range(num_prompts)]


async def _one(client, url: str, payload: Dict[str, Any], chat: bool, res: RequestResult) -> None:
    res.start = time.perf_counter()
    last = res.start
    try:
        async with client.stream("POST", url, json=payload) as resp:
            if resp.status_code != 200:
                res.error = f"HTTP {resp.status_code}: {(await resp.aread())[:200]!r}"
                return
            async for line in resp.aiter_lines():
                if not line.startswith("data:"):
                    continue
                body = line[5:].strip()
                if body == "[DONE]":
                    break
                js = json.loads(body)
                if js.get("usage"):
                    res.prompt_tokens = int(js["usage"].get("prompt_tokens", 0))
                    res.output_tokens = int(js["usage"].get("completion_tokens", 0))
                ch = (js.get("choices") or [None])[0]
                if not ch:
                    continue
                piece = (ch.get("delta") or {}).get("content") if chat else ch.get("text")
                if piece:
                    now = time.perf_counter()
                    if res.ttft == 0.0:
                        res.ttft = now - res.start
                    else:
                        res.itl.append(now - last)
                    last = now
                    res.text += piece
        res.latency = time.perf_counter() - res.start
        res.ok = res.ttft > 0.0 or res.output_tokens > 0
        if not res.ok:
            res.error = "no tokens received"
    except Exception as e:   # noqa: BLE001 - a failed request is a data point, not a crash
        res.error = f"{type(e).__name__}: {e}"


def _stats(xs: List[float], name: str, out: Dict[str, Any], pcts=(99,)) -> None:
    ms = [x * 1000.0 for x in xs]
    out[f"mean_{name}_ms"] = statistics.fmean(ms) if ms else 0.0
    out[f"median_{name}_ms"] = statistics.median(ms) if ms else 0.0
    out[f"std_{name}_ms"] = statistics.pstdev(ms) if len(ms) > 1 else 0.0
    srt = sorted(ms)
    for p in pcts:
        # numpy's default (linear) percentile, which is what the reference's tool reports
        if not srt:
            out[f"p{p}_{name}_ms"] = 0.0
            continue
        k = (len(srt) - 1) * p / 100.0
        lo, hi = math.floor(k), math.ceil(k)
        out[f"p{p}_{name}_ms"] = srt[lo] + (srt[hi] - srt[lo]) * (k - lo)


async def run_benchmark(base_url: str, model: str, requests: List[SampleRequest], *, backend: str = "vllm",
                        max_concurrency: Optional[int] = None, request_rate: float = float("inf"), ignore_eos: bool = False,
                        temperature: float = 0.0, seed: int = 0, tokenizer=None, transport=None, timeout: float = 600.0,
                        extra_body: Optional[Dict[str, Any]] = None) -> Dict[str, Any]:
    """Fire ``requests`` (all at once, or as a Poisson process of ``request_rate`` per second), at most ``max_concurrency``
    in flight; returns the result dictionary ``vllm bench serve --save-result`` writes."""
    import httpx
    chat = backend in ("openai-chat", "chat")
    url = base_url.rstrip("/") + ("/v1/chat/completions" if chat else "/v1/completions")
    sem = asyncio.Semaphore(max_concurrency) if max_concurrency else None
    rng = random.Random(seed)
    results = [RequestResult() for _ in requests]

    async def guarded(client, payload, res):
        if sem is None:
            return await _one(client, url, payload, chat, res)
        async with sem:
            return await _one(client, url, payload, chat, res)

    kw = {"transport": transport} if transport is not None else {}
    t0 = time.perf_counter()
    async with httpx.AsyncClient(timeout=timeout, **kw) as client:
        tasks = []
        for req, res in zip(requests, results):
            payload: Dict[str, Any] = {"model": model, "max_tokens": req.output_len, "temperature": temperature, "stream": True,
                                       "stream_options": {"include_usage": True}}
            if ignore_eos:
                payload["ignore_eos"] = True
            if chat:
                payload["messages"] = [{"role": "user", "content": req.prompt}]
            else:
                payload["prompt"] = req.prompt
            payload.update(extra_body or {})
            tasks.append(asyncio.create_task(guarded(client, payload, res)))
            if math.isfinite(request_rate) and request_rate > 0:
                await asyncio.sleep(rng.expovariate(request_rate))
        await asyncio.gather(*tasks)
    duration = time.perf_counter() - t0
    good = [(q, r) for q, r in zip(requests, results) if r.ok]
    for q, r in good:                                   # token counts: server usage > tokenizer > words
        if not r.output_tokens:
            r.output_tokens = len(tokenizer.encode(r.text)) if tokenizer is not None else max(1, len(r.text.split()))
        if not r.prompt_tokens:
            r.prompt_tokens = q.prompt_len or max(1, len(q.prompt.split()))
    n_in, n_out = sum(r.prompt_tokens for _, r in good), sum(r.output_tokens for _, r in good)
    out: Dict[str, Any] = {
        "backend": backend, "model_id": model, "num_prompts": len(requests), "request_rate": request_rate if math.isfinite(request_rate) else "inf",
        "max_concurrency": max_concurrency, "duration": duration, "completed": len(good), "failed": len(requests) - len(good),
        "total_input_tokens": n_in, "total_output_tokens": n_out,
        "request_throughput": len(good) / duration if duration > 0 else 0.0,
        "output_throughput": n_out / duration if duration > 0 else 0.0,
        "total_token_throughput": (n_in + n_out) / duration if duration > 0 else 0.0,
        "errors": sorted({r.error for r in results if r.error})[:5]}
    _stats([r.ttft for _, r in good], "ttft", out)
    # time per output token after the first: (latency - ttft) / (tokens - 1), one value per request
    _stats([(r.latency - r.ttft) / (r.output_tokens - 1) for _, r in good if r.output_tokens > 1], "tpot", out)
    _stats([x for _, r in good for x in r.itl], "itl", out)
    _stats([r.latency for _, r in good], "e2el", out)
    return out


def format_report(r: Dict[str, Any]) -> str:
    rows = [("Successful requests", r["completed"]), ("Failed requests", r["failed"]), ("Benchmark duration (s)", f"{r['duration']:.2f}"),
            ("Total input tokens", r["total_input_tokens"]), ("Total generated tokens", r["total_output_tokens"]),
            ("Request throughput (req/s)", f"{r['request_throughput']:.2f}"), ("Output token throughput (tok/s)", f"{r['output_throughput']:.2f}"),
            ("Total token throughput (tok/s)", f"{r['total_token_throughput']:.2f}")]
    for name, title in (("ttft", "Time to First Token"), ("tpot", "Time per Output Token (excl. 1st token)"),
                        ("itl", "Inter-token Latency"), ("e2el", "End-to-end Latency")):
        rows += [(f"Mean {name.upper()} (ms)", f"{r[f'mean_{name}_ms']:.2f}"), (f"Median {name.upper()} (ms)", f"{r[f'median_{name}_ms']:.2f}"),
                 (f"P99 {name.upper()} (ms)", f"{r[f'p99_{name}_ms']:.2f}")]
    w = max(len(k) for k, _ in rows) + 2
    return "\n".join(["=" * (w + 12), " Serving Benchmark Result ".center(w + 12, "="), *(f"{k:<{w}}{v}" for k, v in rows), "=" * (w + 12)])


def main(argv=None) -> Dict[str, Any]:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--backend", default="vllm", choices=["vllm", "openai", "openai-chat"])
    ap.add_argument("--base-url", default="http://127.0.0.1:8000")
    ap.add_argument("--model", required=True)
    ap.add_argument("--tokenizer", default=None, help="tokenizer directory (token counts when the server sends no usage)")
    ap.add_argument("--dataset-name", default="random", choices=["sharegpt", "random"])
    ap.add_argument("--dataset-path", default=None)
    ap.add_argument("--num-prompts", type=int, default=128)
    ap.add_argument("--request-rate", type=float, default=float("inf"))
    ap.add_argument("--max-concurrency", type=int, default=None)
    ap.add_argument("--ignore-eos", action="store_true")
    ap.add_argument("--sharegpt-output-len", type=int, default=None)
    ap.add_argument("--random-input-len", type=int, default=256)
    ap.add_argument("--random-output-len", type=int, default=128)
    ap.add_argument("--temperature", type=float, default=0.0)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--save-result", action="store_true")
    ap.add_argument("--result-dir", default=".")
    ap.add_argument("--result-filename", default=None)
    a = ap.parse_args(argv)
    tok = None
    if a.tokenizer:
        from ..data.tokenizers import load_pretrained_tokenizer
        tok = load_pretrained_tokenizer(a.tokenizer, add_pad_token=False)
    if a.dataset_name == "sharegpt":
        if not a.dataset_path:
            ap.error("--dataset-name sharegpt needs --dataset-path")
        reqs = load_sharegpt(a.dataset_path, a.num_prompts, a.sharegpt_output_len, a.seed, tok)
    else:
        reqs = random_requests(a.num_prompts, a.random_input_len, a.random_output_len, a.seed)
    out = asyncio.run(run_benchmark(a.base_url, a.model, reqs, backend=a.backend, max_concurrency=a.max_concurrency,
                                    request_rate=a.request_rate, ignore_eos=a.ignore_eos, temperature=a.temperature, seed=a.seed,
                                    tokenizer=tok))
    print(format_report(out))
    if a.save_result:
        os.makedirs(a.result_dir, exist_ok=True)
        name = a.result_filename or f"{a.backend}-{a.request_rate}qps-{a.model.replace('/', '_')}-{int(time.time())}.json"
        with open(os.path.join(a.result_dir, name), "w") as f:
            json.dump(out, f, indent=1)
    return out


if __name__ == "__main__":
    main()
