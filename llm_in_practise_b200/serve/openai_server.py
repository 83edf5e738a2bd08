"""OpenAI-compatible HTTP front end over ``LLMEngine``.

Same wire format as the reference's server (Scripts/inference/07-deepseek1.5b-api-infr.py:66-161):
``POST /v1/chat/completions`` with ``{model, messages, temperature (0..1, default .7), top_p (default .95),
max_tokens (default 512), stream}`` -> ``{id "chatcmpl-…", object "chat.completion", created, model,
choices[{index, message{role, content}, finish_reason}], usage{prompt_tokens, completion_tokens, total_tokens}}``;
HTTP 503 while no model is loaded. The reference answers ``stream=true`` with HTTP 501; here streaming is
implemented (SSE ``chat.completion.chunk`` events) and ``allow_stream=False`` restores the 501.

Added because the reference's deployment recipes need them from whatever serves the model:
``POST /v1/completions`` (the Llama-Guard wrapper and the Locust files post to it;
Deployment/litellm-proxy/llama-guard-wrapper/model_client.py:35-58, Deployment/Ray/scripts/locustfile.py:4-24),
``GET /v1/models``, ``GET /health`` and a Prometheus ``GET /metrics`` with vLLM's metric names
(LLM_on_Kubernetes/…/05-KEDA-AutoScale/keda-scaledobject.yaml:12-53).
"""
from __future__ import annotations

import json
import time
import uuid
from typing import Any, Dict, List, Optional, Union

from fastapi import FastAPI, HTTPException
from fastapi.responses import PlainTextResponse, StreamingResponse
from pydantic import BaseModel, Field

from ..data.chat import render_chatml
from ..inference.generate import GenerationConfig
from .engine import LLMEngine


class ChatMessage(BaseModel):
    role: str
    content: str


class ChatCompletionRequest(BaseModel):
    model: str
    messages: List[ChatMessage]
    temperature: Optional[float] = Field(0.7, ge=0.0, le=2.0)
    top_p: Optional[float] = Field(0.95, ge=0.0, le=1.0)
    max_tokens: Optional[int] = Field(512, ge=1)
    stream: Optional[bool] = False
    top_k: Optional[int] = 0
    repetition_penalty: Optional[float] = 1.0
    stop: Optional[Union[str, List[str]]] = None    # vLLM / OpenAI stop strings (not in the reference's FastAPI schema)
    ignore_eos: Optional[bool] = False              # vLLM extension (``vllm bench serve --ignore-eos``)
    stream_options: Optional[Dict[str, Any]] = None  # {"include_usage": true}: one more chunk with the token counts


class CompletionRequest(BaseModel):
    model: str
    prompt: Union[str, List[str]]
    temperature: Optional[float] = 1.0
    top_p: Optional[float] = 1.0
    max_tokens: Optional[int] = 16
    stream: Optional[bool] = False
    stop: Optional[Union[str, List[str]]] = None
    ignore_eos: Optional[bool] = False
    stream_options: Optional[Dict[str, Any]] = None


def deepseek_prompt_from_messages(messages: List[ChatMessage]) -> str:
    """``<|user|>\\n…\\n<|assistant|>\\n`` turns, system text prepended (07-…api-infr.py:40-60)."""
    sys_txt, body = "", ""
    for m in messages:
        if m.role == "user":
            body += f"<|user|>\n{m.content}\n"
        elif m.role == "assistant":
            body += f"<|assistant|>\n{m.content}\n"
        elif m.role == "system":
            sys_txt = f"{m.content}\n" + sys_txt
    return sys_txt + body + "<|assistant|>\n"


def chatml_prompt_from_messages(messages: List[ChatMessage]) -> str:
    return render_chatml([m.model_dump() for m in messages], add_generation_prompt=True, space_before_end=False)


PROMPT_STYLES = {"deepseek": deepseek_prompt_from_messages, "chatml": chatml_prompt_from_messages}


def _stops(stop) -> List[str]:
    return [] if not stop else ([stop] if isinstance(stop, str) else [x for x in stop if x])


def cut_at_stop(text: str, stops: List[str]):
    """(text truncated at the earliest stop string, True if one matched)."""
    cut = min((i for i in (text.find(x) for x in stops) if i >= 0), default=-1)
    return (text, False) if cut < 0 else (text[:cut], True)


def render_metrics(engine: LLMEngine) -> str:
    m, name = engine.metrics, engine.model_name
    lab = f'{{model_name="{name}"}}'
    out = []

    def gauge(n, v, help_):
        out.extend([f"# HELP {n} {help_}", f"# TYPE {n} gauge", f"{n}{lab} {v}"])

    def counter(n, v, help_):
        out.extend([f"# HELP {n} {help_}", f"# TYPE {n} counter", f"{n}{lab} {v}"])

    gauge("vllm:num_requests_running", engine.num_running, "Number of requests currently running on GPU.")
    gauge("vllm:num_requests_waiting", engine.num_waiting, "Number of requests waiting to be processed.")
    gauge("vllm:gpu_cache_usage_perc", round(engine.kv_cache_usage(), 6), "GPU KV-cache usage. 1 means 100 percent usage.")
    with m.lock:
        counter("vllm:prompt_tokens_total", m.prompt_tokens, "Number of prefill tokens processed.")
        counter("vllm:generation_tokens_total", m.generation_tokens, "Number of generation tokens processed.")
        counter("vllm:gpu_prefix_cache_queries_total", m.prefix_queries, "Prompt tokens looked up in the prefix cache.")
        counter("vllm:gpu_prefix_cache_hits_total", m.prefix_hits, "Prompt tokens served from the prefix cache.")
        counter("b200:host_prefix_cache_hits_total", m.host_hits, "Prompt tokens restored from the host tier of the prefix cache.")
        counter("b200:host_kv_offloads_total", m.host_offloads, "Cached rows copied to the host tier when they lost their HBM.")
        out.extend(["# HELP vllm:request_success_total Count of successfully processed requests.",
                    "# TYPE vllm:request_success_total counter"])
        for reason, n in sorted(m.requests_finished.items()):
            if reason != "error":
                out.append(f'vllm:request_success_total{{model_name="{name}",finished_reason="{reason}"}} {n}')
        # the canary gate of the reference divides this by request_success_total (09-Canary-Deployment/analysis-template.yaml:24-37)
        counter("vllm:request_failure_total", m.requests_finished.get("error", 0), "Count of requests that failed inside the engine.")
        out.extend(["# HELP vllm:time_to_first_token_seconds Histogram of time to first token in seconds.",
                    "# TYPE vllm:time_to_first_token_seconds histogram"])
        cum = 0
        for b, c in zip(m.TTFT_BUCKETS, m.ttft_counts):
            cum += c
            out.append(f'vllm:time_to_first_token_seconds_bucket{{model_name="{name}",le="{b}"}} {cum}')
        cum += m.ttft_counts[-1]
        out.append(f'vllm:time_to_first_token_seconds_bucket{{model_name="{name}",le="+Inf"}} {cum}')
        out.append(f"vllm:time_to_first_token_seconds_count{lab} {cum}")
        out.append(f"vllm:time_to_first_token_seconds_sum{lab} {m.ttft_sum}")
        counter("vllm:e2e_request_latency_seconds_sum", m.e2e_sum, "Sum of end to end request latency in seconds.")
        counter("vllm:request_queue_time_seconds_sum", m.queue_sum, "Sum of the time requests spent in the waiting queue.")
        counter("vllm:request_queue_time_seconds_count", m.queue_count, "Requests that left the waiting queue.")
    return "\n".join(out) + "\n"


def build_app(engine: Optional[LLMEngine], prompt_style: str = "chatml", allow_stream: bool = True,
              title: str = "llm_in_practise_b200 OpenAI-compatible API") -> FastAPI:
    app = FastAPI(title=title, version="0.1.0")
    render = PROMPT_STYLES[prompt_style]

    def ready() -> LLMEngine:
        if engine is None or engine.tokenizer is None:
            raise HTTPException(status_code=503, detail="Model not loaded. Server is not ready.")
        return engine

    def decode(eng: LLMEngine, ids: List[int]) -> str:
        try:
            return eng.tokenizer.decode(ids, skip_special_tokens=True)
        except TypeError:
            return eng.tokenizer.decode(ids)

    def gen_cfg(temperature, top_p, max_tokens, top_k=0, rep=1.0, ignore_eos=False) -> GenerationConfig:
        t = 0.0 if temperature is None else float(temperature)
        return GenerationConfig(max_new_tokens=int(max_tokens or 16), do_sample=t > 0.0, temperature=max(t, 1e-5),
                                top_p=1.0 if top_p is None else float(top_p), top_k=int(top_k or 0),
                                repetition_penalty=float(rep or 1.0), ignore_eos=bool(ignore_eos))

    def usage_chunk(head, n_p, n_c) -> str:
        return "data: " + json.dumps({**head, "choices": [], "usage": {"prompt_tokens": n_p, "completion_tokens": n_c,
                                                                       "total_tokens": n_p + n_c}}) + "\n\n"

    @app.get("/health")
    @app.get("/healthz")
    async def health():
        ready()
        return {"status": "ok"}

    @app.get("/v1/models")
    async def models():
        eng = ready()
        return {"object": "list", "data": [{"id": eng.model_name, "object": "model", "created": int(time.time()),
                                            "owned_by": "llm_in_practise_b200", "max_model_len": eng.args.max_model_len}]}

    @app.get("/metrics")
    async def metrics():
        return PlainTextResponse(render_metrics(ready()), media_type="text/plain; version=0.0.4")

    @app.post("/v1/chat/completions")
    def chat_completions(req: ChatCompletionRequest):
        eng = ready()
        if req.stream and not allow_stream:
            raise HTTPException(status_code=501, detail="Streaming is not yet implemented.")
        prompt = render(req.messages)
        ids = eng.tokenizer.encode(prompt)
        try:
            r = eng.submit(ids, gen_cfg(req.temperature, req.top_p, req.max_tokens, req.top_k, req.repetition_penalty, req.ignore_eos))
        except ValueError as e:
            raise HTTPException(status_code=400, detail=str(e))
        cid, created = f"chatcmpl-{uuid.uuid4().hex}", int(time.time())
        if req.stream:
            def events():
                head = {"id": cid, "object": "chat.completion.chunk", "created": created, "model": req.model}
                yield "data: " + json.dumps({**head, "choices": [{"index": 0, "delta": {"role": "assistant"}, "finish_reason": None}]}) + "\n\n"
                sent, acc, stops, hit = "", [], _stops(req.stop), False
                hold = max((len(x) for x in stops), default=1) - 1     # characters that could still become a stop string
                for tok in r.tokens():
                    acc.append(tok)
                    text = decode(eng, acc)
                    if text.endswith("�"):          # incomplete UTF-8 sequence: wait for the next token
                        continue
                    text, hit = cut_at_stop(text, stops)
                    safe = text if hit else text[:max(len(text) - hold, len(sent))]
                    delta, sent = safe[len(sent):], safe
                    if delta:
                        yield "data: " + json.dumps({**head, "choices": [{"index": 0, "delta": {"content": delta}, "finish_reason": None}]}, ensure_ascii=False) + "\n\n"
                    if hit:
                        eng.abort(r)
                        break
                if not hit:
                    tail = cut_at_stop(decode(eng, acc), stops)[0][len(sent):]
                    if tail:
                        yield "data: " + json.dumps({**head, "choices": [{"index": 0, "delta": {"content": tail}, "finish_reason": None}]}, ensure_ascii=False) + "\n\n"
                yield "data: " + json.dumps({**head, "choices": [{"index": 0, "delta": {}, "finish_reason": "stop" if hit else r.finish_reason}]}) + "\n\n"
                if (req.stream_options or {}).get("include_usage"):
                    yield usage_chunk(head, len(ids), len(acc))
                yield "data: [DONE]\n\n"
            return StreamingResponse(events(), media_type="text/event-stream")
        r.done.wait()
        if r.error:
            raise HTTPException(status_code=500, detail=r.error)
        text = decode(eng, r.output_ids).replace("<|user|>", "").replace("<|assistant|>", "").strip()
        text, hit = cut_at_stop(text, _stops(req.stop))
        n_p, n_c = len(ids), len(r.output_ids)
        return {"id": cid, "object": "chat.completion", "created": created, "model": req.model,
                "choices": [{"index": 0, "message": {"role": "assistant", "content": text}, "finish_reason": "stop" if hit else r.finish_reason}],
                "usage": {"prompt_tokens": n_p, "completion_tokens": n_c, "total_tokens": n_p + n_c}}

    @app.post("/v1/completions")
    def completions(req: CompletionRequest):
        eng = ready()
        if req.stream and not allow_stream:
            raise HTTPException(status_code=501, detail="Streaming is not yet implemented.")
        prompts = [req.prompt] if isinstance(req.prompt, str) else list(req.prompt)
        cfg = gen_cfg(req.temperature, req.top_p, req.max_tokens, ignore_eos=req.ignore_eos)
        if req.stream and len(prompts) != 1:
            raise HTTPException(status_code=400, detail="streaming takes one prompt per request")
        try:
            reqs = [eng.submit(eng.tokenizer.encode(p), cfg) for p in prompts]
        except ValueError as e:
            raise HTTPException(status_code=400, detail=str(e))
        if req.stream:
            # the form ``vllm bench serve --backend vllm`` reads: text deltas in choices[0].text, then the usage chunk
            r = reqs[0]
            head = {"id": f"cmpl-{uuid.uuid4().hex}", "object": "text_completion", "created": int(time.time()), "model": req.model}

            def events():
                sent, acc, stops, hit = "", [], _stops(req.stop), False
                hold = max((len(x) for x in stops), default=1) - 1
                for tok in r.tokens():
                    acc.append(tok)
                    text = decode(eng, acc)
                    if text.endswith("\ufffd"):          # incomplete UTF-8 sequence: wait for the next token
                        continue
                    text, hit = cut_at_stop(text, stops)
                    safe = text if hit else text[:max(len(text) - hold, len(sent))]
                    delta, sent = safe[len(sent):], safe
                    if delta:
                        yield "data: " + json.dumps({**head, "choices": [{"index": 0, "text": delta, "finish_reason": None, "logprobs": None}]},
                                                    ensure_ascii=False) + "\n\n"
                    if hit:
                        eng.abort(r)
                        break
                tail = "" if hit else cut_at_stop(decode(eng, acc), stops)[0][len(sent):]
                yield "data: " + json.dumps({**head, "choices": [{"index": 0, "text": tail, "logprobs": None,
                                                                  "finish_reason": "stop" if hit else r.finish_reason}]},
                                            ensure_ascii=False) + "\n\n"
                if (req.stream_options or {}).get("include_usage"):
                    yield usage_chunk(head, len(r.prompt_ids), len(acc))
                yield "data: [DONE]\n\n"
            return StreamingResponse(events(), media_type="text/event-stream")
        choices, n_p, n_c = [], 0, 0
        for i, r in enumerate(reqs):
            r.done.wait()
            if r.error:
                raise HTTPException(status_code=500, detail=r.error)
            text, hit = cut_at_stop(decode(eng, r.output_ids), _stops(req.stop))
            choices.append({"index": i, "text": text, "finish_reason": "stop" if hit else r.finish_reason, "logprobs": None})
            n_p += len(r.prompt_ids)
            n_c += len(r.output_ids)
        return {"id": f"cmpl-{uuid.uuid4().hex}", "object": "text_completion", "created": int(time.time()), "model": req.model,
                "choices": choices, "usage": {"prompt_tokens": n_p, "completion_tokens": n_c, "total_tokens": n_p + n_c}}

    return app


def serve(engine: LLMEngine, host: str = "0.0.0.0", port: int = 8000, **kw) -> None:
    import uvicorn
    engine.start()
    try:
        uvicorn.run(build_app(engine, **kw), host=host, port=port)
    finally:
        engine.shutdown()            # tensor- / pipeline-parallel engines: also releases the follower ranks
