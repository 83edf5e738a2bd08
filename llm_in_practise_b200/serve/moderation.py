"""Llama-Guard moderation service: OpenAI ``/v1/moderations`` on top of a completions endpoint.

Behavioural match for the reference wrapper (Deployment/litellm-proxy/llama-guard-wrapper/app.py:22-72,
model_client.py:35-128, openai_moderation_map.py:5-64, schemas.py:4-5). The Llama-Guard system prompt, the category list
and the S1-S14 -> OpenAI category table below are TAKEN VERBATIM from the reference (a guard prompt has to match the
format the model was trained on, and the mapping is a constant table); the function names and signatures follow the
reference's so that its callers port one to one. The engine-backed backend (`engine_backend`) is this repository's own.

* ``POST /v1/moderations`` and ``/moderations`` with ``{"input": text}`` (400 on empty input), ``GET /healthz``,
  ``POST /v1/generate`` -> 501, optional ``X-API-KEY`` check (401) when ``WRAPPER_API_KEY`` is set;
* the text is wrapped in a Llama-3 header prompt whose system part asks for JSON ``{safe, categories[S1..S14],
  explanation}`` and posted to ``MODEL_API_URL + COMPLETIONS_PATH`` with ``max_tokens 256, temperature 0`` —
  up to 3 attempts, linear back-off 0.5 s x attempt on connection errors, HTTP 502 on a non-200 answer or when all
  attempts failed;
* the answer is parsed as JSON first, else as native Llama-Guard output (``unsafe\\nS1 S4``) with keyword
  heuristics (suicide -> S11, explosives -> S9, hate -> S10);
* S-codes map onto the 11 OpenAI categories; ``flagged = not safe``; scores are 0/1.

The completions backend can be a URL (any OpenAI-compatible server, e.g. ``serve.openai_server``) or an in-process
``LLMEngine`` (``backend=engine``), which removes the HTTP hop when guard and wrapper share a GPU box.
"""
from __future__ import annotations

import asyncio
import json
import logging
import os
import re
import uuid
from typing import Any, Awaitable, Callable, Dict, List, Optional

from fastapi import FastAPI, HTTPException, Request
from fastapi.responses import JSONResponse
from pydantic import BaseModel

logger = logging.getLogger("b200.moderation")

OPENAI_CATEGORIES = ["hate", "hate/threatening", "harassment", "harassment/threatening", "self-harm", "self-harm/intent",
                     "self-harm/instructions", "sexual", "sexual/minors", "violence", "violence/graphic"]

S_CODE_TO_OPENAI: Dict[str, List[str]] = {
    "S1": ["violence"], "S2": ["violence"], "S3": ["sexual"], "S4": ["sexual"], "S5": ["harassment"], "S6": ["self-harm"],
    "S7": ["self-harm/intent"], "S8": ["self-harm/instructions"], "S9": ["violence"], "S10": ["hate"],
    "S11": ["self-harm/instructions"], "S12": ["hate/threatening"], "S13": ["harassment/threatening"], "S14": ["sexual/minors"]}

GUARD_SYSTEM_PROMPT = (
    "You are Llama Guard 3, a safety classifier that must map text to a standard S-code taxonomy (S1..S14).\n"
    "You MUST respond in valid JSON ONLY (no additional text) with EXACT fields:\n"
    '  - "safe": boolean,\n'
    '  - "categories": an array of S-codes (e.g. ["S1"] or ["S11"]) - choose the most relevant codes in order of priority (top-3 max),\n'
    '  - "explanation": a short one-line justification why those categories apply.\n'
    "Rules:\n"
    "  1) Always pick S-codes from S1..S14. Up to 3 codes.\n"
    "  2) If the user asks for self-harm or suicide instructions, include S11.\n"
    "  3) If the user asks about making explosives or weapons, include S9 (Indiscriminate Weapons) or S2/S1 as appropriate.\n"
    "  4) If the user expresses hate toward a protected group, include S10.\n"
    "  5) NEVER output anything other than valid JSON.\n"
    "Example output:\n"
    '{"safe": false, "categories": ["S11"], "explanation": "User asks how to commit suicide."}\n')


class ModerationRequest(BaseModel):
    input: str


def build_guard_prompt(user_text: str) -> str:
    return ("<|begin_of_text|><|start_header_id|>system<|end_header_id|>\n" + GUARD_SYSTEM_PROMPT + "\n"
            "<|start_header_id|>user<|end_header_id|>\n" + user_text.strip() + "\n"
            "<|start_header_id|>assistant<|end_header_id|>\n")


def parse_native_scodes(text: str) -> Dict[str, Any]:
    lines = [ln.strip() for ln in (text or "").splitlines() if ln.strip()]
    if not lines:
        return {"safe": True, "categories": [], "explanation": ""}
    safe = "unsafe" not in lines[0].lower()
    codes = [c.upper() for c in re.findall(r"S\d+", " ".join(lines[1:]) or " ".join(lines))]
    if not codes and not safe:
        codes = ["UNSPECIFIED"]
    return {"safe": safe, "categories": codes, "explanation": ""}


_HEURISTICS = ((("suicid", "自杀"), "S11", "Heuristic: contains suicide/self-harm keywords."),
               (("bomb", "炸弹", "explosive", "爆炸"), "S9", "Heuristic: contains weapon/explosive keywords."),
               (("hate", "我讨厌", "黑人", "仇恨", "racist", "杀死他们因为"), "S10", "Heuristic: contains hate-speech keywords."))


def interpret_guard_output(user_text: str, text_out: str, resp: Any = None) -> Dict[str, Any]:
    out = (text_out or "").strip()
    raw = {"text": out, "resp": resp}
    try:
        p = json.loads(out)
        if not isinstance(p, dict):
            raise ValueError
        cats = [c.upper() for c in (p.get("categories") or []) if isinstance(c, str)]
        return {"safe": bool(p.get("safe", True)), "categories": cats, "explanation": p.get("explanation") or "", "raw": raw}
    except Exception:
        pass
    r = parse_native_scodes(out)
    low = user_text.lower()
    for keys, code, why in _HEURISTICS:           # first matching family only
        if any(k in low for k in keys):
            if code not in r["categories"]:
                r["categories"].insert(0, code)
                r["safe"] = False
                r["explanation"] = why
            break
    r["raw"] = raw
    return r


def make_openai_moderation_response(safe: bool, s_codes: List[str], explanation: str = "", raw: Optional[dict] = None,
                                    model: Optional[str] = None) -> Dict[str, Any]:
    cats = {k: False for k in OPENAI_CATEGORIES}
    scores = {k: 0.0 for k in OPENAI_CATEGORIES}
    for s in s_codes or []:
        for c in S_CODE_TO_OPENAI.get(s.upper(), []):
            cats[c], scores[c] = True, 1.0
    applied = {c: (["text"] if cats[c] else []) for c in OPENAI_CATEGORIES}
    return {"id": "modr-" + uuid.uuid4().hex, "model": model or os.getenv("MODEL_NAME", "llama-guard-3-8b"),
            "results": [{"flagged": not safe, "categories": cats, "category_scores": scores,
                         "category_applied_input_types": applied, "explanation": explanation or "", "raw": raw or {}}]}


# ------------------------------------------------------------------------------------------------
# completion back ends
# ------------------------------------------------------------------------------------------------
def http_backend(url: Optional[str] = None, timeout: float = 30.0, retries: int = 2, backoff: float = 0.5,
                 transport=None) -> Callable[[str, str], Awaitable[dict]]:
    import httpx
    url = url or (os.getenv("MODEL_API_URL", "http://llama-guard-vllm:8002").rstrip("/") + os.getenv("COMPLETIONS_PATH", "/v1/completions"))

    async def post(prompt: str, model: str) -> dict:
        payload = {"model": model, "prompt": prompt, "max_tokens": 256, "temperature": 0.0}
        last = None
        async with httpx.AsyncClient(timeout=timeout, transport=transport) as client:
            for attempt in range(retries + 1):
                try:
                    resp = await client.post(url, json=payload)
                except httpx.RequestError as e:
                    last = e
                    logger.warning("model server unreachable (attempt %d/%d): %s", attempt + 1, retries + 1, e)
                    await asyncio.sleep(backoff * (attempt + 1))
                    continue
                if resp.status_code != 200:
                    raise HTTPException(status_code=502, detail=f"Model server error: {resp.status_code} {resp.text}")
                return resp.json()
        raise HTTPException(status_code=502, detail=f"Error contacting model server after retries: {last}")
    return post


def engine_backend(engine) -> Callable[[str, str], Awaitable[dict]]:
    from ..inference.generate import GenerationConfig

    async def post(prompt: str, model: str) -> dict:
        r = engine.submit(engine.tokenizer.encode(prompt), GenerationConfig(max_new_tokens=256, do_sample=False))
        await asyncio.get_running_loop().run_in_executor(None, r.done.wait)
        if r.error:
            raise HTTPException(status_code=502, detail=r.error)
        return {"choices": [{"text": engine.tokenizer.decode(r.output_ids)}]}
    return post


def build_moderation_app(backend: Optional[Callable[[str, str], Awaitable[dict]]] = None, model_name: Optional[str] = None,
                         api_key: Optional[str] = None) -> FastAPI:
    app = FastAPI(title="Llama-Guard-3 OpenAI-Moderation Wrapper")
    backend = backend or http_backend()
    model_name = model_name or os.getenv("MODEL_NAME", "llama-guard-3")
    api_key = os.getenv("WRAPPER_API_KEY", "") if api_key is None else api_key

    @app.middleware("http")
    async def auth(request: Request, call_next):
        if api_key and request.headers.get("X-API-KEY", "") != api_key:
            return JSONResponse(status_code=401, content={"error": "Unauthorized"})
        return await call_next(request)

    @app.post("/v1/moderations")
    @app.post("/moderations")
    async def moderations(req: ModerationRequest):
        text = req.input
        if not text or not text.strip():
            raise HTTPException(status_code=400, detail="input is required")
        try:
            resp = await backend(build_guard_prompt(text), model_name)
        except HTTPException:
            raise
        except Exception as e:
            raise HTTPException(status_code=502, detail=str(e))
        try:
            ch = (resp.get("choices") or [{}])[0]
            out = ch.get("text") or (ch.get("message") or {}).get("content") or ""
        except Exception:
            out = json.dumps(resp)
        r = interpret_guard_output(text, out, resp)
        return JSONResponse(content=make_openai_moderation_response(bool(r["safe"]), r["categories"], r["explanation"], r["raw"], None))

    @app.post("/v1/generate")
    async def generate(request: Request):
        return JSONResponse(status_code=501, content={"error": "generate not implemented in this wrapper. Use /v1/moderations for safety classification."})

    @app.get("/healthz")
    async def healthz():
        return {"status": "ok"}

    return app
