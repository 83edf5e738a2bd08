"""Chat formatting and SFT preprocessing of the Fine-Tuning / Scripts pipelines.

* ``render_chatml``            hand-rendered ChatML ``<|im_start|>{role}\\n{content} <|im_end|>\\n`` — note the space
                               before ``<|im_end|>`` that Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:49-57 emits
                               (``space_before_end=False`` is the deepseek-r1 variant, …/deepseek-r1-0528-qwen3-8b-qlora.dist.py:55).
* ``CHATML_JINJA``             the template of Fine-Tuning/templates/chatml_template.jinja (for ``--chat-template``).
* ``load_self_cognition``      the only dataset shipped in-tree (Fine-Tuning/LLaMA-Factory/self_cognition.jsonl, 108 rows);
                               ``{{NAME}}`` / ``{{AUTHOR}}`` replacement (…deepspeed.py:32-36).
* ``convert_self_cognition_to_alpaca``  Fine-Tuning/LLaMA-Factory/convert_self_cognition_to_alpaca.py:4-31.
* ``build_sft_features``       messages -> ChatML text -> token ids padded/truncated to ``max_length`` (…:66-72) plus the
                               assistant-only ``labels`` column the scripts *intend* (…:74-83); the collator decides
                               whether it is honoured (see data/datasets.py).
* ``deepseek_prompt`` / ``preprocess_generic``  the ``<|user|>\\n…\\n<|assistant|>\\n`` prompt and the schema-sniffing
                               preprocessing of Scripts/fine-tuning/01-deepseek-finetuning-example.py:45-84.
"""
from __future__ import annotations

import json
import os
from typing import Dict, Iterable, List, Optional, Sequence

IM_START, IM_END = "<|im_start|>", "<|im_end|>"

CHATML_JINJA = (
    "{% for message in messages %}"
    "{{ '<|im_start|>' + message['role'] + '\\n' + message['content'] + '<|im_end|>' + '\\n' }}"
    "{% endfor %}"
    "{% if add_generation_prompt %}{{ '<|im_start|>assistant\\n' }}{% endif %}"
)


def render_chatml(messages: Sequence[Dict[str, str]], add_generation_prompt: bool = False,
                  space_before_end: bool = True) -> str:
    sp = " " if space_before_end else ""
    out = "".join(f"{IM_START}{m['role']}\n{m['content']}{sp}{IM_END}\n" for m in messages)
    if add_generation_prompt:
        out += f"{IM_START}assistant\n"
    return out


def build_messages(query: str, response: Optional[str] = None, system: str = "You are a helpful assistant.",
                   history: Optional[Sequence[Sequence[str]]] = None) -> List[Dict[str, str]]:
    msgs = [{"role": "system", "content": system}] if system else []
    for q, a in history or ():
        msgs += [{"role": "user", "content": q}, {"role": "assistant", "content": a}]
    msgs.append({"role": "user", "content": query})
    if response is not None:
        msgs.append({"role": "assistant", "content": response})
    return msgs


def load_self_cognition(path: Optional[str] = None, name: str = "小马", author: str = "马哥教育") -> List[Dict[str, str]]:
    """Rows ``{"query", "response", "tag"}`` with the placeholders filled in."""
    cands = [path, os.environ.get("SELF_COGNITION_JSONL"),
             "/root/reference/Fine-Tuning/LLaMA-Factory/self_cognition.jsonl"]
    rows = []
    for p in cands:
        if p and os.path.exists(p):
            with open(p, encoding="utf-8") as f:
                rows = [json.loads(ln) for ln in f if ln.strip()]
            break
    if not rows:  # offline fallback: a tiny synthetic identity set of the same schema
        rows = [{"query": q, "response": "我是{{NAME}}，由{{AUTHOR}}训练的人工智能助手。", "tag": "zh"}
                for q in ("你是？", "你是谁？", "请介绍一下你自己", "who are you?")]
    out = []
    for r in rows:
        out.append({k: (v.replace("{{NAME}}", name).replace("{{AUTHOR}}", author) if isinstance(v, str) else v)
                    for k, v in r.items()})
    return out


def convert_self_cognition_to_alpaca(rows: Iterable[Dict[str, str]], name_zh="小马", author_zh="马哥教育",
                                     name_en="Pony", author_en="MageEdu") -> List[Dict[str, str]]:
    out = []
    for r in rows:
        zh = r.get("tag", "zh") == "zh"
        nm, au = (name_zh, author_zh) if zh else (name_en, author_en)
        fill = lambda s: s.replace("{{NAME}}", nm).replace("{{AUTHOR}}", au)  # noqa: E731
        out.append({"instruction": fill(r.get("query", "")), "input": "", "output": fill(r.get("response", ""))})
    return out


def build_sft_features(rows: Iterable[Dict[str, str]], tokenizer, max_length: int = 512, system: str = "You are a helpful assistant.",
                       padding: str = "max_length", space_before_end: bool = True) -> List[Dict[str, List[int]]]:
    """One feature dict per row: input_ids (+ attention_mask) and assistant-only labels (-100 elsewhere)."""
    feats = []
    pad_id = getattr(tokenizer, "pad_token_id", 0) or 0
    for r in rows:
        msgs = build_messages(r["query"], r["response"], system)
        text = render_chatml(msgs, space_before_end=space_before_end)
        prefix = render_chatml(msgs[:-1], add_generation_prompt=True, space_before_end=space_before_end)
        ids = tokenizer.encode(text)[:max_length]
        n_prefix = min(len(tokenizer.encode(prefix)), len(ids))
        labels = [-100] * n_prefix + ids[n_prefix:]
        mask = [1] * len(ids)
        if padding == "max_length" and len(ids) < max_length:
            k = max_length - len(ids)
            ids, labels, mask = ids + [pad_id] * k, labels + [-100] * k, mask + [0] * k
        feats.append({"input_ids": ids, "attention_mask": mask, "labels": labels})
    return feats


def deepseek_prompt(user: str, assistant: Optional[str] = None, history: Optional[Sequence[Sequence[str]]] = None) -> str:
    out = ""
    for q, a in history or ():
        out += f"<|user|>\n{q}\n<|assistant|>\n{a}\n"
    out += f"<|user|>\n{user}\n<|assistant|>\n"
    if assistant is not None:
        out += assistant
    return out


def preprocess_generic(example: Dict[str, str]) -> str:
    """Schema sniffing: instruction/input/output, question/answer, or prompt/response rows -> one training text."""
    if "instruction" in example:
        inp = example.get("input") or ""
        q = example["instruction"] + (("\n" + inp) if inp else "")
        return deepseek_prompt(q, example.get("output", ""))
    if "question" in example:
        return deepseek_prompt(example["question"], example.get("answer", ""))
    if "prompt" in example:
        return deepseek_prompt(example["prompt"], example.get("response", example.get("completion", "")))
    if "query" in example:
        return deepseek_prompt(example["query"], example.get("response", ""))
    return str(example.get("text", ""))
