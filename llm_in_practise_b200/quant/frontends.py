"""``GPTQModel`` / ``oneshot`` front-ends and the perplexity checks of the quantisation lesson.

Reference usage: Quantization/GPTQModel/quantize_qwen3_4b_gptq.py:16-46 (load -> quantize -> save),
Quantization/GPTQModel/inference_qwen3_4b_gptq.py:18-68 (load quantised -> chat-template prompt -> generate),
Quantization/LLM-Compressor/GPTQ/quantize_qwen3_4b_gptq.py:19-62 and AWQ/quantize_qwen3_4b_awq.py:17-58 (oneshot),
Quantization/LLM-Compressor/GPTQ/eval_qwen3_4b_gptq.py:30-62 (self-perplexity over greedy continuations).
"""
from __future__ import annotations

import math
import os
from typing import Iterable, List, Optional, Sequence, Tuple, Union

import torch

from ..inference.generate import GenerationConfig, Generator
from ..io.hf import from_pretrained
from ..models.qwen3 import Qwen3ForCausalLM
from .api import AWQModifier, GPTQModifier, QuantizeConfig, _Modifier, tokenize_calibration
from .pipeline import detect_quant_format, load_quantized, quantize_model, save_quantized


class GPTQModel:
    """gptqmodel-style handle: ``GPTQModel.load(path, quantize_config)`` on a dense checkpoint prepares it for
    ``quantize``; ``GPTQModel.load(quantised_dir)`` returns a ready-to-generate model."""

    def __init__(self, model: Qwen3ForCausalLM, quantize_config: Optional[QuantizeConfig], tokenizer=None):
        self.model, self.quantize_config, self.tokenizer = model, quantize_config, tokenizer
        self._gen: Optional[Generator] = None

    @classmethod
    def load(cls, path: str, quantize_config: Optional[QuantizeConfig] = None, tokenizer=None, dtype=torch.bfloat16,
             device=None, trust_remote_code: bool = False, **_) -> "GPTQModel":
        if quantize_config is None and os.path.isdir(path) and detect_quant_format(path) is not None:
            m = load_quantized(path, dtype, device)
            return cls(m, m.quant_config, tokenizer)
        return cls(from_pretrained(path, dtype=dtype, device=device), quantize_config or QuantizeConfig(), tokenizer)

    from_quantized = load

    @property
    def device(self):
        return self.model.embed_tokens.device

    def quantize(self, calibration_dataset, batch_size: int = 1, tokenizer=None, max_seq_length: int = 2048,
                 progress: bool = False) -> "GPTQModel":
        tok = tokenizer or self.tokenizer
        batches = tokenize_calibration(calibration_dataset, tok, max_seq_length, None, batch_size, self.device)
        quantize_model(self.model, batches, self.quantize_config, progress)
        return self

    def save(self, out_dir: str, fmt: str = "gptq") -> None:
        save_quantized(self.model, out_dir, fmt, self.tokenizer)

    save_quantized = save

    def generate(self, input_ids=None, **kw) -> torch.Tensor:
        if self._gen is None:
            self._gen = Generator(self.model, self.tokenizer, kw.pop("max_len", 4096))
        kw.pop("attention_mask", None)
        return self._gen.generate(input_ids.to(self.device), **kw)

    def __call__(self, *a, **kw):
        return self.model(*a, **kw)


def oneshot(model: Union[str, Qwen3ForCausalLM], recipe: Union[_Modifier, Sequence[_Modifier]], dataset=None, tokenizer=None,
            output_dir: Optional[str] = None, max_seq_length: int = 2048, num_calibration_samples: Optional[int] = 512,
            dtype=torch.bfloat16, device=None, save_format: str = "compressed-tensors", progress: bool = False,
            **_) -> Qwen3ForCausalLM:
    """llm-compressor ``oneshot``: apply the recipe's single weight-quantisation modifier with ``dataset`` as the
    calibration set and write a compressed-tensors checkpoint into ``output_dir``."""
    mods = [recipe] if isinstance(recipe, _Modifier) else list(recipe)
    if len(mods) != 1:
        raise NotImplementedError("the reference recipes hold exactly one modifier (GPTQModifier or AWQModifier)")
    mod = mods[0]
    tg = [mod.targets] if isinstance(mod.targets, str) else list(mod.targets)
    if tg != ["Linear"]:
        raise NotImplementedError("targets must be 'Linear' (all decoder projections), as in the reference")
    if "lm_head" not in list(mod.ignore):
        raise NotImplementedError("lm_head is always kept in 16 bit (the reference passes ignore=['lm_head'])")
    if isinstance(model, str):
        model = from_pretrained(model, dtype=dtype, device=device)
    qc = mod.config()
    dev = model.embed_tokens.device
    batches = [] if qc.method == "rtn" else tokenize_calibration(dataset, tokenizer, max_seq_length, num_calibration_samples, 1, dev)
    quantize_model(model, batches, qc, progress)
    if output_dir is not None:
        save_quantized(model, output_dir, save_format, tokenizer)
    return model


# ------------------------------------------------------------------------------------------------
# evaluation
# ------------------------------------------------------------------------------------------------
@torch.no_grad()
def perplexity(model: Qwen3ForCausalLM, batches: Iterable[torch.Tensor]) -> float:
    """Token-level perplexity of given texts (teacher forcing)."""
    was = model.training
    model.eval()
    nll, n = 0.0, 0
    for ids in batches:
        ids = ids.to(model.embed_tokens.device)
        k = ids.numel() - ids.shape[0]
        nll += float(model(ids, labels=ids)["loss"]) * k
        n += k
    model.train(was)
    return math.exp(nll / max(n, 1))


@torch.no_grad()
def self_perplexity(gen: Generator, prompts: Sequence[torch.Tensor], max_tokens: int = 256,
                    eos_token_id=None) -> Tuple[float, float]:
    """The lesson's "eternal PPL": greedy-decode up to ``max_tokens`` per prompt and report mean/std over prompts of
    exp(-mean logprob of the chosen tokens, first token skipped)."""
    vals = []
    for ids in prompts:
        _, lp = gen.generate(ids, GenerationConfig(max_new_tokens=max_tokens, do_sample=False, eos_token_id=eos_token_id),
                             return_logprobs=True)
        lp = lp[0, 1:]
        if lp.numel():
            vals.append(math.exp(-float(lp.mean())))
    if not vals:
        return float("nan"), float("nan")
    t = torch.tensor(vals)
    return float(t.mean()), float(t.std(unbiased=False))
