"""Import namespace for moving a reference script over with its imports changed and little else.

The fine-tuning / quantisation / inference scripts of the reference start with

    from transformers import AutoModelForCausalLM, AutoTokenizer, BitsAndBytesConfig, TrainingArguments, Trainer, \\
        DataCollatorForLanguageModeling, TextIteratorStreamer, set_seed
    from peft import LoraConfig, get_peft_model, TaskType, prepare_model_for_kbit_training, PeftModel
    from gptqmodel import GPTQModel, QuantizeConfig
    from llmcompressor import oneshot; from llmcompressor.modifiers... import GPTQModifier, AWQModifier

(Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:1-14, Quantization/GPTQModel/quantize_qwen3_4b_gptq.py:1-8,
Quantization/LLM-Compressor/*/quantize_*.py, Scripts/inference/06-*.py). Every one of those names is importable from
here and resolves to this framework's implementation:

    from llm_in_practise_b200.compat import AutoModelForCausalLM, AutoTokenizer, BitsAndBytesConfig, ...

``AutoModelForCausalLM.from_pretrained`` reads a local HF checkpoint directory (dense, GPTQ, AutoAWQ or compressed-tensors —
detected from its config) or a preset name; ``AutoTokenizer`` is the real ``transformers`` tokenizer (host-side, not a hot
path). ``tests/test_api_surface_cpu.py`` checks that the keyword arguments and methods the reference's scripts use on
these objects exist.
"""
from __future__ import annotations

import enum
import os
from typing import Optional

import torch

from .data.datasets import DataCollatorForLanguageModeling, DataCollatorWithPadding, set_seed  # noqa: F401
from .inference.generate import GenerationConfig, TextIteratorStreamer  # noqa: F401
from .io.hf import BitsAndBytesConfig  # noqa: F401
from .peft import LoraConfig, PeftModel, get_peft_model, prepare_model_for_kbit_training  # noqa: F401
from .quant import AWQModifier, GPTQModel, GPTQModifier, QuantizationModifier, QuantizeConfig, oneshot  # noqa: F401
from .trainer import EvalPrediction, Trainer, TrainingArguments  # noqa: F401


class TaskType(str, enum.Enum):
    """``peft.TaskType``: ``LoraConfig(task_type=TaskType.CAUSAL_LM)``."""
    CAUSAL_LM = "CAUSAL_LM"
    SEQ_CLS = "SEQ_CLS"
    SEQ_2_SEQ_LM = "SEQ_2_SEQ_LM"
    TOKEN_CLS = "TOKEN_CLS"
    FEATURE_EXTRACTION = "FEATURE_EXTRACTION"


class AutoModelForCausalLM:
    @staticmethod
    def from_pretrained(pretrained_model_name_or_path: str, *args, **kw):
        """Local checkpoint directory or preset name -> ``Qwen3ForCausalLM``. ``torch_dtype`` / ``dtype``, ``device_map``
        (``"auto"`` / ``None`` / ``{"": i}``: one device per process here), ``quantization_config``, ``rope_scaling``,
        ``use_cache``, ``trust_remote_code``, ``low_cpu_mem_usage`` are taken as transformers takes them."""
        from .io.hf import from_pretrained
        from .quant import detect_quant_format, load_quantized
        path = pretrained_model_name_or_path
        dtype = kw.pop("torch_dtype", None) or kw.pop("dtype", None) or torch.bfloat16
        if dtype == "auto":
            dtype = torch.bfloat16
        dm = kw.pop("device_map", None)
        device = kw.pop("device", None)
        if device is None and isinstance(dm, dict) and dm:
            d = next(iter(dm.values()))
            device = torch.device("cuda", d) if isinstance(d, int) else torch.device(d)
        if os.path.isdir(path) and detect_quant_format(path) is not None:
            return load_quantized(path, dtype=dtype, device=device)
        return from_pretrained(path, dtype=dtype, device=device, **kw)


class AutoModelForSequenceClassification:
    @staticmethod
    def from_pretrained(pretrained_model_name_or_path: str, num_labels: int = 2, **kw):
        """The encoder + classification head of HF_Basics/trainer_demo.py:61-64. BERT weights are not converted: the model is
        built with the vocabulary of the tokenizer found at the path (or 30522) and trained from scratch."""
        from .models.classifier import EncoderClassifier
        vocab = 30522
        try:
            vocab = len(AutoTokenizer.from_pretrained(pretrained_model_name_or_path))
        except Exception:
            pass
        dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
        return EncoderClassifier(vocab_size=vocab, num_labels=num_labels, device=dev,
                                 **{k: v for k, v in kw.items() if k in ("d_model", "n_layer", "n_head", "max_len", "dropout")})


class AutoTokenizer:
    @staticmethod
    def from_pretrained(pretrained_model_name_or_path: str, use_fast: bool = True, trust_remote_code: bool = False,
                        padding_side: Optional[str] = None, **_):
        from .data.tokenizers import load_pretrained_tokenizer
        tok = load_pretrained_tokenizer(pretrained_model_name_or_path, add_pad_token=False)
        if padding_side:
            tok.padding_side = padding_side
        return tok


# ---- torch.distributed / deepspeed names of the hand-written training loops (LLM_Distributed_Trainning/PyTorch/**) ----------
from .data.datasets import DistributedSampler  # noqa: E402,F401
from .parallel.ddp import DistributedDataParallel  # noqa: E402,F401
from .parallel.fsdp import (CPUOffload, CPUOffloadPolicy, FSDP, MixedPrecision, MixedPrecisionPolicy, ShardingStrategy,  # noqa: E402,F401
                            StateDictOptions, fully_shard, get_state_dict, init_device_mesh, transformer_auto_wrap_policy)

DDP = DistributedDataParallel
FullyShardedDataParallel = FSDP


class DeepSpeedCPUAdam(torch.optim.AdamW):
    """``deepspeed.ops.adam.DeepSpeedCPUAdam(model.parameters(), lr=…)`` (…/deepspeed_basics/GPTLike_wikitext2_ZeRO-Offload.py:
    27,228). Handed to ``deepspeed.initialize`` it only carries the hyper-parameters: the engine keeps the fp32 masters and the
    AdamW state in pinned host memory and steps them with the C++ host optimizer (``csrc/host/cpu_adam.cpp``) when the config
    says ``offload_optimizer: cpu``. Used on its own it is torch's AdamW."""

    def __init__(self, model_params, lr: float = 1e-3, bias_correction: bool = True, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0, amsgrad: bool = False, adamw_mode: bool = True, fp32_optimizer_states: bool = True):
        super().__init__(model_params, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay if adamw_mode else 0.0, amsgrad=amsgrad)


class _DeepSpeedOpsAdam:
    DeepSpeedCPUAdam = DeepSpeedCPUAdam


class _DeepSpeedOps:
    adam = _DeepSpeedOpsAdam


class deepspeed:  # noqa: N801 - spelled like the module it stands in for
    """``import deepspeed`` for the ZeRO training loops: ``deepspeed.init_distributed()``, ``engine, optimizer, _, scheduler =
    deepspeed.initialize(args=…, model=…, optimizer=…, model_parameters=…, config=…)`` (the engine has ``backward`` / ``step`` /
    ``save_checkpoint`` / ``load_checkpoint``), ``deepspeed.ops.adam.DeepSpeedCPUAdam``."""
    ops = _DeepSpeedOps

    @staticmethod
    def init_distributed(dist_backend: Optional[str] = None, **kw):
        from .parallel.zero import init_distributed
        return init_distributed(dist_backend, **kw)

    @staticmethod
    def initialize(*a, **kw):
        from .parallel.zero import ds_initialize
        return ds_initialize(*a, **kw)

    @staticmethod
    def shutdown() -> None:
        from .parallel.dist import cleanup_distributed
        cleanup_distributed()


# ---- ``from ray import serve`` / ``from ray.serve.llm import LLMConfig, build_openai_app`` (Deployment/Ray/serve_run_examples) ----
from .serve.deploy import LLMConfig, build_openai_app  # noqa: E402,F401


class serve:  # noqa: N801
    """``serve.run(llm_app, route_prefix="/")``; ``serve deploy file.yaml`` is ``recipes.tools serve-deploy file.yaml``."""

    @staticmethod
    def run(app, route_prefix: str = "/", blocking: bool = True, host: str = "0.0.0.0", port: int = 8000, **kw):
        from .serve.deploy import run
        return run(app, host=host, port=port, route_prefix=route_prefix, blocking=blocking, **kw)


# ---- ``import evaluate`` (HF_Basics/trainer_demo.py:66-75, accelerate_demo.py:58-66): the classification metrics ----------------
class _Metric:
    def __init__(self, name: str):
        if name not in ("accuracy", "f1", "precision", "recall"):
            raise ValueError(f"metric {name!r} is not provided (accuracy, f1, precision, recall are)")
        self.name = name

    def compute(self, predictions=None, references=None, average: str = "binary", pos_label: int = 1, **_):
        import numpy as np
        p, r = np.asarray(predictions).reshape(-1), np.asarray(references).reshape(-1)
        if self.name == "accuracy":
            return {"accuracy": float((p == r).mean()) if len(r) else 0.0}
        labels = [pos_label] if average == "binary" else sorted(set(r.tolist()) | set(p.tolist()))
        per, weights = [], []
        for c in labels:
            tp, fp, fn = float(((p == c) & (r == c)).sum()), float(((p == c) & (r != c)).sum()), float(((p != c) & (r == c)).sum())
            prec = tp / (tp + fp) if tp + fp else 0.0
            rec = tp / (tp + fn) if tp + fn else 0.0
            f1 = 2 * prec * rec / (prec + rec) if prec + rec else 0.0
            per.append({"precision": prec, "recall": rec, "f1": f1}[self.name])
            weights.append(float((r == c).sum()))
        if average == "micro" and self.name in ("precision", "recall", "f1"):
            return {self.name: float((p == r).mean())}          # single-label: micro P = R = F1 = accuracy
        if average == "weighted":
            tot = sum(weights) or 1.0
            return {self.name: float(sum(v * w for v, w in zip(per, weights)) / tot)}
        return {self.name: float(sum(per) / max(1, len(per)))}


class evaluate:  # noqa: N801
    @staticmethod
    def load(name: str, *_, **__) -> _Metric:
        return _Metric(name)
