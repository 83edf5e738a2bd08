"""``python -m llm_in_practise_b200.recipes.finetune --recipe NAME --model PATH [overrides]`` — the Qwen3 LoRA / QLoRA
fine-tuning scripts of the reference (``Fine-Tuning/*.py``, ``Scripts/fine-tuning/0{1,3}-*.py``) as presets over one
pipeline:

self-cognition rows -> placeholder substitution -> hand-rendered ChatML -> ``max_length`` padding ->
``DataCollatorForLanguageModeling(mlm=False)`` (labels = input_ids with pad -> -100, the reference's effective loss) ->
``from_pretrained`` (bf16 or NF4 ``BitsAndBytesConfig``) -> ``prepare_model_for_kbit_training`` -> ``get_peft_model`` ->
``TrainingArguments`` -> ``Trainer.train()`` -> ``save_model`` + tokenizer; ``<dir>_interrupted`` on an exception.

Launch the ``*-dist*`` presets with ``torchrun --nproc-per-node N -m llm_in_practise_b200.recipes.finetune …``.
``--model`` may be a local HF checkpoint directory or a preset name (``qwen3-14b`` …: random-init weights, no network).
"""
from __future__ import annotations

import argparse
import json
import os
from typing import Any, Dict

import torch

from ..data.chat import build_sft_features, load_self_cognition
from ..data.datasets import DataCollatorForLanguageModeling, DataCollatorWithPadding
from ..io.hf import BitsAndBytesConfig, from_pretrained
from ..parallel.dist import cleanup_distributed, setup_distributed
from ..peft import LoraConfig, get_peft_model, prepare_model_for_kbit_training
from ..trainer import Trainer, TrainingArguments

_CFG_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "configs")
_QV = ["q_proj", "v_proj"]

PRESETS: Dict[str, Dict[str, Any]] = {
    # Fine-Tuning/qwen3-8b-lora.py:114-204
    "qwen3-8b-lora": dict(model="qwen3-8b", nf4=False, lora=dict(r=16, lora_alpha=32, lora_dropout=0.05, target_modules=["q_proj", "k_proj", "v_proj", "o_proj"]),
                          args=dict(output_dir="./finetuned/qwen3-8b-lora", per_device_train_batch_size=2, gradient_accumulation_steps=4,
                                    num_train_epochs=3, learning_rate=1e-4, bf16=True, optim="adamw_torch", gradient_checkpointing=True,
                                    logging_steps=10, save_strategy="epoch", label_names=["labels"]), dynamic_padding=True),
    # Fine-Tuning/qwen3-8b-lora-dist.py:109-173 (DDP through the Trainer)
    "qwen3-8b-lora-dist": dict(model="qwen3-8b", nf4=False, lora=dict(r=8, lora_alpha=16, lora_dropout=0.1, target_modules=_QV),
                               args=dict(output_dir="./finetuned/qwen3-8b-lora-dist", per_device_train_batch_size=2, gradient_accumulation_steps=2,
                                         num_train_epochs=3, learning_rate=5e-5, weight_decay=0.01, bf16=True, optim="adamw_torch",
                                         gradient_checkpointing=True, gradient_checkpointing_kwargs={"use_reentrant": False},
                                         ddp_backend="nccl", ddp_find_unused_parameters=False, logging_steps=10, save_steps=100, save_total_limit=3)),
    # Fine-Tuning/qwen3-8b-qlora.py:86-169
    "qwen3-8b-qlora": dict(model="qwen3-8b", nf4=True, lora=dict(r=8, lora_alpha=16, lora_dropout=0.1, target_modules=_QV),
                           args=dict(output_dir="./finetuned/qwen3-8b-qlora", per_device_train_batch_size=4, num_train_epochs=3, learning_rate=1e-4,
                                     bf16=True, optim="paged_adamw_8bit", gradient_checkpointing=True, logging_steps=10, save_strategy="epoch")),
    # Fine-Tuning/qwen3-8b-qlora-dist.py
    "qwen3-8b-qlora-dist": dict(model="qwen3-8b", nf4=True, lora=dict(r=8, lora_alpha=16, lora_dropout=0.1, target_modules=_QV),
                                args=dict(output_dir="./finetuned/qwen3-8b-qlora-dist", per_device_train_batch_size=2, gradient_accumulation_steps=2,
                                          num_train_epochs=3, learning_rate=1e-4, bf16=True, optim="paged_adamw_8bit", gradient_checkpointing=True,
                                          ddp_backend="nccl", ddp_find_unused_parameters=False, ddp_timeout=1800, logging_steps=10, save_steps=100,
                                          save_total_limit=3)),
    # Fine-Tuning/qwen3-14b-qlora-dist.py:96-175
    "qwen3-14b-qlora-dist": dict(model="qwen3-14b", nf4=True, lora=dict(r=8, lora_alpha=16, lora_dropout=0.1, target_modules=_QV),
                                 args=dict(output_dir="./finetuned/qwen3-14b-qlora-dist", per_device_train_batch_size=2, gradient_accumulation_steps=2,
                                           num_train_epochs=3, learning_rate=1e-4, bf16=True, optim="paged_adamw_8bit", gradient_checkpointing=True,
                                           ddp_backend="nccl", ddp_find_unused_parameters=False, logging_steps=10, save_steps=100, save_total_limit=3)),
    # Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:95-202 + Fine-Tuning/ds_zero3_config.json — the headline configuration
    "qwen3-14b-qlora-dist-deepspeed": dict(model="qwen3-14b", nf4=True, lora=dict(r=8, lora_alpha=16, lora_dropout=0.1, target_modules=_QV),
                                           args=dict(output_dir="./finetuned/qwen3-14b-qlora-deepspeed", per_device_train_batch_size=2,
                                                     gradient_accumulation_steps=2, num_train_epochs=10, learning_rate=1e-4, bf16=True,
                                                     optim="paged_adamw_8bit", gradient_checkpointing=False, logging_steps=10, save_steps=10,
                                                     save_total_limit=3, deepspeed=os.path.join(_CFG_DIR, "ds_zero3_config.json"))),
    # Fine-Tuning/deepseek-r1-0528-qwen3-8b-qlora.dist.py (no space before <|im_end|>, 5 epochs)
    "deepseek-r1-0528-qwen3-8b-qlora-dist": dict(model="qwen3-8b", nf4=True, space_before_end=False,
                                                 lora=dict(r=8, lora_alpha=16, lora_dropout=0.1, target_modules=_QV),
                                                 args=dict(output_dir="./finetuned/deepseek-r1-0528-qwen3-8b-qlora", per_device_train_batch_size=2,
                                                           gradient_accumulation_steps=2, num_train_epochs=5, learning_rate=1e-4, bf16=True,
                                                           optim="paged_adamw_8bit", gradient_checkpointing=True, ddp_backend="nccl",
                                                           ddp_find_unused_parameters=False, logging_steps=10, save_steps=100, save_total_limit=3)),
    # Scripts/fine-tuning/01-deepseek-finetuning-example.py:96-143 (fp16, warmup 50, bs 4 x accum 8, r 8 / alpha 16 / 0.05 on q,v)
    "deepseek-1.5b-lora": dict(model="qwen3-1.7b", nf4=False, lora=dict(r=8, lora_alpha=16, lora_dropout=0.05, target_modules=_QV),
                               args=dict(output_dir="./deepseek-lora", per_device_train_batch_size=4, gradient_accumulation_steps=8, num_train_epochs=3,
                                         learning_rate=2e-4, fp16=True, warmup_steps=50, logging_steps=10, save_strategy="epoch"), labels_full=True),
}


_ALL_LINEAR = ["q_proj", "k_proj", "v_proj", "o_proj", "gate_proj", "up_proj", "down_proj"]


def preset_from_llamafactory_yaml(path: str) -> Dict[str, Any]:
    """A LLaMA-Factory ``*_lora_sft.yaml`` (Fine-Tuning/LLaMA-Factory/deepseek-r1-0528-qwen3_lora_sft.yaml:1-32:
    ``stage: sft, finetuning_type: lora, lora_target: all, cutoff_len, per_device_train_batch_size,
    gradient_accumulation_steps, learning_rate, num_train_epochs, lr_scheduler_type, warmup_ratio, bf16, output_dir,
    logging_steps, save_steps``) as a preset of this pipeline. LLaMA-Factory's defaults fill what the file leaves out
    (lora_rank 8, lora_alpha 2 x rank, lora_dropout 0); ``quantization_bit: 4`` selects the NF4 base."""
    import yaml
    with open(path, encoding="utf-8") as f:
        text = f.read()
    try:
        y = yaml.safe_load(text) or {}
    except yaml.YAMLError:
        # the shipped recipe is not strict YAML (a scalar `dataset:` followed by list items): read top-level `key: value`
        # lines only, which is all this front end uses
        y = {}
        for ln in text.splitlines():
            ln = ln.split("#", 1)[0].rstrip()
            if not ln or ln[0] in " \t-" or ":" not in ln:
                continue
            k, v = ln.split(":", 1)
            y[k.strip()] = yaml.safe_load(v.strip()) if v.strip() else None
    if y.get("stage", "sft") != "sft" or y.get("finetuning_type", "lora") != "lora":
        raise ValueError(f"{path}: only `stage: sft` with `finetuning_type: lora` is supported")
    tgt = y.get("lora_target", "all")
    targets = _ALL_LINEAR if tgt == "all" else [t.strip() for t in str(tgt).split(",") if t.strip()]
    r = int(y.get("lora_rank", 8))
    args = dict(output_dir=y.get("output_dir", "./finetuned/llamafactory-sft"),
                per_device_train_batch_size=int(y.get("per_device_train_batch_size", 1)),
                gradient_accumulation_steps=int(y.get("gradient_accumulation_steps", 8)),
                learning_rate=float(y.get("learning_rate", 1e-4)), num_train_epochs=float(y.get("num_train_epochs", 3.0)),
                lr_scheduler_type=str(y.get("lr_scheduler_type", "cosine")), warmup_ratio=float(y.get("warmup_ratio", 0.0)),
                bf16=bool(y.get("bf16", False)), fp16=bool(y.get("fp16", False)), logging_steps=int(y.get("logging_steps", 10)),
                save_steps=int(y.get("save_steps", 500)))
    return dict(model=y.get("model_name_or_path", "qwen3-8b"), nf4=int(y.get("quantization_bit", 0) or 0) == 4,
                lora=dict(r=r, lora_alpha=int(y.get("lora_alpha", 2 * r)), lora_dropout=float(y.get("lora_dropout", 0.0)),
                          target_modules=targets),
                args=args, max_length=int(y.get("cutoff_len", 2048)), dynamic_padding=True)


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0], formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--recipe", choices=sorted(PRESETS), default="qwen3-14b-qlora-dist-deepspeed")
    ap.add_argument("--llamafactory_yaml", default=None, help="a LLaMA-Factory *_lora_sft.yaml: replaces --recipe")
    ap.add_argument("--model", default=None, help="HF checkpoint directory or preset name")
    ap.add_argument("--tokenizer", default=None, help="tokenizer directory (defaults to --model; byte tokenizer if none is found)")
    ap.add_argument("--data", default=None, help="self_cognition.jsonl (default: the copy shipped with the reference tree)")
    ap.add_argument("--name", default="小马")
    ap.add_argument("--author", default="马哥教育")
    ap.add_argument("--max_length", type=int, default=512)
    ap.add_argument("--max_steps", type=int, default=-1)
    ap.add_argument("--output_dir", default=None)
    ap.add_argument("--local_rank", type=int, default=-1)
    a = ap.parse_args(argv)
    pre = preset_from_llamafactory_yaml(a.llamafactory_yaml) if a.llamafactory_yaml else PRESETS[a.recipe]
    if a.llamafactory_yaml and "--max_length" not in (argv if argv is not None else __import__("sys").argv):
        a.max_length = pre["max_length"]
    info = setup_distributed(a.local_rank)
    model_path = a.model or pre["model"]
    if a.llamafactory_yaml and a.model is None and not os.path.isdir(str(model_path)):
        from ..models.qwen3 import QWEN3_PRESETS
        if str(model_path) not in QWEN3_PRESETS:
            model_path = "qwen3-8b"                          # the YAML names a checkpoint directory that is not on this box

    # ---- tokenizer ----
    tok = None
    for cand in (a.tokenizer, model_path):
        if cand and os.path.isdir(cand):
            try:
                from ..data.tokenizers import load_pretrained_tokenizer
                tok = load_pretrained_tokenizer(cand, add_pad_token=False)
                if getattr(tok, "pad_token", None) is None:
                    tok.pad_token = tok.eos_token                      # pad = eos, as every FT script does
                break
            except Exception:
                tok = None
    if tok is None:
        from ..data.tokenizers import ByteTokenizer
        tok = ByteTokenizer()

    # ---- data ----
    rows = load_self_cognition(a.data, a.name, a.author)
    feats = build_sft_features(rows, tok, a.max_length, padding="longest" if pre.get("dynamic_padding") else "max_length",
                               space_before_end=pre.get("space_before_end", True))
    pad_id = getattr(tok, "pad_token_id", 0) or 0
    if pre.get("dynamic_padding"):
        collator = DataCollatorWithPadding(tok, pad_token_id=pad_id, pad_to_multiple_of=8)
    else:
        collator = DataCollatorForLanguageModeling(tok, mlm=False, pad_token_id=pad_id)

    # ---- model + adapters ----
    qcfg = BitsAndBytesConfig(load_in_4bit=True, bnb_4bit_compute_dtype=torch.bfloat16, bnb_4bit_quant_type="nf4",
                              bnb_4bit_use_double_quant=True) if pre["nf4"] else None
    model = from_pretrained(model_path, dtype=torch.bfloat16, device=info.device, quantization_config=qcfg, use_cache=False)
    targs = dict(pre["args"])
    if a.output_dir:
        targs["output_dir"] = a.output_dir
    if a.max_steps > 0:
        targs["max_steps"] = a.max_steps
    targs["local_rank"] = info.local_rank
    if not torch.cuda.is_available():
        targs["bf16"], targs["fp16"] = False, False
    if pre["nf4"]:
        model = prepare_model_for_kbit_training(model, use_gradient_checkpointing=targs.get("gradient_checkpointing", False))
    elif targs.get("gradient_checkpointing"):
        model.gradient_checkpointing_enable()
    peft_model = get_peft_model(model, LoraConfig(bias="none", task_type="CAUSAL_LM", **pre["lora"]))
    if info.rank == 0:
        peft_model.print_trainable_parameters()

    args = TrainingArguments(**targs)
    trainer = Trainer(model=peft_model, args=args, train_dataset=feats, data_collator=collator, tokenizer=tok)
    try:
        out = trainer.train()
        trainer.save_model(args.output_dir)
        if trainer.is_world_process_zero() and hasattr(tok, "save_pretrained"):
            tok.save_pretrained(args.output_dir)
        trainer.log_metrics("train", out.metrics)
        trainer.save_metrics("train", out.metrics)
        trainer.save_state()
        return out
    except BaseException:
        trainer.save_model(args.output_dir + "_interrupted")           # qwen3-8b-lora.py:190-204
        raise
    finally:
        cleanup_distributed()


if __name__ == "__main__":
    main()
