"""``python -m llm_in_practise_b200.recipes.pretrain --recipe NAME [overrides]`` — one entry point for every GPT
pretraining script of the reference (launch with ``torchrun --nproc-per-node N -m …`` for the distributed ones).

Each preset names the reference file whose defaults and behaviour it reproduces; every flag of those scripts exists
here under the same name (``--epochs --batch_size --block_size --lr --n_layer --n_head --d_model --dropout
--weight_decay --seed --vocab_size --save_interval --save_dir --clip_grad_norm --latent_dim --num_experts --top_k
--num_shared --rope_theta --ds_config --local_rank --per_device_batch_size``).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from dataclasses import fields

from ..parallel.dist import cleanup_distributed
from ..train.pretrain import PretrainConfig, run_pretraining

_CFG_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "configs")

PRESETS = {
    # LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py:271-339
    "ddp_gpt_wikitext2": dict(strategy="ddp", dataset="wikitext-2-raw-v1", tokenizer="bert", batch_size=16, final_model_path="gpt_like_ddp_wikitext2.pth"),
    # …/ddp_basics/ddp_gpt_wikitext103.py (same, bigger corpus, 32 loader workers)
    "ddp_gpt_wikitext103": dict(strategy="ddp", dataset="wikitext-103-raw-v1", tokenizer="bert", batch_size=16, num_workers=32,
                                final_model_path="gpt_like_ddp_wikitext103.pth"),
    # …/fsdp_basics/fsdp_gpt_wikitext2.py:262-390
    "fsdp_gpt_wikitext2": dict(strategy="fsdp", tokenizer="bert", batch_size=8, save_dir="fsdp_checkpoints"),
    # …/fsdp_basics/fsdp2_gpt_wikitext2.py:255-367
    "fsdp2_gpt_wikitext2": dict(strategy="fsdp2", tokenizer="bert", batch_size=8, save_dir="fsdp2_checkpoints"),
    # …/transformer_basics/GPTLike_wikitext2.py:22-242 (whitespace BPE 3000, post-LN encoder without mask, Adam 1e-3)
    "gptlike_wikitext2_minimal": dict(model="simple", tokenizer="bpe", vocab_size=3000, d_model=128, n_head=4, n_layer=2, block_size=128,
                                      batch_size=32, lr=1e-3, optimizer="adam", weight_decay=0.0, epochs=5),
    # …/transformer_basics/GPTLike_wikitext2_fixed_pe.py:52-415
    "gptlike_wikitext2_fixed_pe": dict(tokenizer="bytelevel-bpe", vocab_size=30000, n_head=8, init="normal", pos="sinusoidal", val_fraction=0.1,
                                       step_lr_size=1, step_lr_gamma=0.95, clip_grad_norm=1.0, save_dir="checkpoints", early_stop_patience=3),
    # …/transformer_basics/GPTLike_wikitext2_learned_pe.py
    "gptlike_wikitext2_learned_pe": dict(tokenizer="bytelevel-bpe", vocab_size=30000, n_head=8, init="normal", pos="learned", val_fraction=0.1,
                                         step_lr_size=1, step_lr_gamma=0.95, clip_grad_norm=1.0, save_dir="checkpoints"),
    # …/transformer_basics/GPTLike_wikitext2_bert_tokenizer.py
    "gptlike_wikitext2_bert_tokenizer": dict(tokenizer="bert", n_head=8, init="normal", lr=1e-4, val_fraction=0.1, clip_grad_norm=1.0,
                                             save_dir="checkpoints"),
    # …/transformer_basics/DeepSeekLike_wikitext2.py (every expert runs on every token, masked)
    "deepseeklike_wikitext2": dict(model="deepseeklike", moe_impl="dense", tokenizer="bytelevel-bpe", vocab_size=30000, n_head=8,
                                   val_fraction=0.1, clip_grad_norm=1.0, save_dir="checkpoints"),
    # …/transformer_basics/DeepSeekLike_spare_MoE_wikitext2.py (index_select / index_add_ dispatch)
    "deepseeklike_sparse_moe_wikitext2": dict(model="deepseeklike", moe_impl="sparse", tokenizer="bytelevel-bpe", vocab_size=30000, n_head=8,
                                              val_fraction=0.1, clip_grad_norm=1.0, save_dir="checkpoints"),
    # LLM_Distributed_Trainning/PyTorch/temp/ddp_gpt_wikitext2.py:79-438 (byte tokenizer, fp16 AMP, no-decay groups, per-rank seeds)
    "ddp_gpt_bytes_amp": dict(strategy="ddp", tokenizer="byte", block_size=128, batch_size=8, amp="fp16", no_decay_groups=True,
                              seed_per_rank=True, save_dir="./checkpoints_ddp", resume=True),
    # …/temp/ddp_gpt_bpe_tokenizer.py / _02.py (rank-0 tokenizer training behind a barrier, resume, early stop)
    "ddp_gpt_bpe_tokenizer": dict(strategy="ddp", tokenizer="bytelevel-bpe", vocab_size=30000, amp="fp16", no_decay_groups=True,
                                  seed_per_rank=True, val_fraction=0.05, save_dir="./checkpoints_ddp_bpe", resume=True, early_stop_patience=3),
    # LLM_Distributed_Trainning/DeepSpeed/GPTLike-Base/GPTLike-Bert-Wikitext2.py:165-246
    "gptlike_base": dict(strategy="single", tokenizer="bert", final_model_path="gpt_like_wikitext2.pth"),
    # …/DeepSpeed/DeepSpeed-GPTLike-ZeRO-{1,2,3,Offload}/*.py + ds_config.json
    "deepspeed_zero1": dict(strategy="zero", tokenizer="bert", ds_config=os.path.join(_CFG_DIR, "ds_zero1.json"), save_dir="checkpoints_zero1"),
    "deepspeed_zero2": dict(strategy="zero", tokenizer="bert", ds_config=os.path.join(_CFG_DIR, "ds_zero2.json"), save_dir="checkpoints_zero2"),
    "deepspeed_zero3": dict(strategy="zero", tokenizer="bert", ds_config=os.path.join(_CFG_DIR, "ds_zero3.json"), save_dir="checkpoints_zero3"),
    "deepspeed_zero_offload": dict(strategy="zero", tokenizer="bert", ds_config=os.path.join(_CFG_DIR, "ds_zero3_offload.json"),
                                   save_dir="checkpoints_offload"),
    # …/DeepSpeed/DeepSpeed-GPTLike-Multihosts/*.py: 12 layers, 16 heads, d 1024, block 512, GPT-2 tokenizer + [PAD] -> 202.6 M params
    "deepspeed_zero2_multihosts": dict(strategy="zero", tokenizer="gpt2", n_layer=12, n_head=16, d_model=1024, block_size=512,
                                       ds_config=os.path.join(_CFG_DIR, "ds_zero2_multihosts.json"), save_dir="checkpoints_multihosts"),
}


def build_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0], formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--recipe", choices=sorted(PRESETS), default="gptlike_base")
    ap.add_argument("--list", action="store_true", help="print the presets and exit")
    for f in fields(PretrainConfig):
        if f.name == "ds_config":
            ap.add_argument("--ds_config", type=str, default=None)
            continue
        t = f.type if isinstance(f.type, type) else None
        if f.type in ("bool", bool):
            ap.add_argument(f"--{f.name}", type=lambda s: s.lower() in ("1", "true", "yes"), default=None)
        else:
            caster = {"int": int, "float": float, "str": str, "Optional[int]": int, "Optional[str]": str, "Optional[float]": float}.get(str(f.type), str)
            ap.add_argument(f"--{f.name}", type=caster, default=None)
    ap.add_argument("--per_device_batch_size", type=int, default=None, help="alias of --batch_size (temp/ddp_gpt_wikitext2.py)")
    # spellings of temp/ddp_gpt_bpe_tokenizer_02.py:55-78 and the DeepSpeed drafts
    ap.add_argument("--deepspeed_config", type=str, default=None, help="alias of --ds_config")
    ap.add_argument("--val_split_ratio", type=float, default=None, help="alias of --val_fraction")
    ap.add_argument("--gradient_accumulation_steps", type=int, default=None, help="alias of --grad_accum")
    ap.add_argument("--eval_interval", type=int, default=None, help="accepted; validation runs once per epoch")
    ap.add_argument("--train_tokenizer", action="store_true", help="train a byte-level BPE on the corpus (tokenizer=bytelevel-bpe)")
    ap.add_argument("--use_pretrained_tokenizer", type=str, default=None, help="name or path of a pretrained tokenizer (tokenizer=<value>)")
    ap.add_argument("--local_rank", type=int, default=None, help="accepted for launcher compatibility")
    return ap


def main(argv=None) -> dict:
    ap = build_parser()
    a = ap.parse_args(argv)
    if a.list:
        print(json.dumps(PRESETS, indent=1, default=str))
        return {}
    kw = dict(PRESETS[a.recipe])
    for f in fields(PretrainConfig):
        v = getattr(a, f.name, None)
        if v is not None:
            kw[f.name] = v
    if a.per_device_batch_size is not None:
        kw["batch_size"] = a.per_device_batch_size
    if a.deepspeed_config:
        kw["ds_config"] = a.deepspeed_config
    if a.val_split_ratio is not None:
        kw["val_fraction"] = a.val_split_ratio
    if a.gradient_accumulation_steps is not None:
        kw["grad_accum"] = a.gradient_accumulation_steps
    if a.train_tokenizer:
        kw["tokenizer"] = "bytelevel-bpe"
    if a.use_pretrained_tokenizer:
        kw["tokenizer"] = a.use_pretrained_tokenizer
    try:
        out = run_pretraining(PretrainConfig(**kw))
    finally:
        cleanup_distributed()
    return out


if __name__ == "__main__":
    main()
