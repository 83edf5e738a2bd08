"""HF checkpoint compatibility for the Qwen family: ``AutoModelForCausalLM.from_pretrained`` / ``save_pretrained``
as the reference uses them (Fine-Tuning/qwen3-8b-lora-dist.py:109-115, Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:95-107,
Scripts/fine-tuning/02-merge-lora-adapter-and-model.py:27-38).

Reads ``config.json`` + ``model.safetensors`` or sharded ``model-0000x-of-0000y.safetensors`` with
``model.safetensors.index.json`` (memory-mapped, one tensor at a time, straight into the fused q|k|v and gate|up
weights), and writes the same layout back with HF parameter names, so checkpoints round-trip with transformers / vLLM.
``BitsAndBytesConfig(load_in_4bit=True, bnb_4bit_quant_type="nf4", bnb_4bit_use_double_quant=True,
bnb_4bit_compute_dtype=bf16)`` quantises each decoder projection to NF4 on the device as it is loaded.
"""
from __future__ import annotations

import json
import os
from dataclasses import dataclass
from typing import Dict, Iterable, Optional, Tuple

import torch

from ..models.qwen3 import QWEN3_PRESETS, Qwen3Config, Qwen3ForCausalLM
from .safetensors import safe_open, save_file


@dataclass
class BitsAndBytesConfig:
    load_in_4bit: bool = False
    bnb_4bit_compute_dtype: torch.dtype = torch.bfloat16
    bnb_4bit_quant_type: str = "nf4"
    bnb_4bit_use_double_quant: bool = False
    bnb_4bit_quant_storage: Optional[torch.dtype] = None
    load_in_8bit: bool = False

    def __post_init__(self):
        if self.load_in_8bit:
            raise NotImplementedError("LLM.int8 is not used by the reference; only 4-bit NF4 is implemented")
        if self.load_in_4bit and self.bnb_4bit_quant_type != "nf4":
            raise NotImplementedError("only bnb_4bit_quant_type='nf4' is implemented")


def config_from_hf(d: dict) -> Qwen3Config:
    mt = d.get("model_type", "qwen3")
    hd = d.get("head_dim") or d["hidden_size"] // d["num_attention_heads"]
    # rope scaling: ``rope_scaling`` (transformers 4) or ``rope_parameters`` (transformers 5, which also holds rope_theta)
    rp = d.get("rope_parameters") or {}
    rs = d.get("rope_scaling") or ({k: v for k, v in rp.items() if k != "rope_theta"} if rp else None)
    if rs and rs.get("rope_type", rs.get("type", "default")) in ("default", None):
        rs = None
    if "rope_theta" not in d and "rope_theta" in rp:
        d = dict(d, rope_theta=rp["rope_theta"])
    return Qwen3Config(
        rope_scaling=rs,
        vocab_size=d["vocab_size"], hidden_size=d["hidden_size"], intermediate_size=d["intermediate_size"],
        num_hidden_layers=d["num_hidden_layers"], num_attention_heads=d["num_attention_heads"],
        num_key_value_heads=d.get("num_key_value_heads", d["num_attention_heads"]), head_dim=hd,
        rms_norm_eps=d.get("rms_norm_eps", 1e-6), rope_theta=d.get("rope_theta", 10000.0),
        max_position_embeddings=d.get("max_position_embeddings", 32768),
        tie_word_embeddings=d.get("tie_word_embeddings", False), model_type=mt,
        attention_bias=d.get("attention_bias", mt == "qwen2"), qk_norm=(mt == "qwen3"))


def config_to_hf(c: Qwen3Config, dtype=torch.bfloat16) -> dict:
    arch = "Qwen3ForCausalLM" if c.model_type == "qwen3" else "Qwen2ForCausalLM"
    return {"architectures": [arch], "model_type": c.model_type, "vocab_size": c.vocab_size, "hidden_size": c.hidden_size,
            "intermediate_size": c.intermediate_size, "num_hidden_layers": c.num_hidden_layers,
            "num_attention_heads": c.num_attention_heads, "num_key_value_heads": c.num_key_value_heads,
            "head_dim": c.head_dim, "rms_norm_eps": c.rms_norm_eps, "rope_theta": c.rope_theta,
            "rope_scaling": c.rope_scaling,
            "max_position_embeddings": c.max_position_embeddings, "tie_word_embeddings": c.tie_word_embeddings,
            "attention_bias": c.attention_bias, "hidden_act": "silu", "torch_dtype": str(dtype).replace("torch.", ""),
            "use_cache": True, "attention_dropout": 0.0}


class _Shards:
    """name -> tensor over one or many safetensors files."""

    def __init__(self, path: str):
        self.files: Dict[str, safe_open] = {}
        idx = os.path.join(path, "model.safetensors.index.json")
        if os.path.exists(idx):
            with open(idx) as f:
                self.map = json.load(f)["weight_map"]
        else:
            single = os.path.join(path, "model.safetensors")
            with safe_open(single) as f:
                self.map = {k: "model.safetensors" for k in f.keys()}
        self.path = path

    def get(self, name: str) -> torch.Tensor:
        fn = self.map[name]
        if fn not in self.files:
            self.files[fn] = safe_open(os.path.join(self.path, fn))
        return self.files[fn].get_tensor(name)

    def has(self, name: str) -> bool:
        return name in self.map


def hf_names(i: int) -> Dict[str, str]:
    p = f"model.layers.{i}."
    return {"q": p + "self_attn.q_proj", "k": p + "self_attn.k_proj", "v": p + "self_attn.v_proj", "o": p + "self_attn.o_proj",
            "gate": p + "mlp.gate_proj", "up": p + "mlp.up_proj", "down": p + "mlp.down_proj",
            "ln1": p + "input_layernorm.weight", "ln2": p + "post_attention_layernorm.weight",
            "qn": p + "self_attn.q_norm.weight", "kn": p + "self_attn.k_norm.weight"}


@torch.no_grad()
def load_hf_weights(model: Qwen3ForCausalLM, path: str, quantization_config: Optional[BitsAndBytesConfig] = None) -> None:
    sh = _Shards(path)
    dev, dt = model.embed_tokens.device, model.embed_tokens.dtype
    put = lambda dst, name: dst.copy_(sh.get(name).to(dev, dt))  # noqa: E731
    put(model.embed_tokens, "model.embed_tokens.weight")
    put(model.norm, "model.norm.weight")
    if model.lm_head is not None:
        put(model.lm_head, "lm_head.weight" if sh.has("lm_head.weight") else "model.embed_tokens.weight")
    q4 = quantization_config is not None and quantization_config.load_in_4bit
    for i, layer in enumerate(model.layers):
        n = hf_names(i)
        put(layer.input_layernorm, n["ln1"])
        put(layer.post_attention_layernorm, n["ln2"])
        if layer.q_norm is not None:
            put(layer.q_norm, n["qn"])
            put(layer.k_norm, n["kn"])
        for lin, parts in ((layer.qkv_proj, ("q", "k", "v")), (layer.o_proj, ("o",)),
                           (layer.gate_up_proj, ("gate", "up")), (layer.down_proj, ("down",))):
            w = torch.cat([sh.get(n[p] + ".weight").to(dev, dt) for p in parts], dim=0)
            if lin.weight is None:
                lin.weight = torch.nn.Parameter(w, requires_grad=False)
            else:
                lin.weight.data = w
            if lin.bias is not None:
                lin.bias.copy_(torch.cat([sh.get(n[p] + ".bias").to(dev, dt) for p in parts], dim=0))
            if q4:
                lin.quantize_nf4(quantization_config.bnb_4bit_use_double_quant)


def from_pretrained(path: str, dtype: torch.dtype = torch.bfloat16, torch_dtype: Optional[torch.dtype] = None,
                    device_map=None, device=None, quantization_config: Optional[BitsAndBytesConfig] = None,
                    low_cpu_mem_usage: bool = True, use_cache: bool = True, trust_remote_code: bool = False, **extra) -> Qwen3ForCausalLM:
    """``AutoModelForCausalLM.from_pretrained`` for Qwen3 / Qwen2-style checkpoints on local disk. ``path`` may also
    be a preset name (``qwen3-14b`` ...) to get random-init weights of that architecture (no network here).
    ``rope_scaling=<dict or None>`` overrides the checkpoint's own setting, as the config kwarg does in transformers
    (the reference passes ``rope_scaling=None`` to switch YaRN off for fine-tuning)."""
    dtype = torch_dtype or dtype
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
    if path in QWEN3_PRESETS:
        model = Qwen3ForCausalLM(QWEN3_PRESETS[path], dtype=dtype, device=device)
        if quantization_config is not None and quantization_config.load_in_4bit:
            model.quantize_nf4(quantization_config.bnb_4bit_use_double_quant)
        return model
    if not os.path.isfile(os.path.join(path, "config.json")):
        raise FileNotFoundError(f"{path!r} is neither a local checkpoint directory (no config.json there) nor one of the presets "
                                f"{sorted(QWEN3_PRESETS)}; hub ids cannot be downloaded here — pass the directory the weights are in")
    with open(os.path.join(path, "config.json")) as f:
        cfg = config_from_hf(json.load(f))
    if "rope_scaling" in extra:
        cfg.rope_scaling = extra["rope_scaling"]
    model = Qwen3ForCausalLM(cfg, dtype=dtype, device=device)
    load_hf_weights(model, path, quantization_config)
    model.name_or_path = path
    return model


@torch.no_grad()
def hf_state_dict(model: Qwen3ForCausalLM) -> Dict[str, torch.Tensor]:
    sd = {"model.embed_tokens.weight": model.embed_tokens.detach(), "model.norm.weight": model.norm.detach()}
    if model.lm_head is not None:
        sd["lm_head.weight"] = model.lm_head.detach()
    for i, layer in enumerate(model.layers):
        n = hf_names(i)
        sd[n["ln1"]], sd[n["ln2"]] = layer.input_layernorm.detach(), layer.post_attention_layernorm.detach()
        if layer.q_norm is not None:
            sd[n["qn"]], sd[n["kn"]] = layer.q_norm.detach(), layer.k_norm.detach()
        for lin, parts in ((layer.qkv_proj, ("q", "k", "v")), (layer.o_proj, ("o",)),
                           (layer.gate_up_proj, ("gate", "up")), (layer.down_proj, ("down",))):
            w = lin.dense_weight()
            off = 0
            for p, size in zip(parts, lin.part_sizes):
                sd[n[p] + ".weight"] = w[off:off + size]
                if lin.bias is not None:
                    sd[n[p] + ".bias"] = lin.bias.detach()[off:off + size]
                off += size
    return sd


def save_pretrained(model: Qwen3ForCausalLM, out_dir: str, max_shard_size: int = 5 * 1024 ** 3, tokenizer=None) -> None:
    os.makedirs(out_dir, exist_ok=True)
    sd = {k: v.cpu().contiguous() for k, v in hf_state_dict(model).items()}
    shards, cur, size = [], {}, 0
    for k, v in sd.items():
        nb = v.numel() * v.element_size()
        if cur and size + nb > max_shard_size:
            shards.append(cur)
            cur, size = {}, 0
        cur[k] = v
        size += nb
    shards.append(cur)
    if len(shards) == 1:
        save_file(shards[0], os.path.join(out_dir, "model.safetensors"), {"format": "pt"})
    else:
        wm, total = {}, 0
        for i, s in enumerate(shards):
            fn = f"model-{i + 1:05d}-of-{len(shards):05d}.safetensors"
            save_file(s, os.path.join(out_dir, fn), {"format": "pt"})
            for k, v in s.items():
                wm[k] = fn
                total += v.numel() * v.element_size()
        with open(os.path.join(out_dir, "model.safetensors.index.json"), "w") as f:
            json.dump({"metadata": {"total_size": total}, "weight_map": wm}, f, indent=2)
    with open(os.path.join(out_dir, "config.json"), "w") as f:
        json.dump(config_to_hf(model.config, model.embed_tokens.dtype), f, indent=2)
    if tokenizer is not None and hasattr(tokenizer, "save_pretrained"):
        tokenizer.save_pretrained(out_dir)
