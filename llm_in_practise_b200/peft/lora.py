"""LoRA / QLoRA adapter management — the peft surface the reference scripts use:

    LoraConfig(task_type, r, lora_alpha, lora_dropout, target_modules, bias, modules_to_save)
    get_peft_model(model, config) -> model with .print_trainable_parameters(), .save_pretrained(dir)
    prepare_model_for_kbit_training(model)
    PeftModel.from_pretrained(base, adapter_dir);  .merge_and_unload()

(Fine-Tuning/qwen3-8b-lora-dist.py:118-130, Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:110-123,
Fine-Tuning/inferences.py:18-26, Scripts/fine-tuning/02-merge-lora-adapter-and-model.py:27-38.)

Adapters are stored on disk in the peft 0.17 layout (``adapter_model.safetensors`` with keys
``base_model.model.model.layers.N.self_attn.q_proj.lora_A.weight`` / ``lora_B.weight`` and
``adapter_config.json``) so peft and ``vllm --enable-lora`` can load them; in memory they live in one
flat fp32 arena and ride inside the fused projections (see ``ops/linear.py``).
"""
from __future__ import annotations

import json
import os
from dataclasses import asdict, dataclass, field
from typing import Dict, Iterable, List, Optional, Sequence, Union

import torch
from torch import nn

from ..ops.linear import FusedLinear
from ..utils.arena import ParamArena

ALL_LINEAR = ["q_proj", "k_proj", "v_proj", "o_proj", "gate_proj", "up_proj", "down_proj"]

# HF module path of each sub-projection inside a decoder layer
_HF_PARENT = {"q_proj": "self_attn", "k_proj": "self_attn", "v_proj": "self_attn", "o_proj": "self_attn",
              "gate_proj": "mlp", "up_proj": "mlp", "down_proj": "mlp"}


@dataclass
class LoraConfig:
    r: int = 8
    lora_alpha: float = 16
    lora_dropout: float = 0.0
    target_modules: Union[Sequence[str], str] = ("q_proj", "v_proj")
    bias: str = "none"
    task_type: str = "CAUSAL_LM"
    modules_to_save: Optional[List[str]] = None
    base_model_name_or_path: Optional[str] = None
    inference_mode: bool = False

    def targets(self) -> List[str]:
        if isinstance(self.target_modules, str):
            if self.target_modules in ("all", "all-linear"):
                return list(ALL_LINEAR)
            return [self.target_modules]
        return list(self.target_modules)

    def to_peft_dict(self) -> dict:
        return {
            "peft_type": "LORA", "task_type": str(self.task_type).split(".")[-1], "r": self.r,
            "lora_alpha": self.lora_alpha, "lora_dropout": self.lora_dropout,
            "target_modules": sorted(self.targets()), "bias": self.bias, "modules_to_save": self.modules_to_save,
            "base_model_name_or_path": self.base_model_name_or_path, "inference_mode": True,
            "fan_in_fan_out": False, "init_lora_weights": True, "use_rslora": False, "use_dora": False,
            "rank_pattern": {}, "alpha_pattern": {}, "layers_to_transform": None, "layers_pattern": None,
            "revision": None, "auto_mapping": None, "megatron_config": None, "megatron_core": "megatron.core",
            "loftq_config": {}, "layer_replication": None, "exclude_modules": None,
        }

    @classmethod
    def from_peft_dict(cls, d: dict) -> "LoraConfig":
        return cls(r=d["r"], lora_alpha=d["lora_alpha"], lora_dropout=d.get("lora_dropout", 0.0),
                   target_modules=list(d["target_modules"]), bias=d.get("bias", "none"),
                   task_type=d.get("task_type", "CAUSAL_LM"), modules_to_save=d.get("modules_to_save"),
                   base_model_name_or_path=d.get("base_model_name_or_path"))


def _fused_linears(model: nn.Module) -> Iterable:
    for name, m in model.named_modules():
        if isinstance(m, FusedLinear):
            yield name, m


def prepare_model_for_kbit_training(model: nn.Module, use_gradient_checkpointing: bool = False, **_) -> nn.Module:
    """Freeze the base model (peft: also upcasts norms to fp32 and enables input grads; here the norm
    kernels already accumulate in fp32 and the LoRA Function roots the graph, so freezing is all
    that is left)."""
    for p in model.parameters():
        p.requires_grad_(False)
    if use_gradient_checkpointing and hasattr(model, "gradient_checkpointing_enable"):
        model.gradient_checkpointing_enable()
    return model


def apply_lora(model: nn.Module, config: LoraConfig, pad_to: int = 2048, alloc=None) -> ParamArena:
    """Attach adapters in place; returns the flat fp32 arena holding every adapter parameter.

    The arena length is a multiple of 2048 (8 ranks x 256-value optimizer blocks) and, when a multi-GPU process
    group already exists, it is placed in symmetric peer memory right away, so a model prepared with
    ``get_peft_model`` can later be handed to the ZeRO engine / Trainer(deepspeed=...) as the reference scripts do."""
    targets = set(config.targets())
    symm_bufs = []
    if alloc is None:
        try:
            from ..parallel.symm import SymmBuffer, symm_available
            import torch.distributed as dist
            if symm_available() and dist.get_world_size() > 1 and next(model.parameters()).is_cuda:
                def alloc(numel, dtype, device):
                    b = SymmBuffer(numel, dtype, device)
                    symm_bufs.append(b)
                    return b.local
        except Exception:
            alloc = None
    for p in model.parameters():
        p.requires_grad_(False)
    dev = next(model.parameters()).device
    arena = ParamArena(dtype=torch.float32, device=dev, pad_to=pad_to)
    touched = []
    for name, lin in _fused_linears(model):
        t = {n: (config.r, float(config.lora_alpha)) for n in lin.part_names if n in targets}
        if t:
            lin.request_lora(arena, t, config.lora_dropout, name)
            touched.append(lin)
    if not touched:
        raise ValueError(f"no module matches target_modules={sorted(targets)}")
    arena.finalize(alloc=alloc)
    for lin in touched:
        lin.materialize_lora()
    if symm_bufs:
        arena._symm_bufs = symm_bufs
    model._lora_arena = arena
    model._lora_config = config
    return arena


class PeftModel(nn.Module):
    """Thin wrapper giving the peft call surface over a model that carries adapters."""

    def __init__(self, model: nn.Module, config: LoraConfig, arena: ParamArena):
        super().__init__()
        self.base_model = model
        self.peft_config = {"default": config}
        self.arena = arena

    def forward(self, *a, **k):
        return self.base_model(*a, **k)

    def __getattr__(self, name):
        try:
            return super().__getattr__(name)
        except AttributeError:
            return getattr(self.base_model, name)

    def get_nb_trainable_parameters(self):
        trainable = sum(p.numel() for p in self.base_model.parameters() if p.requires_grad)
        total = self.base_model.num_parameters() if hasattr(self.base_model, "num_parameters") else \
            sum(p.numel() for p in self.base_model.parameters())
        return trainable, total

    def print_trainable_parameters(self):
        t, a = self.get_nb_trainable_parameters()
        print(f"trainable params: {t:,d} || all params: {a:,d} || trainable%: {100 * t / max(a, 1):.4f}")

    # ---- adapter state dict in peft naming ---------------------------------------------------
    def adapter_state_dict(self) -> Dict[str, torch.Tensor]:
        sd = {}
        for name, lin in _fused_linears(self.base_model):
            if not lin.has_lora:
                continue
            prefix = name.rsplit(".", 1)[0]  # "...layers.N"
            for part in lin._lora_names:
                parent = _HF_PARENT.get(part)
                base = f"base_model.model.model.{prefix.split('model.')[-1] if prefix.startswith('model.') else prefix}"
                key = f"{base}.{parent}.{part}" if parent else f"{base}.{part}"
                sd[f"{key}.lora_A.weight"] = lin.lora_A[part].detach().clone()
                sd[f"{key}.lora_B.weight"] = lin.lora_Bt[part].detach().t().contiguous()
        return sd

    def load_adapter_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        mine = self.adapter_state_dict()
        missing = [k for k in mine if k not in sd]
        if missing:
            raise KeyError(f"adapter checkpoint lacks {len(missing)} tensors, e.g. {missing[0]}")
        with torch.no_grad():
            for name, lin in _fused_linears(self.base_model):
                if not lin.has_lora:
                    continue
                prefix = name.rsplit(".", 1)[0]
                for part in lin._lora_names:
                    parent = _HF_PARENT.get(part)
                    base = f"base_model.model.model.{prefix.split('model.')[-1] if prefix.startswith('model.') else prefix}"
                    key = f"{base}.{parent}.{part}" if parent else f"{base}.{part}"
                    lin.lora_A[part].copy_(sd[f"{key}.lora_A.weight"].to(lin.lora_A[part].device, torch.float32))
                    lin.lora_Bt[part].copy_(sd[f"{key}.lora_B.weight"].to(lin.lora_Bt[part].device, torch.float32).t())
                lin.invalidate_adapter_buffers()

    def save_pretrained(self, save_directory: str, **_) -> None:
        from ..io.safetensors import save_file
        os.makedirs(save_directory, exist_ok=True)
        sd = {k: v.to(torch.float32).cpu() for k, v in self.adapter_state_dict().items()}
        save_file(sd, os.path.join(save_directory, "adapter_model.safetensors"), metadata={"format": "pt"})
        with open(os.path.join(save_directory, "adapter_config.json"), "w") as f:
            json.dump(self.peft_config["default"].to_peft_dict(), f, indent=2, sort_keys=True)

    @classmethod
    def from_pretrained(cls, model: nn.Module, adapter_dir: str, is_trainable: bool = False, **_) -> "PeftModel":
        from ..io.safetensors import load_file
        with open(os.path.join(adapter_dir, "adapter_config.json")) as f:
            cfg = LoraConfig.from_peft_dict(json.load(f))
        path = os.path.join(adapter_dir, "adapter_model.safetensors")
        if not os.path.exists(path):
            path = os.path.join(adapter_dir, "peft_model.safetensors")
        arena = apply_lora(model, cfg)
        pm = cls(model, cfg, arena)
        pm.load_adapter_state_dict(load_file(path))
        if not is_trainable:
            for p in model.parameters():
                p.requires_grad_(False)
            model.eval()
        return pm

    def merge_and_unload(self) -> nn.Module:
        """Fold (alpha/r) B A into the dense base weights and drop the adapters (NF4 bases are
        expanded to bf16 first, as peft does for bnb layers)."""
        with torch.no_grad():
            for _, lin in _fused_linears(self.base_model):
                if not lin.has_lora:
                    continue
                w = lin.dense_weight().float()
                for part in lin._lora_names:
                    lo, hi = lin.part_range(part)
                    w[lo:hi] += lin.lora_delta(part).to(w.device)
                tgt = torch.bfloat16 if lin.nf4 is not None else lin.weight.dtype
                lin.nf4 = None
                lin.weight = nn.Parameter(w.to(tgt), requires_grad=False)
                lin._b_slices = []
                lin._lora_names = []
                for attr in ("lora_A", "lora_Bt"):
                    if hasattr(lin, attr):
                        delattr(lin, attr)
        return self.base_model


def get_peft_model(model: nn.Module, config: LoraConfig, pad_to: int = 1) -> PeftModel:
    arena = apply_lora(model, config, pad_to=pad_to)
    return PeftModel(model, config, arena)
