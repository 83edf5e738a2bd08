"""DeepSpeed-style JSON configuration, including HF's ``"auto"`` placeholders.

Accepts every key the reference's config files use — LLM_Distributed_Trainning/DeepSpeed/*/ds_config.json
(stages 1/2/3/offload, fp16 dynamic loss scaling, AdamW, WarmupLR) and Fine-Tuning/ds_zero3_config.json
(bf16, stage 3, ``"auto"`` bucket sizes, ``stage3_gather_16bit_weights_on_model_save``,
gradient_clipping, wall_clock_breakdown). Unknown keys are kept in ``extra`` and reported, never
silently dropped; ``offload_*: {"device": "cpu"}`` is honoured (pinned-host optimizer state with the
C++ CPU AdamW) rather than ignored.
"""
from __future__ import annotations

import json
from dataclasses import dataclass, field
from typing import Any, Dict, Optional, Union

AUTO = "auto"


def _is_auto(v) -> bool:
    return isinstance(v, str) and v.lower() == AUTO


@dataclass
class ZeroConfig:
    stage: int = 0
    overlap_comm: bool = True
    contiguous_gradients: bool = True
    reduce_scatter: bool = True
    allgather_partitions: bool = True
    reduce_bucket_size: Union[int, str] = int(5e8)
    allgather_bucket_size: Union[int, str] = int(5e8)
    stage3_prefetch_bucket_size: Union[int, str] = int(5e7)
    stage3_param_persistence_threshold: Union[int, str] = int(1e5)
    stage3_max_live_parameters: int = int(1e9)
    stage3_max_reuse_distance: int = int(1e9)
    sub_group_size: int = int(1e9)
    stage3_gather_16bit_weights_on_model_save: bool = False
    offload_optimizer_device: str = "none"
    offload_param_device: str = "none"
    pin_memory: bool = True


@dataclass
class FP16Config:
    enabled: bool = False
    loss_scale: float = 0.0  # 0 = dynamic
    initial_scale_power: int = 16
    loss_scale_window: int = 1000
    hysteresis: int = 2
    min_loss_scale: float = 1.0


@dataclass
class DSConfig:
    zero: ZeroConfig = field(default_factory=ZeroConfig)
    fp16: FP16Config = field(default_factory=FP16Config)
    bf16: bool = False
    optimizer: Optional[Dict[str, Any]] = None
    scheduler: Optional[Dict[str, Any]] = None
    gradient_accumulation_steps: Union[int, str] = 1
    gradient_clipping: Union[float, str] = 0.0
    train_batch_size: Union[int, str, None] = None
    train_micro_batch_size_per_gpu: Union[int, str, None] = None
    steps_per_print: int = 10
    wall_clock_breakdown: bool = False
    extra: Dict[str, Any] = field(default_factory=dict)

    # ------------------------------------------------------------------
    @classmethod
    def from_any(cls, cfg: Union[str, Dict[str, Any], "DSConfig", None]) -> "DSConfig":
        if cfg is None:
            return cls()
        if isinstance(cfg, DSConfig):
            return cfg
        if isinstance(cfg, str):
            with open(cfg) as f:
                cfg = json.load(f)
        d = dict(cfg)
        out = cls()
        z = dict(d.pop("zero_optimization", {}) or {})
        zc = out.zero
        for k in list(z.keys()):
            if k == "offload_optimizer":
                zc.offload_optimizer_device = (z[k] or {}).get("device", "none") or "none"
                zc.pin_memory = bool((z[k] or {}).get("pin_memory", zc.pin_memory))
            elif k == "offload_param":
                zc.offload_param_device = (z[k] or {}).get("device", "none") or "none"
            elif hasattr(zc, k):
                setattr(zc, k, z[k])
            else:
                out.extra["zero_optimization." + k] = z[k]
        f16 = d.pop("fp16", None)
        if f16:
            for k, v in f16.items():
                if hasattr(out.fp16, k):
                    setattr(out.fp16, k, v)
                else:
                    out.extra["fp16." + k] = v
        b16 = d.pop("bf16", None)
        if b16:
            out.bf16 = bool(b16.get("enabled", False)) if isinstance(b16, dict) else bool(b16)
        for k in ("optimizer", "scheduler", "gradient_accumulation_steps", "gradient_clipping", "train_batch_size",
                  "train_micro_batch_size_per_gpu", "steps_per_print", "wall_clock_breakdown"):
            if k in d:
                setattr(out, k, d.pop(k))
        out.extra.update(d)
        return out

    def resolve_auto(self, *, hidden_size: Optional[int] = None, world_size: int = 1,
                     micro_batch: Optional[int] = None, grad_accum: Optional[int] = None,
                     max_grad_norm: Optional[float] = None) -> "DSConfig":
        """Fill "auto" the way HF's DeepSpeed integration does: bucket sizes from ``hidden_size``
        (reduce = h*h, prefetch = 0.9*h*h, persistence = 10*h) and batch sizes from the Trainer arguments."""
        z = self.zero
        h = hidden_size or 4096
        if _is_auto(z.reduce_bucket_size):
            z.reduce_bucket_size = h * h
        if _is_auto(z.allgather_bucket_size):
            z.allgather_bucket_size = h * h
        if _is_auto(z.stage3_prefetch_bucket_size):
            z.stage3_prefetch_bucket_size = int(0.9 * h * h)
        if _is_auto(z.stage3_param_persistence_threshold):
            z.stage3_param_persistence_threshold = 10 * h
        if _is_auto(self.gradient_accumulation_steps) or self.gradient_accumulation_steps is None:
            self.gradient_accumulation_steps = grad_accum or 1
        if _is_auto(self.train_micro_batch_size_per_gpu) or self.train_micro_batch_size_per_gpu is None:
            self.train_micro_batch_size_per_gpu = micro_batch or 1
        if _is_auto(self.train_batch_size) or self.train_batch_size is None:
            self.train_batch_size = int(self.train_micro_batch_size_per_gpu) * int(self.gradient_accumulation_steps) * world_size
        if _is_auto(self.gradient_clipping):
            self.gradient_clipping = max_grad_norm if max_grad_norm is not None else 1.0
        if self.optimizer:
            p = self.optimizer.get("params", {})
            for k, v in list(p.items()):
                if _is_auto(v):
                    p.pop(k)
        return self

    def check_batch(self, world_size: int) -> None:
        tb, mb, ga = self.train_batch_size, self.train_micro_batch_size_per_gpu, self.gradient_accumulation_steps
        if all(isinstance(x, int) for x in (tb, mb, ga)) and tb != mb * ga * world_size:
            raise ValueError(f"train_batch_size {tb} != micro {mb} * accum {ga} * world {world_size}")


def load_accelerate_config(path: str) -> Dict[str, Any]:
    """Parse an ``accelerate`` YAML (Fine-Tuning/multi_hosts.ymal): returns the launcher-relevant fields."""
    import yaml
    with open(path) as f:
        d = yaml.safe_load(f) or {}
    return {
        "distributed_type": d.get("distributed_type", "NO"),
        "machine_rank": int(d.get("machine_rank", 0)),
        "num_machines": int(d.get("num_machines", 1)),
        "num_processes": int(d.get("num_processes", 1)),
        "main_process_ip": d.get("main_process_ip", "127.0.0.1"),
        "main_process_port": int(d.get("main_process_port", 29500)),
        "mixed_precision": d.get("mixed_precision", "no"),
        "compute_environment": d.get("compute_environment", "LOCAL_MACHINE"),
    }


def parse_hostfile(path: str) -> Dict[str, int]:
    """DeepSpeed hostfile: ``hostname slots=N`` per line (LLM_Distributed_Trainning/DeepSpeed/…Multihosts/hostfile)."""
    hosts = {}
    with open(path) as f:
        for ln in f:
            ln = ln.split("#")[0].strip()
            if not ln:
                continue
            parts = ln.split()
            slots = 1
            for p in parts[1:]:
                if p.startswith("slots="):
                    slots = int(p.split("=")[1])
            hosts[parts[0]] = slots
    return hosts
