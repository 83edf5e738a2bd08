"""``TrainingArguments`` — every field the reference passes to HF's class, same names and defaults:
Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:132-172, Fine-Tuning/qwen3-8b-lora-dist.py:137-173,
HF_Basics/trainer_demo.py:86-107, HF_Basics/trainer_evaluate_demo.py:76-88, HF_Basics/accelerate_demo.py:88-90."""
from __future__ import annotations

import json
import os
from dataclasses import asdict, dataclass, field, fields
from typing import Any, Dict, List, Optional, Union


@dataclass
class TrainingArguments:
    output_dir: str = "./output"
    overwrite_output_dir: bool = False
    per_device_train_batch_size: int = 8
    per_device_eval_batch_size: int = 8
    gradient_accumulation_steps: int = 1
    fused_accumulation: bool = True      # run the micro-batches of a step as one layer-major pass (ZeroEngine.train_batch)
    num_train_epochs: float = 3.0
    max_steps: int = -1
    learning_rate: float = 5e-5
    weight_decay: float = 0.0
    adam_beta1: float = 0.9
    adam_beta2: float = 0.999
    adam_epsilon: float = 1e-8
    max_grad_norm: float = 1.0
    warmup_steps: int = 0
    warmup_ratio: float = 0.0
    lr_scheduler_type: str = "linear"
    logging_steps: int = 500
    logging_dir: Optional[str] = None
    logging_first_step: bool = False
    save_steps: int = 500
    save_strategy: str = "steps"            # "no" | "steps" | "epoch"
    save_total_limit: Optional[int] = None
    eval_strategy: str = "no"               # "no" | "steps" | "epoch"
    evaluation_strategy: Optional[str] = None  # older spelling
    eval_steps: Optional[int] = None
    load_best_model_at_end: bool = False
    metric_for_best_model: Optional[str] = None
    greater_is_better: Optional[bool] = None
    optim: str = "adamw_torch"              # "adamw_torch" | "paged_adamw_8bit" | "adamw_8bit"
    bf16: bool = False
    fp16: bool = False
    gradient_checkpointing: bool = False
    gradient_checkpointing_kwargs: Optional[Dict[str, Any]] = None
    ddp_backend: Optional[str] = None
    ddp_find_unused_parameters: Optional[bool] = None
    ddp_timeout: int = 1800
    dataloader_pin_memory: bool = True
    dataloader_num_workers: int = 0
    dataloader_drop_last: bool = False
    local_rank: int = -1
    deepspeed: Optional[Union[str, Dict[str, Any]]] = None
    fsdp: Union[str, List[str], bool] = ""
    fsdp_config: Optional[Dict[str, Any]] = None
    push_to_hub: bool = False
    report_to: Union[str, List[str], None] = "none"
    remove_unused_columns: bool = True
    label_names: Optional[List[str]] = None
    seed: int = 42
    data_seed: Optional[int] = None
    resume_from_checkpoint: Optional[str] = None
    disable_tqdm: bool = True
    save_safetensors: bool = True
    run_name: Optional[str] = None

    def __post_init__(self):
        if self.evaluation_strategy and self.eval_strategy == "no":
            self.eval_strategy = self.evaluation_strategy
        if self.local_rank < 0:
            self.local_rank = int(os.environ.get("LOCAL_RANK", -1))
        if self.logging_dir is None:
            self.logging_dir = os.path.join(self.output_dir, "runs")
        if self.push_to_hub:
            raise ValueError("push_to_hub needs network access, which this environment does not have")
        if self.load_best_model_at_end and self.eval_strategy != self.save_strategy:
            raise ValueError("load_best_model_at_end requires eval_strategy == save_strategy")
        if self.bf16 and self.fp16:
            raise ValueError("choose one of bf16 / fp16")

    @property
    def world_size(self) -> int:
        return int(os.environ.get("WORLD_SIZE", 1))

    @property
    def process_index(self) -> int:
        return int(os.environ.get("RANK", 0))

    @property
    def train_batch_size(self) -> int:
        return self.per_device_train_batch_size

    def to_dict(self) -> Dict[str, Any]:
        return asdict(self)

    def to_json_string(self) -> str:
        return json.dumps(self.to_dict(), indent=2, default=str)


@dataclass
class TrainerState:
    epoch: float = 0.0
    global_step: int = 0
    max_steps: int = 0
    log_history: List[Dict[str, float]] = field(default_factory=list)
    best_metric: Optional[float] = None
    best_model_checkpoint: Optional[str] = None
    total_flos: float = 0.0
    # fields of a ``trainer_state.json`` written by HF transformers that this trainer does not use (``eval_steps``,
    # ``stateful_callbacks``, ``train_batch_size`` ...): kept and written back, so a checkpoint can move both ways
    extra: Dict[str, Any] = field(default_factory=dict)

    def save_to_json(self, path: str) -> None:
        d = asdict(self)
        d.update(d.pop("extra"))
        with open(path, "w") as f:
            json.dump(d, f, indent=2, sort_keys=True)

    @classmethod
    def load_from_json(cls, path: str) -> "TrainerState":
        with open(path) as f:
            d = json.load(f)
        known = {f.name for f in fields(cls)} - {"extra"}
        st = cls(**{k: v for k, v in d.items() if k in known})
        st.extra = {k: v for k, v in d.items() if k not in known}
        st.epoch = float(st.epoch or 0.0)
        return st


@dataclass
class TrainOutput:
    global_step: int
    training_loss: float
    metrics: Dict[str, float]


@dataclass
class EvalPrediction:
    predictions: Any
    label_ids: Any
