"""``Trainer`` — the HF-Trainer-shaped training façade of the reference's finetuning and classification scripts.

    trainer = Trainer(model=model, args=training_args, train_dataset=ds, eval_dataset=..., data_collator=collator,
                      tokenizer=tok, compute_metrics=fn)
    trainer.train(resume_from_checkpoint=...); trainer.evaluate(); trainer.save_model(dir)
    trainer.log_metrics("train", m); trainer.save_metrics("train", m); trainer.save_state()

(Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:175-202, Fine-Tuning/qwen3-8b-lora.py:180-204,
HF_Basics/trainer_demo.py:111-134, HF_Basics/trainer_evaluate_demo.py:108-112.)

The launcher environment decides the parallelism exactly as with HF: world size > 1 means data parallel;
``args.deepspeed`` (path or dict, ``"auto"`` values resolved from the arguments) selects the ZeRO stage;
``args.fsdp="full_shard"`` selects FSDP. Underneath, everything runs through ``parallel.zero.ZeroEngine`` (stage 0 =
DDP semantics) or ``parallel.fsdp.FSDP``: flat arenas, one-launch optimizer, fused peer-memory collectives.
Checkpoints follow the HF layout: ``checkpoint-N/`` with the adapter (or ``model.safetensors``), ``optimizer.pt``,
``scheduler.pt``, ``rng_state.pth``, ``trainer_state.json``, ``training_args.bin``, rotated by ``save_total_limit``.
"""
from __future__ import annotations

import inspect
import json
import logging
import math
import os
import random
import shutil
import time
from typing import Any, Callable, Dict, List, Optional

import numpy as np
import torch
import torch.distributed as dist
from torch import nn
from torch.utils.data import DataLoader

from ..data.datasets import DistributedSampler, set_seed
from ..parallel.dist import get_info, setup_distributed
from .args import EvalPrediction, TrainerState, TrainingArguments, TrainOutput

log = logging.getLogger("llm_in_practise_b200.trainer")


def _get_loss(out):
    if isinstance(out, dict):
        return out["loss"]
    if hasattr(out, "loss"):
        return out.loss
    if isinstance(out, (tuple, list)):
        return out[0]
    return out


def _get_logits(out):
    if isinstance(out, dict):
        return out.get("logits")
    if hasattr(out, "logits"):
        return out.logits
    if isinstance(out, (tuple, list)) and len(out) > 1:
        return out[1]
    return None


class LambdaSchedule:
    """linear / cosine / constant schedules with warm-up, stepped once per optimizer step."""

    def __init__(self, optimizer, kind: str, warmup: int, total: int, base_lr: float):
        self.opt, self.kind, self.warmup, self.total, self.base = optimizer, kind, warmup, max(1, total), base_lr
        self.t = 0
        self._apply()

    def factor(self, t: int) -> float:
        if t < self.warmup:
            return t / max(1, self.warmup)
        p = (t - self.warmup) / max(1, self.total - self.warmup)
        if self.kind == "cosine":
            return max(0.0, 0.5 * (1 + math.cos(math.pi * min(1.0, p))))
        if self.kind.startswith("constant"):
            return 1.0
        return max(0.0, 1.0 - p)

    def _apply(self):
        for g in self.opt.param_groups:
            g["lr"] = self.base * self.factor(self.t)

    def step(self):
        self.t += 1
        self._apply()

    def get_last_lr(self):
        return [g["lr"] for g in self.opt.param_groups]

    def state_dict(self):
        return {"t": self.t}

    def load_state_dict(self, sd):
        self.t = sd["t"]
        self._apply()


class Trainer:
    def __init__(self, model: nn.Module = None, args: TrainingArguments = None, train_dataset=None, eval_dataset=None,
                 data_collator: Optional[Callable] = None, tokenizer=None, processing_class=None,
                 compute_metrics: Optional[Callable] = None, callbacks=None, optimizers=(None, None)):
        self.args = args or TrainingArguments()
        self.info = setup_distributed(self.args.local_rank, timeout_s=self.args.ddp_timeout)
        set_seed(self.args.seed)
        self.model = model
        self.tokenizer = tokenizer or processing_class
        self.train_dataset, self.eval_dataset = train_dataset, eval_dataset
        self.data_collator = data_collator or self._default_collate
        self.compute_metrics = compute_metrics
        self.state = TrainerState()
        self.engine = None
        self.fsdp = None
        self.optimizer, self.lr_scheduler = optimizers
        self._sig = None
        dev = self.info.device
        if next(model.parameters()).device != dev and dev.type == "cuda":
            model.to(dev)
        self.device = next(model.parameters()).device
        if self.args.gradient_checkpointing and hasattr(model, "gradient_checkpointing_enable"):
            model.gradient_checkpointing_enable(**(self.args.gradient_checkpointing_kwargs or {}))
        if self.is_world_process_zero():
            os.makedirs(self.args.output_dir, exist_ok=True)

    # ------------------------------------------------------------------ helpers
    def is_world_process_zero(self) -> bool:
        return self.info.rank == 0

    @staticmethod
    def _default_collate(features):
        out = {}
        for k in features[0]:
            v = [f[k] for f in features]
            out[k] = torch.stack([torch.as_tensor(x) for x in v]) if not isinstance(v[0], str) else v
        if "label" in out and "labels" not in out:
            out["labels"] = out.pop("label")
        return out

    def _filter(self, batch: Dict[str, Any]) -> Dict[str, Any]:
        if not self.args.remove_unused_columns:
            return batch
        if self._sig is None:
            base = getattr(self.model, "base_model", self.model)
            params = inspect.signature(base.forward).parameters
            self._sig = None if any(p.kind == p.VAR_KEYWORD for p in params.values()) else set(params)
            self._sig = self._sig or "any"
        if self._sig == "any":
            return batch
        return {k: v for k, v in batch.items() if k in self._sig}

    def _to_device(self, batch):
        return {k: (v.to(self.device, non_blocking=True) if torch.is_tensor(v) else v) for k, v in batch.items()}

    def _loader(self, ds, batch_size, train: bool):
        sampler = None
        if self.info.world_size > 1:
            sampler = DistributedSampler(ds, self.info.world_size, self.info.rank, shuffle=train,
                                         seed=self.args.data_seed or self.args.seed)
        gen = torch.Generator().manual_seed(self.args.data_seed or self.args.seed)
        return DataLoader(ds, batch_size=batch_size, sampler=sampler, shuffle=(train and sampler is None),
                          collate_fn=self.data_collator, num_workers=self.args.dataloader_num_workers,
                          pin_memory=self.args.dataloader_pin_memory and self.device.type == "cuda",
                          drop_last=self.args.dataloader_drop_last, generator=gen if sampler is None else None)

    # ------------------------------------------------------------------ engine construction
    def _ds_config(self) -> Dict[str, Any]:
        a = self.args
        if a.deepspeed:
            if isinstance(a.deepspeed, str):
                with open(a.deepspeed) as f:
                    cfg = json.load(f)
            else:
                cfg = json.loads(json.dumps(a.deepspeed))
        else:
            cfg = {"zero_optimization": {"stage": 0}}
        cfg.setdefault("gradient_accumulation_steps", "auto")
        cfg.setdefault("gradient_clipping", "auto")
        cfg.setdefault("train_micro_batch_size_per_gpu", "auto")
        cfg.setdefault("train_batch_size", "auto")
        if "optimizer" not in cfg:
            cfg["optimizer"] = {"type": a.optim, "params": {"lr": a.learning_rate, "betas": [a.adam_beta1, a.adam_beta2],
                                                          "eps": a.adam_epsilon, "weight_decay": a.weight_decay}}
        if a.fp16 and "fp16" not in cfg and "bf16" not in cfg:
            cfg["fp16"] = {"enabled": True}
        return cfg

    def _build_engine(self):
        from ..parallel.config import DSConfig
        from ..parallel.zero import ZeroEngine
        a = self.args
        fsdp_on = bool(a.fsdp) and "full_shard" in (a.fsdp if isinstance(a.fsdp, str) else " ".join(a.fsdp))
        base = getattr(self.model, "base_model", self.model)
        if fsdp_on and not hasattr(base, "_lora_arena"):
            from ..parallel import fsdp as F
            from ..models.transformer import TransformerBlock
            classes = {TransformerBlock}
            try:
                from ..models.qwen3 import Qwen3DecoderLayer
                classes.add(Qwen3DecoderLayer)
            except Exception:
                pass
            mp = F.MixedPrecision(torch.bfloat16, torch.bfloat16, torch.bfloat16) if a.bf16 else None
            self.fsdp = F.FSDP(self.model, auto_wrap_policy=F.transformer_auto_wrap_policy(transformer_layer_cls=classes),
                               mixed_precision=mp, sync_module_states=True)
            if self.optimizer is None:
                self.optimizer = torch.optim.AdamW(self.fsdp.parameters(), lr=a.learning_rate, betas=(a.adam_beta1, a.adam_beta2),
                                                   eps=a.adam_epsilon, weight_decay=a.weight_decay)
            return
        cfg = DSConfig.from_any(self._ds_config()).resolve_auto(
            hidden_size=getattr(getattr(base, "config", None), "hidden_size", None), world_size=self.info.world_size,
            micro_batch=a.per_device_train_batch_size, grad_accum=a.gradient_accumulation_steps, max_grad_norm=a.max_grad_norm)
        self.engine = ZeroEngine(base, config=cfg)
        self.optimizer = self.engine.optimizer

    # ------------------------------------------------------------------ training
    def training_step(self, batch: Dict[str, Any]) -> torch.Tensor:
        batch = self._filter(self._to_device(batch))
        if self.engine is not None:
            out = self.engine(**batch)
            loss = _get_loss(out)
            self.engine.backward(loss)
            return loss.detach()
        from ..ops.loss import grad_scale_context
        up = 1.0 / self.args.gradient_accumulation_steps
        with grad_scale_context(up):
            out = self.fsdp(**batch)
        loss = _get_loss(out)
        (loss * up).backward()
        return loss.detach()

    def _optimizer_step(self) -> None:
        if self.engine is not None:
            self.engine.step()
        else:
            if self.args.max_grad_norm and self.args.max_grad_norm > 0:
                self.fsdp.clip_grad_norm_(self.args.max_grad_norm)
            self.optimizer.step()
            self.fsdp.zero_grad()
        if self.lr_scheduler is not None:
            self.lr_scheduler.step()

    def train(self, resume_from_checkpoint: Optional[str] = None, **_) -> TrainOutput:
        a = self.args
        if self.engine is None and self.fsdp is None:
            self._build_engine()
        loader = self._loader(self.train_dataset, a.per_device_train_batch_size, True)
        ga = a.gradient_accumulation_steps
        steps_per_epoch = max(1, len(loader) // ga)
        max_steps = a.max_steps if a.max_steps > 0 else int(math.ceil(a.num_train_epochs * steps_per_epoch))
        epochs = int(math.ceil(max_steps / steps_per_epoch))
        self.state.max_steps = max_steps
        warmup = a.warmup_steps or int(a.warmup_ratio * max_steps)
        if self.lr_scheduler is None:
            self.lr_scheduler = LambdaSchedule(self.optimizer, a.lr_scheduler_type, warmup, max_steps, a.learning_rate)
        resume = resume_from_checkpoint or a.resume_from_checkpoint
        if resume:
            if resume is True:
                resume = self._last_checkpoint()
            if resume:
                self._load_checkpoint(resume)
        model = self.engine.module if self.engine is not None else self.fsdp
        model.train()
        t0 = time.time()
        tot_loss, n_loss, run_loss, run_n = 0.0, 0, torch.zeros((), device=self.device), 0
        start_step = self.state.global_step
        samples = 0
        done = False
        for epoch in range(start_step // steps_per_epoch, epochs):
            if hasattr(loader.sampler, "set_epoch"):
                loader.sampler.set_epoch(epoch)
            pending: List[Dict[str, Any]] = []
            fuse = (self.engine is not None and ga > 1 and getattr(a, "fused_accumulation", True)
                    and self.engine.supports_fused_accumulation())
            for i, batch in enumerate(loader):
                if i // ga >= steps_per_epoch:
                    break
                samples += a.per_device_train_batch_size * self.info.world_size
                if fuse:
                    # the micro-batches of one optimizer step go through the model as ONE layer-major pass
                    # (ZeroEngine.train_batch): same gradients, frozen 4-bit weights gathered and expanded once
                    pending.append(self._filter(self._to_device(batch)))
                    if (i + 1) % ga:
                        continue
                    shapes = {tuple(b["input_ids"].shape) for b in pending}
                    if len(shapes) == 1 and all(set(b) <= {"input_ids", "labels", "attention_mask"} for b in pending):
                        run_loss += self.engine.train_batch([b["input_ids"] for b in pending],
                                                            [b.get("labels", b["input_ids"]) for b in pending]) * ga
                    else:
                        for b in pending:
                            out = self.engine(**b)
                            loss = _get_loss(out)
                            self.engine.backward(loss)
                            run_loss += loss.detach()
                    run_n += ga
                    pending = []
                else:
                    run_loss += self.training_step(batch)
                    run_n += 1
                    if (i + 1) % ga:
                        continue
                self._optimizer_step()
                self.state.global_step += 1
                self.state.epoch = self.state.global_step / steps_per_epoch
                gs = self.state.global_step
                if (a.logging_steps and gs % a.logging_steps == 0) or (a.logging_first_step and gs == 1):
                    mean = float(run_loss.item()) / max(1, run_n)   # the only host sync of the loop
                    tot_loss, n_loss = tot_loss + mean * run_n, n_loss + run_n
                    run_loss.zero_()
                    run_n = 0
                    rec = {"loss": round(mean, 4), "learning_rate": self.lr_scheduler.get_last_lr()[0],
                           "epoch": round(self.state.epoch, 2), "step": gs}
                    gn = self.engine.get_global_grad_norm() if self.engine is not None else None
                    if gn is not None:
                        rec["grad_norm"] = round(gn, 4)
                    self.log(rec)
                if a.eval_strategy == "steps" and a.eval_steps and gs % a.eval_steps == 0:
                    self._evaluate_and_track()
                if a.save_strategy == "steps" and a.save_steps and gs % a.save_steps == 0:
                    self._save_checkpoint()
                if gs >= max_steps:
                    done = True
                    break
            if a.eval_strategy == "epoch":
                self._evaluate_and_track()
            if a.save_strategy == "epoch":
                self._save_checkpoint()
            if done:
                break
        if run_n:
            tot_loss, n_loss = tot_loss + float(run_loss.item()), n_loss + run_n
        if self.device.type == "cuda":
            torch.cuda.synchronize()
        dt = time.time() - t0
        steps = self.state.global_step - start_step
        metrics = {"train_runtime": round(dt, 4), "train_samples_per_second": round(samples / max(dt, 1e-9), 3),
                   "train_steps_per_second": round(steps / max(dt, 1e-9), 3), "train_loss": tot_loss / max(1, n_loss),
                   "epoch": self.state.epoch}
        if a.load_best_model_at_end and self.state.best_model_checkpoint:
            self._load_model_weights(self.state.best_model_checkpoint)
        self.log(metrics)
        return TrainOutput(self.state.global_step, metrics["train_loss"], metrics)

    def log(self, rec: Dict[str, Any]) -> None:
        self.state.log_history.append(rec)
        if self.is_world_process_zero():
            print(rec, flush=True)

    # ------------------------------------------------------------------ evaluation
    @torch.no_grad()
    def evaluate(self, eval_dataset=None, metric_key_prefix: str = "eval") -> Dict[str, float]:
        ds = eval_dataset if eval_dataset is not None else self.eval_dataset
        model = self.fsdp if self.fsdp is not None else (self.engine.module if self.engine is not None else self.model)
        was = model.training
        model.eval()
        loader = self._loader(ds, self.args.per_device_eval_batch_size, False)
        tot, cnt = 0.0, 0
        preds, labels = [], []
        t0 = time.time()
        for batch in loader:
            batch = self._filter(self._to_device(batch))
            out = model(**batch)
            loss = _get_loss(out) if (isinstance(out, dict) and "loss" in out) or hasattr(out, "loss") else None
            n = next(v.shape[0] for v in batch.values() if torch.is_tensor(v))
            if loss is not None:
                tot += float(loss) * n
                cnt += n
            lg = _get_logits(out)
            if self.compute_metrics is not None and lg is not None and "labels" in batch:
                preds.append(lg.float().cpu().numpy())
                labels.append(batch["labels"].cpu().numpy())
        from ..parallel.dist import all_reduce_mean_scalar
        metrics = {}
        if cnt:
            metrics[f"{metric_key_prefix}_loss"] = all_reduce_mean_scalar(tot, cnt, self.device)
        if self.compute_metrics is not None and preds:
            p, l = np.concatenate(preds), np.concatenate(labels)
            if self.info.world_size > 1:
                gathered = [None] * self.info.world_size
                dist.all_gather_object(gathered, (p, l))
                p = np.concatenate([g[0] for g in gathered])[: len(ds)]
                l = np.concatenate([g[1] for g in gathered])[: len(ds)]
            for k, v in (self.compute_metrics(EvalPrediction(p, l)) or {}).items():
                metrics[k if k.startswith(metric_key_prefix) else f"{metric_key_prefix}_{k}"] = float(v)
        dt = time.time() - t0
        metrics[f"{metric_key_prefix}_runtime"] = round(dt, 4)
        metrics[f"{metric_key_prefix}_samples_per_second"] = round(len(ds) / max(dt, 1e-9), 3)
        metrics["epoch"] = self.state.epoch
        self.log(metrics)
        model.train(was)
        return metrics

    def predict(self, test_dataset, metric_key_prefix: str = "test"):
        keep = self.compute_metrics
        m = self.evaluate(test_dataset, metric_key_prefix)
        self.compute_metrics = keep
        return m

    def _evaluate_and_track(self) -> None:
        if self.eval_dataset is None:
            return
        m = self.evaluate()
        key = self.args.metric_for_best_model or "loss"
        key = key if key.startswith("eval_") else "eval_" + key
        if key in m:
            greater = self.args.greater_is_better
            if greater is None:
                greater = not key.endswith("loss")
            better = self.state.best_metric is None or (m[key] > self.state.best_metric if greater else m[key] < self.state.best_metric)
            if better:
                self.state.best_metric = m[key]
                self._pending_best = True

    # ------------------------------------------------------------------ persistence
    def _trainable_model(self):
        return self.model

    def save_model(self, output_dir: Optional[str] = None) -> None:
        out = output_dir or self.args.output_dir
        if self.fsdp is not None:
            sd = self.fsdp.full_state_dict(cpu=True)
        if not self.is_world_process_zero():
            return
        os.makedirs(out, exist_ok=True)
        m = self.model
        if self.fsdp is not None:
            from ..io.safetensors import save_file
            save_file({k: v.contiguous() for k, v in sd.items()}, os.path.join(out, "model.safetensors"), {"format": "pt"})
        elif hasattr(m, "save_pretrained"):
            m.save_pretrained(out)
        else:
            from ..io.safetensors import save_file
            sd = {k: v.detach().cpu().contiguous() for k, v in m.state_dict().items()}
            seen, uniq = {}, {}
            for k, v in sd.items():                      # tied tensors are stored once
                key = (v.data_ptr(), tuple(v.shape))
                if key not in seen:
                    seen[key] = k
                    uniq[k] = v
            save_file(uniq, os.path.join(out, "model.safetensors"), {"format": "pt"})
        if self.tokenizer is not None and hasattr(self.tokenizer, "save_pretrained"):
            self.tokenizer.save_pretrained(out)
        torch.save(self.args, os.path.join(out, "training_args.bin"))

    def _last_checkpoint(self) -> Optional[str]:
        if not os.path.isdir(self.args.output_dir):
            return None
        cks = [d for d in os.listdir(self.args.output_dir) if d.startswith("checkpoint-") and d.split("-")[-1].isdigit()]
        return os.path.join(self.args.output_dir, max(cks, key=lambda d: int(d.split("-")[-1]))) if cks else None

    def _save_checkpoint(self) -> None:
        d = os.path.join(self.args.output_dir, f"checkpoint-{self.state.global_step}")
        self.save_model(d)
        W, R = self.info.world_size, self.info.rank
        if W > 1:
            dist.barrier()          # rank 0 has created the directory
            # ZeRO (stage >= 1) and FSDP keep optimizer state only for this rank's shard: one file per rank, like
            # DeepSpeed's zero_pp_rank_<r>_* files (HF writes the same through engine.save_checkpoint)
            os.makedirs(d, exist_ok=True)
            torch.save({"world_size": W, "rank": R, "state": self.optimizer.state_dict()},
                       os.path.join(d, f"optimizer_rank{R}.pt"))
        if self.is_world_process_zero():
            if W == 1:
                torch.save(self.optimizer.state_dict(), os.path.join(d, "optimizer.pt"))
            torch.save(self.lr_scheduler.state_dict(), os.path.join(d, "scheduler.pt"))
            torch.save({"python": random.getstate(), "numpy": np.random.get_state(), "cpu": torch.get_rng_state(),
                        "cuda": torch.cuda.get_rng_state_all() if torch.cuda.is_available() else None},
                       os.path.join(d, "rng_state.pth"))
            if getattr(self, "_pending_best", False):
                self.state.best_model_checkpoint = d
                self._pending_best = False
            self.state.save_to_json(os.path.join(d, "trainer_state.json"))
            self._rotate()
        if self.info.world_size > 1:
            dist.barrier()

    def _rotate(self) -> None:
        lim = self.args.save_total_limit
        if not lim or lim <= 0:
            return
        cks = sorted((d for d in os.listdir(self.args.output_dir) if d.startswith("checkpoint-") and d.split("-")[-1].isdigit()),
                     key=lambda d: int(d.split("-")[-1]))
        best = os.path.basename(self.state.best_model_checkpoint) if self.state.best_model_checkpoint else None
        while len(cks) > lim:
            victim = next((c for c in cks if c != best), None)
            if victim is None:
                break
            shutil.rmtree(os.path.join(self.args.output_dir, victim), ignore_errors=True)
            cks.remove(victim)

    def _load_model_weights(self, d: str) -> None:
        from ..io.safetensors import load_file
        m = self.model
        if os.path.exists(os.path.join(d, "adapter_model.safetensors")) and hasattr(m, "load_adapter_state_dict"):
            m.load_adapter_state_dict(load_file(os.path.join(d, "adapter_model.safetensors")))
        elif os.path.exists(os.path.join(d, "model.safetensors")):
            tgt = self.fsdp if self.fsdp is not None else m
            tgt.load_state_dict(load_file(os.path.join(d, "model.safetensors")), strict=False)

    def _load_checkpoint(self, d: str) -> None:
        self._load_model_weights(d)
        W, R = self.info.world_size, self.info.rank
        mine = os.path.join(d, f"optimizer_rank{R}.pt")
        if W > 1 and os.path.exists(mine):
            ck = torch.load(mine, map_location="cpu", weights_only=False)
            if ck["world_size"] != W:
                raise RuntimeError(f"{d} holds optimizer shards of a {ck['world_size']}-rank job; this job has {W} ranks "
                                   "(sharded optimizer state cannot be re-partitioned on load)")
            self.optimizer.load_state_dict(ck["state"])
        elif W > 1 and any(f.startswith("optimizer_rank") for f in os.listdir(d)):
            raise RuntimeError(f"{d} has per-rank optimizer files but none for rank {R}: written by a smaller job")
        elif W > 1 and os.path.exists(os.path.join(d, "optimizer.pt")):
            import logging
            logging.getLogger("llm_in_practise_b200.trainer").warning(
                "%s holds a single-process optimizer.pt; with %d ranks the state is sharded, so it is NOT loaded", d, W)
        elif os.path.exists(os.path.join(d, "optimizer.pt")):
            self.optimizer.load_state_dict(torch.load(os.path.join(d, "optimizer.pt"), map_location="cpu", weights_only=False))
        if os.path.exists(os.path.join(d, "scheduler.pt")):
            self.lr_scheduler.load_state_dict(torch.load(os.path.join(d, "scheduler.pt"), weights_only=False))
        if os.path.exists(os.path.join(d, "trainer_state.json")):
            self.state = TrainerState.load_from_json(os.path.join(d, "trainer_state.json"))
        rp = os.path.join(d, "rng_state.pth")
        if os.path.exists(rp):
            r = torch.load(rp, weights_only=False)
            random.setstate(r["python"])
            np.random.set_state(r["numpy"])
            torch.set_rng_state(r["cpu"])
            if r.get("cuda") is not None and torch.cuda.is_available():
                try:
                    torch.cuda.set_rng_state_all(r["cuda"])
                except Exception:
                    pass

    # ------------------------------------------------------------------ metrics files (HF names)
    def log_metrics(self, split: str, metrics: Dict[str, float]) -> None:
        if not self.is_world_process_zero():
            return
        print(f"***** {split} metrics *****")
        w = max(len(k) for k in metrics) if metrics else 0
        for k in sorted(metrics):
            print(f"  {k:<{w}} = {metrics[k]}")

    def save_metrics(self, split: str, metrics: Dict[str, float], combined: bool = True) -> None:
        if not self.is_world_process_zero():
            return
        with open(os.path.join(self.args.output_dir, f"{split}_results.json"), "w") as f:
            json.dump(metrics, f, indent=4, sort_keys=True)
        if combined:
            p = os.path.join(self.args.output_dir, "all_results.json")
            allm = {}
            if os.path.exists(p):
                with open(p) as f:
                    allm = json.load(f)
            allm.update(metrics)
            with open(p, "w") as f:
                json.dump(allm, f, indent=4, sort_keys=True)

    def save_state(self) -> None:
        if self.is_world_process_zero():
            self.state.save_to_json(os.path.join(self.args.output_dir, "trainer_state.json"))
