"""Trainer façade on CPU: LoRA finetune with logging / checkpoint rotation / resume, and a classifier with
evaluation, compute_metrics and load_best_model_at_end (the two Trainer usages of the reference)."""
import json
import os

import numpy as np
import torch

from llm_in_practise_b200.data import DataCollatorForLanguageModeling, SyntheticTokenDataset
from llm_in_practise_b200.models.classifier import EncoderClassifier
from llm_in_practise_b200.models.qwen3 import QWEN3_PRESETS, Qwen3ForCausalLM
from llm_in_practise_b200.peft import LoraConfig, get_peft_model
from llm_in_practise_b200.trainer import Trainer, TrainingArguments


class _Rows(torch.utils.data.Dataset):
    def __init__(self, n, L, V, seed=0, labels=False):
        g = torch.Generator().manual_seed(seed)
        self.x = torch.randint(1, V, (n, L), generator=g)
        self.y = (self.x[:, 0] % 2) if labels else None

    def __len__(self):
        return len(self.x)

    def __getitem__(self, i):
        d = {"input_ids": self.x[i].tolist()}
        if self.y is not None:
            d["label"] = int(self.y[i])
        return d


def test_lora_trainer_checkpoints_and_resume(tmp_path):
    torch.manual_seed(0)
    model = Qwen3ForCausalLM(QWEN3_PRESETS["qwen3-tiny"], dtype=torch.float32)
    pm = get_peft_model(model, LoraConfig(r=4, lora_alpha=8, target_modules=["q_proj", "v_proj"]))
    ds = _Rows(32, 16, 1024)
    out = str(tmp_path / "out")
    args = TrainingArguments(output_dir=out, per_device_train_batch_size=2, gradient_accumulation_steps=2, num_train_epochs=2,
                             learning_rate=5e-3, logging_steps=2, save_steps=4, save_total_limit=2, optim="paged_adamw_8bit",
                             report_to="none", deepspeed={"zero_optimization": {"stage": 3}, "gradient_clipping": 1.0,
                                                          "gradient_accumulation_steps": "auto",
                                                          "train_micro_batch_size_per_gpu": "auto", "train_batch_size": "auto"})
    tr = Trainer(model=pm, args=args, train_dataset=ds, data_collator=DataCollatorForLanguageModeling(pad_token_id=0))
    res = tr.train()
    assert res.global_step == 16
    losses = [r["loss"] for r in tr.state.log_history if "loss" in r]
    assert losses[-1] < losses[0]
    cks = sorted(d for d in os.listdir(out) if d.startswith("checkpoint-"))
    assert cks == ["checkpoint-12", "checkpoint-16"]            # rotation kept the last two
    for f in ("adapter_model.safetensors", "adapter_config.json", "optimizer.pt", "scheduler.pt", "rng_state.pth",
              "trainer_state.json", "training_args.bin"):
        assert os.path.exists(os.path.join(out, "checkpoint-16", f)), f
    tr.save_model(out)
    tr.log_metrics("train", res.metrics); tr.save_metrics("train", res.metrics); tr.save_state()
    assert json.load(open(os.path.join(out, "train_results.json")))["train_loss"] > 0
    assert res.metrics["train_samples_per_second"] > 0
    # resume: a fresh trainer picks up step 16 and has nothing left to do for the same schedule
    model2 = Qwen3ForCausalLM(QWEN3_PRESETS["qwen3-tiny"], dtype=torch.float32)
    pm2 = get_peft_model(model2, LoraConfig(r=4, lora_alpha=8, target_modules=["q_proj", "v_proj"]))
    tr2 = Trainer(model=pm2, args=args, train_dataset=ds, data_collator=DataCollatorForLanguageModeling(pad_token_id=0))
    r2 = tr2.train(resume_from_checkpoint=True)
    assert r2.global_step == 16
    assert torch.allclose(pm2.arena.data, pm.arena.data)


def test_classifier_trainer_eval_and_best_model(tmp_path):
    torch.manual_seed(0)
    model = EncoderClassifier(vocab_size=64, d_model=32, n_layer=1, n_head=2, max_len=16, dropout=0.0)
    train, ev = _Rows(64, 8, 64, 1, labels=True), _Rows(32, 8, 64, 2, labels=True)

    def metrics(p):
        return {"accuracy": float((np.argmax(p.predictions, -1) == p.label_ids).mean())}

    args = TrainingArguments(output_dir=str(tmp_path / "cls"), per_device_train_batch_size=8, per_device_eval_batch_size=8,
                             num_train_epochs=3, learning_rate=5e-3, weight_decay=0.01, eval_strategy="epoch",
                             save_strategy="epoch", load_best_model_at_end=True, metric_for_best_model="accuracy",
                             logging_steps=4, warmup_steps=2)
    tr = Trainer(model=model, args=args, train_dataset=train, eval_dataset=ev, compute_metrics=metrics)
    tr.train()
    m = tr.evaluate()
    assert "eval_accuracy" in m and "eval_loss" in m
    assert m["eval_accuracy"] >= 0.8 and abs(m["eval_accuracy"] - tr.state.best_metric) < 1e-9   # best checkpoint was reloaded
    assert tr.state.best_model_checkpoint is not None


def _two_rank_resume(rank, world, out):
    """2 ranks (gloo), ZeRO stage 2: optimizer state is sharded, so resume must load THIS rank's moments."""
    from llm_in_practise_b200.parallel.dist import setup_distributed
    setup_distributed()

    def make():
        torch.manual_seed(0)
        model = Qwen3ForCausalLM(QWEN3_PRESETS["qwen3-tiny"], dtype=torch.float32)
        return get_peft_model(model, LoraConfig(r=4, lora_alpha=8, target_modules=["q_proj", "v_proj"]))

    ds = _Rows(32, 16, 1024)

    def args(epochs, d):
        return TrainingArguments(output_dir=os.path.join(out, d), per_device_train_batch_size=2, gradient_accumulation_steps=1,
                                 num_train_epochs=epochs, learning_rate=5e-3, logging_steps=100, save_steps=4, save_total_limit=4,
                                 lr_scheduler_type="constant",
                                 report_to="none", deepspeed={"zero_optimization": {"stage": 2}, "gradient_clipping": 1.0,
                                                              "gradient_accumulation_steps": "auto",
                                                              "train_micro_batch_size_per_gpu": "auto", "train_batch_size": "auto"})
    coll = DataCollatorForLanguageModeling(pad_token_id=0)
    # uninterrupted: 2 epochs = 16 steps
    pm = make()
    Trainer(model=pm, args=args(2, "a"), train_dataset=ds, data_collator=coll).train()
    # interrupted after 1 epoch (8 steps, checkpoint-8), resumed by a fresh trainer for the second epoch
    pm1 = make()
    Trainer(model=pm1, args=args(1, "b"), train_dataset=ds, data_collator=coll).train()
    assert os.path.exists(os.path.join(out, "b", "checkpoint-8", f"optimizer_rank{rank}.pt"))
    pm2 = make()
    r = Trainer(model=pm2, args=args(2, "b"), train_dataset=ds, data_collator=coll).train(resume_from_checkpoint=True)
    assert r.global_step == 16
    assert torch.allclose(pm2.arena.data, pm.arena.data, rtol=1e-5, atol=1e-6), float((pm2.arena.data - pm.arena.data).abs().max())


def test_two_rank_resume_restores_each_ranks_optimizer_shard(tmp_path):
    from dist_utils import run_distributed
    run_distributed(_two_rank_resume, 2, (str(tmp_path),))
