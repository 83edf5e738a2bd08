"""A reference fine-tuning script moved over by changing its imports: the body below follows
Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py (load tokenizer + 4-bit model, prepare for k-bit training, LoRA r=8 on q/v,
``Trainer`` with the script's arguments, save the adapter, reload it with ``PeftModel.from_pretrained``, generate) with every
name imported from ``llm_in_practise_b200.compat`` instead of transformers / peft."""
import json
import os

import pytest
import torch


def _tiny_checkpoint(tmp):
    """A local 'pretrained' directory: tiny Qwen3 weights + a real tokenizer.json / tokenizer_config.json."""
    tokenizers = pytest.importorskip("tokenizers")
    transformers = pytest.importorskip("transformers")
    from tokenizers import Tokenizer, models, pre_tokenizers, trainers
    from llm_in_practise_b200.models.qwen3 import Qwen3Config, Qwen3ForCausalLM
    tk = Tokenizer(models.BPE(unk_token="<unk>"))
    tk.pre_tokenizer = pre_tokenizers.ByteLevel()
    corpus = ["Instruction: say hello. Response: hello there.", "Instruction: count to three. Response: one two three."] * 30
    tk.train_from_iterator(iter(corpus), trainers.BpeTrainer(vocab_size=320, special_tokens=["<unk>", "<|endoftext|>"], show_progress=False))
    fast = transformers.PreTrainedTokenizerFast(tokenizer_object=tk, eos_token="<|endoftext|>", unk_token="<unk>")
    fast.save_pretrained(tmp)
    cfg = Qwen3Config(vocab_size=320, hidden_size=128, intermediate_size=256, num_hidden_layers=2, num_attention_heads=4,
                      num_key_value_heads=2, head_dim=32, max_position_embeddings=256, tie_word_embeddings=True)
    torch.manual_seed(0)
    Qwen3ForCausalLM(cfg, dtype=torch.float32, device="cpu", init_std=0.05).save_pretrained(tmp)
    return corpus


def test_qlora_script_with_only_its_imports_changed(tmp_path):
    from llm_in_practise_b200.compat import (AutoModelForCausalLM, AutoTokenizer, BitsAndBytesConfig, DataCollatorForLanguageModeling,
                                             LoraConfig, PeftModel, TaskType, Trainer, TrainingArguments, get_peft_model,
                                             prepare_model_for_kbit_training, set_seed)
    model_checkpoint = str(tmp_path / "Qwen3-tiny")
    corpus = _tiny_checkpoint(model_checkpoint)
    output_dir = str(tmp_path / "out")
    set_seed(42)
    tokenizer = AutoTokenizer.from_pretrained(model_checkpoint, trust_remote_code=True, use_fast=True)
    if tokenizer.pad_token is None:
        tokenizer.pad_token = tokenizer.eos_token
    model = AutoModelForCausalLM.from_pretrained(
        model_checkpoint, device_map=None, torch_dtype=torch.float32,
        quantization_config=BitsAndBytesConfig(load_in_4bit=True, bnb_4bit_compute_dtype=torch.float32, bnb_4bit_quant_type="nf4",
                                               bnb_4bit_use_double_quant=True),
        rope_scaling=None, use_cache=False, trust_remote_code=True)
    model = prepare_model_for_kbit_training(model, use_gradient_checkpointing=True)
    lora_config = LoraConfig(r=8, lora_alpha=16, target_modules=["q_proj", "v_proj"], lora_dropout=0.05, bias="none",
                             task_type=TaskType.CAUSAL_LM)
    model = get_peft_model(model, lora_config)
    model.print_trainable_parameters()
    rows = [{"input_ids": tokenizer(t, truncation=True, max_length=64)["input_ids"]} for t in corpus[:16]]
    training_args = TrainingArguments(
        output_dir=output_dir, per_device_train_batch_size=2, gradient_accumulation_steps=2, learning_rate=2e-3, num_train_epochs=2,
        logging_steps=1, save_steps=4, save_total_limit=2, bf16=False, optim="paged_adamw_8bit", lr_scheduler_type="cosine",
        warmup_ratio=0.03, gradient_checkpointing=True, report_to="none", ddp_find_unused_parameters=False,
        deepspeed={"zero_optimization": {"stage": 3}, "train_micro_batch_size_per_gpu": "auto", "gradient_accumulation_steps": "auto",
                   "train_batch_size": "auto", "bf16": {"enabled": "auto"}})
    trainer = Trainer(model=model, args=training_args, train_dataset=rows,
                      data_collator=DataCollatorForLanguageModeling(tokenizer=tokenizer, mlm=False))
    result = trainer.train()
    losses = [r["loss"] for r in trainer.state.log_history if "loss" in r]
    assert result.global_step == 8 and losses[-1] < losses[0]
    if trainer.is_world_process_zero():
        model.save_pretrained(output_dir)
        tokenizer.save_pretrained(output_dir)
    cfg = json.load(open(os.path.join(output_dir, "adapter_config.json")))
    assert cfg["r"] == 8 and cfg["task_type"] == "CAUSAL_LM" and sorted(cfg["target_modules"]) == ["q_proj", "v_proj"]
    # Fine-Tuning/inferences.py: base model + adapter directory -> chat
    base = AutoModelForCausalLM.from_pretrained(model_checkpoint, torch_dtype=torch.float32, device_map="auto")
    tuned = PeftModel.from_pretrained(base, output_dir)
    tuned.eval()
    inputs = tokenizer("Instruction: say hello. Response:", return_tensors="pt").to(tuned.device)
    out = tuned.generate(**inputs, max_new_tokens=6, do_sample=True, top_p=0.9, temperature=0.7, seed=0,
                         eos_token_id=tokenizer.eos_token_id, pad_token_id=tokenizer.pad_token_id)
    assert out.shape[0] == 1 and inputs["input_ids"].shape[1] < out.shape[1] <= inputs["input_ids"].shape[1] + 6
    assert isinstance(tokenizer.decode(out[0], skip_special_tokens=True), str)


def test_deepspeed_style_loop_with_only_its_imports_changed(tmp_path):
    """The ZeRO training loop of LLM_Distributed_Trainning/PyTorch/deepspeed_basics/*.py: ``deepspeed.init_distributed()``,
    ``engine, optimizer, _, _ = deepspeed.initialize(args=args, model=model, optimizer=optimizer, config=args.ds_config)``,
    ``logits = engine(x)`` -> criterion -> ``engine.backward(loss)`` -> ``engine.step()``, ``engine.save_checkpoint``; the
    ZeRO-Offload variant builds ``DeepSpeedCPUAdam`` first."""
    import argparse
    from llm_in_practise_b200.compat import deepspeed
    from llm_in_practise_b200.models.gpt import GPTLike
    ds_config = {"train_micro_batch_size_per_gpu": 4, "gradient_accumulation_steps": 1, "gradient_clipping": 1.0,
                 "zero_optimization": {"stage": 2, "offload_optimizer": {"device": "cpu", "pin_memory": True}},
                 "optimizer": {"type": "AdamW", "params": {"lr": 1e-2, "weight_decay": 0.01}}}
    path = tmp_path / "ds_config.json"
    path.write_text(json.dumps(ds_config))
    args = argparse.Namespace(ds_config=str(path), local_rank=0, lr=1e-2)
    deepspeed.init_distributed()
    torch.manual_seed(0)
    model = GPTLike(vocab_size=64, block_size=16, n_layer=2, n_head=2, d_model=32, dropout=0.0)
    optimizer = deepspeed.ops.adam.DeepSpeedCPUAdam(model.parameters(), lr=args.lr)
    engine, optimizer, _, _ = deepspeed.initialize(args=args, model=model, optimizer=optimizer, config=args.ds_config)
    criterion = torch.nn.CrossEntropyLoss()
    g = torch.Generator().manual_seed(1)
    x = torch.randint(0, 64, (4, 16), generator=g)
    y = torch.roll(x, -1, 1)
    losses = []
    for _ in range(12):
        logits = engine(x)
        logits = logits["logits"] if isinstance(logits, dict) else logits
        loss = criterion(logits.reshape(-1, 64).float(), y.reshape(-1))
        engine.backward(loss)
        engine.step()
        losses.append(float(loss.detach()))
    assert losses[-1] < 0.7 * losses[0], losses
    engine.save_checkpoint(str(tmp_path / "ck"), tag="epoch_1")
    assert os.path.exists(tmp_path / "ck" / "latest")
    deepspeed.shutdown()


def test_evaluate_shim_matches_sklearn():
    sk = pytest.importorskip("sklearn.metrics")
    import numpy as np
    from llm_in_practise_b200.compat import evaluate
    rng = np.random.default_rng(0)
    y, p = rng.integers(0, 3, 200), rng.integers(0, 3, 200)
    assert evaluate.load("accuracy").compute(predictions=p, references=y)["accuracy"] == pytest.approx(sk.accuracy_score(y, p))
    for avg in ("macro", "weighted", "micro"):
        for name, fn in (("f1", sk.f1_score), ("precision", sk.precision_score), ("recall", sk.recall_score)):
            got = evaluate.load(name).compute(predictions=p, references=y, average=avg)[name]
            assert got == pytest.approx(fn(y, p, average=avg, zero_division=0)), (name, avg)
    yb, pb = (y > 0).astype(int), (p > 0).astype(int)
    assert evaluate.load("f1").compute(predictions=pb, references=yb)["f1"] == pytest.approx(sk.f1_score(yb, pb))
    with pytest.raises(ValueError):
        evaluate.load("bleu")
