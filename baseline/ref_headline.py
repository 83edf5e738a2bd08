"""Reference arm of the headline benchmark: a torch-only restatement of
Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py (+ Fine-Tuning/ds_zero3_config.json).

The stock script cannot run in this image: it imports peft, bitsandbytes, deepspeed and accelerate, none
of which is installed or in /opt/wheelhouse, and /root/reference is not a pip-installable package. What
*is* in the image is the rest of that script's stack: torch 2.11 (cuBLAS, SDPA, NCCL) and transformers 5.5
(``Qwen3ForCausalLM``, the class ``AutoModelForCausalLM`` resolves to). This file rebuilds the script's
training step on exactly those libraries and nothing from ``llm_in_practise_b200``:

* model     HF ``Qwen3ForCausalLM`` with the Qwen3-14B shape, random init, bf16, ``use_cache=False``, HF's default
            attention (sdpa); gradient checkpointing off (script line 160)
* adapters  LoRA r 8 / alpha 16 / dropout 0.1 on q_proj and v_proj (script lines 113-120) written as a plain
            ``nn.Module`` (peft is absent): fp32 A (kaiming) and B (zeros) around the frozen bf16 ``nn.Linear``
* step      micro-batch 2 x 512 tokens, gradient accumulation 2, bf16 autocast, HF causal-LM loss, clip 1.0,
            ``torch.optim.AdamW`` (fused) on the adapter parameters
* parallel  N > 1: both torch-only data-parallel forms are timed and the FASTER one is the reported value —
            ``fsdp2``: ``fully_shard`` per decoder layer + root (the ZeRO-3 equivalent: parameters sharded, all-gathered
            per layer in forward and backward, adapter gradients reduce-scattered), and ``ddp``: replicated base,
            ``DistributedDataParallel`` all-reduce of the adapter gradients (what the script degenerates to if DeepSpeed
            leaves the quantised base replicated — SURVEY.md section 6)

Deviations from the stock script (all forced by the missing packages): the frozen base is bf16 instead of bitsandbytes
NF4 (no dequantisation kernel in the step, 4x the bytes in the ZeRO-3 gather), the optimizer is fp32 AdamW instead of
``paged_adamw_8bit``, and the partitioning engine is torch FSDP2 / DDP instead of DeepSpeed. The JSON line says so
(``reference_class``, ``same_config: false``, ``deviations``).
"""
from __future__ import annotations

import json
import math
import os
import statistics
import subprocess
import threading
import time

METRIC = "Qwen3-14B QLoRA ZeRO-3 train tok/s"

SHAPES = {
    "qwen3-14b": dict(hidden_size=5120, intermediate_size=17408, num_hidden_layers=40, num_attention_heads=40,
                      num_key_value_heads=8),
    "qwen3-8b": dict(hidden_size=4096, intermediate_size=12288, num_hidden_layers=36, num_attention_heads=32,
                     num_key_value_heads=8),
    "qwen3-tiny": dict(vocab_size=1024, hidden_size=256, intermediate_size=512, num_hidden_layers=2,
                       num_attention_heads=4, num_key_value_heads=2, head_dim=64, max_position_embeddings=512),
}


class _Clocks:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, idx):
        self.idx, self.proc, self.lines = idx, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms",
                                          "200", "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL,
                                         text=True)
        except OSError:
            return
        threading.Thread(target=lambda: [self.lines.append(l.strip()) for l in self.proc.stdout], daemon=True).start()

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def build_model(name, layers, dev, dtype):
    """Random-init HF Qwen3 of the named shape, created directly on ``dev`` in ``dtype``."""
    import torch
    from transformers import Qwen3Config, Qwen3ForCausalLM
    kw = dict(vocab_size=151936, head_dim=128, max_position_embeddings=40960, rms_norm_eps=1e-6, rope_theta=1e6,
              tie_word_embeddings=False, use_cache=False)
    kw.update(SHAPES[name])
    if layers:
        kw["num_hidden_layers"] = layers
    cfg = Qwen3Config(**kw)
    old = torch.get_default_dtype()
    torch.set_default_dtype(dtype)
    try:
        with torch.device(dev):
            model = Qwen3ForCausalLM(cfg)
    finally:
        torch.set_default_dtype(old)
    model.to(dtype)
    model.config.use_cache = False
    return model, cfg


def add_lora(model, r=8, alpha=16, dropout=0.1, targets=("q_proj", "v_proj")):
    """peft ``LoraConfig(r, lora_alpha, lora_dropout, target_modules)`` + ``get_peft_model`` by hand."""
    import torch
    from torch import nn

    class LoraLinear(nn.Module):
        def __init__(self, base: nn.Linear):
            super().__init__()
            self.base_layer = base
            dev = base.weight.device
            self.lora_A = nn.Parameter(torch.empty(r, base.in_features, dtype=torch.float32, device=dev))
            self.lora_B = nn.Parameter(torch.zeros(base.out_features, r, dtype=torch.float32, device=dev))
            nn.init.kaiming_uniform_(self.lora_A, a=math.sqrt(5))
            self.drop = nn.Dropout(dropout)
            self.scaling = alpha / r

        def forward(self, x):
            y = self.base_layer(x)
            z = nn.functional.linear(nn.functional.linear(self.drop(x), self.lora_A.to(x.dtype)), self.lora_B.to(x.dtype))
            return y + z * self.scaling

    for p in model.parameters():
        p.requires_grad_(False)
    n = 0
    for mod in list(model.modules()):
        for t in targets:
            child = getattr(mod, t, None)
            if isinstance(child, nn.Linear):
                setattr(mod, t, LoraLinear(child))
                n += 1
    # prepare_model_for_kbit_training / enable_input_require_grads: let gradients flow from the embedding output
    model.get_input_embeddings().register_forward_hook(lambda m, i, o: o.requires_grad_(True))
    return [p for p in model.parameters() if p.requires_grad], n


def run(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    use_cuda = torch.cuda.is_available() and os.environ.get("REF_DEVICE", "cuda") != "cpu"
    if not use_cuda and os.environ.get("REF_DEVICE") != "cpu":
        print(json.dumps({"impl": "reference", "unavailable": "no CUDA device"}))
        return 0
    dev = torch.device("cuda", local_rank) if use_cuda else torch.device("cpu")
    if use_cuda:
        torch.cuda.set_device(dev)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29512")
        if use_cuda:
            dist.init_process_group("nccl", device_id=dev)
        else:
            dist.init_process_group("gloo")
    dtype = torch.bfloat16
    B, S, GA = args.micro_batch, args.seq_len, args.grad_accum
    tokens_per_step = B * S * GA * world
    torch.manual_seed(1234)
    t0 = time.time()
    model, cfg = build_model(args.model, args.layers, dev, dtype)
    params, n_wrapped = add_lora(model)
    model.train()
    build_s = time.time() - t0

    gen = torch.Generator().manual_seed(100 + rank)
    n_host = max(4, min(16, args.steps + args.warmup)) * GA
    host = [torch.randint(0, cfg.vocab_size, (B, S), generator=gen) for _ in range(n_host)]
    if use_cuda:
        host = [h.pin_memory() for h in host]
    devb = [h.to(dev) for h in host]

    def sync_all():
        if world > 1:
            dist.barrier()
        if use_cuda:
            torch.cuda.synchronize()

    def timer():
        if use_cuda:
            return torch.cuda.Event(enable_timing=True)

        class T:
            def record(self):
                self.t = time.perf_counter()

            def elapsed_time(self, o):
                return (o.t - self.t) * 1e3
        return T()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def measure(wrapped, opt, no_sync):
        import contextlib

        def steps(n, e2e):
            k, last = 0, None
            for _ in range(n):
                for m in range(GA):
                    ids = host[k % n_host].to(dev, non_blocking=True) if e2e else devb[k % n_host]
                    k += 1
                    ctx = no_sync() if (no_sync is not None and m < GA - 1) else contextlib.nullcontext()
                    with ctx:
                        with torch.autocast(dev.type, dtype=dtype):
                            out = wrapped(input_ids=ids, labels=ids)
                        (out.loss / GA).backward()
                torch.nn.utils.clip_grad_norm_(params, 1.0)
                opt.step()
                opt.zero_grad(set_to_none=True)
                if e2e:
                    last = float(out.loss.item())
            return last

        steps(args.warmup, False)
        sync_all()
        clocks = _Clocks(local_rank)
        clocks.start()
        e0, e1 = timer(), timer()
        sync_all()
        e0.record()
        steps(args.steps, False)
        e1.record()
        sync_all()
        ms = max_over_ranks(e0.elapsed_time(e1)) / args.steps
        ck = clocks.stop()
        steps(1, True)
        sync_all()
        f0, f1 = timer(), timer()
        f0.record()
        last = steps(args.steps, True)
        f1.record()
        sync_all()
        ms_e2e = max_over_ranks(f0.elapsed_time(f1)) / args.steps
        return {"ms_per_step": ms, "ms_e2e": ms_e2e, "clocks": ck, "last_loss": last}

    def make_opt():
        kw = dict(lr=1e-4, weight_decay=0.0)
        if use_cuda:
            kw["fused"] = True
        return torch.optim.AdamW([p for p in model.parameters() if p.requires_grad], **kw)

    results = {}
    want = os.environ.get("REF_PARALLEL", "both")
    if world == 1:
        results["single"] = measure(model, make_opt(), None)
    else:
        if want in ("both", "ddp"):
            from torch.nn.parallel import DistributedDataParallel as DDP
            ddp = DDP(model, device_ids=[local_rank] if use_cuda else None, find_unused_parameters=False)
            results["ddp"] = measure(ddp, make_opt(), ddp.no_sync)
            del ddp
        if want in ("both", "fsdp2"):
            from torch.distributed.fsdp import fully_shard
            for layer in model.model.layers:
                fully_shard(layer)
            fully_shard(model)
            params[:] = [p for p in model.parameters() if p.requires_grad]
            results["fsdp2"] = measure(model, make_opt(), None)

    if rank == 0:
        best = min(results, key=lambda k: results[k]["ms_per_step"])
        r = results[best]
        ms = r["ms_per_step"]
        out = {
            "metric": METRIC, "value": round(tokens_per_step / ms * 1e3, 1), "unit": "tokens/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms, 3), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic token ids, random-init weights",
            "impl": "reference", "reference_class": "torch-only restatement", "same_config": False,
            "deviations": ["frozen base is bf16 (bitsandbytes NF4 is not installable here)",
                           "torch.optim.AdamW fp32 (fused) instead of paged_adamw_8bit",
                           ("torch FSDP2 fully_shard per decoder layer" if best == "fsdp2" else
                            "torch DistributedDataParallel" if best == "ddp" else "single process")
                           + " instead of DeepSpeed ZeRO-3",
                           "LoRA written as a plain nn.Module (peft is not installable here)"],
            "stack": {"torch": torch.__version__, "transformers": __import__("transformers").__version__,
                      "attention": getattr(cfg, "_attn_implementation", "sdpa")},
            "config": {"model": args.model + ("" if not args.layers else f" (DEBUG {args.layers} layers: INVALID)"),
                       "global_batch": B * GA * world, "micro_batch_per_gpu": B, "grad_accum": GA, "seq_len": S,
                       "parallelism": f"{best}-dp{world}", "base": "bf16 (frozen)",
                       "lora": f"r8 alpha16 dropout0.1 q_proj,v_proj ({n_wrapped} wrapped linears)",
                       "optimizer": "torch AdamW fp32", "l2": "weights (29.5 GB bf16) exceed the 126 MB L2: no flush needed",
                       "tokens_per_step": tokens_per_step},
            "clocks": r["clocks"], "build_s": round(build_s, 1),
            "e2e": {"value": round(tokens_per_step / r["ms_e2e"] * 1e3, 1), "unit": "tokens/s",
                    "h2d_bytes_per_step": B * S * 8 * GA, "d2h_bytes_per_step": 4, "ms_per_step": round(r["ms_e2e"], 3),
                    "last_loss": r["last_loss"]},
            "variants": {k: {"ms_per_step": round(v["ms_per_step"], 3), "value": round(tokens_per_step / v["ms_per_step"] * 1e3, 1),
                             "e2e_value": round(tokens_per_step / v["ms_e2e"] * 1e3, 1)} for k, v in results.items()},
        }
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0
