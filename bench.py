#!/usr/bin/env python
"""Headline benchmark (BASELINE.json): Qwen3-14B QLoRA (NF4 base, LoRA r=8 on q_proj/v_proj, dropout 0.1,
8-bit AdamW) ZeRO-3 training throughput in tokens/s at 1/2/4/8 B200.

Configuration = the reference's Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py + ds_zero3_config.json:
per-device micro-batch 2 x 512 tokens, gradient_accumulation_steps 2, bf16 compute, gradient clipping 1.0,
no activation checkpointing. One "step" is one optimizer step (2 micro-batches). Synthetic token ids,
random-init weights of the Qwen3-14B architecture (no network for datasets or checkpoints).

    python bench.py --gpus 1 --steps 10 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus 8 --steps 10 --warmup 3
    python bench.py --impl reference ...     # the unmodified reference stack, if it can be installed

Prints ONE JSON line on rank 0 (see the contract in the repo README / DESIGN.md).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Qwen3-14B QLoRA ZeRO-3 train tok/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="qwen3-14b")
    ap.add_argument("--micro-batch", type=int, default=2)
    ap.add_argument("--seq-len", type=int, default=512)
    ap.add_argument("--grad-accum", type=int, default=2)
    ap.add_argument("--layers", type=int, default=0, help="debug only: override the layer count (result is then marked invalid)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--sequential-accum", action="store_true",
                    help="run the micro-batches one after the other (round-1 schedule) instead of one layer-major pass")
    ap.add_argument("--local_rank", type=int, default=-1)
    return ap.parse_args()


# --------------------------------------------------------------------------------------------------
class ClockSampler:
    """Samples SM clocks / throttle reasons with nvidia-smi while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.proc = None
        self.lines = []
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.idx)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        def pump():
            for ln in self.proc.stdout:
                self.lines.append(ln.strip())
        self.thread = threading.Thread(target=pump, daemon=True)
        self.thread.start()

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def reference_arm(args):
    """``--impl reference``. The stock headline script (Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py) needs peft +
    bitsandbytes + deepspeed + accelerate, which are neither installed nor in /opt/wheelhouse, and /root/reference has no
    setup.py / pyproject.toml to install. What runs instead is ``baseline/ref_headline.py``: the same training step on the
    part of the reference's stack that IS in the image (HF transformers ``Qwen3ForCausalLM`` + torch AdamW + torch
    FSDP2 / DDP over NCCL). It imports nothing from this repository's package and labels itself
    ``"reference_class": "torch-only restatement", "same_config": false`` with the list of deviations."""
    import importlib.util
    path = os.path.join(ROOT, "baseline", "ref_headline.py")
    try:
        spec = importlib.util.spec_from_file_location("ref_headline", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        import transformers  # noqa: F401  (the restatement needs HF's Qwen3)
    except Exception as e:  # pragma: no cover - depends on the image
        print(json.dumps({"impl": "reference", "unavailable": f"torch-only restatement cannot start: {type(e).__name__}: {e}"}))
        return 0
    return mod.run(args)


# --------------------------------------------------------------------------------------------------
def main():
    args = parse_args()
    if args.impl == "reference":
        return reference_arm(args)

    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        print(json.dumps({"metric": METRIC, "error": "no CUDA device"}))
        return 1
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        dist.init_process_group("nccl", device_id=dev)

    from llm_in_practise_b200.models.qwen3 import QWEN3_PRESETS, Qwen3ForCausalLM, flops_per_token
    from llm_in_practise_b200.ops import _native as nv
    from llm_in_practise_b200.peft import LoraConfig
    from llm_in_practise_b200.parallel.zero import initialize

    nv.lib()  # fail loudly if the kernels are not built
    cfg = QWEN3_PRESETS[args.model]
    if args.layers:
        import dataclasses
        cfg = dataclasses.replace(cfg, num_hidden_layers=args.layers)
    torch.manual_seed(1234)  # same base weights on every rank
    t_build = time.time()
    model = Qwen3ForCausalLM(cfg, dtype=torch.bfloat16, device=dev)
    model.quantize_nf4(double_quant=True)
    torch.cuda.empty_cache()
    lora_cfg = LoraConfig(r=8, lora_alpha=16, lora_dropout=0.1, target_modules=["q_proj", "v_proj"])
    ds_config = {
        "bf16": {"enabled": True},
        "zero_optimization": {"stage": 3, "overlap_comm": True, "contiguous_gradients": True,
                              "reduce_bucket_size": "auto", "stage3_prefetch_bucket_size": "auto",
                              "stage3_param_persistence_threshold": "auto",
                              "stage3_gather_16bit_weights_on_model_save": True},
        "gradient_accumulation_steps": args.grad_accum, "gradient_clipping": 1.0,
        "train_micro_batch_size_per_gpu": args.micro_batch, "train_batch_size": "auto",
        "optimizer": {"type": "paged_adamw_8bit", "params": {"lr": 1e-4, "weight_decay": 0.0}},
    }
    engine = initialize(model=model, lora_config=lora_cfg, config=ds_config)
    build_s = time.time() - t_build

    B, S, GA = args.micro_batch, args.seq_len, args.grad_accum
    tokens_per_step = B * S * GA * world
    gen = torch.Generator().manual_seed(100 + rank)
    n_host = max(4, min(16, args.steps + args.warmup)) * GA
    host_batches = [torch.randint(0, cfg.vocab_size, (B, S), generator=gen).pin_memory() for _ in range(n_host)]
    dev_batches = [b.to(dev) for b in host_batches]

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def run_steps(n, e2e):
        """n optimizer steps. e2e: inputs come from pinned host memory and the loss is read back each step."""
        k = 0
        loss_val = None
        for _ in range(n):
            mbs = []
            for _m in range(GA):
                mbs.append(host_batches[k % n_host].to(dev, non_blocking=True) if e2e else dev_batches[k % n_host])
                k += 1
            if args.sequential_accum:
                for ids in mbs:
                    loss = engine.train_micro_batch(ids, ids)
            else:
                loss = engine.train_batch(mbs)      # the GA micro-batches of the step, one layer-major pass
            engine.step()
            if e2e:
                loss_val = float(loss.item())  # D2H read of the step's result
        return loss_val

    # ---- warm-up (untimed) ----
    run_steps(args.warmup, e2e=False)
    sync_all()

    # ---- device-timed region ----
    engine.time_skew = world > 1
    engine.reset_comm_timers()
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = nv.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    ev0.record()
    t_host0 = time.perf_counter()
    run_steps(args.steps, e2e=False)
    host_enqueue_ms = (time.perf_counter() - t_host0) * 1e3 / args.steps   # how long Python needs to ISSUE a step's launches
    ev1.record()
    sync_all()
    ms = ev0.elapsed_time(ev1)
    launches = nv.launch_count - launches0
    clocks = sampler.stop()
    exposed = engine.exposed_comm_ms()
    stats = engine.comm_stats() if hasattr(engine, "comm_stats") else {}
    engine.time_skew = False

    # ---- end-to-end region (pinned-host inputs each step, loss read back each step) ----
    e2e = None
    if not args.no_e2e:
        run_steps(1, e2e=True)
        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        last_loss = run_steps(args.steps, e2e=True)
        e1.record()
        sync_all()
        ms_e2e = e0.elapsed_time(e1)
    else:
        ms_e2e = None
        last_loss = None

    def max_over_ranks(x):
        if world == 1 or x is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    ms = max_over_ranks(ms)
    ms_e2e = max_over_ranks(ms_e2e)
    exposed = max_over_ranks(exposed)
    for k_ in ("zero_step_ms", "step_skew_ms"):
        if k_ in stats:
            stats[k_] = round(max_over_ranks(stats[k_]), 4)

    if rank == 0:
        ms_per_step = ms / args.steps
        value = tokens_per_step / (ms_per_step / 1e3)
        out = {
            "metric": METRIC, "value": round(value, 1), "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms_per_step, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic token ids, random-init weights",
            "impl": "b200",
            "config": {"model": args.model + ("" if not args.layers else f" (DEBUG {args.layers} layers: INVALID)"),
                       "global_batch": B * GA * world, "micro_batch_per_gpu": B, "grad_accum": GA, "seq_len": S,
                       "parallelism": f"zero3-dp{world}", "base": "NF4 double-quant",
                       "accumulation": "sequential micro-batches" if args.sequential_accum else
                       f"the {GA} micro-batches of a step run as one layer-major pass (same gradients, weights gathered/expanded once)", "lora": "r8 alpha16 dropout0.1 q_proj,v_proj",
                       "optimizer": "8-bit blockwise AdamW", "l2": "weights (7.5 GB NF4) exceed the 126 MB L2: no flush needed",
                       "tokens_per_step": tokens_per_step},
            "clocks": clocks,
            "gpu_launches": launches,
            # wall time the host needs to issue one step's launches (no synchronisation inside): close to ms_per_step means
            # the step is bound by the Python / ctypes launch path rather than by the GPU
            "host_enqueue_ms_per_step": round(host_enqueue_ms, 3),
            "model_tflops_per_gpu": round(flops_per_token(cfg, S) * tokens_per_step / world / (ms_per_step / 1e3) / 1e12, 1),
            "build_s": round(build_s, 1),
        }
        if ms_e2e is not None:
            h2d = B * S * 8 * GA  # int64 token ids, one copy per micro-batch (labels alias the ids)
            out["e2e"] = {"value": round(tokens_per_step / (ms_e2e / args.steps / 1e3), 1), "unit": "tokens/s",
                          "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4, "ms_per_step": round(ms_e2e / args.steps, 3),
                          "last_loss": last_loss}
        if stats:
            out["comm"] = stats
        # second half of BASELINE.json's metric: exposed communication per optimizer step, max over ranks
        out["exposed_comm_ms"] = None if exposed is None else round(exposed, 3)
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
