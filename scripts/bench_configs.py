"""Secondary training configurations of BASELINE.md section 2 (rows 2 and 3), timed like bench.py (CUDA events, barrier +
synchronize on both sides, max over ranks, synthetic ids, random-init weights):

  --config gptlike-ddp    GPTLike-66M (6 layers, 12 heads, d 768, block 256, vocab 30522: the model of
                          LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py:86-165, the bandwidth-bound DDP
                          case; 251.7 MiB of fp32 gradients per step) under the bucketed, overlapped DDP wrapper, bf16
                          working weights; reports train tokens/s and the exposed all-reduce time per step
  --config minigpt2-ddp   llm-demo/minigpt2/model.py:5-72 (4 layers, 4 heads, d 128, seq 256) under the same wrapper
  --config lora8b-4k      Qwen3-8B LoRA bf16 (r 8, alpha 16, dropout 0.1 on q_proj, v_proj;
                          Fine-Tuning/qwen3-8b-lora-dist.py:118-173), 4096-token sequences, micro-batch 1 x accum 2,
                          data parallel through the ZeRO engine at stage 0 (= DDP) or --stage 2 / 3

  torchrun --nproc-per-node N --master-addr 127.0.0.1 scripts/bench_configs.py --config gptlike-ddp
"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from llm_in_practise_b200.parallel.dist import cleanup_distributed, setup_distributed

ap = argparse.ArgumentParser()
ap.add_argument("--config", required=True, choices=["gptlike-ddp", "minigpt2-ddp", "lora8b-4k"])
ap.add_argument("--steps", type=int, default=8)
ap.add_argument("--warmup", type=int, default=3)
ap.add_argument("--batch", type=int, default=0)
ap.add_argument("--stage", type=int, default=0)
ap.add_argument("--layers", type=int, default=0)
ap.add_argument("--graph", action="store_true", help="gptlike / minigpt2 on one GPU: the whole step replayed from a CUDA graph (utils/graph_step.py)")
ap.add_argument("--wrapper", default="ddp", choices=["ddp", "fsdp"], help="gptlike / minigpt2: DDP (reference ddp_basics) or FSDP FULL_SHARD (fsdp_basics)")
a = ap.parse_args()
info = setup_distributed()
rank, world, dev = info.rank, info.world_size, info.device
torch.manual_seed(1234)


def sync():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def timed(step_fn):
    for _ in range(a.warmup):
        step_fn()
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        step_fn()
    e1.record()
    sync()
    t = torch.tensor([e0.elapsed_time(e1) / a.steps], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


out = {"config": a.config, "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "dtype": "bf16", "data": "synthetic token ids, random-init weights"}
if a.config in ("gptlike-ddp", "minigpt2-ddp"):
    from llm_in_practise_b200.parallel.ddp import DistributedDataParallel as DDP
    if a.config == "gptlike-ddp":
        from llm_in_practise_b200.models.gpt import GPTLike
        B, S, V = a.batch or 16, 256, 30522
        model = GPTLike(V, S, 6, 12, 768, dropout=0.1, dtype=torch.bfloat16, device=dev)
    else:
        from llm_in_practise_b200.models.minigpt import Config, MiniGPT2
        cfg = Config(vocab_size=4096)
        B, S, V = a.batch or 32, cfg.seq_len, 4096
        model = MiniGPT2(cfg, causal=True, dtype=torch.bfloat16, device=dev)
    if a.wrapper == "fsdp":
        # LLM_Distributed_Trainning/PyTorch/fsdp_basics/fsdp_gpt_wikitext2.py:278-312: FULL_SHARD, one unit per block, bf16
        from llm_in_practise_b200.models.gpt import TransformerBlock
        from llm_in_practise_b200.parallel.fsdp import (FullyShardedDataParallel as FSDP, MixedPrecision, ShardingStrategy,
                                                        transformer_auto_wrap_policy)
        ddp = FSDP(model, auto_wrap_policy=transformer_auto_wrap_policy(transformer_layer_cls={TransformerBlock}),
                   sharding_strategy=ShardingStrategy.FULL_SHARD,
                   mixed_precision=MixedPrecision(param_dtype=torch.bfloat16, reduce_dtype=torch.bfloat16, buffer_dtype=torch.bfloat16))
        opt = torch.optim.AdamW(ddp.parameters(), lr=3e-4, weight_decay=0.01, fused=True)
    else:
        ddp = DDP(model, mixed_precision=torch.bfloat16)
        opt = ddp.make_optimizer(lr=3e-4, weight_decay=0.01)
    gen = torch.Generator(device=dev).manual_seed(7 + rank)
    xs = [torch.randint(0, V, (B, S), device=dev, generator=gen) for _ in range(4)]
    k = [0]

    def step():
        x = xs[k[0] % 4]
        k[0] += 1
        o = ddp(x, x)
        loss = o[1] if isinstance(o, tuple) else o
        loss.backward()
        opt.step()
        ddp.zero_grad()
    if a.graph:
        if world != 1 or a.wrapper != "ddp":
            raise SystemExit("--graph: single process, DDP wrapper only")
        from llm_in_practise_b200.utils.graph_step import GraphedTrainStep

        def loss_fn(x):
            o = ddp(x, x)
            return o[1] if isinstance(o, tuple) else o
        gstep = GraphedTrainStep(loss_fn, opt, [xs[0]], zero_grad=ddp.zero_grad)

        def step():                                          # noqa: F811 - the graph-replayed step replaces the eager one
            x = xs[k[0] % 4]
            k[0] += 1
            gstep(x)
        out["cuda_graph"] = True
    ms = timed(step)
    n_params = sum(p.numel() for p in model.parameters())
    torch.cuda.synchronize()
    if a.wrapper == "fsdp":
        exposed = [ddp.exposed_ms() / (a.steps + a.warmup)] if ddp.use_symm else []
        out["wrapper"], out["units"] = "fsdp-full-shard", len(ddp.units)
        out["peak_mem_gb"] = round(torch.cuda.max_memory_allocated() / 2 ** 30, 2)
    else:
        exposed = ddp.exposed_ms[-a.steps:]
        out["exposed_events"] = len(ddp._exposed_events)
    out.update({"metric": f"{a.config} train tok/s", "ms_per_step": round(ms, 3), "value": round(B * S * world / ms * 1e3, 1), "unit": "tokens/s",
                "params_m": round(n_params / 1e6, 2), "grad_bytes_per_step": n_params * 4, "batch_per_gpu": B, "seq_len": S,
                "exposed_allreduce_ms": round(sum(exposed) / max(len(exposed), 1), 3) if exposed else None,
                "comm": "nvlink-peer-fused" if getattr(ddp, "use_symm", False) else ("single" if world == 1 else "nccl")})
else:
    import dataclasses
    from llm_in_practise_b200.models.qwen3 import QWEN3_PRESETS, Qwen3ForCausalLM
    from llm_in_practise_b200.parallel.zero import initialize
    from llm_in_practise_b200.peft import LoraConfig
    cfg = QWEN3_PRESETS["qwen3-8b"]
    if a.layers:
        cfg = dataclasses.replace(cfg, num_hidden_layers=a.layers)
    model = Qwen3ForCausalLM(cfg, dtype=torch.bfloat16, device=dev)
    B, S, GA = a.batch or 1, 4096, 2
    eng = initialize(model=model, lora_config=LoraConfig(r=8, lora_alpha=16, lora_dropout=0.1, target_modules=["q_proj", "v_proj"]),
                     config={"bf16": {"enabled": True}, "zero_optimization": {"stage": a.stage}, "gradient_accumulation_steps": GA,
                             "gradient_clipping": 1.0, "train_micro_batch_size_per_gpu": B,
                             "optimizer": {"type": "AdamW", "params": {"lr": 5e-5, "weight_decay": 0.01}}})
    gen = torch.Generator(device=dev).manual_seed(7 + rank)
    xs = [torch.randint(0, cfg.vocab_size, (B, S), device=dev, generator=gen) for _ in range(4)]
    k = [0]

    def step():
        for _ in range(GA):
            x = xs[k[0] % 4]
            k[0] += 1
            eng.train_micro_batch(x, x)
        eng.step()
    ms = timed(step)
    out.update({"metric": "Qwen3-8B LoRA bf16 train tok/s (4k-token sequences)" + (f" [{a.layers} layers: INVALID]" if a.layers else ""),
                "ms_per_step": round(ms, 3), "value": round(B * S * GA * world / ms * 1e3, 1), "unit": "tokens/s", "micro_batch_per_gpu": B,
                "grad_accum": GA, "seq_len": S, "parallelism": f"zero{a.stage}-dp{world}" if a.stage else f"ddp{world}",
                "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2 ** 30, 1)})
if rank == 0:
    print(json.dumps(out))
cleanup_distributed()
