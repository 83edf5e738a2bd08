"""Multi-GPU correctness cases (run under torchrun on >= 2 GPUs; tests/test_yy_multigpu.py drives them and compares the JSON
lines). Every case trains a small model for a few optimizer steps on fixed synthetic data and prints the per-step losses
and a checksum / dump of the trainable parameters.

  --case zero   ZeRO engine, LoRA on a small NF4 Qwen3, --stage 0 (all-reduce, K-C1) / 1 / 2 / 3 (fused zero_step, K-C2),
                --opt adamw|adamw8bit. B200_COMM=nccl selects the torch.distributed path of the same engine (the oracle).
  --case ddp    bucketed DDP wrapper (--impl ours) or torch.nn.parallel.DistributedDataParallel around the SAME model
                (--impl torch), GPTLike 4 layers, fp32.
  --case fsdp   FullyShardedDataParallel (FULL_SHARD, bf16 reduce) on GPTLike; B200_COMM=nccl = NCCL collectives.
  --case dense3 ZeRO-3 of a dense GPTLike through the engine (parameters partitioned); prints peak memory.
  --skew        rank (step % world) sleeps ~10 ms on the device before its backward: results must not change.
"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from llm_in_practise_b200.parallel.dist import cleanup_distributed, setup_distributed

ap = argparse.ArgumentParser()
ap.add_argument("--case", default="", choices=["", "zero", "ddp", "fsdp", "dense3"])
ap.add_argument("--suite", default="", choices=["", "multi", "single"],
                help="multi: every variant the tests compare, one after the other in ONE process group (a torchrun start "
                     "per variant would cost more GPU time than the variants); single: the 1-GPU counterparts")
ap.add_argument("--stage", type=int, default=2)
ap.add_argument("--opt", default="adamw")
ap.add_argument("--impl", default="ours")
ap.add_argument("--comm", default="", help="symm | nccl (sets B200_COMM for this variant)")
ap.add_argument("--skew", action="store_true")
ap.add_argument("--synthetic", action="store_true", help="zero case: seeded random gradients instead of training (collective in isolation)")
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--dump", default="", help="file (single case) or directory (suite) for the parameter dumps")
cli = ap.parse_args()
info = setup_distributed()
rank, world = info.rank, info.world_size
dev = torch.device("cuda", info.local_rank)
TOTAL = int(os.environ.get("CHECK_TOTAL_RANKS", world))     # data is generated for TOTAL virtual ranks


def gptlike(dtype=torch.float32, layers=4):
    from llm_in_practise_b200.models.gpt import GPTLike
    torch.manual_seed(3)
    return GPTLike(2048, 128, layers, 4, 256, dropout=0.0, dtype=dtype, device=dev)


def run_case(a, dump=""):
    if a.comm:
        os.environ["B200_COMM"] = a.comm
    out = {"case": a.case, "world": world, "comm": os.environ.get("B200_COMM", "symm"), "skew": a.skew, "stage": a.stage,
           "opt": a.opt, "impl": a.impl}

    def maybe_skew(step):
        if a.skew and world > 1 and step % world == rank:
            torch.cuda._sleep(20_000_000)

    gen = torch.Generator().manual_seed(99)
    losses = []
    if a.case == "zero":
        from llm_in_practise_b200.models.qwen3 import Qwen3Config, Qwen3ForCausalLM
        from llm_in_practise_b200.parallel.zero import initialize
        from llm_in_practise_b200.peft import LoraConfig
        cfg = Qwen3Config(vocab_size=2048, hidden_size=1024, intermediate_size=2048, num_hidden_layers=4, num_attention_heads=8,
                          num_key_value_heads=4, head_dim=128, max_position_embeddings=512)
        torch.manual_seed(7)
        model = Qwen3ForCausalLM(cfg, dtype=torch.bfloat16, device=dev, init_std=0.05)
        model.quantize_nf4(True)
        eng = initialize(model=model, lora_config=LoraConfig(r=8, lora_alpha=16, lora_dropout=0.0, target_modules=["q_proj", "v_proj"]),
                         config={"bf16": {"enabled": True}, "zero_optimization": {"stage": a.stage}, "gradient_accumulation_steps": 2,
                                 "gradient_clipping": 1.0, "train_micro_batch_size_per_gpu": 2,
                                 "optimizer": {"type": "paged_adamw_8bit" if a.opt == "adamw8bit" else "AdamW",
                                               "params": {"lr": 1e-3, "weight_decay": 0.0}}})
        data = torch.randint(0, cfg.vocab_size, (a.steps, 2, TOTAL, 2, 256), generator=gen)   # [step, micro, vrank, B, S]
        if getattr(a, "synthetic", False):
            # the collective in isolation: every rank's gradient arena is a seeded random tensor (identical in the runs
            # that are compared), so the only difference between the fused kernel and NCCL is the order of the fp32 adds
            for step in range(a.steps):
                g = torch.Generator(device=dev).manual_seed(1000 * step + rank)
                eng.arena.grad.copy_(torch.randn(eng.arena.numel, device=dev, generator=g) * 1e-2)
                eng.micro_steps = eng.ga * (step + 1)
                eng.step()
                losses.append(0.0)
        for step in range(0 if getattr(a, "synthetic", False) else a.steps):
            for m in range(2):
                for v in [v for v in range(TOTAL) if v % world == rank]:
                    ids = data[step, m, v].to(dev)
                    if m == 1:
                        maybe_skew(step)
                    loss = eng.train_micro_batch(ids, ids)
                losses.append(float(loss))
            eng.step()
        params = eng.arena.data.detach().float().cpu()
        out["fused"] = bool(eng.use_symm)
    elif a.case == "ddp":
        model = gptlike()
        data = torch.randint(0, 2048, (a.steps, TOTAL, 4, 128), generator=gen)
        if a.impl == "ours":
            from llm_in_practise_b200.parallel.ddp import DistributedDataParallel as DDP
            ddp = DDP(model, bucket_cap_mb=1.0)
            opt = ddp.make_optimizer(lr=1e-3, weight_decay=0.0)
            out["fused"] = bool(getattr(ddp, "use_symm", False))
            out["buckets"] = len(ddp.buckets)
        else:
            ddp = torch.nn.parallel.DistributedDataParallel(model, device_ids=[info.local_rank], bucket_cap_mb=1.0) if world > 1 else model
            opt = torch.optim.AdamW(model.parameters(), lr=1e-3, weight_decay=0.0)
        for step in range(a.steps):
            x = data[step, rank % TOTAL].to(dev)
            o = ddp(x, x)
            loss = o[1] if isinstance(o, tuple) else o
            maybe_skew(step)
            loss.backward()
            opt.step()
            ddp.zero_grad() if a.impl == "ours" else opt.zero_grad()
            losses.append(float(loss))
        params = torch.cat([p.detach().float().reshape(-1).cpu() for p in model.parameters()])
    elif a.case == "fsdp":
        from llm_in_practise_b200.models.gpt import TransformerBlock
        from llm_in_practise_b200.parallel.fsdp import (FullyShardedDataParallel as FSDP, MixedPrecision, ShardingStrategy,
                                                        transformer_auto_wrap_policy)
        model = gptlike(torch.bfloat16)
        f = FSDP(model, auto_wrap_policy=transformer_auto_wrap_policy(transformer_layer_cls={TransformerBlock}),
                 sharding_strategy=ShardingStrategy.FULL_SHARD,
                 mixed_precision=MixedPrecision(param_dtype=torch.bfloat16, reduce_dtype=torch.bfloat16, buffer_dtype=torch.bfloat16),
                 sync_module_states=True)
        opt = torch.optim.AdamW(f.parameters(), lr=1e-3, weight_decay=0.0)
        data = torch.randint(0, 2048, (a.steps, TOTAL, 4, 128), generator=gen)
        for step in range(a.steps):
            x = data[step, rank % TOTAL].to(dev)
            o = f(x, x)
            loss = o[1] if isinstance(o, tuple) else o
            maybe_skew(step)
            loss.backward()
            f.clip_grad_norm_(1.0)
            opt.step()
            f.zero_grad()
            losses.append(float(loss))
        sd = f.full_state_dict(cpu=True)
        params = torch.cat([v.reshape(-1) for v in sd.values()])
        out["fused"] = bool(f.use_symm)
        out["units"] = len(f.units)
        out["exposed_ms"] = round(f.exposed_ms(), 3) if f.use_symm else None
    else:
        from llm_in_practise_b200.parallel.zero import initialize
        import gc
        gc.collect()
        torch.cuda.empty_cache()
        base_mem = torch.cuda.memory_allocated()        # whatever earlier variants of a suite left behind
        # 1 rank: the replicated baseline (fp32 arena: parameters + gradients + AdamW moments = 16 B / parameter)
        model = gptlike(torch.bfloat16 if world > 1 else torch.float32, layers=8)
        n_params = sum(p.numel() for p in model.parameters())
        eng = initialize(model=model, config={"bf16": {"enabled": world > 1}, "zero_optimization": {"stage": a.stage},
                                              "gradient_accumulation_steps": 1, "gradient_clipping": 1.0,
                                              "train_micro_batch_size_per_gpu": 4,
                                              "optimizer": {"type": "AdamW", "params": {"lr": 1e-3, "weight_decay": 0.0}}})
        data = torch.randint(0, 2048, (a.steps, TOTAL, 4, 128), generator=gen)
        for step in range(a.steps):
            x = data[step, rank % TOTAL].to(dev)
            o = eng(x, x)
            loss = o[1] if isinstance(o, tuple) else o
            eng.backward(loss)
            eng.step()
            losses.append(float(loss))
        torch.cuda.synchronize()
        sd = eng.consolidated_state_dict()
        params = torch.cat([v.float().reshape(-1).cpu() for v in sd.values()])
        out["params"] = n_params
        del sd
        gc.collect()
        out["persistent_bytes_per_rank"] = int(torch.cuda.memory_allocated() - base_mem)   # after the step: masters + state (+ replicas)
        out["dense_sharded"] = eng._fsdp is not None

    out["losses"] = losses
    out["abs_sum"] = float(params.abs().sum())
    if dump and rank == 0:
        torch.save(params, dump)
    # every rank must hold the same parameters
    chk = torch.tensor([out["abs_sum"]], device=dev, dtype=torch.float64)
    mx, mn = chk.clone(), chk.clone()
    if world > 1:
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(mn, op=dist.ReduceOp.MIN)
    out["ranks_identical"] = bool(float(mx) - float(mn) <= 1e-9 * abs(float(mx)))
    if rank == 0:
        print(json.dumps(out), flush=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    return out


def V(**kw):
    d = dict(case="zero", stage=2, opt="adamw", impl="ours", comm="symm", skew=False, steps=3, synthetic=False)
    d.update(kw)
    return argparse.Namespace(**d)


SUITES = {
    "multi": {
        "grad0_symm": V(stage=0, synthetic=True), "grad0_nccl": V(stage=0, comm="nccl", synthetic=True),
        "grad2_symm": V(stage=2, synthetic=True), "grad2_nccl": V(stage=2, comm="nccl", synthetic=True),
        "grad3_8bit_symm": V(stage=3, opt="adamw8bit", synthetic=True),
        "grad3_8bit_nccl": V(stage=3, opt="adamw8bit", comm="nccl", synthetic=True),
        "zero2_symm": V(stage=2), "zero2_nccl": V(stage=2, comm="nccl"), "zero2_skew": V(stage=2, skew=True),
        "zero3_symm": V(stage=3),
        "ddp_ours": V(case="ddp", steps=4), "ddp_torch": V(case="ddp", impl="torch", steps=4),
        "ddp_skew": V(case="ddp", steps=4, skew=True),
        "fsdp_symm": V(case="fsdp", steps=4), "fsdp_nccl": V(case="fsdp", steps=4, comm="nccl"),
        "fsdp_skew": V(case="fsdp", steps=4, skew=True),
        "dense3": V(case="dense3", stage=3),
    },
    "single": {"zero3_symm": V(stage=3), "dense3": V(case="dense3", stage=3)},
}
if cli.suite:
    if cli.dump and rank == 0:
        os.makedirs(cli.dump, exist_ok=True)
    for name, v in SUITES[cli.suite].items():
        try:
            r = run_case(v, os.path.join(cli.dump, name + ".pt") if cli.dump else "")
        except Exception as e:      # keep going: the tests report which variant failed
            import traceback
            if rank == 0:
                print(json.dumps({"variant": name, "error": f"{type(e).__name__}: {e}", "trace": traceback.format_exc()[-1500:]}), flush=True)
            raise
        if rank == 0:
            print(json.dumps({"variant": name, **r}), flush=True)
else:
    run_case(cli, cli.dump)
sys.stdout.flush()
cleanup_distributed()
os._exit(0)      # the work is done and printed: skip interpreter teardown (symmetric-memory handles outliving the process group abort there)
