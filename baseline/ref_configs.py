"""Reference arm of BASELINE config 2: the reference's OWN model code under torch DistributedDataParallel.

``baseline/_ref/`` holds unmodified copies of two reference files (``__graft_entry__.build()`` copies them from
/root/reference; they are git-ignored and travel to the GPU box with the snapshot):

    ddp_gpt_wikitext2.py   LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py  (``GPTLike``, 65.97 M parameters)
    minigpt2_model.py      llm-demo/minigpt2/model.py                                         (``Config``, ``MiniGPT``)

The models are imported from those copies and trained exactly as ``ddp_gpt_wikitext2.py:271-308`` does:
``DistributedDataParallel(model, device_ids=[local_rank], output_device=local_rank, find_unused_parameters=False)``,
``optim.AdamW(lr=3e-4)``, ``nn.CrossEntropyLoss``, ``logits -> loss -> zero_grad -> backward -> step``, fp32, no AMP,
no clipping. Only the data is synthetic (``[B, 256]`` random ids: wikitext is not reachable offline and the minigpt2
corpus yields zero samples). ``--amp bf16`` additionally times the same loop under
``torch.autocast`` (not what the reference does; it is the stronger baseline for a bf16 framework).

Nothing from ``llm_in_practise_b200`` is imported. Timing = CUDA events, barrier + synchronize on both sides, max over
ranks, like ``scripts/bench_configs.py``.

    torchrun --nproc-per-node N --master-addr 127.0.0.1 baseline/ref_configs.py --config gptlike-ddp
"""
import argparse
import importlib.util
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))


def _load(fname, modname):
    path = os.path.join(HERE, "_ref", fname)
    if not os.path.exists(path):
        print(json.dumps({"impl": "reference", "unavailable": f"{path} is missing (run __graft_entry__.build() where /root/reference exists)"}))
        sys.exit(0)
    spec = importlib.util.spec_from_file_location(modname, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", required=True, choices=["gptlike-ddp", "minigpt2-ddp"])
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--amp", default="none", choices=["none", "bf16"])
    ap.add_argument("--tf32", type=int, default=0, help="1: torch.backends.cuda.matmul.allow_tf32 (the reference leaves it off)")
    a = ap.parse_args()

    import torch
    import torch.distributed as dist
    import torch.nn as nn
    import torch.optim as optim

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    cuda = torch.cuda.is_available()
    dev = torch.device("cuda", local_rank) if cuda else torch.device("cpu")
    if cuda:
        torch.cuda.set_device(dev)
        torch.backends.cuda.matmul.allow_tf32 = bool(a.tf32)
        torch.backends.cudnn.allow_tf32 = bool(a.tf32)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29513")
        dist.init_process_group("nccl" if cuda else "gloo")
    torch.manual_seed(1234)

    if a.config == "gptlike-ddp":
        ref = _load("ddp_gpt_wikitext2.py", "ref_ddp_gpt_wikitext2")
        B, S, V = a.batch or 16, 256, 30522
        # the reference feeds block[:-1] -> 255 positions; 256 keeps the attention tiles aligned and matches our arm
        model = ref.GPTLike(vocab_size=V, block_size=S, n_layer=6, n_head=12, d_model=768, dropout=0.1)
    else:
        ref = _load("minigpt2_model.py", "ref_minigpt2_model")
        cfg = ref.Config()
        cfg.vocab_size = 4096
        B, S, V = a.batch or 32, cfg.seq_len, cfg.vocab_size
        model = ref.MiniGPT(cfg)
    model.to(dev)
    if world > 1:
        train = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local_rank], output_device=local_rank,
                                                          find_unused_parameters=False) if cuda else \
            torch.nn.parallel.DistributedDataParallel(model)
    else:
        train = model
    optimizer = optim.AdamW(train.parameters(), lr=3e-4)
    criterion = nn.CrossEntropyLoss()
    gen = torch.Generator(device=dev).manual_seed(7 + rank)
    xs = [torch.randint(0, V, (B, S), device=dev, generator=gen) for _ in range(4)]
    k = [0]
    train.train()

    def step():
        x = xs[k[0] % 4]
        k[0] += 1
        if a.amp == "bf16":
            with torch.autocast(dev.type, dtype=torch.bfloat16):
                logits = train(x)
        else:
            logits = train(x)
        loss = criterion(logits.view(-1, V).float(), x.view(-1))
        optimizer.zero_grad()
        loss.backward()
        optimizer.step()

    def sync():
        if world > 1:
            dist.barrier()
        if cuda:
            torch.cuda.synchronize()

    for _ in range(a.warmup):
        step()
    sync()
    if cuda:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.steps):
            step()
        e1.record()
        sync()
        ms = e0.elapsed_time(e1) / a.steps
    else:
        import time
        t0 = time.perf_counter()
        for _ in range(a.steps):
            step()
        ms = (time.perf_counter() - t0) * 1e3 / a.steps
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t)
    if rank == 0:
        n_params = sum(p.numel() for p in model.parameters())
        print(json.dumps({"impl": "reference", "reference_class": "unmodified reference model code under torch DDP (NCCL)",
                          "config": a.config, "metric": f"{a.config} train tok/s", "n_gpus": world, "steps": a.steps,
                          "warmup": a.warmup, "dtype": "fp32" if a.amp == "none" else "bf16 autocast", "tf32": bool(a.tf32),
                          "data": "synthetic token ids, random-init weights", "ms_per_step": round(ms, 3),
                          "value": round(B * S * world / ms * 1e3, 1), "unit": "tokens/s", "params_m": round(n_params / 1e6, 2),
                          "batch_per_gpu": B, "seq_len": S, "torch": torch.__version__}))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
