"""K-C1 all-reduce over NVLink: peer-pointer two-shot kernel vs the NVLS (multimem, in-switch reduction) kernel vs NCCL.

    torchrun --nproc-per-node N scripts/bench_allreduce.py [--sizes-mib 1,8,64,252] [--check-only]

Correctness: every variant against the NCCL fp32 sum of the same data (fp32: rel <= 1e-6; bf16: rel <= 1e-2).
Timing: CUDA events on the launching stream, max over ranks; `busbw_frac` = 2 (W-1)/W * bytes / time / 900 GB/s, the
ring-equivalent bus bandwidth NCCL reports, as a fraction of one NVLink 5 direction (what ncclAllReduce under DDP,
LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py:271-279, is bound by)."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from llm_in_practise_b200.parallel.symm import PeerFlags, SymmBuffer, allreduce_, allreduce_bf16_nvls_, nvls_enabled
from llm_in_practise_b200.parallel.dist import cleanup_distributed, setup_distributed

ap = argparse.ArgumentParser()
ap.add_argument("--sizes-mib", default="1,8,64,252")
ap.add_argument("--reps", type=int, default=10)
ap.add_argument("--check-only", action="store_true")
a = ap.parse_args()
info = setup_distributed()
rank, world, dev = info.rank, info.world_size, info.device
sizes = [int(float(s) * (1 << 20)) for s in a.sizes_mib.split(",")]
nmax = max(sizes) // 4
buf = SymmBuffer(nmax, torch.float32, dev)
low = SymmBuffer(nmax, torch.bfloat16, dev)
b16 = SymmBuffer(nmax * 2, torch.bfloat16, dev)
flags = PeerFlags(dev)
have_nvls = bool(buf.multicast_ptr)
out = {"world": world, "nvls_available": have_nvls, "rows": [], "ok": True}


def timed(fn, reps):
    for _ in range(3):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def rel(x, y):
    return float((x.float() - y.float()).norm() / y.float().norm().clamp_min(1e-30))


for nbytes in sizes:
    n = nbytes // 4 // 8 * 8
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    src = torch.randn(n, device=dev, generator=g)
    want = src.clone()
    dist.all_reduce(want)
    want_avg = want / world
    row = {"mib": round(nbytes / 2 ** 20, 1)}
    # ---- correctness ----
    for name, env in (("peer", "0"), ("nvls", "1")):
        if name == "nvls" and not have_nvls:
            continue
        os.environ["B200_NVLS"] = env
        buf.local[:n].copy_(src)
        allreduce_(buf, flags, 0, n, 1.0 / world, lowp=low)
        torch.cuda.synchronize()
        row[f"{name}_fp32_rel"] = rel(buf.local[:n], want_avg)
        row[f"{name}_lowp_rel"] = rel(low.local[:n], want_avg)
        out["ok"] &= row[f"{name}_fp32_rel"] < 1e-6 and row[f"{name}_lowp_rel"] < 1e-2
    if have_nvls:
        os.environ["B200_NVLS"] = "1"
        b16.local[:n].copy_(src)
        allreduce_bf16_nvls_(b16, flags, 0, n)
        torch.cuda.synchronize()
        row["nvls_bf16_rel"] = rel(b16.local[:n], want)
        out["ok"] &= row["nvls_bf16_rel"] < 1e-2
    # ---- timing ----
    if not a.check_only:
        bus = 2.0 * (world - 1) / world
        def frac(ms, bytes_):
            return round(bus * bytes_ / (ms * 1e-3) / 900e9, 3)
        os.environ["B200_NVLS"] = "0"
        ms = timed(lambda: allreduce_(buf, flags, 0, n, 1.0 / world), a.reps)
        row["peer_fp32_ms"], row["peer_fp32_busbw_frac"] = round(ms, 4), frac(ms, n * 4)
        if have_nvls:
            os.environ["B200_NVLS"] = "1"
            ms = timed(lambda: allreduce_(buf, flags, 0, n, 1.0 / world), a.reps)
            row["nvls_fp32_ms"], row["nvls_fp32_busbw_frac"] = round(ms, 4), frac(ms, n * 4)
            ms = timed(lambda: allreduce_bf16_nvls_(b16, flags, 0, n), a.reps)
            row["nvls_bf16_ms"], row["nvls_bf16_busbw_frac"] = round(ms, 4), frac(ms, n * 2)
        t32 = src.clone()
        ms = timed(lambda: dist.all_reduce(t32), a.reps)
        row["nccl_fp32_ms"], row["nccl_fp32_busbw_frac"] = round(ms, 4), frac(ms, n * 4)
        t16 = src.to(torch.bfloat16)
        ms = timed(lambda: dist.all_reduce(t16), a.reps)
        row["nccl_bf16_ms"], row["nccl_bf16_busbw_frac"] = round(ms, 4), frac(ms, n * 2)
    out["rows"].append(row)
os.environ.pop("B200_NVLS", None)
ok = torch.tensor([1 if out["ok"] else 0], device=dev)
dist.all_reduce(ok, op=dist.ReduceOp.MIN)
out["ok"] = bool(int(ok))
if rank == 0:
    print(json.dumps(out), flush=True)
torch.cuda.synchronize()
dist.barrier()
sys.stdout.flush()
cleanup_distributed()
os._exit(0)
