"""K-C5: row-parallel tcgen05 GEMM with the all-reduce inside the launch (`TensorParallel.gemm_allreduce`) against
GEMM + NCCL all-reduce of the same operands, and timed against the unfused path (GEMM, then the NVLS all-reduce kernel).

    torchrun --nproc-per-node N scripts/check_gemm_allreduce.py [--time]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from llm_in_practise_b200.ops.gemm import gemm
from llm_in_practise_b200.parallel.dist import cleanup_distributed, setup_distributed
from llm_in_practise_b200.parallel.tp import TensorParallel

ap = argparse.ArgumentParser()
ap.add_argument("--time", action="store_true")
a = ap.parse_args()
info = setup_distributed()
rank, world, dev = info.rank, info.world_size, info.device
H = 5120
tp = TensorParallel(H, max_tokens=8192, device=dev)
out = {"world": world, "nvls": bool(tp.fused and tp.lowp.multicast_ptr), "rows": [], "ok": True}
shapes = [(80, 1024, 512), (80, 1024, 1024), (100, 4096, 2048), (512, 4096, 4096 // world), (4096, 4096, 4096 // world), (300, 5120, 1024), (130, 1024, 256), (2048, 5120, 17408 // world // 64 * 64)]
for M, N, k in shapes:
    g = torch.Generator(device=dev).manual_seed(1000 * rank + M)
    x = torch.randn(M, k, device=dev, generator=g).to(torch.bfloat16)
    w = (torch.randn(N, k, device=dev, generator=g) * k ** -0.5).to(torch.bfloat16)
    want = x.float() @ w.float().t()
    dist.all_reduce(want)
    row = {"M": M, "N": N, "k_local": k, "fusable": bool(tp.can_fuse_gemm(x, N))}
    if row["fusable"]:
        y = tp.gemm_allreduce(x, w)
        y2 = tp.gemm_allreduce(x, w)            # same buffer again: the epochs / tickets of a second launch
        torch.cuda.synchronize()
        row["rel"] = float((y.float() - want).norm() / want.norm())
        row["repeat_identical"] = bool(torch.equal(y, y2))
        chk = y.float().abs().sum().double().reshape(1)
        mx, mn = chk.clone(), chk.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(mn, op=dist.ReduceOp.MIN)
        row["ranks_identical"] = bool(float(mx) == float(mn))
        out["ok"] &= row["rel"] < 1e-2 and row["repeat_identical"] and row["ranks_identical"]
        if a.time:
            def timed(fn, reps=20):
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
                return round(float(t) * 1e3, 1)
            row["fused_us"] = timed(lambda: tp.gemm_allreduce(x, w))
            row["gemm_then_nvls_allreduce_us"] = timed(lambda: tp.all_reduce(gemm(x, w)))
            row["gemm_only_us"] = timed(lambda: gemm(x, w))
            os.environ["B200_NVLS"] = "0"
            row["gemm_then_peer_allreduce_us"] = timed(lambda: tp.all_reduce(gemm(x, w)))
            os.environ.pop("B200_NVLS")
            yb = gemm(x, w)
            row["gemm_then_nccl_us"] = timed(lambda: dist.all_reduce(gemm(x, w, out=yb)))
    out["rows"].append(row)
ok = torch.tensor([1 if out["ok"] else 0], device=dev)
dist.all_reduce(ok, op=dist.ReduceOp.MIN)
out["ok"] = bool(int(ok))
out["fused_calls"] = tp.fused_gemm_calls
if rank == 0:
    print(json.dumps(out), flush=True)
torch.cuda.synchronize()
dist.barrier()
sys.stdout.flush()
cleanup_distributed()
os._exit(0)
