"""One-shot tensor-parallel all-reduce (b200_allreduce_oneshot) against NCCL, eagerly and replayed from a CUDA graph, with
injected rank skew (device-side sleeps before eager calls, host-side sleeps before replays).
  torchrun --nproc-per-node 2 --master-addr 127.0.0.1 scripts/check_tp_oneshot.py"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from llm_in_practise_b200.parallel.dist import cleanup_distributed, setup_distributed
from llm_in_practise_b200.parallel.tp import TensorParallel

info = setup_distributed()
rank, world, dev = info.rank, info.world_size, info.device
tp = TensorParallel(1024, max_tokens=256, device=dev)
assert tp.fused and tp.oneshot
g = torch.Generator(device=dev).manual_seed(11 + rank)
worst = 0.0
for it, (rows, dt) in enumerate([(1, torch.bfloat16), (8, torch.bfloat16), (64, torch.float16), (3, torch.bfloat16), (8, torch.bfloat16)]):
    y = torch.randn(rows, 1024, device=dev, generator=g).to(dt)
    want = y.float().clone()
    dist.all_reduce(want)
    if it % world == rank:
        torch.cuda._sleep(20_000_000)                     # injected rank skew (~10 ms on the device): peers must wait on the flags
    got = tp.all_reduce(y)
    worst = max(worst, float((got.float() - want).abs().max() / want.abs().max()))
# graph: three dependent all-reduces per replay, new inputs every replay
x = torch.zeros(8, 1024, device=dev, dtype=torch.bfloat16)


def body():
    a = tp.all_reduce(x)
    b = tp.all_reduce(a * 0.5)
    return tp.all_reduce(b + x)


body()
torch.cuda.synchronize()
graph = torch.cuda.CUDAGraph()
with torch.cuda.graph(graph):
    out = body()
gworst = 0.0
for it in range(6):
    x.copy_(torch.randn(8, 1024, device=dev, generator=g).to(torch.bfloat16))
    if it % world == rank:
        time.sleep(0.02)                                  # host-side skew between the ranks' replays
    graph.replay()
    xs = x.float().clone()
    dist.all_reduce(xs)                                   # a = sum x
    want = (xs.bfloat16().float() * 0.5).bfloat16().float() * world          # b = sum of identical a/2 on every rank
    want = want.bfloat16().float() * world + xs          # c = sum over ranks of (b + x_r)
    gworst = max(gworst, float((out.float() - want).abs().max() / want.abs().max()))
same = out.float().clone()
dist.all_reduce(same, op=dist.ReduceOp.MAX)
identical = bool(torch.equal(same, out.float()))         # every rank holds bitwise the same result
if rank == 0:
    print(json.dumps({"world": world, "eager_rel_err": worst, "graph_rel_err": gworst, "ranks_identical": identical}))
cleanup_distributed()
