"""Small-M GEMV kernels (csrc/gemv.cu) on the Qwen3-8B decode shapes: time and effective weight bandwidth, CUDA events,
256 MiB L2 flush between iterations.   [B200_GEMV_MINB=2|3|4] python scripts/bench_gemv.py [--w4]"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from llm_in_practise_b200.ops.gemm import gemv16

ap = argparse.ArgumentParser()
ap.add_argument("--w4", action="store_true")
a = ap.parse_args()
dev = "cuda"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(fns, reps=5):
    """``fns``: the same product on enough distinct weight copies to exceed the 126 MB L2 (every call streams its weights
    from HBM). One CUDA graph holds the whole sequence, so the time per call is kernel time, not the launch path."""
    for f in fns:
        f()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for f in fns:
            f()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) / len(fns))
    ts.sort()
    return ts[len(ts) // 2]


shapes = [("qkv", 6144, 4096), ("o", 4096, 4096), ("gate_up", 24576, 4096), ("down", 4096, 12288), ("lm_head", 151936, 4096)]
for name, N, K in shapes:
    if a.w4 and name == "lm_head":
        continue
    nbytes = (N * K // 2 + N * K // 128 * 8) if a.w4 else N * K * 2
    copies = max(2, min(64, (320 << 20) // nbytes + 1))
    ws = []
    for _ in range(copies):
        if a.w4:
            from llm_in_practise_b200.quant.formats import quantize_rtn
            ws.append(quantize_rtn(torch.randn(N, K, device=dev, dtype=torch.bfloat16), 128, True))
        else:
            ws.append(torch.randn(N, K, device=dev, dtype=torch.bfloat16))
    for M in (1, 8, 16, 32):
        x = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
        fns = [(lambda q=q: q.gemv(x)) for q in ws] if a.w4 else [(lambda w=w: gemv16(x, w)) for w in ws]
        if fns[0]() is None:
            continue
        ms = timed(fns)
        print(json.dumps({"weights": "w4" if a.w4 else "bf16", "shape": name, "N": N, "K": K, "M": M, "us": round(ms * 1e3, 2),
                          "weight_TBps": round(nbytes / ms / 1e9, 3), "copies": copies,
                          "post": os.environ.get("B200_GEMV_POST", "1")}), flush=True)
    del ws
