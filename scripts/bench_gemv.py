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


def timed(fn, reps=12):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


shapes = [("qkv", 6144, 4096), ("o", 4096, 4096), ("gate_up", 24576, 4096), ("down", 4096, 12288), ("lm_head", 151936, 4096)]
for name, N, K in shapes:
    if a.w4:
        from llm_in_practise_b200.quant.formats import quantize_rtn
        from llm_in_practise_b200.ops.w4 import w4_gemv
        if name == "lm_head":
            continue
        q = quantize_rtn(torch.randn(N, K, device=dev, dtype=torch.bfloat16), 128, True)
        nbytes = N * K // 2 + N * K // 128 * 4
    else:
        w = torch.randn(N, K, device=dev, dtype=torch.bfloat16)
        nbytes = N * K * 2
    for M in (1, 8, 16, 32):
        x = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
        if a.w4:
            fn = lambda: q.gemv(x) if hasattr(q, "gemv") else w4_gemv(x, q)
        else:
            fn = lambda: gemv16(x, w)
        if fn() is None:
            continue
        ms = timed(fn)
        print(json.dumps({"weights": "w4" if a.w4 else "bf16", "shape": name, "N": N, "K": K, "M": M, "us": round(ms * 1e3, 2),
                          "weight_TBps": round(nbytes / ms / 1e9, 3), "minb": os.environ.get("B200_GEMV_MINB", "default")}), flush=True)
