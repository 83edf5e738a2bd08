"""Fused NF4-dequant GEMM vs (dequant kernel + dense tcgen05 GEMM) vs cuBLAS on a pre-expanded bf16 weight.
Qwen3-14B projection shapes at M = 1024 tokens; CUDA-event timed, L2 flushed between runs."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from llm_in_practise_b200.ops import gemm as G
from llm_in_practise_b200.ops.w4 import NF4Weight

dev = "cuda"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

def timeit(fn, iters=10):
    for _ in range(3): fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]

M = int(os.environ.get("M", 1024))
shapes = [("qkv", 7168, 5120), ("o", 5120, 5120), ("gate_up", 34816, 5120), ("down", 5120, 17408)]
rows = []
only = os.environ.get("ONLY")
for name, out_f, in_f in shapes:
    if only and name != only:
        continue
    w = torch.randn(out_f, in_f, device=dev, dtype=torch.bfloat16) * 0.02
    q = NF4Weight.quantize(w, True)
    del w
    ws = torch.empty(out_f, in_f, device=dev, dtype=torch.bfloat16)
    wst = torch.empty(in_f, out_f, device=dev, dtype=torch.bfloat16)
    x = torch.randn(M, in_f, device=dev, dtype=torch.bfloat16)
    dy = torch.randn(M, out_f, device=dev, dtype=torch.bfloat16)
    y = torch.empty(M, out_f, device=dev, dtype=torch.bfloat16)
    dx = torch.empty(M, in_f, device=dev, dtype=torch.bfloat16)
    fl = 2.0 * M * out_f * in_f
    for tag, fused, unfused, cublas in (
        ("fwd", lambda: G.gemm_w4(x, q, out=y, force=True), lambda: G.gemm(x, q.dequantize(out=ws), out=y), lambda: torch.matmul(x, ws.t(), out=y)),
        ("dgrad", lambda: G.gemm_w4(dy, q, b_mn=True, out=dx, force=True), lambda: G.gemm(dy, q.dequantize(out=wst, transpose=True), out=dx),
         lambda: torch.matmul(dy, ws, out=dx))):
        tf, tu, tc = timeit(fused), timeit(unfused), timeit(cublas)
        rows.append({"shape": f"{name} {tag}", "M": M, "out": out_f, "in": in_f, "fused_ms": round(tf, 4), "dequant+gemm_ms": round(tu, 4),
                     "cublas_bf16_only_ms": round(tc, 4), "fused_tflops": round(fl / tf / 1e9, 1), "unfused_tflops": round(fl / tu / 1e9, 1),
                     "cublas_tflops": round(fl / tc / 1e9, 1)})
        print(rows[-1], flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(rows, open("gpurun_out/gemm_w4_bench.json", "w"), indent=1)
