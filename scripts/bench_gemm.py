"""GEMM micro-benchmark: Qwen3-14B training shapes, CUDA-event timed, L2 flushed between runs.
Prints achieved TFLOP/s of the tcgen05 kernel and of torch.matmul (cuBLAS) for the same product."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from llm_in_practise_b200.ops import gemm as G

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

shapes = [("qkv fwd", 1024, 7168, 5120, False), ("o fwd", 1024, 5120, 5120, False), ("gate_up fwd", 1024, 34816, 5120, False),
          ("down fwd", 1024, 5120, 17408, False), ("lm_head fwd", 1024, 151936, 5120, False),
          ("gate_up dgrad", 1024, 5120, 34816, True), ("down dgrad", 1024, 17408, 5120, True),
          ("lm_head dgrad", 1024, 5120, 151936, True), ("qkv dgrad", 1024, 5120, 7168, True),
          ("8k x 8k x 8k", 8192, 8192, 8192, False), ("4k tok gate_up 8B", 8192, 24576, 4096, False)]
# the layer-major schedule runs both micro-batches of a step through a layer together: M = 2048; with an NF4 base the
# dgrad reads W^T K-major (expanded transposed), so those rows are b_mn = False
shapes += [("qkv fwd M2048", 2048, 7168, 5120, False), ("o fwd M2048", 2048, 5120, 5120, False),
           ("gate_up fwd M2048", 2048, 34816, 5120, False), ("down fwd M2048", 2048, 5120, 17408, False),
           ("gate_up dgrad(K-major) M2048", 2048, 5120, 34816, False), ("down dgrad(K-major) M2048", 2048, 17408, 5120, False),
           ("lm_head fwd M2048", 2048, 151936, 5120, False), ("lm_head dgrad M2048", 2048, 5120, 151936, True)]
if len(sys.argv) > 1:
    shapes = [s_ for s_ in shapes if any(k in s_[0] for k in sys.argv[1:])]
rows = []
for name, M, N, K, bmn in shapes:
    a = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
    b = torch.randn(K, N, device=dev, dtype=torch.bfloat16) if bmn else torch.randn(N, K, device=dev, dtype=torch.bfloat16)
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    t_mine = timeit(lambda: G.gemm(a, b, b_mn=bmn, out=out))
    t_ref = timeit(lambda: torch.matmul(a, b if bmn else b.t(), out=out))
    fl = 2.0 * M * N * K
    rows.append({"shape": name, "M": M, "N": N, "K": K, "b_mn": bmn, "b200_ms": round(t_mine, 4), "cublas_ms": round(t_ref, 4),
                 "b200_tflops": round(fl / t_mine / 1e9, 1), "cublas_tflops": round(fl / t_ref / 1e9, 1)})
    print(rows[-1], flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(rows, open(f"gpurun_out/gemm_bench_ctas{os.environ.get('B200_GEMM_CTAS','auto')}.json", "w"), indent=1)
