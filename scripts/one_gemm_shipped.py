"""A few launches of the shipped tcgen05 GEMM on two Qwen3-14B shapes (gate_up forward: classic tiles; o_proj forward at
M = 1024: stream-K) for an ncu capture."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from llm_in_practise_b200.ops.gemm import gemm
torch.manual_seed(0)
shapes = [(1024, 34816, 5120), (1024, 5120, 5120)]
ops = []
for M, N, K in shapes:
    ops.append((torch.randn(M, K, device="cuda", dtype=torch.bfloat16), torch.randn(N, K, device="cuda", dtype=torch.bfloat16) / K ** 0.5))
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    for x, w in ops:
        gemm(x, w)
torch.cuda.synchronize()
print("ok")
