"""One GEMM shape, a few launches: the target of `ncu --set full` captures."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from llm_in_practise_b200.ops import gemm as G
M, N, K = (int(x) for x in (sys.argv[1:4] if len(sys.argv) > 3 else (1024, 34816, 5120)))
bmn = len(sys.argv) > 4 and sys.argv[4] == "bmn"
a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
b = torch.randn(K, N, device="cuda", dtype=torch.bfloat16) if bmn else torch.randn(N, K, device="cuda", dtype=torch.bfloat16)
out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for _ in range(6):
    G.gemm(a, b, b_mn=bmn, out=out)
torch.cuda.synchronize()
print("ok", float(out.float().abs().mean()))
