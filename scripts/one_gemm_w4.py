"""One fused NF4 GEMM shape, a few launches: the target of `ncu --set full` captures."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from llm_in_practise_b200.ops import gemm as G
from llm_in_practise_b200.ops.w4 import NF4Weight
M, N, K = (int(x) for x in (sys.argv[1:4] if len(sys.argv) > 3 else (1024, 34816, 5120)))
bmn = len(sys.argv) > 4 and sys.argv[4] == "bmn"
rows, kw = (K, N) if bmn else (N, K)
q = NF4Weight.quantize(torch.randn(rows, kw, device="cuda", dtype=torch.bfloat16) * 0.02, True)
a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for _ in range(4):
    assert G.gemm_w4(a, q, b_mn=bmn, out=out, force=True) is not None
torch.cuda.synchronize()
print("ok", float(out.float().abs().mean()))
