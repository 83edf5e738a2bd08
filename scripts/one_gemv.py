"""A few 4-bit GEMV launches on the Qwen3-8B gate_up / qkv decode shapes (for `ncu -k regex:gemv_mma`)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from llm_in_practise_b200.quant.formats import quantize_rtn
dev = "cuda"
qs = [quantize_rtn(torch.randn(N, 4096, device=dev, dtype=torch.bfloat16), 128, True) for N in (6144, 24576)]
x = torch.randn(8, 4096, device=dev, dtype=torch.bfloat16)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for it in range(6):
    flush.zero_()
    for q in qs:
        y = q.gemv(x)
torch.cuda.synchronize()
print("ok", float(y.float().abs().mean()))
