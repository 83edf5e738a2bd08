"""One forward + backward of the tcgen05 flash kernels at Qwen3-8B 1 x 4096 (for an ncu capture)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from llm_in_practise_b200.ops import flash as FL
B, S, H, Hkv, D = 1, 4096, 32, 8, 128
qkv = torch.randn(B * S, (H + 2 * Hkv) * D, device="cuda", dtype=torch.bfloat16).requires_grad_()
go = torch.randn(B * S, H * D, device="cuda", dtype=torch.bfloat16)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    o = FL.flash_attention(qkv, B, S, H, Hkv, D, True, D ** -0.5)
    torch.autograd.grad(o, qkv, go)
torch.cuda.synchronize()
print("ok")
