"""Small-M GEMV kernels (dense 16-bit, 4-bit integer codes) over the odd projection shapes tensor parallelism produces
(N = 384, 512, 640, ...; K = 128 ... 1024) against fp32 matmul."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from llm_in_practise_b200.ops.linear import FusedLinear
from llm_in_practise_b200.quant.formats import quantize_rtn
from llm_in_practise_b200.ops import w4 as W4
torch.manual_seed(0)
bad = []
rows = 0
for M in (1, 2, 8):
    for N in (128, 384, 512, 640, 1024, 1536, 3072):
        for K in (128, 256, 512, 1024):
            lin = FusedLinear(K, [("p", N)], dtype=torch.bfloat16, device="cuda", init_std=0.05)
            x = torch.randn(M, K, device="cuda").bfloat16()
            want = x.float() @ lin.weight.data.float().t()
            with torch.no_grad():
                got = lin(x).float()
            e = float((got - want).abs().max() / want.abs().max())
            rows += 1
            if not e < 2e-2:
                bad.append(("dense", M, N, K, e))
            q = quantize_rtn(lin.weight.data, 128, True)
            wref = W4._w4_dequant_ref(q.packed, q.scale, q.zs, N, K, 128, q._lut(), torch.float32)
            lin.set_nf4(q)
            with torch.no_grad():
                got = lin(x).float()
            want = x.float() @ wref.t()
            e = float((got - want).abs().max() / want.abs().max())
            rows += 1
            if not e < 2e-2:
                bad.append(("int4", M, N, K, e))
print(json.dumps({"cases": rows, "bad": bad}))
