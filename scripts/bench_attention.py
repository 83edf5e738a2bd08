"""Attention forward / backward timing: flash (csrc/flash_attn.cu) vs the tiled GEMM + softmax path, CUDA events, L2 flushed
between iterations. Shapes: the headline (Qwen3-14B: B 2, S 512, 40/8 heads, D 128), the 4k LoRA config (Qwen3-8B: 32/8) and
GPTLike (12 heads, D 64, S 256).   python scripts/bench_attention.py"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from llm_in_practise_b200.ops.attention import attention

dev = "cuda"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(fn, reps=10):
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


rows = []
for name, B, S, H, Hkv, D in (("qwen3-14b s512", 2, 512, 40, 8, 128), ("qwen3-8b s4096", 1, 4096, 32, 8, 128), ("gptlike s256", 16, 256, 12, 12, 64)):
    qkv = torch.randn(B * S, (H + 2 * Hkv) * D, device=dev, dtype=torch.bfloat16, requires_grad=True)
    flops_f = 4.0 * B * H * S * S * D / 2
    for impl in ("flash", "tiled"):
        out = attention(qkv, B, S, H, Hkv, D, causal=True, impl=impl)
        g = torch.randn_like(out)
        with torch.no_grad():
            tf = timed(lambda: attention(qkv, B, S, H, Hkv, D, causal=True, impl=impl))

        def fb():
            o = attention(qkv, B, S, H, Hkv, D, causal=True, impl=impl)
            o.backward(g)
            qkv.grad = None
        tfb = timed(fb)
        rows.append({"shape": name, "impl": impl, "fwd_ms": round(tf, 4), "fwd_bwd_ms": round(tfb, 4),
                     "fwd_tflops": round(flops_f / tf / 1e9, 1), "fwd_bwd_tflops": round(3.5 * flops_f / tfb / 1e9, 1)})
        print(json.dumps(rows[-1]), flush=True)
