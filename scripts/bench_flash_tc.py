"""Forward flash attention: tcgen05 kernel (csrc/flash_tc.cu) vs the mma.sync kernel (csrc/flash_attn.cu) vs torch SDPA,
CUDA events, L2 flushed between iterations, causal, head_dim 128, bf16."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from llm_in_practise_b200.ops import flash as FL

dev = "cuda"
BASE = "--no-baselines" not in sys.argv
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


for name, B, S, H, Hkv in (("qwen3-14b 4x512", 4, 512, 40, 8), ("qwen3-8b 1x4096", 1, 4096, 32, 8), ("qwen3-8b 2x4096", 2, 4096, 32, 8),
                           ("qwen3-8b 1x16384", 1, 16384, 32, 8)):
    D = 128
    qkv = torch.randn(B * S, (H + 2 * Hkv) * D, device=dev, dtype=torch.bfloat16)
    fl = 4.0 * B * H * S * S * D / 2
    row = {"shape": name}
    for tc in ((True, False) if BASE else (True,)):
        FL._TC = tc
        t = timed(lambda: FL._fwd(qkv, B, S, H, Hkv, D, True, D ** -0.5, True))
        row["tcgen05_ms" if tc else "mma_sync_ms"] = round(t, 4)
        row["tcgen05_tflops" if tc else "mma_sync_tflops"] = round(fl / t / 1e9, 1)
    # forward + backward through the autograd function (delta + dK/dV pass + dQ pass, or the mma.sync kernel)
    qg = qkv.clone().requires_grad_()
    go = torch.randn(B * S, H * D, device=dev, dtype=torch.bfloat16)
    def fb():
        o = FL.flash_attention(qg, B, S, H, Hkv, D, True, D ** -0.5)
        torch.autograd.grad(o, qg, go)
    for tc in ((True, False) if BASE else (True,)):
        FL._TC = tc
        FL._TC_BWD = tc
        t = timed(fb, 6)
        row["tcgen05_fwd_bwd_ms" if tc else "mma_sync_fwd_bwd_ms"] = round(t, 4)
        row["tcgen05_fwd_bwd_tflops" if tc else "mma_sync_fwd_bwd_tflops"] = round(3.5 * fl / t / 1e9, 1)
    FL._TC = FL._TC_BWD = True
    if not BASE:
        print(json.dumps(row), flush=True)
        continue
    t4 = qkv.view(B, S, H + 2 * Hkv, D)
    q, k, v = t4[:, :, :H].transpose(1, 2), t4[:, :, H:H + Hkv].transpose(1, 2), t4[:, :, H + Hkv:].transpose(1, 2)
    t = timed(lambda: torch.nn.functional.scaled_dot_product_attention(q, k, v, is_causal=True, enable_gqa=True))
    row["torch_sdpa_ms"], row["torch_sdpa_tflops"] = round(t, 4), round(fl / t / 1e9, 1)
    print(json.dumps(row), flush=True)
