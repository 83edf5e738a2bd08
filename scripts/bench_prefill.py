"""Serving prefill cost: one prompt per pass vs several prompts packed into one pass (Generator._prefill_packed), Qwen3-8B
architecture, random-init bf16 weights (or --w4), CUDA events.   python scripts/bench_prefill.py [--w4] [--len 259]"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from llm_in_practise_b200.inference.generate import Generator, KVCache
from llm_in_practise_b200.models.qwen3 import QWEN3_PRESETS, Qwen3ForCausalLM

ap = argparse.ArgumentParser()
ap.add_argument("--w4", action="store_true")
ap.add_argument("--len", type=int, default=259)
ap.add_argument("--model", default="qwen3-8b")
a = ap.parse_args()
dev = torch.device("cuda")
cfg = QWEN3_PRESETS[a.model]
model = Qwen3ForCausalLM(cfg, dtype=torch.bfloat16, device=dev)
if a.w4:
    from llm_in_practise_b200.quant.formats import quantize_rtn
    for layer in model.layers:
        for lin in layer.fused_linears().values():
            lin.set_nf4(quantize_rtn(lin.weight.data, 128, True))
    torch.cuda.empty_cache()
gen = Generator(model, max_len=2048)
cache = KVCache(cfg.num_hidden_layers, 16, cfg.num_key_value_heads, 2048, cfg.head_dim, torch.bfloat16, dev)
g = torch.Generator().manual_seed(0)


def timed(fn, reps=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


with torch.no_grad():
    for n in (1, 2, 4, 8):
        prompts = [torch.randint(0, cfg.vocab_size, (a.len + 7 * i,), generator=g).tolist() for i in range(n)]
        toks = sum(len(p) for p in prompts)
        slots = list(range(n))

        def one_by_one():
            for p, s in zip(prompts, slots):
                gen._prefill(torch.tensor([p], device=dev), cache, torch.tensor([s], device=dev))
        t1 = timed(one_by_one)
        t2 = timed(lambda: gen._prefill_packed(prompts, cache, slots))
        t3 = timed(lambda: gen._prefill_packed(prompts, cache, slots, scratch_slot=15))
        print(json.dumps({"weights": "w4a16" if a.w4 else "bf16", "prompts": n, "tokens": toks, "per_prompt_ms": round(t1, 3),
                          "packed_ms": round(t2, 3), "packed_graph_ms": round(t3, 3), "packed_graph_tok_per_s": round(toks / t3 * 1e3, 1)}), flush=True)
