"""Cold vs warm time-to-first-token with automatic prefix caching (the reference's APC table: 1200 -> 200 ms etc. on an
RTX 3090, LLM_on_Kubernetes/Inference_Platfrom/README.md:2376-2380): Qwen3-8B architecture, random bf16 weights, a
`--prefix`-token shared system prompt followed by a short user turn. Cold = first request (nothing cached), warm = the
following requests, whose prefix is served from the cache row of the first. Also a long cold prompt with and without
chunked prefill while 8 sequences are decoding (inter-token latency of the running sequences).

  python scripts/bench_ttft.py [--model qwen3-8b] [--prefix 2048] [--suffix 32]"""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from llm_in_practise_b200.inference import GenerationConfig
from llm_in_practise_b200.models.qwen3 import QWEN3_PRESETS, Qwen3ForCausalLM
from llm_in_practise_b200.serve import EngineArgs, LLMEngine

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="qwen3-8b")
ap.add_argument("--prefix", type=int, default=2048)
ap.add_argument("--suffix", type=int, default=32)
ap.add_argument("--layers", type=int, default=0)
a = ap.parse_args()
dev = torch.device("cuda", 0)
cfg = QWEN3_PRESETS[a.model]
if a.layers:
    import dataclasses
    cfg = dataclasses.replace(cfg, num_hidden_layers=a.layers)
torch.manual_seed(0)
model = Qwen3ForCausalLM(cfg, dtype=torch.bfloat16, device=dev)
for p in model.parameters():
    p.requires_grad_(False)
g = torch.Generator().manual_seed(1)
system = torch.randint(1000, 100000, (a.prefix,), generator=g).tolist()
users = [torch.randint(1000, 100000, (a.suffix,), generator=g).tolist() for _ in range(6)]


def ttft(eng, prompt):
    torch.cuda.synchronize()
    r = eng.submit(prompt, GenerationConfig(max_new_tokens=4, do_sample=False))
    while not r.done.is_set():
        eng.step()
    return r.ttft * 1e3, r.cached_tokens


out = {"model": a.model, "prefix_tokens": a.prefix, "suffix_tokens": a.suffix}
for apc in (True, False):
    eng = LLMEngine(model, None, EngineArgs(max_num_seqs=8, max_model_len=a.prefix + 512, enable_prefix_caching=apc,
                                            max_prefill_tokens=4096, warmup=True))
    eng.warmup()
    ttft(eng, users[5])                                    # engine warm, caches still cold for the system prompt
    cold, _ = ttft(eng, system + users[0])
    warm = [ttft(eng, system + u) for u in users[1:4]]
    out["apc_on" if apc else "apc_off"] = {"cold_ttft_ms": round(cold, 2), "next_ttft_ms": [round(w[0], 2) for w in warm],
                                           "cached_tokens": [w[1] for w in warm]}
    del eng
    torch.cuda.empty_cache()
print(json.dumps(out))
