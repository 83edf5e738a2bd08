"""Serving benchmark in the shape of the reference's published table (`vllm bench serve`, ShareGPT-like prompts, output
length 256, 512 prompts, max-concurrency 8..256; LLM_on_Kubernetes/Inference_Platfrom/README.md:1504-1511, quoted in
BASELINE.md section 1): mean / P99 TTFT, mean / P99 inter-token latency, requests/s and output tokens/s per concurrency.

The engine is driven in-process (closed loop: a new request is submitted whenever one finishes, so `concurrency` requests
are always in flight), Qwen3-8B architecture with random-init bf16 weights (the reference serves Qwen3-8B bf16), prompt
lengths drawn from a log-normal around 220 tokens (ShareGPT's typical turn length), every request generates exactly 256
tokens (no EOS with random weights — the same as vllm's `--ignore-eos` fixed output length).

  python scripts/bench_serving.py [--model qwen3-8b] [--concurrency 8 16 32 64] [--num-prompts 512] [--w4]
"""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from llm_in_practise_b200.inference import GenerationConfig
from llm_in_practise_b200.models.qwen3 import QWEN3_PRESETS, Qwen3ForCausalLM
from llm_in_practise_b200.serve import EngineArgs, LLMEngine

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="qwen3-8b")
ap.add_argument("--concurrency", type=int, nargs="+", default=[8, 16, 32, 64])
ap.add_argument("--num-prompts", type=int, default=512)
ap.add_argument("--output-len", type=int, default=256)
ap.add_argument("--max-model-len", type=int, default=2048)
ap.add_argument("--w4", action="store_true", help="W4A16 (RTN g128) weights instead of bf16")
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
if a.w4:
    from llm_in_practise_b200.quant.formats import quantize_rtn
    for layer in model.layers:
        for lin in layer.fused_linears().values():
            lin.set_nf4(quantize_rtn(lin.weight.data, 128, True))
    torch.cuda.empty_cache()

rng = np.random.default_rng(0)
lens = np.clip(rng.lognormal(np.log(220), 0.6, a.num_prompts).astype(int), 16, a.max_model_len - a.output_len - 8)
prompts = [rng.integers(0, cfg.vocab_size, int(n)).tolist() for n in lens]
gcfg = GenerationConfig(max_new_tokens=a.output_len, do_sample=False, eos_token_id=[])
rows = []
for C in a.concurrency:
    eng = LLMEngine(model, None, EngineArgs(max_num_seqs=C, max_model_len=a.max_model_len, max_prefill_per_step=8), model_name=a.model)
    eng.warmup()
    # warm-up: a few requests through prefill and decode (graph capture, lazily built buffers)
    for r in eng.generate(prompts[:min(C, 8)], GenerationConfig(max_new_tokens=8, do_sample=False, eos_token_id=[])):
        assert r.error is None, r.error
    torch.cuda.synchronize()
    pending = list(prompts)
    live, done = [], []
    t0 = time.perf_counter()
    while pending or live:
        while pending and len(live) < C:
            live.append(eng.submit(pending.pop(), gcfg))
        eng.step()
        still = []
        for r in live:
            (done if r.done.is_set() else still).append(r)
        live = still
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    ttft = np.array([r.ttft for r in done]) * 1e3
    itl = np.concatenate([np.diff(r.token_times) for r in done if len(r.token_times) > 1]) * 1e3
    out_tok = sum(len(r.output_ids) for r in done)
    row = {"concurrency": C, "num_prompts": len(done), "mean_ttft_ms": round(float(ttft.mean()), 1), "p99_ttft_ms": round(float(np.percentile(ttft, 99)), 1),
           "mean_itl_ms": round(float(itl.mean()), 2), "p99_itl_ms": round(float(np.percentile(itl, 99)), 2),
           "req_per_s": round(len(done) / wall, 2), "output_tok_per_s": round(out_tok / wall, 1), "wall_s": round(wall, 1)}
    rows.append(row)
    print(json.dumps(row), flush=True)
    del eng
    torch.cuda.empty_cache()
os.makedirs("gpurun_out", exist_ok=True)
json.dump({"model": a.model, "weights": "W4A16 RTN g128" if a.w4 else "bf16", "mean_prompt_len": float(lens.mean()), "output_len": a.output_len,
           "rows": rows}, open(f"gpurun_out/serving_{'w4' if a.w4 else 'bf16'}.json", "w"), indent=1)
