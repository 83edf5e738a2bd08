"""Reference arm of BASELINE config 5 and of the serving table: vLLM (the engine every serving recipe of the reference
uses — ``vllm serve … --tensor-parallel-size T`` Quantization/README.md:175, Deployment/Ray/serve_run_examples/deepseek.py:26-34)
on the same box, same model shape, random weights (``load_format="dummy"``), no tokenizer (token-id prompts).

Two measurements, both mirrored by a script of this repository:

  --mode generate   prompt 512, batch B: prefill time (max_tokens=1) and TPOT ((t(gen+1) - t(1)) / gen)      <-> scripts/bench_infer_tp.py
  --mode serve      512 prompts x 256 output tokens pushed through the engine with max_num_seqs = concurrency  <-> scripts/bench_serving.py
                    (offline ``LLM.generate``: vLLM's continuous batching with the concurrency cap; output tok/s)

bf16 weights by default; ``--quant w4a16`` writes a compressed-tensors W4A16 ``quantization_config`` (group 128, symmetric,
pack-quantized) into the dummy checkpoint's config.json so vLLM allocates and runs its own int4 (Marlin-class) kernels.
Wall-clock timed around ``generate`` (vLLM runs its engine in another process, CUDA events cannot bracket it); every
number is the best of ``--reps`` runs after a warm-up. Nothing from ``llm_in_practise_b200`` is imported.

    python baseline/ref_vllm.py --mode generate --tp 1 --batch 8
"""
import argparse
import json
import os
import tempfile
import time

SHAPES = {
    "qwen3-8b": dict(hidden_size=4096, intermediate_size=12288, num_hidden_layers=36, num_attention_heads=32,
                     num_key_value_heads=8),
    "qwen3-14b": dict(hidden_size=5120, intermediate_size=17408, num_hidden_layers=40, num_attention_heads=40,
                      num_key_value_heads=8),
}


def write_config(model, quant, layers):
    d = tempfile.mkdtemp(prefix="ref_vllm_")
    cfg = {"architectures": ["Qwen3ForCausalLM"], "model_type": "qwen3", "vocab_size": 151936, "head_dim": 128,
           "max_position_embeddings": 40960, "rms_norm_eps": 1e-6, "rope_theta": 1000000.0, "tie_word_embeddings": False,
           "hidden_act": "silu", "attention_bias": False, "torch_dtype": "bfloat16", "bos_token_id": 151643,
           "eos_token_id": 151645, "use_cache": True}
    cfg.update(SHAPES[model])
    if layers:
        cfg["num_hidden_layers"] = layers
    if quant == "w4a16":
        cfg["quantization_config"] = {
            "quant_method": "compressed-tensors", "format": "pack-quantized", "quantization_status": "compressed",
            "ignore": ["lm_head"], "global_compression_ratio": None,
            "config_groups": {"group_0": {"targets": ["Linear"], "input_activations": None, "output_activations": None,
                                          "weights": {"num_bits": 4, "type": "int", "symmetric": True, "strategy": "group",
                                                      "group_size": 128, "dynamic": False, "actorder": None,
                                                      "observer": "minmax", "observer_kwargs": {}, "block_structure": None}}}}
    with open(os.path.join(d, "config.json"), "w") as f:
        json.dump(cfg, f)
    with open(os.path.join(d, "generation_config.json"), "w") as f:
        json.dump({"do_sample": False}, f)
    return d


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="generate", choices=["generate", "serve"])
    ap.add_argument("--model", default="qwen3-8b")
    ap.add_argument("--quant", default="w4a16", choices=["bf16", "w4a16"])
    ap.add_argument("--tp", type=int, default=1)
    ap.add_argument("--batch", type=int, nargs="+", default=[1, 8])
    ap.add_argument("--prompt", type=int, default=512)
    ap.add_argument("--gen", type=int, default=64)
    ap.add_argument("--concurrency", type=int, nargs="+", default=[8, 32, 128])
    ap.add_argument("--num-prompts", type=int, default=512)
    ap.add_argument("--output-len", type=int, default=256)
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--layers", type=int, default=0)
    a = ap.parse_args()
    try:
        import numpy as np
        from vllm import LLM, SamplingParams
    except Exception as e:
        print(json.dumps({"impl": "reference", "unavailable": f"vllm import failed: {type(e).__name__}: {e}"}))
        return 0
    path = write_config(a.model, a.quant, a.layers)
    rng = np.random.default_rng(0)

    def engine(max_num_seqs):
        return LLM(model=path, load_format="dummy", skip_tokenizer_init=True, dtype="bfloat16", tensor_parallel_size=a.tp,
                   max_model_len=2048, max_num_seqs=max_num_seqs, gpu_memory_utilization=0.85, enable_prefix_caching=False,
                   seed=0)

    def prompts(n, lens):
        return [{"prompt_token_ids": rng.integers(1000, 150000, size=int(L)).tolist()} for L in lens[:n]]

    def run(llm, ps, max_tokens):
        sp = SamplingParams(max_tokens=max_tokens, temperature=0.0, ignore_eos=True, detokenize=False)
        t0 = time.perf_counter()
        llm.generate(ps, sp, use_tqdm=False)
        return time.perf_counter() - t0

    rows = []
    try:
        if a.mode == "generate":
            llm = engine(max(a.batch))
            for B in a.batch:
                ps = prompts(B, [a.prompt] * B)
                run(llm, ps, 4)
                t1 = min(run(llm, prompts(B, [a.prompt] * B), 1) for _ in range(a.reps))
                tg = min(run(llm, prompts(B, [a.prompt] * B), a.gen + 1) for _ in range(a.reps))
                tpot = (tg - t1) / a.gen
                rows.append({"batch": B, "prompt_len": a.prompt, "gen_len": a.gen, "prefill_ms": round(t1 * 1e3, 3),
                             "prefill_tokens_per_s": round(B * a.prompt / t1, 1), "tpot_ms": round(tpot * 1e3, 3),
                             "decode_tokens_per_s": round(B / tpot, 1)})
        else:
            lens = np.clip(rng.lognormal(mean=5.4, sigma=0.6, size=a.num_prompts), 16, 1536).astype(int)
            for c in a.concurrency:
                llm = engine(c)
                run(llm, prompts(min(c, 32), lens), 8)
                t = run(llm, prompts(a.num_prompts, lens), a.output_len)
                rows.append({"concurrency": c, "num_prompts": a.num_prompts, "output_len": a.output_len,
                             "output_tokens_per_s": round(a.num_prompts * a.output_len / t, 1),
                             "requests_per_s": round(a.num_prompts / t, 2), "wall_s": round(t, 2)})
                del llm
                import gc
                gc.collect()
    except Exception as e:
        print(json.dumps({"impl": "reference", "engine": "vllm", "unavailable": f"{type(e).__name__}: {str(e)[:300]}", "rows": rows}))
        return 0
    import vllm
    print(json.dumps({"impl": "reference", "engine": f"vllm {vllm.__version__}", "mode": a.mode, "model": a.model + (f" ({a.layers} layers)" if a.layers else ""),
                      "quant": a.quant, "tp": a.tp, "weights": "dummy (random)", "timing": "wall clock around LLM.generate", "rows": rows}))
    return 0


if __name__ == "__main__":
    main()
