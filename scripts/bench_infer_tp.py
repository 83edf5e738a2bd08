"""W4A16 inference benchmark with tensor parallelism: Qwen3-8B architecture, random-init weights,
RTN W4A16 (group 128, symmetric: the GPTQ/compressed-tensors kernel layout), TP = WORLD_SIZE.

  torchrun --nproc-per-node T --master-addr 127.0.0.1 scripts/bench_infer_tp.py [--model qwen3-8b] [--batch 8] [--prompt 512] [--gen 64]

Reports prefill tokens/s and decode tokens/s / TPOT, CUDA-event timed, max over ranks, and first checks on a small model
that the TP logits equal the unsharded ones."""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from llm_in_practise_b200.inference.generate import Generator, KVCache
from llm_in_practise_b200.models.qwen3 import QWEN3_PRESETS, Qwen3Config, Qwen3ForCausalLM
from llm_in_practise_b200.parallel.dist import cleanup_distributed, setup_distributed
from llm_in_practise_b200.parallel.tp import TensorParallel, shard_qwen3
from llm_in_practise_b200.quant.formats import quantize_rtn

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="qwen3-8b")
ap.add_argument("--batch", type=int, default=8)
ap.add_argument("--prompt", type=int, default=512)
ap.add_argument("--gen", type=int, default=64)
ap.add_argument("--layers", type=int, default=0)
a = ap.parse_args()
info = setup_distributed()
rank, world, dev = info.rank, info.world_size, info.device


def quantize_all(m):
    for layer in m.layers:
        for lin in layer.fused_linears().values():
            lin.set_nf4(quantize_rtn(lin.weight.data, 128, True))
    torch.cuda.empty_cache()


# ---- correctness on a small model -------------------------------------------------------------
small = Qwen3Config(vocab_size=1024, hidden_size=1024, intermediate_size=2048, num_hidden_layers=2, num_attention_heads=8,
                    num_key_value_heads=8 if world > 4 else 4, head_dim=128, max_position_embeddings=512)
torch.manual_seed(3)
ref = Qwen3ForCausalLM(small, dtype=torch.bfloat16, device=dev, init_std=0.05)
quantize_all(ref)
ids = torch.randint(0, 1024, (2, 40), device=dev, generator=torch.Generator(device=dev).manual_seed(5))
_forced = []   # the unsharded run's greedy tokens, replayed in the sharded run: a near-tie in the argmax would otherwise
               # send the two runs down different continuations and compare unrelated logits


def decode_logits(model):
    """prefill 40 tokens, then 3 decode steps through the (fused-tail) decode path: logits of the last step"""
    g = Generator(model, max_len=64)
    cache = KVCache(small.num_hidden_layers, 2, model.config.num_key_value_heads, 64, small.head_dim, torch.bfloat16, dev)
    lg = g._prefill(ids, cache)
    record = not _forced
    for i in range(3):
        if record:
            _forced.append(lg.argmax(-1).reshape(-1).clone())
        lg = g._decode(_forced[i], cache)
    return lg.float()


with torch.no_grad():
    want = ref(ids)["logits"].float()
    want_dec = decode_logits(ref)
tp = shard_qwen3(ref, max_tokens=a.batch * a.prompt)
with torch.no_grad():
    got = ref(ids)["logits"].float()
    got_dec = decode_logits(ref)
err = float((got - want).abs().max() / want.abs().max())
assert err < 3e-2, f"TP logits differ: {err}"
err_dec = float((got_dec - want_dec).abs().max() / want_dec.abs().max())
assert err_dec < 5e-2, f"TP decode logits differ: {err_dec}"
del ref
torch.cuda.empty_cache()

# ---- the benchmark -----------------------------------------------------------------------------
cfg = QWEN3_PRESETS[a.model]
if a.layers:
    import dataclasses
    cfg = dataclasses.replace(cfg, num_hidden_layers=a.layers)
torch.manual_seed(1)
model = Qwen3ForCausalLM(cfg, dtype=torch.bfloat16, device=dev)
quantize_all(model)
tp = shard_qwen3(model, None, max_tokens=a.batch * a.prompt)
gen = Generator(model, max_len=a.prompt + a.gen + 8)
B, S = a.batch, a.prompt
ids = torch.randint(0, cfg.vocab_size, (B, S), device=dev)


def sync():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def timed(fn, reps):
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    sync()
    t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


with torch.no_grad():
    cache = KVCache(cfg.num_hidden_layers, B, model.config.num_key_value_heads, S + a.gen + 8, cfg.head_dim, torch.bfloat16, dev)
    for _ in range(2):
        gen._prefill(ids, cache)
    t_prefill = timed(lambda: gen._prefill(ids, cache), 3)
    tok = torch.randint(0, cfg.vocab_size, (B,), device=dev)

    pos_t = torch.zeros(B, dtype=torch.int32, device=dev)
    slots_t = torch.arange(B, dtype=torch.int32, device=dev)

    def decode_run():
        cache.len = S
        for _ in range(a.gen):
            pos_t.fill_(cache.len)
            gen.decode_graphed(tok, cache, pos_t, slots_t)     # CUDA graph at TP 1, eager under TP
            cache.len += 1
    decode_run()
    t_decode = timed(decode_run, 2) / a.gen
if rank == 0:
    print(json.dumps({"model": a.model + (f" ({a.layers} layers)" if a.layers else ""), "quant": "W4A16 RTN g128 sym (random-init)", "tp": world,
                      "batch": B, "prompt_len": S, "gen_len": a.gen, "tp_logits_rel_err_small_model": err, "tp_decode_logits_rel_err_small_model": err_dec,
                      "prefill_ms": round(t_prefill, 3), "prefill_tokens_per_s": round(B * S / t_prefill * 1e3, 1),
                      "tpot_ms": round(t_decode, 3), "decode_tokens_per_s": round(B / t_decode * 1e3, 1),
                      "allreduce": "nvlink-peer-fused" if (tp.world > 1 and tp.fused) else ("none" if tp.world == 1 else "nccl")}))
cleanup_distributed()
