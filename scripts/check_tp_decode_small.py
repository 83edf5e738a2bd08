"""TP decode logits of a tiny model whose row-parallel layers are left with very short contractions per rank (K = 128 / 256,
what TP 8 does to a 1024-wide model), fused tail on and off, against the unsharded model."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from llm_in_practise_b200.inference.generate import Generator, KVCache
from llm_in_practise_b200.models.qwen3 import Qwen3Config, Qwen3ForCausalLM
from llm_in_practise_b200.parallel.dist import cleanup_distributed, setup_distributed
from llm_in_practise_b200.parallel.tp import shard_qwen3
from llm_in_practise_b200.quant.formats import quantize_rtn
import llm_in_practise_b200.inference.generate as G

info = setup_distributed()
rank, world, dev = info.rank, info.world_size, info.device
heads = world            # one head per rank: o_proj K = 128 per rank
cfg = Qwen3Config(vocab_size=1024, hidden_size=128 * heads, intermediate_size=256 * world, num_hidden_layers=2,
                  num_attention_heads=heads, num_key_value_heads=heads, head_dim=128, max_position_embeddings=512)
out = {"world": world}
for quant in (False, True):
    torch.manual_seed(3)
    ref = Qwen3ForCausalLM(cfg, dtype=torch.bfloat16, device=dev, init_std=0.05)
    if quant:
        for layer in ref.layers:
            for lin in (layer.qkv_proj, layer.o_proj, layer.gate_up_proj, layer.down_proj):
                lin.set_nf4(quantize_rtn(lin.weight.data, 128, True))
    ids = torch.randint(0, 1024, (2, 40), device=dev, generator=torch.Generator(device=dev).manual_seed(5))

    forced = []          # the reference run's greedy tokens, replayed in the sharded runs (a near-tie in the argmax would
                         # otherwise send the two runs down different continuations and compare unrelated logits)

    def decode_logits(model):
        g = Generator(model, max_len=64)
        cache = KVCache(cfg.num_hidden_layers, 2, model.config.num_key_value_heads, 64, cfg.head_dim, torch.bfloat16, dev)
        lg = g._prefill(ids, cache)
        record = not forced
        for i in range(3):
            if record:
                forced.append(lg.argmax(-1).reshape(-1).clone())
            lg = g._decode(forced[i], cache)
        return lg.float()

    def prefill_logits(model):
        g = Generator(model, max_len=64)
        cache = KVCache(cfg.num_hidden_layers, 2, model.config.num_key_value_heads, 64, cfg.head_dim, torch.bfloat16, dev)
        return g._prefill(ids, cache).float()

    with torch.no_grad():
        want = decode_logits(ref)
        want_p = prefill_logits(ref)
    tp = shard_qwen3(ref, max_tokens=256)
    tag = "w4" if quant else "bf16"
    for gemm_ar in (True, False):
        tp.fuse_gemm = gemm_ar
        with torch.no_grad():
            got_p = prefill_logits(ref)
        out[f"{tag}_prefill_gemmar{int(gemm_ar)}_rel"] = float((got_p - want_p).abs().max() / want_p.abs().max())
        for fused in ("1", "0"):
            os.environ["B200_FUSED_DECODE"] = fused
            with torch.no_grad():
                got = decode_logits(ref)
            out[f"{tag}_decode_gemmar{int(gemm_ar)}_fused{fused}_rel"] = float((got - want).abs().max() / want.abs().max())
    del ref
if rank == 0:
    print(json.dumps(out), flush=True)
torch.cuda.synchronize(); dist.barrier(); sys.stdout.flush()
cleanup_distributed()
os._exit(0)
