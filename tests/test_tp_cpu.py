"""Tensor-parallel sharding on CPU/gloo (2 processes): dense and 4-bit models give the single-process logits and tokens."""
import pytest
import torch
from dist_utils import run_distributed


def _model(quant):
    from llm_in_practise_b200.models.qwen3 import Qwen3Config, Qwen3ForCausalLM
    from llm_in_practise_b200.quant.formats import quantize_rtn
    cfg = Qwen3Config(vocab_size=256, hidden_size=256, intermediate_size=512, num_hidden_layers=2, num_attention_heads=4,
                      num_key_value_heads=2, head_dim=64, max_position_embeddings=128)
    torch.manual_seed(0)
    m = Qwen3ForCausalLM(cfg, dtype=torch.float32, device="cpu", init_std=0.05)
    if quant == "nf4":
        m.quantize_nf4(True)
    elif quant == "w4":
        for layer in m.layers:
            for lin in layer.fused_linears().values():
                lin.set_nf4(quantize_rtn(lin.weight.data, 128, False))
    return m


def _worker(rank, world, quant, tmp):
    from llm_in_practise_b200.inference import Generator
    from llm_in_practise_b200.parallel.dist import setup_distributed
    from llm_in_practise_b200.parallel.tp import shard_qwen3
    setup_distributed()
    m = _model(quant)
    tp = shard_qwen3(m)
    assert tp.world == 2 and m.config.num_attention_heads == 2 and m.layers[0].qkv_proj.out_features == (4 + 2 + 2) * 64 // 2
    ids = torch.arange(12).view(1, 12) % 256
    with torch.no_grad():
        logits = m(ids)["logits"]
    out = Generator(m, max_len=64).generate(ids, max_new_tokens=5, do_sample=False)
    if rank == 0:
        torch.save({"logits": logits, "out": out, "calls": tp.calls}, f"{tmp}/tp.pt")


@pytest.mark.parametrize("quant", ["dense", "nf4", "w4"])
def test_tp2_matches_single(quant, tmp_path):
    from llm_in_practise_b200.inference import Generator
    run_distributed(_worker, 2, (quant, str(tmp_path)))
    got = torch.load(tmp_path / "tp.pt")
    m = _model(quant)
    ids = torch.arange(12).view(1, 12) % 256
    with torch.no_grad():
        ref = m(ids)["logits"]
    assert torch.allclose(got["logits"], ref, atol=2e-4, rtol=1e-3), (got["logits"] - ref).abs().max()
    assert torch.equal(got["out"], Generator(m, max_len=64).generate(ids, max_new_tokens=5, do_sample=False))
    assert got["calls"] > 0


def _serve_worker(rank, world, tmp, block_size=0):
    """Tensor-parallel serving: rank 0 schedules (continuous batching, chunked prefill, prefix cache), rank 1 follows."""
    from llm_in_practise_b200.inference import GenerationConfig
    from llm_in_practise_b200.parallel.dist import setup_distributed
    from llm_in_practise_b200.serve import EngineArgs
    from llm_in_practise_b200.serve.tp_engine import build_tp_engine
    setup_distributed()
    eng = build_tp_engine(_model("w4"), None, EngineArgs(max_num_seqs=3, max_model_len=96, max_prefill_tokens=24, warmup=False,
                                                          block_size=block_size))
    if rank != 0:
        eng.serve_forever()
        return
    torch.manual_seed(4)
    shared = torch.randint(0, 256, (32,)).tolist()
    prompts = [shared + [5, 6, 7], torch.randint(0, 256, (40,)).tolist(), shared + [9]]
    reqs = eng.generate(prompts, GenerationConfig(max_new_tokens=4, do_sample=False))
    torch.save({"out": [r.output_ids for r in reqs], "cached": [r.cached_tokens for r in reqs], "prompts": prompts}, f"{tmp}/serve.pt")
    eng.shutdown()


@pytest.mark.parametrize("block_size", [0, 16])
def test_tp2_serving_engine_matches_single_process(tmp_path, block_size):
    """block_size 16: the paged KV cache under tensor parallelism — rank 0's block-table updates are announced and applied
    to the follower's slice of the cache."""
    from llm_in_practise_b200.inference import Generator
    run_distributed(_serve_worker, 2, (str(tmp_path), block_size))
    got = torch.load(tmp_path / "serve.pt")
    g = Generator(_model("w4"), max_len=96)
    want = [g.generate(torch.tensor([p]), max_new_tokens=4, do_sample=False)[0, len(p):].tolist() for p in got["prompts"]]
    assert got["out"] == want
    assert max(got["cached"]) == 32            # the third prompt reused the first one's prefix


def _offload_worker(rank, world, tmp):
    """Host tier of the prefix cache under tensor parallelism: every rank keeps its own kv-head slice of an off-loaded row."""
    from llm_in_practise_b200.inference import GenerationConfig
    from llm_in_practise_b200.parallel.dist import setup_distributed
    from llm_in_practise_b200.serve import EngineArgs
    from llm_in_practise_b200.serve.tp_engine import build_tp_engine
    setup_distributed()
    eng = build_tp_engine(_model("dense"), None, EngineArgs(max_num_seqs=1, max_model_len=96, warmup=False, kv_offload_gb=0.01))
    if rank != 0:
        eng.serve_forever()
        assert len(eng.gen._host_kv) >= 1 and next(iter(eng.gen._host_kv.values()))[0][0].shape[0] == 1     # one local kv head
        return
    torch.manual_seed(4)
    shared = torch.randint(0, 256, (32,)).tolist()
    prompts = [shared + [5, 6, 7], torch.randint(0, 256, (40,)).tolist(), shared + [9]]
    outs, cached = [], []
    for p in prompts:                                  # one row: each prompt evicts the previous one's
        r = eng.generate([p], GenerationConfig(max_new_tokens=4, do_sample=False))[0]
        outs.append(r.output_ids)
        cached.append(r.cached_tokens)
    torch.save({"out": outs, "cached": cached, "prompts": prompts, "host_hits": eng.metrics.host_hits}, f"{tmp}/offload.pt")
    eng.shutdown()


def test_tp2_host_tier_restores_prefix_on_every_rank(tmp_path):
    from llm_in_practise_b200.inference import Generator
    run_distributed(_offload_worker, 2, (str(tmp_path),))
    got = torch.load(tmp_path / "offload.pt")
    g = Generator(_model("dense"), max_len=96)
    want = [g.generate(torch.tensor([p]), max_new_tokens=4, do_sample=False)[0, len(p):].tolist() for p in got["prompts"]]
    assert got["out"] == want
    assert got["cached"] == [0, 0, 32] and got["host_hits"] == 32
