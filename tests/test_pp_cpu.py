"""Pipeline-parallel inference on CPU / gloo: stages (and stages x tensor-parallel ranks) give the single-process tokens,
through the plain generator and through the serving engine (continuous batching, chunked prefill, prefix cache, paged KV)."""
import pytest
import torch
from dist_utils import run_distributed


def _model(layers=3, quant="dense"):
    from llm_in_practise_b200.models.qwen3 import Qwen3Config, Qwen3ForCausalLM
    from llm_in_practise_b200.quant.formats import quantize_rtn
    cfg = Qwen3Config(vocab_size=256, hidden_size=256, intermediate_size=512, num_hidden_layers=layers, num_attention_heads=4,
                      num_key_value_heads=2, head_dim=64, max_position_embeddings=128)
    torch.manual_seed(0)
    m = Qwen3ForCausalLM(cfg, dtype=torch.float32, device="cpu", init_std=0.05)
    if quant == "w4":
        for layer in m.layers:
            for lin in layer.fused_linears().values():
                lin.set_nf4(quantize_rtn(lin.weight.data, 128, False))
    return m


def test_partition_layers():
    from llm_in_practise_b200.parallel.pp import partition_layers
    assert [list(r) for r in partition_layers(5, 2)] == [[0, 1, 2], [3, 4]]
    assert [len(r) for r in partition_layers(40, 8)] == [5] * 8
    assert [len(r) for r in partition_layers(7, 3)] == [3, 2, 2]
    with pytest.raises(ValueError):
        partition_layers(2, 3)


def _gen_worker(rank, world, tmp, stages, tp):
    from llm_in_practise_b200.parallel.dist import setup_distributed
    from llm_in_practise_b200.parallel.pp import PipelineGenerator, shard_qwen3_pipeline
    setup_distributed()
    m = _model(4 if tp > 1 else 3)
    n_all = len(m.layers)
    st = shard_qwen3_pipeline(m, stages, tp)
    assert st.stages == stages and st.stage == rank // tp and len(m.layers) == len(st.layers) == m.config.num_hidden_layers
    assert (m.lm_head is None) == (not st.last)
    if tp > 1:
        assert m.config.num_attention_heads == 4 // tp and m.tp.world == tp
    ids = torch.arange(12).view(1, 12) % 256
    g = PipelineGenerator(m, max_len=64)
    out = g.generate(ids, max_new_tokens=5, do_sample=False)
    sampled = g.generate(ids, max_new_tokens=5, do_sample=True, temperature=0.8, top_k=20, seed=3)
    torch.save({"out": out, "sampled": sampled, "layers": list(st.layers), "n_all": n_all, "sends": st.sends},
               f"{tmp}/pp{rank}.pt")


@pytest.mark.parametrize("stages,tp", [(2, 1), (3, 1), (2, 2)])
def test_pipeline_generate_matches_single_process(tmp_path, stages, tp):
    from llm_in_practise_b200.inference import Generator
    world = stages * tp
    run_distributed(_gen_worker, world, (str(tmp_path), stages, tp))
    got = [torch.load(tmp_path / f"pp{r}.pt") for r in range(world)]
    m = _model(4 if tp > 1 else 3)
    ids = torch.arange(12).view(1, 12) % 256
    g = Generator(m, max_len=64)
    want = g.generate(ids, max_new_tokens=5, do_sample=False)
    want_s = g.generate(ids, max_new_tokens=5, do_sample=True, temperature=0.8, top_k=20, seed=3)
    for r in got:                                  # every rank ends with the same tokens (the logits are broadcast)
        assert torch.equal(r["out"], want)
        assert torch.equal(r["sampled"], want_s)
    covered = sorted(i for r in got[::tp] for i in r["layers"])
    assert covered == list(range(got[0]["n_all"]))                  # every layer lives on exactly one stage
    assert got[0]["sends"] > 0 and got[-1]["sends"] == 0            # the last stage only receives


def _serve_worker(rank, world, tmp, block_size, stages, tp):
    from llm_in_practise_b200.inference import GenerationConfig
    from llm_in_practise_b200.parallel.dist import setup_distributed
    from llm_in_practise_b200.serve import EngineArgs
    from llm_in_practise_b200.serve.tp_engine import build_pp_engine
    setup_distributed()
    eng = build_pp_engine(_model(4, "w4"), None, EngineArgs(max_num_seqs=3, max_model_len=96, max_prefill_tokens=24, warmup=False,
                                                            block_size=block_size),
                          pipeline_parallel_size=stages, tensor_parallel_size=tp)
    assert len(eng.cache.k) == 4 // stages                          # the KV cache only covers the stage's layers
    if rank != 0:
        eng.serve_forever()
        return
    torch.manual_seed(4)
    shared = torch.randint(0, 256, (32,)).tolist()
    prompts = [shared + [5, 6, 7], torch.randint(0, 256, (40,)).tolist(), shared + [9]]
    reqs = eng.generate(prompts, GenerationConfig(max_new_tokens=4, do_sample=False))
    torch.save({"out": [r.output_ids for r in reqs], "cached": [r.cached_tokens for r in reqs], "prompts": prompts}, f"{tmp}/serve.pt")
    eng.shutdown()


@pytest.mark.parametrize("block_size,stages,tp", [(0, 2, 1), (16, 2, 1), (16, 2, 2)])
def test_pipeline_serving_engine_matches_single_process(tmp_path, block_size, stages, tp):
    from llm_in_practise_b200.inference import Generator
    run_distributed(_serve_worker, stages * tp, (str(tmp_path), block_size, stages, tp))
    got = torch.load(tmp_path / "serve.pt")
    g = Generator(_model(4, "w4"), max_len=96)
    want = [g.generate(torch.tensor([p]), max_new_tokens=4, do_sample=False)[0, len(p):].tolist() for p in got["prompts"]]
    assert got["out"] == want
    assert max(got["cached"]) == 32            # the third prompt reused the first one's prefix on every stage
