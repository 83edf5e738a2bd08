"""ZeRO engine semantics on CPU/gloo with 2 processes: every stage must produce the same parameters
as single-process training on the concatenated batch (multi-process CPU tests for
wrapper semantics)."""
import copy

import pytest
import torch

from dist_utils import run_distributed


def _make(seed=0):
    from llm_in_practise_b200.models.qwen3 import QWEN3_PRESETS, Qwen3ForCausalLM
    torch.manual_seed(seed)
    return Qwen3ForCausalLM(QWEN3_PRESETS["qwen3-tiny"], dtype=torch.float32)


def _cfg(stage, ga=1, opt="AdamW"):
    return {"zero_optimization": {"stage": stage, "reduce_bucket_size": "auto"}, "gradient_accumulation_steps": ga,
            "gradient_clipping": 1.0, "train_micro_batch_size_per_gpu": 2, "train_batch_size": "auto",
            "optimizer": {"type": opt, "params": {"lr": 1e-2, "weight_decay": 0.0}}}


def _worker(rank, world, stage, tmp):
    from llm_in_practise_b200.parallel.dist import setup_distributed
    from llm_in_practise_b200.parallel.zero import initialize
    from llm_in_practise_b200.peft import LoraConfig
    setup_distributed()
    model = _make()
    eng = initialize(model=model, lora_config=LoraConfig(r=4, lora_alpha=8, target_modules=["q_proj", "v_proj"]),
                     config=_cfg(stage, ga=2))
    g = torch.Generator().manual_seed(7)
    data = torch.randint(0, 1024, (8, 2, 16), generator=g)  # 8 micro-batches of [2, 16]
    k = 0
    for step in range(2):
        for m in range(2):
            ids = data[(k * world + rank) % 8]
            k += 1
            eng.train_micro_batch(ids, ids)
        eng.step()
    torch.save(eng.arena.data.clone(), f"{tmp}/rank{rank}.pt")
    eng.save_checkpoint(f"{tmp}/ckpt", tag="t")


@pytest.mark.parametrize("stage", [0, 1, 2, 3])
def test_zero_stages_match_single_process(stage, tmp_path):
    run_distributed(_worker, 2, (stage, str(tmp_path)))
    a = torch.load(tmp_path / "rank0.pt")
    b = torch.load(tmp_path / "rank1.pt")
    assert torch.equal(a, b), "replicas diverged"
    # single-process oracle: same micro-batches, accumulation 4 (= 2 ranks x 2)
    from llm_in_practise_b200.parallel.zero import initialize
    from llm_in_practise_b200.peft import LoraConfig
    model = _make()
    cfg = _cfg(0, ga=4)
    cfg["train_batch_size"] = "auto"
    eng = initialize(model=model, lora_config=LoraConfig(r=4, lora_alpha=8, target_modules=["q_proj", "v_proj"]), config=cfg)
    g = torch.Generator().manual_seed(7)
    data = torch.randint(0, 1024, (8, 2, 16), generator=g)
    k = 0
    for step in range(2):
        for m in range(2):
            for r in range(2):
                ids = data[(k * 2 + r) % 8]
                eng.train_micro_batch(ids, ids)
            k += 1
        eng.step()
    n = min(a.numel(), eng.arena.numel)
    assert torch.allclose(a[:n], eng.arena.data[:n], atol=2e-5), (a[:n] - eng.arena.data[:n]).abs().max()
    assert (tmp_path / "ckpt" / "t" / "mp_rank_00_model_states.pt").exists()
    assert (tmp_path / "ckpt" / "t" / "zero_pp_rank_1_mp_rank_00_optim_states.pt").exists()
    assert (tmp_path / "ckpt" / "latest").read_text() == "t"


def test_train_batch_equals_sequential_micro_batches():
    """One layer-major pass over the G micro-batches of a step gives the gradients of G ``backward(loss / G)`` calls,
    also when the micro-batches have different numbers of non-ignored targets."""
    import copy
    from llm_in_practise_b200.models.qwen3 import QWEN3_PRESETS, Qwen3ForCausalLM
    from llm_in_practise_b200.parallel.zero import initialize
    from llm_in_practise_b200.peft import LoraConfig

    def engine():
        torch.manual_seed(0)
        m = Qwen3ForCausalLM(QWEN3_PRESETS["qwen3-tiny"], dtype=torch.float32)
        return initialize(model=m, lora_config=LoraConfig(r=4, lora_alpha=8, lora_dropout=0.0, target_modules=["q_proj", "v_proj"]),
                          config={"zero_optimization": {"stage": 3}, "gradient_accumulation_steps": 2, "gradient_clipping": 0.0,
                                  "train_micro_batch_size_per_gpu": 2, "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}}})
    g = torch.Generator().manual_seed(1)
    ids = [torch.randint(0, 1024, (2, 24), generator=g) for _ in range(2)]
    labels = [t.clone() for t in ids]
    labels[1][:, :10] = -100                      # the second micro-batch has fewer valid targets
    a, b = engine(), engine()
    # B needs a non-zero adapter B matrix or every gradient of A is zero
    for e in (a, b):
        with torch.no_grad():
            e.arena.data.copy_(torch.randn(e.arena.numel, generator=torch.Generator().manual_seed(5)) * 0.05)
        for mod in e.module.modules():
            if hasattr(mod, "invalidate_adapter_buffers"):
                mod.invalidate_adapter_buffers()
    la = [a.train_micro_batch(x, y) for x, y in zip(ids, labels)]
    lb = b.train_batch(ids, labels)
    assert b.supports_fused_accumulation() and b.micro_steps == 2 and b.is_gradient_accumulation_boundary()
    assert abs(float(lb) - float(sum(la) / 2)) < 1e-5
    assert torch.allclose(a.arena.grad, b.arena.grad, rtol=1e-4, atol=1e-6), float((a.arena.grad - b.arena.grad).abs().max())
    a.step(); b.step()
    assert torch.allclose(a.arena.data, b.arena.data, rtol=1e-4, atol=1e-6)
