"""Parity with the reference's model stack: a tiny random Qwen3 built by the installed ``transformers`` is saved
with its own ``save_pretrained``; our loader reads those safetensors and must reproduce its logits and loss; our
``save_pretrained`` must be loadable by transformers again (HF safetensors in, HF safetensors out)."""
import pytest
import torch

transformers = pytest.importorskip("transformers")


def _hf_tiny(tmp, bias=False):
    from transformers import Qwen3Config, Qwen3ForCausalLM
    cfg = Qwen3Config(vocab_size=512, hidden_size=128, intermediate_size=256, num_hidden_layers=2, num_attention_heads=4,
                      num_key_value_heads=2, head_dim=32, max_position_embeddings=128, rope_theta=10000.0,
                      tie_word_embeddings=False, attention_bias=bias)
    torch.manual_seed(0)
    m = Qwen3ForCausalLM(cfg).eval()
    with torch.no_grad():                       # make the norms non-trivial
        for n, p in m.named_parameters():
            if "norm" in n:
                p.add_(0.1 * torch.randn_like(p))
    m.save_pretrained(tmp, safe_serialization=True)
    return m


def _compare(device, dtype, tmp, tol):
    from llm_in_practise_b200.io.hf import from_pretrained, save_pretrained
    hf = _hf_tiny(tmp)
    ours = from_pretrained(tmp, dtype=dtype, device=torch.device(device)).eval()
    ids = torch.randint(0, 512, (2, 24))
    with torch.no_grad():
        ref = hf(ids, labels=ids)
        out = ours(ids.to(device))
        loss = ours(ids.to(device), labels=ids.to(device))["loss"]
    err = (out["logits"].float().cpu() - ref.logits).abs().max() / ref.logits.abs().max()
    assert err < tol, err
    assert abs(float(loss) - float(ref.loss)) < tol * 10
    return ours, hf, ids


def test_hf_qwen3_logits_and_roundtrip_cpu(tmp_path):
    ours, hf, ids = _compare("cpu", torch.float32, str(tmp_path / "hf"), 1e-4)
    from llm_in_practise_b200.io.hf import save_pretrained
    from transformers import AutoModelForCausalLM
    save_pretrained(ours, str(tmp_path / "ours"))
    back = AutoModelForCausalLM.from_pretrained(str(tmp_path / "ours"), torch_dtype=torch.float32).eval()
    with torch.no_grad():
        assert torch.allclose(back(ids).logits, hf(ids).logits, atol=1e-4)


@pytest.mark.gpu
def test_hf_qwen3_logits_gpu_kernels(tmp_path):
    _compare("cuda", torch.bfloat16, str(tmp_path / "hf"), 4e-2)


@pytest.mark.parametrize("scaling", [
    {"rope_type": "yarn", "factor": 4.0, "original_max_position_embeddings": 32},
    {"rope_type": "yarn", "factor": 4.0, "original_max_position_embeddings": 32, "attention_factor": 0.8782, "beta_fast": 16, "beta_slow": 2},
    {"rope_type": "linear", "factor": 2.0},
    {"rope_type": "llama3", "factor": 8.0, "original_max_position_embeddings": 32, "low_freq_factor": 1.0, "high_freq_factor": 4.0},
])
def test_rope_scaling_matches_transformers(tmp_path, scaling):
    """A checkpoint whose config carries ``rope_scaling`` (DeepSeek-R1-0528-Qwen3-8B ships YaRN): same logits as transformers
    with the scaling on, and ``from_pretrained(..., rope_scaling=None)`` — what the reference's fine-tune script passes —
    equals transformers with the scaling removed."""
    from transformers import Qwen3Config, Qwen3ForCausalLM
    from llm_in_practise_b200.io.hf import from_pretrained
    kw = dict(vocab_size=512, hidden_size=128, intermediate_size=256, num_hidden_layers=2, num_attention_heads=4,
              num_key_value_heads=2, head_dim=32, max_position_embeddings=128, rope_theta=10000.0, tie_word_embeddings=False)
    cfg = Qwen3Config(**kw, rope_scaling=dict(scaling))
    torch.manual_seed(0)
    hf = Qwen3ForCausalLM(cfg).eval()
    path = str(tmp_path / "hf")
    hf.save_pretrained(path, safe_serialization=True)
    ids = torch.randint(0, 512, (2, 96))                               # beyond the original 32 positions
    ours = from_pretrained(path, dtype=torch.float32, device=torch.device("cpu")).eval()
    assert ours.config.rope_scaling is not None
    with torch.no_grad():
        want, got = hf(ids).logits, ours(ids)["logits"]
    assert (got - want).abs().max() / want.abs().max() < 1e-4
    plain_hf = Qwen3ForCausalLM(Qwen3Config(**kw)).eval()
    plain_hf.load_state_dict(hf.state_dict())
    off = from_pretrained(path, dtype=torch.float32, device=torch.device("cpu"), rope_scaling=None).eval()
    assert off.config.rope_scaling is None
    with torch.no_grad():
        want_off, got_off = plain_hf(ids).logits, off(ids)["logits"]
    assert (got_off - want_off).abs().max() / want_off.abs().max() < 1e-4
    assert (want_off - want).abs().max() / want.abs().max() > 1e-3       # the scaling is not a no-op at these positions
