"""Qwen3 + LoRA / QLoRA end to end: CPU logic test, and native-kernel vs PyTorch-path parity on GPU."""
import os

import pytest
import torch

from llm_in_practise_b200.models.qwen3 import QWEN3_PRESETS, Qwen3ForCausalLM
from llm_in_practise_b200.optim import FusedAdamW, PagedAdamW8bit, clip_grad_norm_
from llm_in_practise_b200.peft import LoraConfig, PeftModel, get_peft_model


def _tiny(dtype, device):
    torch.manual_seed(0)
    return Qwen3ForCausalLM(QWEN3_PRESETS["qwen3-tiny"], dtype=dtype, device=device)


def test_lora_training_reduces_loss_cpu(tmp_path):
    m = _tiny(torch.float32, "cpu")
    pm = get_peft_model(m, LoraConfig(r=4, lora_alpha=8, target_modules=["q_proj", "v_proj"]))
    trainable, total = pm.get_nb_trainable_parameters()
    cfg = m.config
    # r * (in + out) per targeted projection, per layer
    expect = cfg.num_hidden_layers * 4 * (2 * cfg.hidden_size + cfg.num_attention_heads * cfg.head_dim
                                          + cfg.num_key_value_heads * cfg.head_dim)
    assert trainable == expect
    ids = torch.randint(0, cfg.vocab_size, (2, 32))
    opt = FusedAdamW(pm.arena, lr=1e-2, weight_decay=0.0)
    losses = []
    for _ in range(8):
        opt.zero_grad()
        loss = pm(ids, labels=ids)["loss"]
        loss.backward()
        clip_grad_norm_(pm.arena, 1.0, opt)
        opt.step()
        losses.append(loss.item())
    assert losses[-1] < losses[0]
    # adapter round trip in peft layout
    pm.save_pretrained(str(tmp_path / "adapter"))
    m2 = _tiny(torch.float32, "cpu")
    pm2 = PeftModel.from_pretrained(m2, str(tmp_path / "adapter"))
    with torch.no_grad():
        a = pm(ids, labels=ids)["loss"]
        b = pm2(ids, labels=ids)["loss"]
    assert torch.allclose(a, b, atol=1e-5)
    # merge_and_unload keeps the function
    merged = pm2.merge_and_unload()
    with torch.no_grad():
        c = merged(ids, labels=ids)["loss"]
    assert torch.allclose(b, c, atol=1e-2)  # merged weights are stored in bf16


def test_lora_param_count_matches_reference_numbers():
    # r=8 on q_proj,v_proj -> 3 833 856 (Qwen3-8B) and 5 242 880 (Qwen3-14B)
    for name, want in [("qwen3-8b", 3833856), ("qwen3-14b", 5242880)]:
        c = QWEN3_PRESETS[name]
        per_layer = 8 * (2 * c.hidden_size + c.num_attention_heads * c.head_dim + c.num_key_value_heads * c.head_dim)
        assert per_layer * c.num_hidden_layers == want


@pytest.mark.gpu
@pytest.mark.parametrize("nf4", [False, True])
def test_native_matches_torch_path(nf4):
    dev = "cuda"
    m = _tiny(torch.bfloat16, dev)
    if nf4:
        m.quantize_nf4()
    pm = get_peft_model(m, LoraConfig(r=8, lora_alpha=16, lora_dropout=0.0, target_modules=["q_proj", "v_proj"]))
    with torch.no_grad():
        pm.arena.data.normal_(std=0.05)
    for _, lin in m.named_modules():
        if hasattr(lin, "invalidate_adapter_buffers"):
            lin.invalidate_adapter_buffers()
    ids = torch.randint(0, m.config.vocab_size, (2, 64), device=dev)
    pm.arena.zero_grad()
    loss = pm(ids, labels=ids)["loss"]
    loss.backward()
    g_native = pm.arena.grad.clone()
    os.environ["B200_FORCE_TORCH"] = "1"
    try:
        pm.arena.zero_grad()
        loss_ref = pm(ids, labels=ids)["loss"]
        loss_ref.backward()
        g_ref = pm.arena.grad.clone()
    finally:
        del os.environ["B200_FORCE_TORCH"]
    assert abs(loss.item() - loss_ref.item()) < 3e-2 * abs(loss_ref.item())
    cos = torch.nn.functional.cosine_similarity(g_native, g_ref, dim=0).item()
    assert cos > 0.98, cos


@pytest.mark.gpu
def test_qlora_step_8bit_optimizer_gpu():
    dev = "cuda"
    m = _tiny(torch.bfloat16, dev).quantize_nf4()
    pm = get_peft_model(m, LoraConfig(r=8, lora_alpha=16, lora_dropout=0.1, target_modules=["q_proj", "v_proj"]))
    opt = PagedAdamW8bit(pm.arena, lr=5e-3, weight_decay=0.0)
    ids = torch.randint(0, m.config.vocab_size, (2, 64), device=dev)
    losses = []
    for _ in range(10):
        opt.zero_grad()
        loss = pm(ids, labels=ids)["loss"]
        loss.backward()
        clip_grad_norm_(pm.arena, 1.0, opt)
        opt.step()
        losses.append(loss.item())
    assert losses[-1] < losses[0]


def test_transformers_call_surface_on_the_model(tmp_path):
    """What the reference's scripts call on a model object: ``.device``, ``.generate(**inputs, …)``, ``.save_pretrained``,
    ``.enable_input_require_grads()`` — on the bare model and through a PEFT wrapper."""
    import torch
    from llm_in_practise_b200.inference import Generator
    from llm_in_practise_b200.io.hf import from_pretrained
    from llm_in_practise_b200.models.qwen3 import QWEN3_PRESETS, Qwen3ForCausalLM
    from llm_in_practise_b200.peft import LoraConfig, get_peft_model
    torch.manual_seed(0)
    m = Qwen3ForCausalLM(QWEN3_PRESETS["qwen3-tiny"], dtype=torch.float32, device="cpu").eval()
    assert m.device == torch.device("cpu")
    inputs = {"input_ids": torch.arange(1, 9).view(1, 8), "attention_mask": torch.ones(1, 8, dtype=torch.long)}
    out = m.generate(**inputs, max_new_tokens=5, do_sample=False)
    assert out.shape == (1, 13) and torch.equal(out, Generator(m, max_len=64).generate(inputs["input_ids"], max_new_tokens=5, do_sample=False))
    assert m.generate(inputs["input_ids"], max_length=11, do_sample=False).shape == (1, 11)
    sampled = m.generate(**inputs, max_new_tokens=4, do_sample=True, temperature=0.7, top_p=0.9, seed=1)
    assert sampled.shape == (1, 12)
    with pytest.raises(NotImplementedError):
        m.generate(input_ids=inputs["input_ids"], attention_mask=torch.tensor([[0, 1, 1, 1, 1, 1, 1, 1]]), max_new_tokens=2)
    m.save_pretrained(str(tmp_path / "m"), max_shard_size="5GB")
    back = from_pretrained(str(tmp_path / "m"), dtype=torch.float32, device=torch.device("cpu")).eval()
    with torch.no_grad():
        assert torch.allclose(back(inputs["input_ids"])["logits"], m(inputs["input_ids"])["logits"], atol=1e-5)
    pm = get_peft_model(m, LoraConfig(r=2, lora_alpha=4, target_modules=["q_proj", "v_proj"]))
    pm.enable_input_require_grads()
    assert pm.device == torch.device("cpu") and pm.generate(**inputs, max_new_tokens=3, do_sample=False).shape == (1, 11)


def test_streaming_thread_flow_of_the_webui_script():
    """Scripts/inference/06-…streaming-infr.py:55-85: ``Thread(target=model.generate, kwargs=dict(**inputs, streamer=streamer, …))``
    and ``for new_text in streamer``. The consumer also terminates when generation fails inside the thread."""
    import threading
    import torch
    from llm_in_practise_b200.compat import TextIteratorStreamer
    from llm_in_practise_b200.data.tokenizers import ByteTokenizer
    from llm_in_practise_b200.models.qwen3 import QWEN3_PRESETS, Qwen3ForCausalLM
    tok = ByteTokenizer()
    torch.manual_seed(0)
    m = Qwen3ForCausalLM(QWEN3_PRESETS["qwen3-tiny"], dtype=torch.float32, device="cpu").eval()
    ids = torch.tensor([tok.encode("hello")])
    streamer = TextIteratorStreamer(tok, skip_prompt=True, skip_special_tokens=True, timeout=30)
    kw = dict(input_ids=ids, attention_mask=torch.ones_like(ids), streamer=streamer, max_new_tokens=8, do_sample=False)
    t = threading.Thread(target=m.generate, kwargs=kw)
    t.start()
    pieces = list(streamer)
    t.join(30)
    want = m.generate(ids, max_new_tokens=8, do_sample=False)[0, ids.shape[1]:].tolist()
    assert not t.is_alive() and "".join(pieces) == tok.decode(want)

    class Broken(ByteTokenizer):
        def decode(self, ids, **kw):
            raise RuntimeError("injected tokenizer failure")
    s2 = TextIteratorStreamer(Broken(), skip_prompt=True, timeout=30)
    t2 = threading.Thread(target=lambda: pytest.raises(RuntimeError, m.generate, ids, streamer=s2, max_new_tokens=4, do_sample=False))
    t2.start()
    assert list(s2) == []                                   # the iterator ends instead of waiting for ever
    t2.join(30)
    assert not t2.is_alive()


def test_streamer_emits_whole_characters_only():
    """Byte-level tokens cut multi-byte characters; the streamer holds back until a character is complete."""
    import torch
    from llm_in_practise_b200.data.tokenizers import ByteTokenizer
    from llm_in_practise_b200.inference import TextIteratorStreamer
    tok = ByteTokenizer()
    text = "马哥教育\nok 云原生"
    st = TextIteratorStreamer(tok, timeout=5)
    for b in tok.encode(text):
        st.put(torch.tensor([b]))
    st.end()
    pieces = list(st)
    assert "".join(pieces) == text and all("�" not in p for p in pieces)
    assert pieces[:4] == ["马", "哥", "教", "育"]                 # three bytes each, one piece per character
