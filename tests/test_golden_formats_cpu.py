"""Checkpoint formats pinned against the real libraries that are in the image:

* compressed-tensors ``pack-quantized``: the ``compressed_tensors`` package itself quantises / packs / unpacks a weight, and
  this framework must read its tensors and produce tensors it reads back, to the bit.
* GPTQ ``qweight`` / ``qzeros`` and AutoAWQ ``qweight`` / ``qzeros`` word layouts: vLLM's own packing helpers
  (``gptq_pack`` / ``pack_cols`` / ``awq_pack``, the layouts its loaders consume) produce the golden words.
* safetensors: files written here are read by the ``safetensors`` package and vice versa, bf16 and metadata included.

* PEFT adapters: tensor names parse with vLLM's PEFT name parser and resolve to Linears of the real ``transformers`` Qwen3,
  ``adapter_config.json`` passes vLLM's ``PEFTHelper``.

bitsandbytes, peft, gptqmodel and llm-compressor themselves are not installable here: the NF4 nibble order / nested absmax and
the GPTQ v1 zero-point offset stay matched from documentation.
"""
import json

import pytest
import torch

from llm_in_practise_b200.quant.formats import (from_awq, from_compressed_tensors, from_gptq, quantize_rtn, to_awq,
                                                to_compressed_tensors, to_gptq)


def _weight(n=24, k=256, seed=0):
    torch.manual_seed(seed)
    return torch.randn(n, k) * 0.05


@pytest.mark.parametrize("symmetric", [True, False])
def test_compressed_tensors_pack_quantized_both_ways(symmetric):
    ct = pytest.importorskip("compressed_tensors")
    from compressed_tensors.compressors.pack_quantized.base import PackedQuantizationCompressor
    from compressed_tensors.quantization import QuantizationArgs, QuantizationScheme
    from compressed_tensors.quantization.utils import calculate_qparams
    args = QuantizationArgs(num_bits=4, type="int", symmetric=symmetric, strategy="group", group_size=128)
    scheme = QuantizationScheme(targets=["Linear"], weights=args)
    w = _weight()
    N, K = w.shape
    g = w.view(N, K // 128, 128)
    scale, zp = calculate_qparams(g.amin(-1), g.amax(-1), args)           # the library's own scale / zero-point rule
    sd = {"weight": w, "weight_scale": scale, "weight_zero_point": zp}
    golden = PackedQuantizationCompressor.compress(sd, scheme)              # what llm-compressor writes to disk
    assert golden["weight_packed"].dtype == torch.int32 and tuple(golden["weight_packed"].shape) == (N, K // 8)
    assert ("weight_zero_point" in golden) == (not symmetric)
    want = PackedQuantizationCompressor.decompress(golden, scheme)["weight"].float()
    # 1. their tensors, our reader
    ours = from_compressed_tensors({k: v for k, v in golden.items() if k != "weight"}, 128)
    got = ours.dequantize(torch.float32)
    assert torch.allclose(got, want, atol=1e-6), (got - want).abs().max()
    assert float((want - w).abs().max()) < float(scale.max()) * 0.51 + 1e-6             # and it is a sane quantisation of w
    # 2. our tensors, their reader
    q = quantize_rtn(w, 128, symmetric)
    mine = to_compressed_tensors(q, symmetric)
    assert mine["weight_packed"].dtype == torch.int32 and tuple(mine["weight_packed"].shape) == (N, K // 8)
    back = PackedQuantizationCompressor.decompress(dict(mine), scheme)["weight"].float()
    ref = q.dequantize(torch.float32)
    tol = float(q.scale.max()) * 2 ** -7                                                # weight_scale is stored as bf16
    assert torch.allclose(back, ref, atol=16 * tol), (back - ref).abs().max()
    # 3. word for word: packing their integer codes with our packer gives their words
    repacked = to_compressed_tensors(ours, symmetric)
    assert torch.equal(repacked["weight_packed"], golden["weight_packed"])
    if not symmetric:
        assert torch.equal(repacked["weight_zero_point"], golden["weight_zero_point"])
    assert ct.__version__


def _vllm_quant_utils():
    pytest.importorskip("vllm")
    try:
        from vllm.model_executor.layers.quantization.utils import quant_utils
    except Exception as e:                                         # a CPU-only import problem is not this repo's failure
        pytest.skip(f"vllm quant_utils not importable here: {type(e).__name__}")
    return quant_utils


def test_gptq_words_match_vllm_packing():
    qu = _vllm_quant_utils()
    q = quantize_rtn(_weight(32, 256, 1), 128, False)
    t = to_gptq(q)
    codes_kn = q.codes().t().contiguous().to(torch.int32)                              # [K, N]
    assert torch.equal(t["qweight"], qu.gptq_pack(codes_kn, 4, q.K, q.N))               # 8 rows of K per word, LSB first
    zeros_gn = (q.zero_points().t().contiguous() - 1) & 15                              # v1 checkpoints store zero - 1
    assert torch.equal(t["qzeros"], qu.pack_cols(zeros_gn.to(torch.int32), 4, q.K // 128, q.N))
    assert torch.equal(qu.unpack_cols(t["qzeros"], 4, q.K // 128, q.N), zeros_gn.to(torch.int32))
    back = from_gptq(t, 128)
    assert torch.equal(back.codes(), q.codes())
    assert torch.allclose(back.dequantize(torch.float32), q.dequantize(torch.float32), atol=float(q.scale.max()) * 16 * 2 ** -10)


def test_awq_words_match_vllm_packing_and_roundtrip(tmp_path):
    qu = _vllm_quant_utils()
    q = quantize_rtn(_weight(32, 256, 2), 128, False)
    t = to_awq(q)
    codes_kn = q.codes().t().contiguous().to(torch.int32)
    assert tuple(t["qweight"].shape) == (q.K, q.N // 8) and tuple(t["qzeros"].shape) == (q.K // 128, q.N // 8)
    assert torch.equal(t["qweight"], qu.awq_pack(codes_kn, 4, q.K, q.N))                # interleaved column order 0 2 4 6 1 3 5 7
    assert torch.equal(t["qzeros"], qu.awq_pack(q.zero_points().t().contiguous().to(torch.int32), 4, q.K // 128, q.N))
    back = from_awq(t, 128)
    assert torch.equal(back.codes(), q.codes()) and torch.equal(back.zero_points(), q.zero_points())
    # a whole checkpoint in the AutoAWQ layout is detected and loaded
    from llm_in_practise_b200.models.qwen3 import QWEN3_PRESETS, Qwen3ForCausalLM
    from llm_in_practise_b200.quant import QuantizeConfig, detect_quant_format, load_quantized, quantize_model, save_quantized
    torch.manual_seed(0)
    m = Qwen3ForCausalLM(QWEN3_PRESETS["qwen3-tiny"], dtype=torch.float32, device="cpu")
    quantize_model(m, [], QuantizeConfig(4, 64 if m.config.hidden_size % 128 else 128, False, False, method="rtn"))
    out = tmp_path / "tiny-AWQ"
    save_quantized(m, str(out), fmt="awq")
    cfg = json.load(open(out / "config.json"))["quantization_config"]
    assert cfg["quant_method"] == "awq" and cfg["version"] == "gemm" and cfg["zero_point"] is True
    assert detect_quant_format(str(out)) == "awq"
    m2 = load_quantized(str(out), dtype=torch.float32, device="cpu")
    ids = torch.arange(10).view(1, 10)
    with torch.no_grad():
        a, b = m(ids)["logits"], m2(ids)["logits"]
    assert torch.allclose(a, b, atol=2e-3, rtol=1e-3), (a - b).abs().max()              # scales travel as fp16


def test_safetensors_interoperates_with_the_real_package(tmp_path):
    st = pytest.importorskip("safetensors.torch")
    from llm_in_practise_b200.io.safetensors import load_file, save_file
    torch.manual_seed(3)
    tensors = {"a.weight": torch.randn(5, 7).to(torch.bfloat16), "b": torch.arange(12, dtype=torch.int32).view(3, 4),
               "c": torch.randn(2, 3, 4).half(), "scalar": torch.tensor(3.5), "empty": torch.zeros(0, 4),
               "u8": torch.arange(6, dtype=torch.uint8), "i64": torch.tensor([2 ** 40, -5])}
    ours, theirs = tmp_path / "ours.safetensors", tmp_path / "theirs.safetensors"
    save_file(tensors, str(ours), {"format": "pt", "note": "x"})
    got = st.load_file(str(ours))                                                       # the real reader on our file
    assert set(got) == set(tensors)
    for k, v in tensors.items():
        assert got[k].dtype == v.dtype and got[k].shape == v.shape and torch.equal(got[k], v), k
    from safetensors import safe_open
    with safe_open(str(ours), framework="pt") as f:
        assert f.metadata() == {"format": "pt", "note": "x"}
    st.save_file(tensors, str(theirs), metadata={"format": "pt"})                       # our reader on the real writer's file
    mine = load_file(str(theirs))
    for k, v in tensors.items():
        assert mine[k].dtype == v.dtype and mine[k].shape == v.shape and torch.equal(mine[k], v), k


def test_peft_adapter_is_what_vllm_and_transformers_expect(tmp_path):
    """``adapter_model.safetensors`` / ``adapter_config.json`` written here: every tensor name parses with vLLM's PEFT name
    parser and resolves to a Linear of the real ``transformers`` Qwen3 (so ``vllm --enable-lora`` / peft find their targets),
    the config passes vLLM's ``PEFTHelper``, A / B shapes are peft's ([r, in] and [out, r])."""
    transformers = pytest.importorskip("transformers")
    pytest.importorskip("vllm")
    try:
        from vllm.lora.peft_helper import PEFTHelper
        from vllm.lora.utils import parse_fine_tuned_lora_name
    except Exception as e:
        pytest.skip(f"vllm.lora not importable here: {type(e).__name__}")
    st = pytest.importorskip("safetensors.torch")
    from llm_in_practise_b200.models.qwen3 import QWEN3_PRESETS, Qwen3ForCausalLM
    from llm_in_practise_b200.peft import LoraConfig, get_peft_model
    torch.manual_seed(0)
    cfg = QWEN3_PRESETS["qwen3-tiny"]
    m = Qwen3ForCausalLM(cfg, dtype=torch.float32, device="cpu")
    pm = get_peft_model(m, LoraConfig(r=4, lora_alpha=8, lora_dropout=0.05, target_modules=["q_proj", "v_proj", "down_proj"],
                                      task_type="CAUSAL_LM"))
    out = tmp_path / "adapter"
    pm.save_pretrained(str(out))
    helper = PEFTHelper.from_local_dir(str(out), max_position_embeddings=cfg.max_position_embeddings)
    assert helper.r == 4 and helper.lora_alpha == 8 and set(helper.target_modules) == {"q_proj", "v_proj", "down_proj"}
    assert helper.vllm_lora_scaling_factor == 2.0
    hf = transformers.Qwen3ForCausalLM(transformers.Qwen3Config(
        vocab_size=cfg.vocab_size, hidden_size=cfg.hidden_size, intermediate_size=cfg.intermediate_size,
        num_hidden_layers=cfg.num_hidden_layers, num_attention_heads=cfg.num_attention_heads,
        num_key_value_heads=cfg.num_key_value_heads, head_dim=cfg.head_dim, max_position_embeddings=cfg.max_position_embeddings))
    linears = {n: mod for n, mod in hf.named_modules() if isinstance(mod, torch.nn.Linear)}
    sd = st.load_file(str(out / "adapter_model.safetensors"))
    seen = set()
    for name, t in sd.items():
        assert name.startswith("base_model.model.model.layers."), name
        module, is_a = parse_fine_tuned_lora_name(name)
        assert module in linears, f"{name} -> {module} is not a Linear of transformers' Qwen3"
        lin = linears[module]
        assert tuple(t.shape) == ((4, lin.in_features) if is_a else (lin.out_features, 4)), (name, t.shape)
        seen.add((module.rsplit(".", 1)[-1], is_a))
    assert seen == {(p, a) for p in ("q_proj", "v_proj", "down_proj") for a in (True, False)}
    assert len(sd) == cfg.num_hidden_layers * 3 * 2


def test_nf4_codebook_equals_its_definition():
    """The 16 NF4 levels are not free parameters: QLoRA defines them as normal quantiles (8 on the positive side, 7 on the
    negative, plus an exact zero) scaled to [-1, 1]. Recomputing them with scipy gives the table the kernels use."""
    stats = pytest.importorskip("scipy.stats")
    import numpy as np
    from llm_in_practise_b200.ops.w4 import NF4_CODE, create_dynamic_map
    offset = 0.9677083                                                    # the probability mass cut-off of the construction
    pos = stats.norm.ppf(np.linspace(offset, 0.5, 9)[:-1])
    neg = -stats.norm.ppf(np.linspace(offset, 0.5, 8)[:-1])
    levels = np.sort(np.concatenate([pos, [0.0], neg]))
    levels /= levels.max()
    assert levels.shape == (16,) and np.abs(levels - NF4_CODE.numpy()).max() < 5e-7
    # the 8-bit dynamic code of the nested absmax / 8-bit Adam state: 256 sorted levels in [-1, 1] with exact 0 and 1
    for signed in (True, False):
        code = create_dynamic_map(signed=signed)
        assert code.numel() == 256 and bool((code[1:] >= code[:-1]).all()) and float(code.max()) == 1.0
        assert 0.0 in code.tolist() and (float(code.min()) < 0) == signed


def test_reference_tokenizer_json_keeps_its_ids(tmp_path):
    """A ``tokenizer.json`` produced the way the reference's training scripts produce it (HF tokenizers, ByteLevel BPE, the
    four special tokens) loads through ``BPETokenizer.from_file`` / ``load_or_train_bpe`` and encodes to the same ids."""
    tokenizers = pytest.importorskip("tokenizers")
    from tokenizers import Tokenizer, models, pre_tokenizers, trainers
    from llm_in_practise_b200.data.tokenizers import BPETokenizer, HFTokenizerFile, load_or_train_bpe
    texts = ["the quick brown fox jumps over the lazy dog", "pack my box with five dozen liquor jugs", "the fox and the dog"] * 20
    tk = Tokenizer(models.BPE(unk_token="[UNK]"))
    tk.pre_tokenizer = pre_tokenizers.ByteLevel()
    tk.train_from_iterator(iter(texts), trainer=trainers.BpeTrainer(vocab_size=300, special_tokens=["[PAD]", "[UNK]", "[CLS]", "[SEP]"],
                                                                      show_progress=False))
    path = tmp_path / "tokenizer.json"
    tk.save(str(path))
    mine = BPETokenizer.from_file(str(path))
    assert isinstance(mine, HFTokenizerFile) and mine.vocab_size == tk.get_vocab_size()
    s = "the lazy fox packs five boxes"
    assert mine.encode(s) == tk.encode(s).ids
    assert mine.encode_batch(texts[:3]) == [e.ids for e in tk.encode_batch(texts[:3])]
    assert mine(texts[:2], truncation=True, max_length=4)["input_ids"] == [e.ids[:4] for e in tk.encode_batch(texts[:2])]
    assert mine.pad_token_id == 0 and mine.token_to_id("[SEP]") == 3 == mine.eos_token_id
    assert mine.decode(mine.encode(s)) == tk.decode(tk.encode(s).ids)        # (the reference sets no decoder: pieces joined by spaces)
    # the cache-or-train helper finds the file and does not retrain over it
    again = load_or_train_bpe(str(path), lambda: (_ for _ in ()).throw(AssertionError("must not retrain")), vocab_size=300)
    assert again.encode(s) == tk.encode(s).ids
    # this framework's own file format still round-trips
    own = BPETokenizer.train_from_iterator(texts, vocab_size=280)
    own.save(str(tmp_path / "bpe_tokenizer.json"))
    assert BPETokenizer.from_file(str(tmp_path / "bpe_tokenizer.json")).encode(s) == own.encode(s)
    assert tokenizers.__version__


def test_trainer_state_json_moves_between_hf_and_this_trainer(tmp_path):
    """``checkpoint-N/trainer_state.json``: a file written by the real ``transformers.TrainerState`` resumes here (its extra
    fields survive a save), and a file written here loads with transformers."""
    tc = pytest.importorskip("transformers.trainer_callback")
    from llm_in_practise_b200.trainer.args import TrainerState
    hf = tc.TrainerState(epoch=1.5, global_step=30, max_steps=100, logging_steps=10, eval_steps=20, save_steps=20, train_batch_size=2,
                         num_train_epochs=3, total_flos=1.5e12, best_metric=0.42, best_model_checkpoint="out/checkpoint-20",
                         log_history=[{"loss": 2.5, "learning_rate": 1e-4, "epoch": 0.5, "step": 10}, {"eval_loss": 2.1, "epoch": 1.0, "step": 20}])
    theirs, ours = str(tmp_path / "hf.json"), str(tmp_path / "ours.json")
    hf.save_to_json(theirs)
    st = TrainerState.load_from_json(theirs)
    assert (st.global_step, st.epoch, st.max_steps, st.best_metric, st.best_model_checkpoint) == (30, 1.5, 100, 0.42, "out/checkpoint-20")
    assert st.log_history == hf.log_history and st.extra["train_batch_size"] == 2 and st.extra["eval_steps"] == 20
    st.global_step, st.epoch = 40, 2.0
    st.log_history.append({"loss": 1.9, "epoch": 2.0, "step": 40})
    st.save_to_json(ours)
    back = tc.TrainerState.load_from_json(ours)                              # transformers reads what this trainer wrote
    assert back.global_step == 40 and back.epoch == 2.0 and back.train_batch_size == 2 and len(back.log_history) == 3
    fresh = str(tmp_path / "fresh.json")
    TrainerState(epoch=0.25, global_step=5, max_steps=20, log_history=[{"loss": 3.0, "step": 5}]).save_to_json(fresh)
    assert tc.TrainerState.load_from_json(fresh).global_step == 5
