"""llm-demo counterparts (minigpt / minigpt2: config 1 of BASELINE.md), the data utilities, the GPTQModel-style front end and
the CPU AdamW of ZeRO-Offload — the pieces that only benchmarks exercised so far."""
import os

import torch


def test_minigpt_trains_generates_and_loads_reference_checkpoints(tmp_path):
    """llm-demo/minigpt/train.py + generate.py: char vocabulary of the 61-character corpus, 16-token sliding windows, AdamW
    1e-3, clip 1.0; the checkpoint layout {"model_state", "char2idx", "config"} round-trips; greedy decoding continues the
    prompt; with reference_quirks the state dict carries nn.TransformerDecoderLayer's parameter names, so a reference
    .pth loads (and vice versa)."""
    from llm_in_practise_b200.models import minigpt as M
    path = str(tmp_path / "mg.pth")
    model, c2i, hist = M.train_minigpt(epochs=12, save_path=path, seed=0, device="cpu", log=None)
    assert len(c2i) == len(set(M.MAGE_TEXT)) and hist[-1] < 0.5 * hist[0]
    m2, c2 = M.load_minigpt(path)
    assert c2 == c2i
    out = M.generate_greedy(m2, c2, M.MAGE_TEXT[:4], max_len=12)
    assert out.startswith(M.MAGE_TEXT[:4]) and len(out) == 16 and set(out) <= set(c2i)
    # the reference's module tree: same parameter names and shapes
    ref = torch.nn.TransformerDecoderLayer(d_model=64, nhead=2, dropout=0.1)
    ref_names = {f"transformer.0.{k}": v.shape for k, v in ref.state_dict().items()}
    sd = model.state_dict()
    hit = [k for k in sd if any(k.endswith(r.split("transformer.0.")[1]) for r in ref_names)]
    assert hit, list(sd)[:8]
    # the corrected variant (causal, batch-first) trains too
    _, _, hist_fixed = M.train_minigpt(epochs=6, save_path=None, seed=0, device="cpu", reference_quirks=False, log=None)
    assert hist_fixed[-1] < hist_fixed[0]


def test_minigpt2_dataset_model_tester_and_recipe(tmp_path):
    """llm-demo/minigpt2/model.py + test_model.py: Config constants, TextDataset writes vocab_size, MiniGPT2 returns the
    loss when targets are given, GPTTester rebuilds the model from the checkpoint, left-pads prompts and samples at
    temperature 0.8; the recipe CLI drives all of it."""
    from llm_in_practise_b200.models import minigpt as M
    from llm_in_practise_b200.recipes import minigpt as R
    cfg = M.Config(seq_len=32, epochs=1, batch_size=4)
    assert (cfg.n_layer, cfg.n_head, cfg.n_embd, cfg.lr, cfg.weight_decay) == (4, 4, 128, 3e-4, 0.1)
    ds = M.TextDataset(M.MAGE_TEXT * 6, cfg)
    assert cfg.vocab_size == len(set(M.MAGE_TEXT)) and len(ds) > 0
    x, y = ds[0]
    assert x.shape == y.shape == (32,) and torch.equal(x[1:], y[:-1])
    model = M.MiniGPT2(cfg)
    logits = model(x[None])
    loss = model(x[None], y[None])
    assert logits.shape == (1, 32, cfg.vocab_size) and loss.dim() == 0 and torch.isfinite(loss)
    path = str(tmp_path / "mg2.pth")
    M.save_minigpt2(model, ds, path)
    t = M.GPTTester(path)
    assert t.test_output_shape()
    torch.manual_seed(0)
    s = t.generate(M.MAGE_TEXT[:3], max_len=10)
    assert s.startswith(M.MAGE_TEXT[:3]) and len(s) == 13
    # CLI: train (minigpt), generate, train2, test2
    p1, p2 = str(tmp_path / "a.pth"), str(tmp_path / "b.pth")
    hist = R.main(["train", "--epochs", "2", "--path", p1, "--device", "cpu"])
    assert len(hist) == 2
    assert R.main(["generate", "--path", p1, "--start", M.MAGE_TEXT[:2]]).startswith(M.MAGE_TEXT[:2])
    R.main(["train2", "--epochs", "1", "--path", p2, "--repeat", "5", "--device", "cpu"])
    R.main(["test2", "--path", p2, "--start", M.MAGE_TEXT[:2]])
    assert os.path.exists(p2)


def test_data_utilities():
    """data/datasets.py: deterministic seeding per rank, TokenizedDataset windows (ddp_gpt_wikitext2.py:54-82), the
    DistributedSampler contract (disjoint shards covering the set, reshuffled by set_epoch), the two collators."""
    from llm_in_practise_b200.data.datasets import (DataCollatorForLanguageModeling, DataCollatorWithPadding, DistributedSampler,
                                                    SyntheticTokenDataset, TokenizedDataset, set_seed, synthetic_corpus)
    from llm_in_practise_b200.data.tokenizers import ByteTokenizer
    set_seed(7, rank=1)
    a = torch.rand(3)
    set_seed(7, rank=1)
    assert torch.equal(a, torch.rand(3))
    set_seed(7, rank=0)
    assert not torch.equal(a, torch.rand(3))
    lines = synthetic_corpus(50, seed=1)
    assert len(lines) == 50 and lines == synthetic_corpus(50, seed=1)
    ds = TokenizedDataset(lines, ByteTokenizer(), block_size=16, max_length=64)
    x, y = ds[0]
    assert x.shape == y.shape == (15,) and torch.equal(x[1:], y[:-1])          # a block of 16 ids -> 15 (input, next-token) pairs
    syn = SyntheticTokenDataset(10, 8, 100, seed=0, shifted=True)
    assert len(syn) == 10 and int(syn[3][0].max()) < 100 and torch.equal(syn[3][0][1:], syn[3][1][:-1])
    assert set(SyntheticTokenDataset(2, 8, 100)[0]) == {"input_ids", "labels"}
    shards = []
    for r in range(3):
        s = DistributedSampler(syn, num_replicas=3, rank=r, shuffle=True, seed=5)
        s.set_epoch(0)
        shards.append(list(s))
    assert len({len(s) for s in shards}) == 1 and set().union(*shards) == set(range(10))
    s0 = DistributedSampler(syn, num_replicas=3, rank=0, shuffle=True, seed=5)
    s0.set_epoch(1)
    assert list(s0) != shards[0]
    feats = [{"input_ids": [5, 6, 7, 0], "attention_mask": [1, 1, 1, 0]}, {"input_ids": [8, 9, 0, 0], "attention_mask": [1, 1, 0, 0]}]
    b = DataCollatorForLanguageModeling(pad_token_id=0)(feats)
    assert b["labels"].tolist() == [[5, 6, 7, -100], [8, 9, -100, -100]]           # every non-pad token is a label (mlm=False)
    p = DataCollatorWithPadding(pad_token_id=0, pad_to_multiple_of=8)([{"input_ids": [1, 2, 3]}, {"input_ids": [4]}])
    assert p["input_ids"].shape == (2, 8) and p["attention_mask"].sum().item() == 4


def test_gptqmodel_front_end_and_perplexity(tmp_path):
    """quant/frontends.py: the GPTQModel-shaped API of Quantization/GPTQModel/quantize_qwen3_4b_gptq.py (QuantizeConfig ->
    quantize(calibration) -> save -> load -> generate) and the perplexity helpers of the evaluation scripts."""
    from llm_in_practise_b200.data.tokenizers import ByteTokenizer
    from llm_in_practise_b200.io.hf import from_pretrained
    from llm_in_practise_b200.inference import Generator
    from llm_in_practise_b200.quant import GPTQModel, QuantizeConfig, detect_quant_format
    from llm_in_practise_b200.quant.frontends import perplexity, self_perplexity
    torch.manual_seed(0)
    model = from_pretrained("qwen3-tiny", dtype=torch.float32, device="cpu")
    tok = ByteTokenizer()
    batches = [torch.randint(0, model.config.vocab_size, (2, 24)) for _ in range(2)]
    ppl0 = perplexity(model, batches)
    gm = GPTQModel(model, QuantizeConfig(bits=4, group_size=32, desc_act=False), tok)
    gm.quantize(["hello world, this is calibration text " * 3] * 3, batch_size=1, max_seq_length=32)
    out = str(tmp_path / "gptq")
    gm.save(out)
    assert detect_quant_format(out) == "gptq"
    gm2 = GPTQModel.load(out, dtype=torch.float32, device="cpu")
    ppl1 = perplexity(gm2.model, batches)
    assert ppl0 > 1.0 and ppl1 > 1.0 and abs(ppl1 - ppl0) / ppl0 < 0.5        # 4-bit weights move a random model's PPL a little
    ids = torch.randint(0, model.config.vocab_size, (1, 6))
    y = gm2.generate(input_ids=ids, max_new_tokens=4, do_sample=False)
    assert y.shape == (1, 10)
    mean, std = self_perplexity(Generator(gm2.model, tok, 64), [ids, ids.flip(1)], max_tokens=5)
    assert mean == mean and mean >= 1.0 and std >= 0.0


def test_cpu_adamw_matches_torch():
    """optim/cpu_adam.py (DeepSpeedCPUAdam's role under ZeRO-Offload, DeepSpeed-GPTLike-ZeRO-Offload.py:27,228): the native host
    AdamW over a parameter arena equals torch.optim.AdamW step for step; state round-trips."""
    from llm_in_practise_b200.optim.cpu_adam import CPUAdamW
    from llm_in_practise_b200.utils.arena import ParamArena
    torch.manual_seed(0)
    arena = ParamArena(device="cpu")
    arena.request("w", (257,))
    arena.request("b", (31,))
    arena.finalize()
    arena.data.normal_()
    ref = torch.nn.Parameter(arena.data.clone())
    opt = CPUAdamW(arena, lr=1e-2, weight_decay=0.05)
    ropt = torch.optim.AdamW([ref], lr=1e-2, weight_decay=0.05)
    for _ in range(5):
        g = torch.randn(arena.numel)
        arena.grad.copy_(g)
        ref.grad = g.clone()
        opt.step()
        ropt.step()
    assert torch.allclose(arena.data, ref.data, atol=1e-6)
    sd = opt.state_dict()
    opt2 = CPUAdamW(arena, lr=1e-2, weight_decay=0.05)
    opt2.load_state_dict(sd)
    g = torch.randn(arena.numel)
    arena.grad.copy_(g)
    ref.grad = g.clone()
    opt2.step()
    ropt.step()
    assert torch.allclose(arena.data, ref.data, atol=1e-6)
