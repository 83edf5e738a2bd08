"""Recipes: the pretraining runner under every strategy (2 gloo processes) and the tool sub-commands, on CPU."""
import json
import os

import pytest
import torch
from dist_utils import run_distributed


def _run(rank, world, strategy, tmp):
    from llm_in_practise_b200.recipes.pretrain import PRESETS
    from llm_in_practise_b200.train import PretrainConfig, run_pretraining
    ds = None
    if strategy.startswith("zero"):
        ds = {"train_micro_batch_size_per_gpu": 4, "gradient_accumulation_steps": 1, "zero_optimization": {"stage": int(strategy[-1])},
              "optimizer": {"type": "AdamW", "params": {"lr": 1e-3, "weight_decay": 0.0}}}
        strategy = "zero"
    cfg = PretrainConfig(strategy=strategy, ds_config=ds, tokenizer="byte", block_size=32, n_layer=2, n_head=2, d_model=32, dropout=0.0,
                         epochs=2, batch_size=4, lr=1e-3, max_lines=120, log_every=0, save_dir=os.path.join(tmp, "ck"), val_fraction=0.1,
                         clip_grad_norm=1.0, no_decay_groups=True)
    out = run_pretraining(cfg)
    h = out["history"]
    assert len(h) == 2 and h[1]["train_loss"] < h[0]["train_loss"], h
    if rank == 0:
        json.dump([r["train_loss"] for r in h], open(os.path.join(tmp, "loss.json"), "w"))


@pytest.mark.parametrize("strategy", ["ddp", "fsdp", "fsdp2", "zero2", "zero3"])
def test_pretrain_strategies(strategy, tmp_path):
    run_distributed(_run, 2, (strategy, str(tmp_path)))
    assert os.path.exists(tmp_path / "loss.json")
    assert os.listdir(tmp_path / "ck")


def test_preset_table_is_consistent():
    from dataclasses import fields
    from llm_in_practise_b200.recipes.pretrain import PRESETS, build_parser
    from llm_in_practise_b200.train import PretrainConfig
    names = {f.name for f in fields(PretrainConfig)}
    for k, v in PRESETS.items():
        assert set(v) <= names, (k, set(v) - names)
        PretrainConfig(**v)
        if isinstance(v.get("ds_config"), str):
            assert os.path.exists(v["ds_config"]), v["ds_config"]
    a = build_parser().parse_args(["--recipe", "ddp_gpt_wikitext2", "--epochs", "1", "--n_head", "4", "--resume", "true"])
    assert a.epochs == 1 and a.n_head == 4 and a.resume is True


def test_tools_roundtrip(tmp_path):
    from llm_in_practise_b200.io.hf import from_pretrained, save_pretrained
    from llm_in_practise_b200.quant import detect_quant_format
    from llm_in_practise_b200.recipes import finetune, infer, tools
    d = str(tmp_path)
    save_pretrained(from_pretrained("qwen3-tiny", dtype=torch.float32, device="cpu"), d + "/base")
    finetune.main(["--recipe", "qwen3-8b-qlora", "--model", d + "/base", "--max_steps", "2", "--max_length", "48", "--output_dir", d + "/ft"])
    assert os.path.exists(d + "/ft/adapter_model.safetensors")
    tools.main(["merge", "--base", d + "/base", "--adapter", d + "/ft", "--out", d + "/merged"])
    tools.main(["quantize", "--model", d + "/merged", "--method", "awq", "--out", d + "/awq", "--num_calibration_samples", "3",
                "--max_seq_length", "48"])
    assert detect_quant_format(d + "/awq") == "compressed-tensors"
    ppl = tools.main(["eval-ppl", "--model", d + "/awq", "--num_samples", "2", "--max_tokens", "6"])
    assert ppl == ppl and ppl > 1.0
    infer.main(["once", "--model", d + "/awq", "--max_new_tokens", "4"])


def test_finetune_from_llamafactory_yaml(tmp_path):
    """A LLaMA-Factory `*_lora_sft.yaml` drives the same fine-tuning pipeline: lora_target all -> the seven Qwen3 linears,
    cutoff_len, cosine schedule; the (not strictly valid) recipe shipped with the reference parses too."""
    import json
    from llm_in_practise_b200.recipes import finetune
    ref = "/root/reference/Fine-Tuning/LLaMA-Factory/deepseek-r1-0528-qwen3_lora_sft.yaml"
    if os.path.exists(ref):
        q = finetune.preset_from_llamafactory_yaml(ref)
        assert q["args"]["gradient_accumulation_steps"] == 8 and q["args"]["lr_scheduler_type"] == "cosine" and q["max_length"] == 2048
        assert len(q["lora"]["target_modules"]) == 7 and q["lora"]["r"] == 8 and q["lora"]["lora_alpha"] == 16
    d = str(tmp_path)
    y = d + "/sft.yaml"
    with open(y, "w") as f:
        f.write("model_name_or_path: qwen3-tiny\nstage: sft\nfinetuning_type: lora\nlora_target: q_proj,v_proj\nlora_rank: 4\n"
                "cutoff_len: 40\noutput_dir: %s/out\nper_device_train_batch_size: 2\ngradient_accumulation_steps: 1\n"
                "learning_rate: 1.0e-3\nnum_train_epochs: 1.0\nlr_scheduler_type: cosine\nwarmup_ratio: 0.1\nbf16: true\n" % d)
    finetune.main(["--llamafactory_yaml", y, "--max_steps", "2"])
    assert os.path.exists(d + "/out/adapter_model.safetensors")
    cfg = json.load(open(d + "/out/adapter_config.json"))
    assert cfg["r"] == 4 and cfg["lora_alpha"] == 8 and sorted(cfg["target_modules"]) == ["q_proj", "v_proj"]


def test_webui_path_with_a_stand_in_gradio(tmp_path, monkeypatch):
    """`recipes.infer webui` (Scripts/inference/05/06-…webui-infr.py: Blocks + Chatbot + Textbox + three sliders + clear
    button on 0.0.0.0:7860). gradio is not in this image, so a minimal stand-in module records what the recipe builds and
    drives the registered streaming callback once."""
    import sys
    import types
    from llm_in_practise_b200.io.hf import from_pretrained, save_pretrained
    from llm_in_practise_b200.recipes import infer
    d = str(tmp_path)
    save_pretrained(from_pretrained("qwen3-tiny", dtype=torch.float32, device="cpu"), d + "/base")
    rec = {"components": [], "launch": None, "submit": None, "clear": None}

    class _Comp:
        def __init__(self, kind, *a, **k):
            self.kind, self.args, self.kw = kind, a, k
            rec["components"].append(self)

        def submit(self, fn, inputs, outputs):
            rec["submit"] = (fn, inputs, outputs)

        def click(self, fn, inputs, outputs):
            rec["clear"] = (fn, inputs, outputs)

    class _Blocks:
        def __init__(self, **k):
            self.kw = k

        def __enter__(self):
            return self

        def __exit__(self, *a):
            return False

        def launch(self, **k):
            rec["launch"] = k

    gr = types.ModuleType("gradio")
    gr.Blocks = _Blocks
    for kind in ("Chatbot", "Textbox", "Slider", "Button"):
        setattr(gr, kind, (lambda kind: (lambda *a, **k: _Comp(kind, *a, **k)))(kind))
    monkeypatch.setitem(sys.modules, "gradio", gr)
    infer.main(["webui", "--model", d + "/base", "--max_new_tokens", "16"])
    assert rec["launch"] == {"server_name": "0.0.0.0", "server_port": 7860}
    kinds = [c.kind for c in rec["components"]]
    assert kinds.count("Slider") == 3 and "Chatbot" in kinds and "Textbox" in kinds and "Button" in kinds
    respond, inputs, outputs = rec["submit"]
    assert len(inputs) == 5 and outputs.kind == "Chatbot"
    frames = list(respond("hello", [("hi", "there")], 0.7, 0.9, 4))
    assert frames and frames[-1][0] == ("hi", "there") and frames[-1][-1][0] == "hello" and isinstance(frames[-1][-1][1], str)
    assert rec["clear"][0]() == []


def test_serve_parallel_flags_need_matching_torchrun(capsys):
    """``serve --tensor-parallel-size 2 --pipeline-parallel-size 2`` outside a 4-rank torchrun is refused with the launch line."""
    from llm_in_practise_b200.recipes import infer
    with pytest.raises(SystemExit):
        infer.main(["serve", "--model", "qwen3-tiny", "--tensor-parallel-size", "2", "--pipeline-parallel-size", "2"])
    assert "torchrun --nproc-per-node 4" in capsys.readouterr().err


def test_serve_accepts_the_vllm_flags_of_the_compose_files(monkeypatch):
    """``vllm serve``'s flags as the reference passes them map onto EngineArgs; ``--quantization bitsandbytes`` quantises at load."""
    from llm_in_practise_b200.recipes import infer
    from llm_in_practise_b200.serve import openai_server
    got = {}
    monkeypatch.setattr(openai_server, "serve", lambda eng, host, port, **kw: got.update(eng=eng, host=host, port=port, kw=kw))
    infer.main(["serve", "--model", "qwen3-tiny", "--quantization", "bitsandbytes", "--no-enable-prefix-caching", "--max-num-seqs", "3",
                "--max-model-len", "128", "--served-model-name", "tiny-nf4", "--block-size", "16", "--enforce-eager",
                "--gpu-memory-utilization", "0.9", "--port", "8123"])
    eng = got["eng"]
    a = eng.args
    assert (a.max_num_seqs, a.max_model_len, a.block_size, a.enable_prefix_caching, a.enable_chunked_prefill, a.warmup) == \
        (3, 128, 16, False, True, False)
    assert a.gpu_memory_utilization == 0.0                # an explicit --max-num-seqs wins over the memory share
    assert eng.model_name == "tiny-nf4" and eng.paged and got["port"] == 8123
    assert eng.model.layers[0].qkv_proj.nf4 is not None   # the decoder projections were quantised to NF4
    with pytest.raises(SystemExit):
        infer.main(["serve", "--model", "qwen3-tiny", "--quantization", "awq"])          # not an AWQ checkpoint


def test_tools_passthrough_commands_see_their_own_flags(capsys):
    """``tools bench-serve --base-url …`` / ``tools loadtest --host …``: flags go to the sub-tool's parser (argparse's REMAINDER
    would have refused a leading option)."""
    from llm_in_practise_b200.recipes import tools
    out = tools.main(["bench-serve", "--base-url", "http://127.0.0.1:9", "--model", "x", "--num-prompts", "2",
                      "--random-input-len", "4", "--random-output-len", "2"])
    assert out["completed"] == 0 and out["failed"] == 2 and out["num_prompts"] == 2       # nothing listens on port 9
    assert "Serving Benchmark Result" in capsys.readouterr().out
    with pytest.raises(SystemExit) as e:
        tools.main(["loadtest", "--help"])
    assert e.value.code == 0 and "--profile" in capsys.readouterr().out


def test_stream_command_exits_cleanly():
    """``recipes.infer stream`` in its own process: the generation thread is joined before the interpreter shuts down (it used
    to be killed inside native code: 'terminate called without an active exception', exit status 134)."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "llm_in_practise_b200.recipes.infer", "stream", "--model", "qwen3-tiny",
                        "--max_new_tokens", "6", "--prompt", "hi"], cwd=root, capture_output=True, timeout=300)
    assert r.returncode == 0, (r.returncode, r.stderr[-400:])
    assert b"terminate called" not in r.stderr
