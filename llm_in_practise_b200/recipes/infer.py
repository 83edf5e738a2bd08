"""Inference entry points of the reference as sub-commands:

``python -m llm_in_practise_b200.recipes.infer <cmd> --model PATH [...]``

* ``once``      one prompt -> one completion; ``<|user|>\\n…\\n<|assistant|>\\n`` prompt, ``max_new_tokens 256, temperature 0.7,
                top_p 0.95`` (Scripts/inference/01..03-*.py:8-30)
* ``console``   multi-turn REPL keeping the history in the prompt (Scripts/inference/04-…multisession-infr.py:27-76)
* ``chat``      ChatML chat with an optional LoRA adapter directory, ``top_p 0.9, temperature 0.7`` (Fine-Tuning/inferences.py:18-83)
* ``stream``    token streaming to stdout through ``TextIteratorStreamer`` and a generation thread
                (Scripts/inference/06-…streaming-infr.py:55-85)
* ``webui``     the Gradio chat UI (sliders, clear button, 0.0.0.0:7860) when gradio is importable (05/06-…webui-infr.py); otherwise
                prints how to use ``serve`` + any OpenAI-compatible UI
* ``serve``     OpenAI-compatible server on 0.0.0.0:8000 (Scripts/inference/07-deepseek1.5b-api-infr.py:66-174)
* ``quantized`` chat against a GPTQ / AWQ (compressed-tensors) checkpoint, ``max_new_tokens 512, temperature 0.7, top_p 0.9,
                repetition_penalty 1.1`` (Quantization/GPTQModel/inference_qwen3_4b_gptq.py:29-115,
                Quantization/LLM-Compressor/*/inference_qwen3_4b_*.py:6-52)
"""
from __future__ import annotations

import argparse
import os
import sys

import torch

from ..data.chat import deepseek_prompt
from ..inference import GenerationConfig, Generator, chat, stream_generate
from ..io.hf import from_pretrained


def _load(path: str, adapter: str = None, dtype=torch.bfloat16):
    from ..quant import detect_quant_format, load_quantized
    if os.path.isdir(path) and detect_quant_format(path) is not None:
        model = load_quantized(path, dtype=dtype)
    else:
        model = from_pretrained(path, dtype=dtype)
    if adapter:
        from ..peft import PeftModel
        model = PeftModel.from_pretrained(model, adapter)
    tok = None
    for cand in (adapter, path):
        if cand and os.path.isdir(cand):
            try:
                from ..data.tokenizers import load_pretrained_tokenizer
                tok = load_pretrained_tokenizer(cand, add_pad_token=False)
                break
            except Exception:
                continue
    if tok is None:
        from ..data.tokenizers import ByteTokenizer
        tok = ByteTokenizer()
    return model, tok


def _ids(tok, text, device):
    return torch.tensor([tok.encode(text)], device=device)


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("cmd", choices=["once", "console", "chat", "stream", "webui", "serve", "quantized"])
    ap.add_argument("--model", required=True)
    ap.add_argument("--adapter", default=None)
    ap.add_argument("--prompt", default="你好，请介绍一下你自己。")
    ap.add_argument("--system", default="You are a helpful assistant.")
    ap.add_argument("--max_new_tokens", type=int, default=256)
    ap.add_argument("--temperature", type=float, default=0.7)
    ap.add_argument("--top_p", type=float, default=0.95)
    ap.add_argument("--host", default="0.0.0.0")
    ap.add_argument("--port", type=int, default=8000)
    ap.add_argument("--max_model_len", "--max-model-len", dest="max_model_len", type=int, default=4096)
    ap.add_argument("--max_num_seqs", "--max-num-seqs", dest="max_num_seqs", type=int, default=None)
    # the vLLM server flags the reference's compose files and manifests pass (Deployment/Basics-02/docker-compose.yml:56-65,
    # LLM_on_Kubernetes/Inference_Platfrom/06-GPU-Timeslicing/vllm-statefulset.yaml:65-82)
    ap.add_argument("--gpu-memory-utilization", "--gpu_memory_utilization", dest="gpu_mem", type=float, default=0.0,
                    help="serve: size the KV pool from this share of HBM when --max-num-seqs is not given")
    ap.add_argument("--enable-prefix-caching", dest="prefix_caching", action=argparse.BooleanOptionalAction, default=True)
    ap.add_argument("--enable-chunked-prefill", dest="chunked_prefill", action=argparse.BooleanOptionalAction, default=True)
    ap.add_argument("--served-model-name", dest="served_name", default=None)
    ap.add_argument("--quantization", default=None, choices=["awq", "gptq", "compressed-tensors", "bitsandbytes"],
                    help="serve: awq / gptq / compressed-tensors checkpoints are detected from their config; bitsandbytes "
                         "quantises a 16-bit checkpoint to NF4 while loading")
    ap.add_argument("--kv-offloading-size", "--kv-offload-gb", dest="kv_offload_gb", type=float, default=0.0,
                    help="serve: GiB of pinned host memory for cached prefixes that lose their HBM row (LMCache-style tier)")
    ap.add_argument("--enforce-eager", dest="enforce_eager", action="store_true", help="serve: no CUDA-graph warm-up at start")
    ap.add_argument("--prompt_style", default="chatml", choices=["chatml", "deepseek"])
    ap.add_argument("--dtype", default="bfloat16", choices=["bfloat16", "float16"])
    ap.add_argument("--tensor-parallel-size", "--tensor_parallel_size", dest="tp", type=int, default=1,
                    help="serve: ranks that share every layer (vllm serve --tensor-parallel-size; launch under torchrun)")
    ap.add_argument("--pipeline-parallel-size", "--pipeline_parallel_size", dest="pp", type=int, default=1,
                    help="serve: pipeline stages (engine_kwargs.pipeline_parallel_size of the Ray Serve recipe)")
    ap.add_argument("--block-size", "--block_size", dest="block_size", type=int, default=0, help="serve: paged KV cache block (0 = one row per sequence)")
    a = ap.parse_args(argv)
    dtype = getattr(torch, a.dtype)
    if a.tp * a.pp > 1:
        from ..parallel.dist import setup_distributed
        info = setup_distributed()                      # one process per GPU; sets the device the checkpoint is loaded onto
        if a.cmd != "serve" or info.world_size != a.tp * a.pp:
            ap.error(f"--tensor-parallel-size x --pipeline-parallel-size = {a.tp * a.pp} needs `serve` under "
                     f"`torchrun --nproc-per-node {a.tp * a.pp}` (this run has {info.world_size} rank(s))")
    model, tok = _load(a.model, a.adapter, dtype)
    if a.quantization == "bitsandbytes":
        getattr(model, "base_model", model).quantize_nf4(True)
    elif a.quantization is not None:
        from ..quant import detect_quant_format
        if not (os.path.isdir(a.model) and detect_quant_format(a.model)):
            ap.error(f"--quantization {a.quantization}: {a.model} is not a quantised checkpoint (no quantization_config found)")
    dev = getattr(model, "base_model", model).embed_tokens.device
    eos = [t for t in (getattr(tok, "eos_token_id", None),) if t is not None]
    gen = Generator(model, tok, a.max_model_len)

    if a.cmd == "once":
        ids = _ids(tok, deepseek_prompt(a.prompt), dev)
        out = gen.generate(ids, max_new_tokens=a.max_new_tokens, do_sample=True, temperature=a.temperature, top_p=a.top_p, eos_token_id=eos)
        print(tok.decode(out[0, ids.shape[1]:].tolist()))
    elif a.cmd == "console":
        history = []
        print("输入 exit 退出。")
        while True:
            try:
                q = input("用户: ").strip()
            except EOFError:
                break
            if q.lower() in ("exit", "quit"):
                break
            ids = _ids(tok, deepseek_prompt(q, history=history), dev)
            out = gen.generate(ids, max_new_tokens=a.max_new_tokens, do_sample=True, temperature=a.temperature, top_p=a.top_p, eos_token_id=eos)
            ans = tok.decode(out[0, ids.shape[1]:].tolist()).strip()
            print("助手:", ans)
            history.append((q, ans))
    elif a.cmd in ("chat", "quantized"):
        kw = dict(repetition_penalty=1.1, top_p=0.9) if a.cmd == "quantized" else dict(top_p=0.9)
        history = []
        for q in (a.prompt, "你是谁开发的？"):                      # single- and multi-turn smoke prompts
            ans, history = chat(gen, tok, q, history, a.system, max_new_tokens=a.max_new_tokens if a.cmd == "chat" else 512,
                                temperature=a.temperature, **kw)
            print(f"用户: {q}\n助手: {ans}\n")
    elif a.cmd == "stream":
        ids = _ids(tok, deepseek_prompt(a.prompt), dev)
        for piece in stream_generate(gen, ids, tok, max_new_tokens=a.max_new_tokens, do_sample=True, temperature=a.temperature,
                                     top_p=a.top_p, eos_token_id=eos):
            sys.stdout.write(piece)
            sys.stdout.flush()
        print()
    elif a.cmd == "webui":
        try:
            import gradio as gr
        except ImportError:
            print("gradio is not installed in this image. Start `recipes.infer serve` and point any OpenAI-compatible chat UI "
                  f"(Open-WebUI, AnythingLLM, …) at http://{a.host}:{a.port}/v1 instead.")
            return

        def respond(message, chat_history, temperature, top_p, max_tokens):
            ids = _ids(tok, deepseek_prompt(message, history=chat_history), dev)
            acc = ""
            for piece in stream_generate(gen, ids, tok, max_new_tokens=int(max_tokens), do_sample=True, temperature=temperature,
                                         top_p=top_p, eos_token_id=eos):
                acc += piece
                yield chat_history + [(message, acc)]

        with gr.Blocks(title="llm_in_practise_b200") as demo:
            bot = gr.Chatbot()
            msg = gr.Textbox(label="输入")
            t = gr.Slider(0.1, 1.5, value=a.temperature, label="temperature")
            p = gr.Slider(0.1, 1.0, value=a.top_p, label="top_p")
            n = gr.Slider(16, 1024, value=a.max_new_tokens, step=16, label="max_new_tokens")
            msg.submit(respond, [msg, bot, t, p, n], bot)
            gr.Button("清空").click(lambda: [], None, bot)
        demo.launch(server_name="0.0.0.0", server_port=7860)
    elif a.cmd == "serve":
        from ..serve import EngineArgs, LLMEngine
        from ..serve.openai_server import serve
        eargs = EngineArgs(max_num_seqs=a.max_num_seqs or 16, max_model_len=a.max_model_len, block_size=a.block_size,
                           enable_prefix_caching=a.prefix_caching, enable_chunked_prefill=a.chunked_prefill,
                           gpu_memory_utilization=(a.gpu_mem if a.max_num_seqs is None else 0.0), warmup=not a.enforce_eager,
                           kv_offload_gb=a.kv_offload_gb)
        name = a.served_name or os.path.basename(a.model.rstrip("/"))
        if a.tp * a.pp > 1:
            # SPMD: rank 0 schedules and answers HTTP, the others repeat its model calls on their shard / stage
            from ..serve.tp_engine import build_pp_engine, build_tp_engine
            base = getattr(model, "base_model", model)
            eng = (build_pp_engine(base, tok, eargs, name, a.pp, a.tp) if a.pp > 1 else build_tp_engine(base, tok, eargs, name))
            if eng.rank != 0:
                eng.serve_forever()
                return
        else:
            eng = LLMEngine(model, tok, eargs, model_name=name)
        serve(eng, a.host, a.port, prompt_style=a.prompt_style)


if __name__ == "__main__":
    main()
