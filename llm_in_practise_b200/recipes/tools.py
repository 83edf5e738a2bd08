"""Small command-line tools of the reference, one sub-command each:

``python -m llm_in_practise_b200.recipes.tools <cmd> [...]``

* ``merge``      merge a LoRA adapter into its base and save a plain HF checkpoint; auto-detects ``adapter_model.safetensors`` /
                 ``peft_model.safetensors`` vs a full fine-tune (Scripts/fine-tuning/02-merge-lora-adapter-and-model.py:27-38,
                 04-…-02.py:35-48; ``llamafactory-cli export`` in Fine-Tuning/LLaMA-Factory/README.md:240-246)
* ``quantize``   GPTQ / AWQ / RTN W4A16 quantisation with the calibration recipe of the lesson: first 128 rows,
                 ``instruction\\ninput\\noutput`` (Quantization/GPTQModel/quantize_qwen3_4b_gptq.py:16-48,
                 Quantization/LLM-Compressor/{GPTQ,AWQ}/quantize_qwen3_4b_*.py, Quantization/LoRA-AWQ/quantize-…-awq.py)
* ``eval-ppl``   the "self perplexity" check: greedy 256 tokens per prompt, mean chosen-token log-prob, pass bar 9.0 vs 8.19
                 (Quantization/LLM-Compressor/*/eval_qwen3_4b_*.py:30-81)
* ``convert-self-cognition``   JSONL -> Alpaca JSON with the per-language placeholder substitution
                 (Fine-Tuning/LLaMA-Factory/convert_self_cognition_to_alpaca.py:4-48)
* ``env``        device name, compute capability, CUDA / torch versions (HF_Basics/env_test.py)
* ``loadtest``   the Locust profiles against an OpenAI-compatible endpoint (Deployment/Ray/scripts/locustfile*.py)
* ``moderation`` the Llama-Guard moderation wrapper service (Deployment/litellm-proxy/llama-guard-wrapper/app.py)
* ``bench-serve`` the HTTP serving benchmark with ``vllm bench serve``'s flags and result keys (serve/bench_client.py;
                 LLM_on_Kubernetes/Inference_Platfrom/README.md:1361-1375)
* ``serve-deploy`` ``serve deploy <file>.yaml`` of the Ray Serve recipes: every application of the YAML at its route prefix
                 (Deployment/Ray/serve_deploy_examples/*.yaml; serve/deploy.py::deploy_yaml)
* ``preload``    mirror a model directory into the node-local cache behind a ``.ready`` marker
                 (LLM_on_Kubernetes/Inference_Platfrom/02-Preloader/model-preloader-daemonset.yaml:44-78)
* ``autoscale``  poll replicas' ``/metrics`` and print the replica count the KEDA ScaledObject would set
                 (…/05-KEDA-AutoScale/keda-scaledobject.yaml:12-53)
* ``canary``     run the canary steps and the Prometheus smoke analysis against a canary replica's ``/metrics``
                 (…/09-Canary-Deployment/vllm-rollout.yaml:22-33, analysis-template.yaml:8-53)
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch


def _calibration_rows(path, n):
    """JSON / JSONL rows with instruction / input / output (alpaca-gpt4-data-zh layout); synthetic text if no file."""
    rows = []
    if path and os.path.exists(path):
        with open(path, encoding="utf-8") as f:
            txt = f.read().strip()
        data = json.loads(txt) if txt.startswith("[") else [json.loads(ln) for ln in txt.splitlines() if ln.strip()]
        for r in data:
            instr = r.get("instruction") or r.get("prompt") or r.get("query") or ""
            outp = r.get("output") or r.get("completion") or r.get("response") or r.get("text") or ""
            t = (instr + "\n" + (r.get("input") or "") + "\n" + outp).strip() or " ".join(str(v) for v in r.values())
            rows.append(t)
    else:
        from ..data.datasets import synthetic_corpus
        rows = synthetic_corpus(max(8 * n, 1024), seed=1)
    return [r for r in rows if len(r.strip()) >= 16]


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    if argv and argv[0] in ("loadtest", "bench-serve"):       # pass-through commands: their own parsers see every flag
        if argv[0] == "loadtest":
            from ..serve.loadtest import main as lt
            return lt(argv[1:])
        from ..serve.bench_client import main as bench_main
        return bench_main(argv[1:])
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    sub = ap.add_subparsers(dest="cmd", required=True)
    m = sub.add_parser("merge")
    m.add_argument("--base", required=True)
    m.add_argument("--adapter", required=True)
    m.add_argument("--out", required=True)
    q = sub.add_parser("quantize")
    q.add_argument("--model", required=True)
    q.add_argument("--method", choices=["gptq", "awq", "rtn"], default="gptq")
    q.add_argument("--frontend", choices=["gptqmodel", "llmcompressor"], default="llmcompressor")
    q.add_argument("--scheme", default="W4A16")
    q.add_argument("--out", required=True)
    q.add_argument("--calib", default=None)
    q.add_argument("--num_calibration_samples", type=int, default=128)
    q.add_argument("--max_seq_length", type=int, default=2048)
    e = sub.add_parser("eval-ppl")
    e.add_argument("--model", required=True)
    e.add_argument("--calib", default=None)
    e.add_argument("--num_samples", type=int, default=128)
    e.add_argument("--max_tokens", type=int, default=256)
    c = sub.add_parser("convert-self-cognition")
    c.add_argument("--input", default=None)
    c.add_argument("--output", required=True)
    sub.add_parser("env")
    l = sub.add_parser("loadtest")
    l.add_argument("rest", nargs=argparse.REMAINDER)
    mod = sub.add_parser("moderation")
    mod.add_argument("--port", type=int, default=int(os.getenv("WRAPPER_PORT", "8099")))
    hc = sub.add_parser("healthcheck", help="node / cluster health check (single process, or every rank of a torchrun job)")
    hc.add_argument("--object-mib", type=int, default=16)
    ro = sub.add_parser("router", help="LiteLLM-style model router / load balancer (serve/router.py)")
    ro.add_argument("--config", required=True, help="YAML in the reference's litellm-config-router-lb.yaml shape")
    ro.add_argument("--port", type=int, default=4000)
    ro.add_argument("--master-key", default=os.getenv("LITELLM_MASTER_KEY"))
    sd = sub.add_parser("serve-deploy", help="serve a Ray Serve LLM YAML (serve/deploy.py)")
    sd.add_argument("config")
    sd.add_argument("--host", default="0.0.0.0")
    sd.add_argument("--port", type=int, default=8000)
    bs = sub.add_parser("bench-serve", help="HTTP serving benchmark (serve/bench_client.py)")
    bs.add_argument("rest", nargs=argparse.REMAINDER)
    pl = sub.add_parser("preload", help="model preloader (serve/platform.preload_model)")
    pl.add_argument("--source", required=True)
    pl.add_argument("--target", required=True)
    pl.add_argument("--require", nargs="*", default=["config.json", "tokenizer.json"])
    au = sub.add_parser("autoscale", help="KEDA scaling rule over /metrics endpoints (serve/platform.KedaScaler)")
    au.add_argument("--config", default=None, help="ScaledObject YAML (default: configs/keda_scaledobject.yaml)")
    au.add_argument("--metrics-url", nargs="+", required=True)
    au.add_argument("--replicas", type=int, default=1, help="current replica count")
    au.add_argument("--polls", type=int, default=0, help="stop after this many polls (0 = run until interrupted)")
    ca = sub.add_parser("canary", help="canary steps + smoke analysis (serve/platform.CanaryRollout)")
    ca.add_argument("--metrics-url", nargs="+", required=True, help="/metrics of the canary replicas (and of a GPU exporter)")
    ca.add_argument("--interval", type=float, default=60.0)
    a = ap.parse_args(argv)

    if a.cmd in ("autoscale", "canary"):
        import time
        import urllib.request
        from ..serve import platform as P
        scrape = lambda: [urllib.request.urlopen(u, timeout=5).read().decode() for u in a.metrics_url]
        if a.cmd == "autoscale":
            import yaml
            path = a.config or os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "configs", "keda_scaledobject.yaml")
            with open(path, encoding="utf-8") as f:
                k = P.KedaScaler.from_scaled_object(yaml.safe_load(f))
            n, i = a.replicas, 0
            while True:
                n = k.tick(n, scrape)
                print(json.dumps({"replicas": n, "fallback": k.in_fallback, **k.last_values}), flush=True)
                i += 1
                if a.polls and i >= a.polls:
                    return n
                time.sleep(k.polling_interval_s)
        ro_ = P.CanaryRollout()
        while True:
            ph = ro_.advance(lambda: scrape() + [P.gpu_memory_metrics_text()])
            print(json.dumps({"step": ro_.index, "weight": ro_.weight, "phase": ph}), flush=True)
            if ph != "Progressing":
                return ph
            time.sleep(a.interval)
    if a.cmd == "serve-deploy":
        from ..serve import deploy
        app = deploy.deploy_yaml(a.config)
        return None if app is None else deploy.run(app, a.host, a.port)
    if a.cmd == "bench-serve":
        from ..serve.bench_client import main as bench_main
        return bench_main(a.rest)
    if a.cmd == "preload":
        from ..serve.platform import preload_model
        rep = preload_model(a.source, a.target, a.require)
        print(json.dumps(rep))
        return rep
    if a.cmd == "env":
        print("torch", torch.__version__, "cuda", torch.version.cuda, "available", torch.cuda.is_available())
        if torch.cuda.is_available():
            for i in range(torch.cuda.device_count()):
                p = torch.cuda.get_device_properties(i)
                print(f"GPU {i}: {p.name}, sm_{p.major}{p.minor}, {p.total_memory / 2 ** 30:.0f} GiB, {p.multi_processor_count} SMs")
        from ..ops import _native as nv
        state = "NOT BUILT" if not nv.lib_path().exists() else ("loadable" if nv.available() else "built (no CUDA device here: CPU paths run)")
        print("libb200ops:", nv.lib_path(), state)
        return
    if a.cmd == "loadtest":
        from ..serve.loadtest import main as lt
        return lt(a.rest)
    if a.cmd == "moderation":
        import uvicorn
        from ..serve.moderation import build_moderation_app
        return uvicorn.run(build_moderation_app(), host="0.0.0.0", port=a.port, log_level="info")
    if a.cmd == "healthcheck":
        from ..serve.deploy import cluster_healthcheck, cluster_healthcheck_distributed
        if int(os.environ.get("WORLD_SIZE", "1")) > 1:
            rep = cluster_healthcheck_distributed(a.object_mib)
        else:
            rep = cluster_healthcheck(object_mib=a.object_mib)
        if rep is not None:
            print(json.dumps(rep, indent=1, default=str))
        return rep
    if a.cmd == "router":
        import uvicorn
        import yaml
        from ..serve.router import Router, build_router_app
        with open(a.config, encoding="utf-8") as f:
            cfg = yaml.safe_load(f)
        interval = float((cfg.get("router_settings") or {}).get("health_check_interval", 0))
        return uvicorn.run(build_router_app(Router.from_config(cfg), interval, a.master_key), host="0.0.0.0", port=a.port, log_level="info")
    if a.cmd == "convert-self-cognition":
        from ..data.chat import convert_self_cognition_to_alpaca, load_self_cognition
        rows = load_self_cognition(a.input, "{{NAME}}", "{{AUTHOR}}")
        with open(a.output, "w", encoding="utf-8") as f:
            json.dump(convert_self_cognition_to_alpaca(rows), f, ensure_ascii=False, indent=2)
        print(f"{len(rows)} rows -> {a.output}")
        return

    from ..io.hf import from_pretrained, save_pretrained
    from .infer import _load
    if a.cmd == "merge":
        has_adapter = any(os.path.exists(os.path.join(a.adapter, n)) for n in ("adapter_model.safetensors", "peft_model.safetensors", "adapter_model.bin"))
        if not has_adapter:                       # a full fine-tune: nothing to merge, just re-save
            model, tok = _load(a.adapter)
        else:
            model, tok = _load(a.base, a.adapter)
            model = model.merge_and_unload()
        save_pretrained(model, a.out, tokenizer=tok)
        print("merged checkpoint written to", a.out)
        return
    if a.cmd == "quantize":
        from ..quant import AWQModifier, GPTQModel, GPTQModifier, QuantizationModifier, QuantizeConfig, oneshot
        model, tok = _load(a.model)
        rows = _calibration_rows(a.calib, a.num_calibration_samples)[: a.num_calibration_samples]
        if a.frontend == "gptqmodel":
            gm = GPTQModel(model, QuantizeConfig(bits=4, group_size=128, desc_act=False, method=a.method), tok)
            gm.quantize(rows, batch_size=1, max_seq_length=a.max_seq_length, progress=True)
            gm.save(a.out)
        else:
            Mod = {"gptq": GPTQModifier, "awq": AWQModifier, "rtn": QuantizationModifier}[a.method]
            oneshot(model=model, recipe=[Mod(targets="Linear", scheme=a.scheme, ignore=["lm_head"])], tokenizer=tok,
                    dataset=[{"text": t} for t in rows], output_dir=a.out, max_seq_length=a.max_seq_length,
                    num_calibration_samples=len(rows), progress=True)
        if hasattr(tok, "save_pretrained"):
            tok.save_pretrained(a.out)
        print("quantised checkpoint written to", a.out)
        return
    if a.cmd == "eval-ppl":
        from ..data.chat import build_messages, render_chatml
        from ..inference import Generator
        from ..quant import self_perplexity
        model, tok = _load(a.model)
        rows = _calibration_rows(a.calib, 2 * a.num_samples)[a.num_samples: 2 * a.num_samples]      # held-out rows [128:256]
        dev = model.embed_tokens.device
        prompts = [torch.tensor([tok.encode(render_chatml(build_messages(r.split("\n")[0]), add_generation_prompt=True, space_before_end=False))],
                                device=dev) for r in rows]
        eos = getattr(tok, "eos_token_id", None)
        mean, std = self_perplexity(Generator(model, tok, 4096), prompts, a.max_tokens, [eos] if eos is not None else None)
        print(f"self perplexity: {mean:.3f} ± {std:.3f}   (reference bar: unquantised ≈ 8.19, pass < 9.0)")
        return mean


if __name__ == "__main__":
    main()
