"""Layer-by-layer calibration driver, quantised checkpoint IO (GPTQ and compressed-tensors layouts, one tensor set
per HF sub-projection so vLLM / transformers can read them), ``GPTQModel`` and ``oneshot`` front-ends, and the
perplexity evaluations of the lesson (Quantization/LLM-Compressor/*/eval_qwen3_4b_*.py:30-62)."""
from __future__ import annotations

import json
import math
import os
from dataclasses import asdict
from typing import Dict, Iterable, List, Optional, Sequence

import torch

from ..io.hf import _Shards, config_from_hf, config_to_hf, from_pretrained, hf_names
from ..io.safetensors import save_file
from ..models.qwen3 import Qwen3ForCausalLM
from ..ops import elementwise as ew
from .algorithms import ActivationStats, gptq_quantize
from .api import AWQModifier, GPTQModifier, QuantizeConfig, _awq_fold_layer, _Modifier, tokenize_calibration
from .formats import (W4Weight, from_awq, from_compressed_tensors, from_gptq, pack_codes, quantize_rtn, to_awq,
                      to_compressed_tensors, to_gptq)

_PARTS = (("qkv_proj", ("q", "k", "v")), ("o_proj", ("o",)), ("gate_up_proj", ("gate", "up")), ("down_proj", ("down",)))


@torch.no_grad()
def quantize_model(model: Qwen3ForCausalLM, batches: Sequence[torch.Tensor], qc: QuantizeConfig,
                   progress: bool = False) -> Qwen3ForCausalLM:
    """In-place W4A16 quantisation of every decoder projection (lm_head / embeddings / norms stay 16-bit)."""
    was = model.training
    model.eval()
    dev = model.embed_tokens.device
    states = []
    for ids in batches:
        ids = ids.to(dev)
        B, S = ids.shape
        pos = torch.arange(S, device=dev, dtype=torch.int32).repeat(B)
        states.append([ew.embedding(ids.reshape(-1), model.embed_tokens), None, pos, B, S])
    smax = max((s[4] for s in states), default=1)
    cos, sin = model.rope(dev, smax)
    if qc.method != "rtn" and not states:
        raise ValueError(f"{qc.method} needs calibration data: every row was empty or shorter than 2 tokens")
    need_h = qc.method == "gptq"
    for li, layer in enumerate(model.layers):
        lins = layer.fused_linears()
        if qc.method != "rtn":
            stats = {k: ActivationStats(l.in_features, dev, need_h) for k, l in lins.items()}
            hooks = [l.register_forward_pre_hook(lambda mod, a, s=stats[k]: s.add(a[0])) for k, l in lins.items()]
            for st in states:
                layer(st[0], st[1], cos, sin, st[2], st[3], st[4])
            for h in hooks:
                h.remove()
            if qc.method == "awq":
                _awq_fold_layer(layer, stats, qc)
        for k, lin in lins.items():
            if lin.weight is None:
                continue                                  # already quantised
            w = lin.weight.data
            if qc.method == "gptq":
                q = gptq_quantize(w, stats[k].H, qc.group_size, qc.sym, qc.damp_percent)
            else:
                q = quantize_rtn(w, qc.group_size, qc.sym)
            lin.set_nf4(q)
        for st in states:
            st[0], st[1] = layer(st[0], st[1], cos, sin, st[2], st[3], st[4])
        if progress:
            print(f"[quant:{qc.method}] layer {li + 1}/{len(model.layers)}", flush=True)
    model.quant_config = qc
    model.train(was)
    return model


# ------------------------------------------------------------------------------------------------
# checkpoint IO
# ------------------------------------------------------------------------------------------------
def _split_rows(q: W4Weight, lo: int, hi: int) -> W4Weight:
    kp, kg = q.K // 2, q.K // q.group
    return W4Weight(q.packed.view(q.N, kp)[lo:hi].reshape(-1).contiguous(), q.scale.view(q.N, kg)[lo:hi].reshape(-1).contiguous(),
                    None if q.zs is None else q.zs.view(q.N, kg)[lo:hi].reshape(-1).contiguous(), hi - lo, q.K, q.group, q.lut)


def _cat_rows(parts: List[W4Weight]) -> W4Weight:
    p0 = parts[0]
    zs = None if p0.zs is None else torch.cat([p.zs for p in parts])
    return W4Weight(torch.cat([p.packed for p in parts]), torch.cat([p.scale for p in parts]), zs,
                    sum(p.N for p in parts), p0.K, p0.group, p0.lut)


def save_quantized(model: Qwen3ForCausalLM, out_dir: str, fmt: str = "gptq", tokenizer=None) -> None:
    """fmt: ``gptq`` (qweight/qzeros/scales/g_idx + quantize_config.json), ``compressed-tensors``
    (weight_packed/weight_scale/weight_zero_point/weight_shape + quantization_config in config.json) or ``awq`` (AutoAWQ
    gemm layout: qweight/qzeros/scales, the form of the published ``*-AWQ`` checkpoints)."""
    if fmt not in ("gptq", "compressed-tensors", "awq"):
        raise ValueError(f"unknown checkpoint format {fmt!r}")
    qc: QuantizeConfig = getattr(model, "quant_config", QuantizeConfig())
    os.makedirs(out_dir, exist_ok=True)
    dt = model.embed_tokens.dtype
    sd = {"model.embed_tokens.weight": model.embed_tokens.detach(), "model.norm.weight": model.norm.detach()}
    if model.lm_head is not None:
        sd["lm_head.weight"] = model.lm_head.detach()
    for i, layer in enumerate(model.layers):
        n = hf_names(i)
        sd[n["ln1"]], sd[n["ln2"]] = layer.input_layernorm.detach(), layer.post_attention_layernorm.detach()
        if layer.q_norm is not None:
            sd[n["qn"]], sd[n["kn"]] = layer.q_norm.detach(), layer.k_norm.detach()
        lins = layer.fused_linears()
        for key, parts in _PARTS:
            lin = lins[key]
            off = 0
            for p, size in zip(parts, lin.part_sizes):
                if isinstance(lin.nf4, W4Weight):
                    sub = _split_rows(lin.nf4, off, off + size)
                    tensors = to_gptq(sub) if fmt == "gptq" else to_awq(sub) if fmt == "awq" else to_compressed_tensors(sub, qc.sym)
                    for tn, tv in tensors.items():
                        sd[f"{n[p]}.{tn}"] = tv
                else:
                    sd[n[p] + ".weight"] = lin.dense_weight()[off:off + size]
                if lin.bias is not None:
                    sd[n[p] + ".bias"] = lin.bias.detach()[off:off + size]
                off += size
    save_file({k: v.cpu().contiguous() for k, v in sd.items()}, os.path.join(out_dir, "model.safetensors"), {"format": "pt"})
    cfg = config_to_hf(model.config, dt)
    if fmt == "gptq":
        qd = {"bits": 4, "group_size": qc.group_size, "desc_act": False, "sym": qc.sym, "damp_percent": qc.damp_percent,
              "quant_method": "gptq", "checkpoint_format": "gptq", "lm_head": False}
        cfg["quantization_config"] = qd
        with open(os.path.join(out_dir, "quantize_config.json"), "w") as f:
            json.dump(qd, f, indent=2)
    elif fmt == "awq":
        cfg["quantization_config"] = {"quant_method": "awq", "bits": 4, "group_size": qc.group_size, "zero_point": True,
                                      "version": "gemm", "modules_to_not_convert": None}
    else:
        cfg["quantization_config"] = {
            "quant_method": "compressed-tensors", "format": "pack-quantized", "ignore": ["lm_head"],
            "config_groups": {"group_0": {"targets": ["Linear"], "weights": {
                "num_bits": 4, "type": "int", "symmetric": qc.sym, "strategy": "group", "group_size": qc.group_size}}}}
    with open(os.path.join(out_dir, "config.json"), "w") as f:
        json.dump(cfg, f, indent=2)
    if tokenizer is not None and hasattr(tokenizer, "save_pretrained"):
        tokenizer.save_pretrained(out_dir)


def detect_quant_format(path: str) -> Optional[str]:
    with open(os.path.join(path, "config.json")) as f:
        q = json.load(f).get("quantization_config")
    if q is None:
        return "gptq" if os.path.exists(os.path.join(path, "quantize_config.json")) else None
    return {"gptq": "gptq", "compressed-tensors": "compressed-tensors", "awq": "awq"}.get(q.get("quant_method"))


@torch.no_grad()
def load_quantized(path: str, dtype=torch.bfloat16, device=None) -> Qwen3ForCausalLM:
    """Read a GPTQ, AutoAWQ or compressed-tensors W4A16 checkpoint; the sub-projections of each fused linear are re-joined
    into one kernel-layout ``W4Weight`` (row concatenation — scales are per output row, so nothing is requantised)."""
    fmt = detect_quant_format(path)
    if fmt is None:
        return from_pretrained(path, dtype=dtype, device=device)
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
    with open(os.path.join(path, "config.json")) as f:
        raw = json.load(f)
    qd = raw["quantization_config"] if "quantization_config" in raw else json.load(open(os.path.join(path, "quantize_config.json")))
    if fmt == "gptq":
        group, sym = qd.get("group_size", 128), qd.get("sym", True)
        if qd.get("desc_act"):
            raise NotImplementedError("desc_act=True checkpoints are not supported")
    elif fmt == "awq":
        group, sym = qd.get("group_size", qd.get("q_group_size", 128)), not qd.get("zero_point", True)
        if int(qd.get("bits", qd.get("w_bit", 4))) != 4 or str(qd.get("version", "gemm")).lower() != "gemm":
            raise NotImplementedError("AutoAWQ checkpoints are read in the 4-bit 'gemm' layout only")
    else:
        wq = next(iter(qd["config_groups"].values()))["weights"]
        group, sym = wq.get("group_size", 128), wq.get("symmetric", True)
    model = Qwen3ForCausalLM(config_from_hf(raw), dtype=dtype, device=device)
    sh = _Shards(path)
    put = lambda dst, name: dst.copy_(sh.get(name).to(device, dtype))  # noqa: E731
    put(model.embed_tokens, "model.embed_tokens.weight")
    put(model.norm, "model.norm.weight")
    if model.lm_head is not None:
        put(model.lm_head, "lm_head.weight" if sh.has("lm_head.weight") else "model.embed_tokens.weight")
    keys = {"gptq": ("qweight", "qzeros", "scales", "g_idx"), "awq": ("qweight", "qzeros", "scales")}.get(
        fmt, ("weight_packed", "weight_scale", "weight_zero_point", "weight_shape"))
    read = {"gptq": from_gptq, "awq": from_awq}.get(fmt, from_compressed_tensors)
    for i, layer in enumerate(model.layers):
        n = hf_names(i)
        put(layer.input_layernorm, n["ln1"])
        put(layer.post_attention_layernorm, n["ln2"])
        if layer.q_norm is not None:
            put(layer.q_norm, n["qn"])
            put(layer.k_norm, n["kn"])
        lins = layer.fused_linears()
        for key, parts in _PARTS:
            lin = lins[key]
            subs = []
            for p in parts:
                if sh.has(f"{n[p]}.{keys[0]}"):
                    t = {k: sh.get(f"{n[p]}.{k}").to(device) for k in keys if sh.has(f"{n[p]}.{k}")}
                    subs.append(read(t, group))
                else:                                     # an ignored (dense) projection inside a fused group
                    subs.append(sh.get(n[p] + ".weight").to(device, dtype))
            if all(isinstance(s, W4Weight) for s in subs):
                lin.set_nf4(_cat_rows(subs))
            else:
                dense = [s.dequantize(dtype) if isinstance(s, W4Weight) else s for s in subs]
                lin.weight.data = torch.cat(dense, 0)
            if lin.bias is not None:
                lin.bias.copy_(torch.cat([sh.get(n[p] + ".bias").to(device, dtype) for p in parts], 0))
    model.quant_config = QuantizeConfig(4, group, False, sym)
    model.name_or_path = path
    return model
