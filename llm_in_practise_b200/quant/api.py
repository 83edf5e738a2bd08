"""Model-level W4A16 post-training quantisation with the front-ends the reference's lesson uses.

* ``QuantizeConfig`` / ``GPTQModel.load(path, quantize_config).quantize(texts, batch_size).save(dir)`` and
  ``GPTQModel.load(dir)`` -> ``generate`` (Quantization/GPTQModel/quantize_qwen3_4b_gptq.py:16-46,
  inference_qwen3_4b_gptq.py:18-68).
* ``oneshot(model, recipe=[GPTQModifier | AWQModifier(targets="Linear", scheme="W4A16" | "W4A16_SYM",
  ignore=["lm_head"])], dataset, tokenizer, output_dir, max_seq_length, num_calibration_samples)``
  (Quantization/LLM-Compressor/{GPTQ,AWQ}/quantize_qwen3_4b_*.py, Quantization/LoRA-AWQ/quantize-…-awq.py).

Calibration walks the decoder one layer at a time: the hidden states of every calibration batch are kept on the
device, the four fused projections of the layer are hooked (Hessian X^T X through the tcgen05 tf32 GEMM for GPTQ,
mean |x| for AWQ), quantised, and the layer is re-run with its 4-bit weights to produce the next layer's inputs —
peak extra memory is one layer's Hessians (1.6 GB for Qwen3-14B), not the model's.
"""
from __future__ import annotations

import json
import os
from dataclasses import asdict, dataclass, field
from typing import Dict, Iterable, List, Optional, Sequence, Union

import torch

from ..io.hf import _Shards, config_from_hf, config_to_hf, from_pretrained, hf_names
from ..io.safetensors import save_file
from ..models.qwen3 import Qwen3ForCausalLM
from ..ops import elementwise as ew
from ..ops.linear import FusedLinear
from .algorithms import ActivationStats, awq_search_scale, gptq_quantize
from .formats import (W4Weight, from_awq, from_compressed_tensors, from_gptq, pack_codes, quantize_rtn, to_compressed_tensors,
                      to_gptq)


@dataclass
class QuantizeConfig:
    bits: int = 4
    group_size: int = 128
    desc_act: bool = False
    sym: bool = True
    damp_percent: float = 0.01
    method: str = "gptq"            # gptq | awq | rtn

    def __post_init__(self):
        if self.bits != 4:
            raise NotImplementedError("the reference quantises to 4 bits only (W4A16)")
        if self.desc_act:
            raise NotImplementedError("desc_act=True is not used by the reference (it passes desc_act=False)")


@dataclass
class _Modifier:
    targets: Union[str, Sequence[str]] = "Linear"
    scheme: str = "W4A16"
    ignore: Sequence[str] = field(default_factory=lambda: ["lm_head"])
    group_size: int = 128
    dampening_frac: float = 0.01
    method = "rtn"

    def config(self) -> QuantizeConfig:
        if self.scheme not in ("W4A16", "W4A16_SYM", "W4A16_ASYM"):
            raise NotImplementedError(f"scheme {self.scheme}")
        # llm-compressor: the preset W4A16 is symmetric for GPTQ; the AWQ lesson uses it as the asymmetric one
        sym = self.scheme == "W4A16_SYM" or (self.scheme == "W4A16" and self.method == "gptq")
        return QuantizeConfig(4, self.group_size, False, sym, self.dampening_frac, self.method)


@dataclass
class GPTQModifier(_Modifier):
    method = "gptq"


@dataclass
class AWQModifier(_Modifier):
    method = "awq"


@dataclass
class QuantizationModifier(_Modifier):
    method = "rtn"


# ------------------------------------------------------------------------------------------------
# calibration data
# ------------------------------------------------------------------------------------------------
def tokenize_calibration(dataset, tokenizer, max_seq_length: int = 2048, num_samples: Optional[int] = None,
                         batch_size: int = 1, device=None) -> List[torch.Tensor]:
    """Strings, ``{"text": ...}`` rows or pre-tokenised id lists -> list of [b, S] id tensors (batches hold
    equal-length rows; with batch_size 1 nothing is padded, as in the reference's ``batch_size=1``)."""
    rows = []
    for i, r in enumerate(dataset):
        if num_samples is not None and i >= num_samples:
            break
        if isinstance(r, dict):
            r = r.get("input_ids", r.get("text"))
        if isinstance(r, str):
            r = tokenizer.encode(r)
        r = list(r)[:max_seq_length]
        if len(r) >= 2:
            rows.append(r)
    out = []
    if batch_size == 1:
        return [torch.tensor([r], device=device) for r in rows]
    rows.sort(key=len)
    for i in range(0, len(rows), batch_size):
        chunk = rows[i:i + batch_size]
        L = min(len(r) for r in chunk)
        out.append(torch.tensor([r[:L] for r in chunk], device=device))
    return out


# ------------------------------------------------------------------------------------------------
# AWQ scale folding for one Qwen decoder layer
# ------------------------------------------------------------------------------------------------
@torch.no_grad()
def _awq_fold_layer(layer, stats: Dict[str, ActivationStats], qc: QuantizeConfig) -> None:
    cfg = layer.cfg
    H, Hkv, D = cfg.num_attention_heads, cfg.num_key_value_heads, cfg.head_dim
    g = H // Hkv
    mean = lambda s: s.abs_mean / max(s.n, 1)  # noqa: E731

    # o_proj <- v_proj rows; the scale has to be shared by the g query heads reading one kv head
    st = stats["o_proj"]
    am = mean(st).view(Hkv, g, D).mean(1).reshape(-1)                       # [Hkv*D]
    expand = lambda s: s.view(Hkv, 1, D).expand(Hkv, g, D).reshape(-1)      # noqa: E731
    wo = layer.o_proj.weight.data
    s_kv = _search_shared(wo, am, st.sample, qc, expand)
    s = expand(s_kv)
    wo.mul_(s.to(wo.dtype))
    v0, v1 = layer.qkv_proj.part_range("v_proj")
    layer.qkv_proj.weight.data[v0:v1].div_(s_kv.to(wo.dtype).unsqueeze(1))
    if layer.qkv_proj.bias is not None:
        layer.qkv_proj.bias.data[v0:v1].div_(s_kv.to(wo.dtype))

    # down_proj <- up_proj rows
    st = stats["down_proj"]
    wd = layer.down_proj.weight.data
    s = awq_search_scale([wd], mean(st), st.sample, qc.group_size, qc.sym)
    wd.mul_(s.to(wd.dtype))
    u0, u1 = layer.gate_up_proj.part_range("up_proj")
    layer.gate_up_proj.weight.data[u0:u1].div_(s.to(wd.dtype).unsqueeze(1))

    # qkv_proj <- input_layernorm, gate_up_proj <- post_attention_layernorm
    for lin, norm, key in ((layer.qkv_proj, layer.input_layernorm, "qkv_proj"),
                           (layer.gate_up_proj, layer.post_attention_layernorm, "gate_up_proj")):
        st = stats[key]
        w = lin.weight.data
        s = awq_search_scale([w], mean(st), st.sample, qc.group_size, qc.sym)
        w.mul_(s.to(w.dtype))
        norm.data.div_(s.to(norm.dtype))


def _search_shared(w, am_kv, x, qc: QuantizeConfig, expand, n_grid: int = 20):
    """AWQ grid search where the per-channel scale is tied across GQA groups."""
    Wf = w.float()
    ref = x @ Wf.t()
    am = am_kv.clamp_min(1e-8)
    best, best_s = None, torch.ones_like(am)
    for i in range(n_grid):
        s_kv = am.pow(i / n_grid)
        s_kv = s_kv / (s_kv.max() * s_kv.min()).sqrt()
        s = expand(s_kv)
        wq = quantize_rtn(Wf * s, qc.group_size, qc.sym).dequantize(torch.float32) / s
        err = ((x @ wq.t()) - ref).pow(2).mean().item()
        if best is None or err < best:
            best, best_s = err, s_kv
    return best_s
