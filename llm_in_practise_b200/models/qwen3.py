"""Qwen3 (dense) decoder on the sm_100a kernels — the model family behind every Fine-Tuning/*.py and
Quantization/*.py script of the reference (loaded there with ``AutoModelForCausalLM``).

Architecture (checked against transformers' Qwen3Attention / Qwen3MLP / Qwen3RMSNorm): GQA with
bias-free q/k/v/o projections, per-head RMSNorm on q and k *before* RoPE (rotate-half layout),
SwiGLU MLP ``down(silu(gate(x)) * up(x))``, RMSNorm eps 1e-6, head_dim 128, vocab 151 936.

B200-first layout: q|k|v and gate|up are single fused projections (one tcgen05 GEMM each, NF4 or
bf16 base, LoRA folded in as an extra k-block); the residual add is fused into the following
RMSNorm; q/k-norm + RoPE is one kernel; attention reads q, k, v in place from the fused projection;
lm_head + cross-entropy is fused and chunked so no fp32 [tokens, vocab] tensor exists.
HF parameter names are restored on save/load (``io/hf.py``).
"""
from __future__ import annotations

from dataclasses import dataclass, asdict
from typing import Dict, Optional, Tuple

import torch
from torch import nn
from torch.utils.checkpoint import checkpoint

from ..ops import elementwise as ew
from ..ops.attention import attention
from ..ops.linear import FusedLinear
from ..ops.loss import linear_cross_entropy, micro_batch_row_scale
from ..ops.norm import rms_norm
from ..ops.gemm import gemm


@dataclass
class Qwen3Config:
    vocab_size: int = 151936
    hidden_size: int = 4096
    intermediate_size: int = 12288
    num_hidden_layers: int = 36
    num_attention_heads: int = 32
    num_key_value_heads: int = 8
    head_dim: int = 128
    rms_norm_eps: float = 1e-6
    rope_theta: float = 1000000.0
    rope_scaling: Optional[dict] = None      # HF ``rope_scaling`` / ``rope_parameters`` (linear, yarn, llama3); None = plain RoPE
    max_position_embeddings: int = 40960
    tie_word_embeddings: bool = False
    model_type: str = "qwen3"
    attention_bias: bool = False   # Qwen2-family checkpoints (DeepSeek-R1-Distill-Qwen-1.5B) have q/k/v biases ...
    qk_norm: bool = True           # ... and no per-head q/k RMSNorm

    def to_dict(self):
        return asdict(self)

    @classmethod
    def from_dict(cls, d: dict) -> "Qwen3Config":
        keys = cls.__dataclass_fields__.keys()
        return cls(**{k: v for k, v in d.items() if k in keys})


# Public checkpoint shapes (recalled, re-read config.json when real weights are used).
QWEN3_PRESETS: Dict[str, Qwen3Config] = {
    "qwen3-14b": Qwen3Config(hidden_size=5120, intermediate_size=17408, num_hidden_layers=40,
                             num_attention_heads=40, num_key_value_heads=8),
    "qwen3-8b": Qwen3Config(hidden_size=4096, intermediate_size=12288, num_hidden_layers=36,
                            num_attention_heads=32, num_key_value_heads=8),
    "qwen3-4b": Qwen3Config(hidden_size=2560, intermediate_size=9728, num_hidden_layers=36,
                            num_attention_heads=32, num_key_value_heads=8, tie_word_embeddings=True),
    "qwen3-1.7b": Qwen3Config(hidden_size=2048, intermediate_size=6144, num_hidden_layers=28,
                              num_attention_heads=16, num_key_value_heads=8, tie_word_embeddings=True),
    "qwen3-tiny": Qwen3Config(vocab_size=1024, hidden_size=256, intermediate_size=512, num_hidden_layers=2,
                              num_attention_heads=4, num_key_value_heads=2, head_dim=64,
                              max_position_embeddings=512),
}


class Qwen3DecoderLayer(nn.Module):
    def __init__(self, cfg: Qwen3Config, idx: int, dtype, device):
        super().__init__()
        h, D = cfg.hidden_size, cfg.head_dim
        H, Hkv, F = cfg.num_attention_heads, cfg.num_key_value_heads, cfg.intermediate_size
        self.cfg = cfg
        self.idx = idx
        self.input_layernorm = nn.Parameter(torch.ones(h, dtype=dtype, device=device))
        self.post_attention_layernorm = nn.Parameter(torch.ones(h, dtype=dtype, device=device))
        self.q_norm = nn.Parameter(torch.ones(D, dtype=dtype, device=device)) if cfg.qk_norm else None
        self.k_norm = nn.Parameter(torch.ones(D, dtype=dtype, device=device)) if cfg.qk_norm else None
        self.qkv_proj = FusedLinear(h, [("q_proj", H * D), ("k_proj", Hkv * D), ("v_proj", Hkv * D)], bias=cfg.attention_bias,
                                    dtype=dtype, device=device)
        self.o_proj = FusedLinear(H * D, [("o_proj", h)], dtype=dtype, device=device)
        self.gate_up_proj = FusedLinear(h, [("gate_proj", F), ("up_proj", F)], dtype=dtype, device=device)
        self.down_proj = FusedLinear(F, [("down_proj", h)], dtype=dtype, device=device)
        for j, lin in enumerate((self.qkv_proj, self.o_proj, self.gate_up_proj, self.down_proj)):
            lin.layer_uid = idx * 4 + j

    def fused_linears(self):
        return {"qkv_proj": self.qkv_proj, "o_proj": self.o_proj, "gate_up_proj": self.gate_up_proj,
                "down_proj": self.down_proj}

    def forward(self, delta, residual, cos, sin, pos, B, S):
        cfg = self.cfg
        x, residual = rms_norm(delta, self.input_layernorm, cfg.rms_norm_eps, residual)
        qkv = self.qkv_proj(x)
        qkv = ew.qknorm_rope(qkv, self.q_norm, self.k_norm, cos, sin, pos, cfg.num_attention_heads,
                             cfg.num_key_value_heads, cfg.head_dim, False, cfg.rms_norm_eps)
        ctx = attention(qkv, B, S, cfg.num_attention_heads, cfg.num_key_value_heads, cfg.head_dim, causal=True)
        a = self.o_proj(ctx)
        x, residual = rms_norm(a, self.post_attention_layernorm, cfg.rms_norm_eps, residual)
        gu = self.gate_up_proj(x)
        hmid = ew.swiglu(gu)
        return self.down_proj(hmid), residual


class Qwen3ForCausalLM(nn.Module):
    """``forward(input_ids, labels=None)`` -> dict(loss=..., logits=... (only when labels is None))."""

    def __init__(self, cfg: Qwen3Config, dtype=torch.bfloat16, device=None, init_std: float = 0.02):
        super().__init__()
        self.config = cfg
        self.dtype_ = dtype
        self.embed_tokens = nn.Parameter(torch.empty(cfg.vocab_size, cfg.hidden_size, dtype=dtype, device=device))
        nn.init.normal_(self.embed_tokens, std=init_std)
        self.layers = nn.ModuleList([Qwen3DecoderLayer(cfg, i, dtype, device) for i in range(cfg.num_hidden_layers)])
        self.norm = nn.Parameter(torch.ones(cfg.hidden_size, dtype=dtype, device=device))
        if cfg.tie_word_embeddings:
            self.lm_head = None
        else:
            self.lm_head = nn.Parameter(torch.empty(cfg.vocab_size, cfg.hidden_size, dtype=dtype, device=device))
            nn.init.normal_(self.lm_head, std=init_std)
        self.gradient_checkpointing = False
        self._rope: Optional[Tuple[torch.Tensor, torch.Tensor]] = None
        self.ce_chunk_tokens = 4096

    # -------------------------------------------------------------------------------------
    def head_weight(self) -> torch.Tensor:
        return self.embed_tokens if self.lm_head is None else self.lm_head

    def gradient_checkpointing_enable(self, **_):
        self.gradient_checkpointing = True

    def gradient_checkpointing_disable(self):
        self.gradient_checkpointing = False

    # ---- the rest of the transformers call surface the reference's scripts use on a model -------------------------------
    @property
    def device(self) -> torch.device:
        return self.embed_tokens.device

    def enable_input_require_grads(self) -> None:
        """transformers needs this so that checkpointed blocks of a frozen base still get a gradient path to the adapters
        (Fine-Tuning/qwen3-8b-qlora.py calls it after ``prepare_model_for_kbit_training``); the fused LoRA linears here take
        their gradients from their own autograd functions, so there is nothing to switch on."""
        self.input_require_grads = True

    def save_pretrained(self, save_directory: str, **kw) -> None:
        from ..io.hf import save_pretrained
        shard = kw.get("max_shard_size")
        if isinstance(shard, str):                       # "5GB", "500MB" as transformers spells it
            units = {"KB": 10 ** 3, "MB": 10 ** 6, "GB": 10 ** 9, "KIB": 2 ** 10, "MIB": 2 ** 20, "GIB": 2 ** 30}
            u = next((s for s in sorted(units, key=len, reverse=True) if shard.upper().endswith(s)), None)
            shard = int(float(shard[:-len(u)]) * units[u]) if u else int(shard)
        save_pretrained(self, save_directory, **({"max_shard_size": shard} if shard else {}), tokenizer=kw.get("tokenizer"))

    @torch.no_grad()
    def generate(self, input_ids: Optional[torch.Tensor] = None, attention_mask: Optional[torch.Tensor] = None,
                 max_length: Optional[int] = None, streamer=None, **kw) -> torch.Tensor:
        """``model.generate(**tokenizer(text, return_tensors="pt"), max_new_tokens=…, do_sample=…, temperature=…, top_p=…)``
        as the inference scripts call it (Scripts/inference/01-*.py:20-30, Fine-Tuning/inferences.py:60-75): runs
        ``inference.Generator`` (KV cache, CUDA-graph decode) and returns prompt + continuation ids. Padded batches are not
        taken (the scripts generate one prompt at a time)."""
        from ..inference.generate import Generator
        if input_ids is None:
            input_ids = kw.pop("inputs")
        if attention_mask is not None and not bool(attention_mask.bool().all()):
            raise NotImplementedError("generate() takes unpadded prompts (attention_mask must be all ones)")
        if max_length is not None and "max_new_tokens" not in kw:
            kw["max_new_tokens"] = max(1, int(max_length) - input_ids.shape[1])
        need = input_ids.shape[1] + int(kw.get("max_new_tokens", 256))
        g = self.__dict__.get("_generator")
        if g is None or g.max_len < need:
            g = Generator(self, None, max(need, min(self.config.max_position_embeddings, 4096)))
            self.__dict__["_generator"] = g
        return g.generate(input_ids.to(self.device), streamer=streamer, **kw)

    def rope(self, device, need: int):
        if self._rope is None or self._rope[0].device != device or self._rope[0].shape[0] < need:
            n = max(need, min(self.config.max_position_embeddings, 8192))
            self._rope = ew.rope_tables(n, self.config.head_dim, self.config.rope_theta, device, self.config.rope_scaling,
                                        self.config.max_position_embeddings)
        return self._rope

    def quantize_nf4(self, double_quant: bool = True) -> "Qwen3ForCausalLM":
        """QLoRA base: every decoder projection becomes NF4; embeddings, norms and lm_head stay bf16
        (``BitsAndBytesConfig(load_in_4bit, nf4, double_quant)`` semantics)."""
        for layer in self.layers:
            for lin in layer.fused_linears().values():
                if lin.nf4 is None:
                    lin.quantize_nf4(double_quant)
        return self

    def hidden_states(self, input_ids: torch.Tensor, position_ids: Optional[torch.Tensor] = None) -> torch.Tensor:
        B, S = input_ids.shape
        dev = input_ids.device
        if position_ids is None:
            pos = torch.arange(S, device=dev, dtype=torch.int32).repeat(B)
        else:
            pos = position_ids.reshape(-1).to(torch.int32)
        cos, sin = self.rope(dev, S if position_ids is None else int(pos.max().item()) + 1)
        delta = ew.embedding(input_ids.reshape(-1), self.embed_tokens)
        residual = None
        for layer in self.layers:
            if self.gradient_checkpointing and self.training and torch.is_grad_enabled():
                delta, residual = checkpoint(layer, delta, residual, cos, sin, pos, B, S, use_reentrant=False)
            else:
                delta, residual = layer(delta, residual, cos, sin, pos, B, S)
        hidden, _ = rms_norm(delta, self.norm, self.config.rms_norm_eps, residual)
        return hidden  # [B*S, h]

    def forward(self, input_ids: torch.Tensor, labels: Optional[torch.Tensor] = None,
                position_ids: Optional[torch.Tensor] = None, attention_mask=None, n_valid: Optional[int] = None,
                micro_batches: int = 0, **_) -> dict:
        """``micro_batches`` = G > 0: the batch is G gradient-accumulation micro-batches stacked along dim 0; the loss
        is the mean over micro-batches of each one's mean token loss (identical to G separate passes with loss / G)
        and the normalisation is computed on the device (no host read-back)."""
        B, S = input_ids.shape
        hidden = self.hidden_states(input_ids, position_ids)
        if labels is None:
            logits = gemm(hidden, self.head_weight())
            return {"logits": logits.view(B, S, -1)}
        # HF causal-LM loss: predict token t+1 from position t; ignore_index = -100
        targets = torch.full_like(labels, -100)
        targets[:, :-1] = labels[:, 1:]
        rs = micro_batch_row_scale(targets, micro_batches) if micro_batches > 0 else None
        loss = linear_cross_entropy(hidden, self.head_weight(), targets.reshape(-1), -100, n_valid,
                                    self.ce_chunk_tokens, row_scale=rs)
        return {"loss": loss}

    # -------------------------------------------------------------------------------------
    def num_parameters(self, only_trainable: bool = False) -> int:
        n = 0
        for p in self.parameters():
            if only_trainable and not p.requires_grad:
                continue
            n += p.numel()
        if not only_trainable:
            for layer in self.layers:
                for lin in layer.fused_linears().values():
                    if lin.nf4 is not None:
                        n += lin.out_features * lin.in_features
        return n


def flops_per_token(cfg: Qwen3Config, seq_len: int, lora: bool = True) -> float:
    """Model FLOPs per trained token (forward + backward; frozen base => no weight gradients)."""
    h, D, H, Hkv, F, L = (cfg.hidden_size, cfg.head_dim, cfg.num_attention_heads, cfg.num_key_value_heads,
                          cfg.intermediate_size, cfg.num_hidden_layers)
    lin = h * (H + 2 * Hkv) * D + H * D * h + 3 * h * F
    per_layer = 2 * lin + 2 * 2 * H * D * seq_len / 2  # causal attention: half the square
    head = 2 * h * cfg.vocab_size
    fwd = L * per_layer + head
    return fwd * (2.0 if lora else 3.0)
