"""Pipeline (layer) parallelism for inference, optionally combined with tensor parallelism inside every stage.

The reference asks for it in one place: ``engine_kwargs.pipeline_parallel_size: 2`` of the Ray Serve recipe
(Deployment/Ray/serve_deploy_examples/qwen3_app_pipeline_parallel.yaml:21-31), which makes vLLM put half of the decoder
layers on each of two Ray workers. Here the layout is SPMD like the tensor-parallel engine (serve/tp_engine.py): one
process per GPU, ``world = stages x tp``, rank ``= stage * tp + tp_rank``.

* ``shard_qwen3_pipeline`` keeps a contiguous slice of the decoder layers on every stage (balanced, earlier stages take the
  remainder), drops the rest, drops the lm_head on every stage but the last and shrinks ``config.num_hidden_layers`` to
  the local count, so the KV cache a ``Generator`` / ``LLMEngine`` allocates only covers local layers. With ``tp > 1`` the
  stage's layers are then sharded Megatron-style over the ranks of the stage (``parallel/tp.shard_qwen3`` on the stage's
  sub-group).
* ``PipelineGenerator`` runs the same four model calls as ``Generator`` (prefill, prefill piece, packed prefill, decode).
  Every stage runs its slice; the residual stream (the pair ``delta, residual`` the fused norms work on, so the result
  is bit-identical to one process) travels to the same tensor-parallel rank of the next stage as ONE point-to-point
  message; the last stage applies the final norm and the lm_head and broadcasts the logits, so every rank returns the
  same tensor and rank 0 (the scheduler) samples.

A call occupies one stage at a time (no micro-batch interleaving): pipeline stages here buy capacity — a model that
does not fit one GPU's 180 GB — not throughput; inside one NVSwitch node tensor parallelism is the faster use of the same
GPUs, which is why ``serve/deploy.py`` only builds stages when the YAML asks for them.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, List, Optional

import torch
import torch.distributed as dist
from torch import nn

from ..inference.generate import Generator, KVCache, _decode_attn
from ..ops import elementwise as ew
from ..ops.gemm import gemm
from ..ops.norm import rms_norm


def partition_layers(n_layers: int, stages: int) -> List[range]:
    """Contiguous, balanced layer ranges; the first ``n_layers % stages`` stages hold one layer more."""
    if not 1 <= stages <= n_layers:
        raise ValueError(f"{stages} pipeline stages need at least as many decoder layers (model has {n_layers})")
    base, extra = divmod(n_layers, stages)
    out, a = [], 0
    for s in range(stages):
        b = a + base + (1 if s < extra else 0)
        out.append(range(a, b))
        a = b
    return out


@dataclass
class PipelineStage:
    stage: int
    stages: int
    tp: int
    layers: range                       # global indices of the decoder layers this stage holds
    prev_rank: Optional[int]            # global rank of the same tensor-parallel rank one stage earlier / later
    next_rank: Optional[int]
    logits_src: int                     # global rank whose logits every rank receives (last stage, tensor-parallel rank 0)
    tp_group: Any = None
    sends: int = 0

    @property
    def first(self) -> bool:
        return self.stage == 0

    @property
    def last(self) -> bool:
        return self.stage == self.stages - 1

    def send(self, delta: torch.Tensor, res: torch.Tensor) -> None:
        dist.send(torch.stack([delta, res]).contiguous(), dst=self.next_rank)
        self.sends += 1

    def recv(self, tokens: int, hidden: int, dtype, device):
        buf = torch.empty(2, tokens, hidden, dtype=dtype, device=device)
        dist.recv(buf, src=self.prev_rank)
        return buf[0], buf[1]

    def share_logits(self, logits: torch.Tensor) -> torch.Tensor:
        dist.broadcast(logits, src=self.logits_src)
        return logits


def shard_qwen3_pipeline(model, stages: Optional[int] = None, tp: int = 1, max_tokens: int = 8192) -> PipelineStage:
    """Rewrite a loaded ``Qwen3ForCausalLM`` in place for this rank's place in a ``stages x tp`` grid over the default
    process group. Every rank must call it (sub-groups are created collectively)."""
    m = getattr(model, "base_model", model)
    world = dist.get_world_size() if (dist.is_available() and dist.is_initialized()) else 1
    rank = dist.get_rank() if world > 1 else 0
    stages = stages if stages is not None else world // tp
    if stages * tp != world:
        raise ValueError(f"pipeline stages ({stages}) x tensor-parallel ranks ({tp}) must equal the world size ({world})")
    parts = partition_layers(len(m.layers), stages)
    s, r = divmod(rank, tp)
    tp_group = None
    if tp > 1:
        for g in range(stages):                                   # every rank creates every sub-group, in the same order
            grp = dist.new_group(list(range(g * tp, (g + 1) * tp)))
            if g == s:
                tp_group = grp
    st = PipelineStage(stage=s, stages=stages, tp=tp, layers=parts[s],
                       prev_rank=(rank - tp) if s > 0 else None, next_rank=(rank + tp) if s < stages - 1 else None,
                       logits_src=(stages - 1) * tp, tp_group=tp_group)
    m.layers = nn.ModuleList([m.layers[i] for i in parts[s]])
    m.config.num_hidden_layers = len(m.layers)
    if not st.last and m.lm_head is not None:
        m.lm_head = None                                           # only the last stage projects onto the vocabulary
    if tp > 1:
        from .tp import TensorParallel, shard_qwen3
        shard_qwen3(m, TensorParallel(m.config.hidden_size, max_tokens, m.embed_tokens.device, group=tp_group), max_tokens)
    m.pp = st
    return st


class PipelineGenerator(Generator):
    """``Generator`` whose model is one pipeline stage. Same call surface and results; every rank of the grid makes the
    same calls in the same order (serve/tp_engine.py announces rank 0's calls to the others)."""

    def __init__(self, model, tokenizer=None, max_len: int = 4096):
        super().__init__(model, tokenizer, max_len)
        self.pp: PipelineStage = self.model.pp
        self._graph_prefill = False          # point-to-point messages between the stages: every pass runs eagerly

    def _run_stage(self, ids, pos, store, attend, pick, n_out: int) -> torch.Tensor:
        """The decoder slice of this stage over ``ids`` / ``pos`` [T]. ``store(li, k, v)`` writes a layer's new K/V rows to
        the cache, ``attend(li, qkv)`` returns the attention context, ``pick(h)`` selects the rows of the final hidden
        state the lm_head is applied to (``n_out`` of them)."""
        m, c, st = self.model, self.cfg, self.pp
        T = ids.shape[0]
        dev = m.embed_tokens.device
        H, Hkv, D = c.num_attention_heads, c.num_key_value_heads, c.head_dim
        cos, sin = m.rope(dev, self.max_len)
        if st.first:
            delta, res = ew.embedding(ids, m.embed_tokens), None
        else:
            delta, res = st.recv(T, c.hidden_size, m.embed_tokens.dtype, dev)
        for li, layer in enumerate(m.layers):
            x, res = rms_norm(delta, layer.input_layernorm, c.rms_norm_eps, res)
            qkv = ew.qknorm_rope(layer.qkv_proj(x), layer.q_norm, layer.k_norm, cos, sin, pos, H, Hkv, D, False, c.rms_norm_eps)
            t = qkv.view(T, H + 2 * Hkv, D)
            store(li, t[:, H:H + Hkv], t[:, H + Hkv:])
            ctx = attend(li, qkv)
            x, res = rms_norm(layer.o_proj(ctx), layer.post_attention_layernorm, c.rms_norm_eps, res)
            delta = layer.down_proj(ew.swiglu(layer.gate_up_proj(x)))
        if st.last:
            h, _ = rms_norm(delta, m.norm, c.rms_norm_eps, res)
            logits = gemm(pick(h).contiguous(), m.head_weight(), out_dtype=torch.float32)
        else:
            st.send(delta, res)
            logits = torch.empty(n_out, c.vocab_size, dtype=torch.float32, device=dev)
        return st.share_logits(logits)

    @torch.no_grad()
    def _prefill(self, ids: torch.Tensor, cache: KVCache, slots: Optional[torch.Tensor] = None) -> torch.Tensor:
        from ..ops.attention import attention
        c = self.cfg
        B, S = ids.shape
        dev = ids.device
        H, Hkv, D = c.num_attention_heads, c.num_key_value_heads, c.head_dim
        pos = torch.arange(S, device=dev, dtype=torch.int32).repeat(B)
        srow = (slots.long() if slots is not None else torch.arange(B, device=dev)).repeat_interleave(S)
        pl = pos.long()
        out = self._run_stage(ids.reshape(-1), pos,
                              lambda li, k, v: cache.store(li, srow, pl, k, v),
                              lambda li, qkv: attention(qkv, B, S, H, Hkv, D, causal=True),
                              lambda h: h.view(B, S, -1)[:, -1], B)
        cache.len = S
        return out

    @torch.no_grad()
    def _prefill_chunk(self, ids: torch.Tensor, cache: KVCache, slot: int, pos0: int) -> torch.Tensor:
        from ..ops.attention import attention_with_past
        c = self.cfg
        n_real = ids.shape[0]
        dev = ids.device
        padn = (-(pos0 + n_real)) % 8               # same padding rule as Generator._prefill_chunk (16-byte score rows)
        if padn and pos0 + n_real + padn <= cache.max_len and ids.is_cuda:
            ids = torch.cat([ids, ids.new_zeros(padn)])
        T = ids.shape[0]
        H, Hkv, D = c.num_attention_heads, c.num_key_value_heads, c.head_dim
        pos = torch.arange(pos0, pos0 + T, device=dev, dtype=torch.int32)
        srow = torch.full((T,), slot, device=dev, dtype=torch.long)
        pl = pos.long()

        def attend(li, qkv):
            kq, vq = cache.seq_view(li, slot, pos0 + T)
            return attention_with_past(qkv, kq, vq, pos0, H, Hkv, D)

        return self._run_stage(ids, pos, lambda li, k, v: cache.store(li, srow, pl, k, v), attend,
                               lambda h: h[n_real - 1:n_real], 1)

    @torch.no_grad()
    def _prefill_packed_core(self, ids, pos, row, cu, last, max_seqlen: int, cache: KVCache) -> torch.Tensor:
        from ..ops.flash import flash_attention_varlen
        c = self.cfg
        H, Hkv, D = c.num_attention_heads, c.num_key_value_heads, c.head_dim
        pl = pos.long()
        return self._run_stage(ids, pos, lambda li, k, v: cache.store(li, row, pl, k, v),
                               lambda li, qkv: flash_attention_varlen(qkv, cu, max_seqlen, H, Hkv, D, D ** -0.5),
                               lambda h: h[last], last.shape[0])

    @torch.no_grad()
    def _decode(self, tok: torch.Tensor, cache: KVCache, pos: Optional[torch.Tensor] = None,
                slots: Optional[torch.Tensor] = None) -> torch.Tensor:
        c = self.cfg
        B = tok.shape[0]
        H, Hkv, D = c.num_attention_heads, c.num_key_value_heads, c.head_dim
        if pos is None:
            pos = torch.full((B,), cache.len, device=tok.device, dtype=torch.int32)
        table = getattr(cache, "table", None)
        # decode attention stores the new K/V row itself (one kernel: append + attend)
        out = self._run_stage(tok.reshape(-1), pos, lambda li, k, v: None,
                              lambda li, qkv: _decode_attn(qkv, cache.k[li], cache.v[li], pos, H, Hkv, D, D ** -0.5, slots, table),
                              lambda h: h, B)
        cache.len += 1
        return out

    @torch.no_grad()
    def decode_graphed(self, tok: torch.Tensor, cache: KVCache, pos: torch.Tensor, slots: torch.Tensor) -> torch.Tensor:
        L = cache.len                                   # ``cache.len`` is the caller's to advance (Generator.decode_graphed)
        out = self._decode(tok, cache, pos, slots)
        cache.len = L
        return out
