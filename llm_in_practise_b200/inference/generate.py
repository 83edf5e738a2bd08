"""Autoregressive generation for the Qwen family: KV cache, sampling, streaming, chat.

Covers the HF ``generate`` usage of the reference: ``do_sample / temperature / top_p / top_k / repetition_penalty /
max_new_tokens / eos_token_id`` (Fine-Tuning/inferences.py:50-58, Scripts/inference/01-deepseek1.5b-hf-infr.py:8-30,
Quantization/GPTQModel/inference_qwen3_4b_gptq.py:58-68), token streaming through an iterator while generation runs in
a background thread (``TextIteratorStreamer`` in Scripts/inference/06-…streaming-infr.py:55-85), and the multi-turn
``chat(query, history, system_prompt, max_new_tokens)`` helper (Fine-Tuning/inferences.py:28-83).

Prefill runs the training kernels (tcgen05 GEMMs, tiled causal attention) and stores the rotated K/V into the cache;
each decode step is q|k|v GEMM/GEMV -> qk-norm+RoPE -> ``b200_decode_attn`` (appends to the cache, online softmax) ->
o_proj -> MLP. 4-bit bases (NF4, GPTQ, AWQ) take the fused GEMV path for M <= 8.
"""
from __future__ import annotations

import os

import ctypes
import queue
import threading
from dataclasses import dataclass
from typing import Callable, Iterator, List, Optional, Sequence

import torch

from ..ops import _native as nv
from ..ops import elementwise as ew
from ..ops import fused_decode
from ..ops.attention import attention
from ..ops.gemm import ensure_graph_workspace, gemm, gemv16
from ..ops.norm import rms_norm

vp, i32, f32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
_SPLIT_DECODE = __import__("os").environ.get("B200_DECODE_SPLIT", "1") != "0"


@dataclass
class GenerationConfig:
    max_new_tokens: int = 256
    do_sample: bool = False
    temperature: float = 1.0
    top_p: float = 1.0
    top_k: int = 0
    repetition_penalty: float = 1.0
    eos_token_id: Optional[Sequence[int]] = None
    pad_token_id: int = 0
    ignore_eos: bool = False             # vLLM's sampling parameter: generate max_new_tokens whatever is sampled (benchmarks)


def sample_next(logits: torch.Tensor, cfg: GenerationConfig, prev: Optional[torch.Tensor] = None,
                generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """logits [B, V] (any float dtype) -> next ids [B]; HF logits-processor order: repetition penalty,
    temperature, top-k, top-p, then multinomial (or argmax when not sampling)."""
    lg = logits.float()
    if cfg.repetition_penalty != 1.0 and prev is not None:
        g = lg.gather(1, prev)
        g = torch.where(g < 0, g * cfg.repetition_penalty, g / cfg.repetition_penalty)
        lg = lg.scatter(1, prev, g)
    if not cfg.do_sample or cfg.temperature <= 0:
        return lg.argmax(-1)
    lg = lg / cfg.temperature
    if cfg.top_k and cfg.top_k > 0:
        kth = lg.topk(min(cfg.top_k, lg.shape[-1])).values[:, -1:]
        lg = lg.masked_fill(lg < kth, float("-inf"))
    if cfg.top_p < 1.0:
        sl, si = lg.sort(descending=True)
        cp = sl.softmax(-1).cumsum(-1)
        drop = cp - sl.softmax(-1) > cfg.top_p          # keep the first token that crosses top_p
        sl = sl.masked_fill(drop, float("-inf"))
        lg = torch.full_like(lg, float("-inf")).scatter(1, si, sl)
    return torch.multinomial(lg.softmax(-1), 1, generator=generator).squeeze(1)


class KVCache:
    """Slot pool: sequence ``slot`` owns row ``slot`` of ``k[layer]`` / ``v[layer]`` [slots, Hkv, max_len, D]."""
    paged = False

    def __init__(self, n_layers, B, Hkv, max_len, D, dtype, device):
        self.k = [torch.zeros(B, Hkv, max_len, D, dtype=dtype, device=device) for _ in range(n_layers)]
        self.v = [torch.zeros(B, Hkv, max_len, D, dtype=dtype, device=device) for _ in range(n_layers)]
        self.max_len = max_len
        self.len = 0

    def store(self, li: int, rows: torch.Tensor, pos: torch.Tensor, k: torch.Tensor, v: torch.Tensor) -> None:
        """k / v [T, Hkv, D] of token t -> position pos[t] of sequence rows[t] (both long [T])."""
        self.k[li][rows, :, pos] = k
        self.v[li][rows, :, pos] = v

    def seq_view(self, li: int, slot: int, n: int):
        """(K, V) of sequence ``slot`` as [Hkv, >= n, D] tensors (positions 0 .. n-1 valid)."""
        return self.k[li][slot], self.v[li][slot]


class PagedKVCache(KVCache):
    """vLLM's PagedAttention layout (the engine behind every `vllm serve` deployment of the reference,
    LLM_on_Kubernetes/Inference_Platfrom/README.md:1504-1509): ``k[layer]`` / ``v[layer]`` are
    [num_blocks, Hkv, block_size, D]; ``table`` [max_seqs, max_blocks] int32 maps the logical blocks of every sequence to
    physical ones. Memory is committed per ``block_size`` tokens a sequence actually holds, not per ``max_len``; two
    sequences may list the same physical block (a shared prompt prefix). The block allocator lives with the scheduler
    (serve/engine.py); this class only holds the storage and the table."""
    paged = True

    def __init__(self, n_layers, num_blocks, Hkv, block_size, D, dtype, device, max_seqs, max_len):
        if block_size & (block_size - 1):
            raise ValueError("block_size must be a power of two")
        self.k = [torch.zeros(num_blocks, Hkv, block_size, D, dtype=dtype, device=device) for _ in range(n_layers)]
        self.v = [torch.zeros(num_blocks, Hkv, block_size, D, dtype=dtype, device=device) for _ in range(n_layers)]
        self.block_size, self.num_blocks = block_size, num_blocks
        self.max_len = max_len
        self.max_blocks = -(-max_len // block_size)
        self.table = torch.zeros(max_seqs, self.max_blocks, dtype=torch.int32, device=device)
        self.len = 0

    def set_row(self, slot: int, blocks) -> None:
        """Logical block j of sequence ``slot`` -> physical block blocks[j]."""
        if len(blocks):
            self.table[slot, :len(blocks)] = torch.as_tensor(list(blocks), dtype=torch.int32).to(self.table.device, non_blocking=True)

    def _where(self, rows, pos):
        blk = self.table[rows, pos // self.block_size].long()
        return blk, pos % self.block_size

    def store(self, li, rows, pos, k, v):
        blk, off = self._where(rows, pos)
        self.k[li][blk, :, off] = k
        self.v[li][blk, :, off] = v

    def seq_view(self, li, slot, n):
        nb = -(-n // self.block_size)
        blks = self.table[slot, :nb].long()
        Hkv, D = self.k[li].shape[1], self.k[li].shape[3]
        k = self.k[li].index_select(0, blks).permute(1, 0, 2, 3).reshape(Hkv, nb * self.block_size, D)
        v = self.v[li].index_select(0, blks).permute(1, 0, 2, 3).reshape(Hkv, nb * self.block_size, D)
        return k, v


def _decode_attn(qkv, ck, cv, pos, H, Hkv, D, scale, slots=None, table=None):
    """``table`` (int32 [rows, max_blocks]): ck / cv are a paged cache [blocks, Hkv, block_size, D]."""
    B = qkv.shape[0]
    if nv.use_native(qkv):
        out = torch.empty(B, H * D, dtype=qkv.dtype, device=qkv.device)
        if _SPLIT_DECODE and qkv.dtype in (torch.bfloat16, torch.float16) and D in (64, 128) and H // Hkv in (1, 2, 4, 5, 8):
            # K/V read once per GQA group; enough (sequence, kv head, split) CTAs to fill the 148 SMs about twice
            nsplit = max(1, min(16, -(-296 // (B * Hkv))))
            scratch = torch.empty(B * H * nsplit * (D + 2), dtype=torch.float32, device=qkv.device)
            if table is not None:
                fp = nv.proto("b200_decode_attn_split_paged", [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, vp, i32, i32, f32, i32, i32, vp])
                nv.check(fp(nv.ptr(qkv), nv.ptr(ck), nv.ptr(cv), nv.ptr(pos), nv.ptr(slots), nv.ptr(out), nv.ptr(scratch), B, H, Hkv, D,
                            nv.ptr(table), table.shape[1], ck.shape[2], scale, nsplit, nv.dtype_code(qkv.dtype), nv.stream()),
                         "decode_attn_split_paged")
            else:
                fs = nv.proto("b200_decode_attn_split", [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, f32, i32, i32, vp])
                nv.check(fs(nv.ptr(qkv), nv.ptr(ck), nv.ptr(cv), nv.ptr(pos), nv.ptr(slots), nv.ptr(out), nv.ptr(scratch), B, H, Hkv, D,
                            ck.shape[2], scale, nsplit, nv.dtype_code(qkv.dtype), nv.stream()), "decode_attn_split")
            nv.count_launch(2)
            return out
        if table is not None:
            fp = nv.proto("b200_decode_attn_paged", [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, vp, i32, i32, f32, i32, i32, vp])
            nv.check(fp(nv.ptr(qkv), nv.ptr(ck), nv.ptr(cv), nv.ptr(pos), nv.ptr(slots), nv.ptr(out), B, H, Hkv, D, nv.ptr(table),
                        table.shape[1], ck.shape[2], scale, 0, nv.dtype_code(qkv.dtype), nv.stream()), "decode_attn_paged")
        else:
            fn = nv.proto("b200_decode_attn", [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, f32, i32, i32, vp])
            nv.check(fn(nv.ptr(qkv), nv.ptr(ck), nv.ptr(cv), nv.ptr(pos), nv.ptr(slots), nv.ptr(out), B, H, Hkv, D, ck.shape[2], scale, 0,
                        nv.dtype_code(qkv.dtype), nv.stream()), "decode_attn")
        nv.count_launch()
        return out
    x = qkv.view(B, H + 2 * Hkv, D)
    g = H // Hkv
    rows = (slots if slots is not None else torch.arange(B, device=qkv.device)).long()
    p = pos.long()
    tmax = int(p.max()) + 1
    if table is not None:
        BS = ck.shape[2]
        blk = table[rows, p // BS].long()
        ck[blk, :, p % BS] = x[:, H:H + Hkv]
        cv[blk, :, p % BS] = x[:, H + Hkv:]
        nb = -(-tmax // BS)
        tb = table[rows, :nb].long()                                             # [B, nb]
        k = ck[tb].permute(0, 2, 1, 3, 4).reshape(B, Hkv, nb * BS, D)[:, :, :tmax].float().repeat_interleave(g, 1)
        v = cv[tb].permute(0, 2, 1, 3, 4).reshape(B, Hkv, nb * BS, D)[:, :, :tmax].float().repeat_interleave(g, 1)
    else:
        ck[rows, :, p] = x[:, H:H + Hkv]
        cv[rows, :, p] = x[:, H + Hkv:]
        k = ck[rows, :, :tmax].float().repeat_interleave(g, 1)
        v = cv[rows, :, :tmax].float().repeat_interleave(g, 1)
    s = torch.einsum("bhd,bhtd->bht", x[:, :H].float(), k) * scale
    s = s.masked_fill(torch.arange(tmax, device=qkv.device)[None, None, :] > p[:, None, None], float("-inf"))
    return torch.einsum("bht,bhtd->bhd", s.softmax(-1), v).reshape(B, H * D).to(qkv.dtype)


class Generator:
    """Stateful runner around a ``Qwen3ForCausalLM`` (dense, NF4 or W4A16 base; merged or live LoRA)."""

    def __init__(self, model, tokenizer=None, max_len: int = 4096):
        self.model = getattr(model, "base_model", model)
        self.tok = tokenizer
        self.max_len = max_len
        self.cfg = self.model.config
        self._graphs = {}
        self._caches = {}
        self._graph_prefill = os.environ.get("B200_GRAPH_PREFILL", "1") != "0"

    @torch.no_grad()
    def _prefill(self, ids: torch.Tensor, cache: KVCache, slots: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Run the prompt(s) [B, S], store K/V in cache rows ``slots`` (default 0..B-1), return last-token logits."""
        m, c = self.model, self.cfg
        B, S = ids.shape
        dev = ids.device
        H, Hkv, D = c.num_attention_heads, c.num_key_value_heads, c.head_dim
        cos, sin = m.rope(dev, self.max_len)
        pos = torch.arange(S, device=dev, dtype=torch.int32).repeat(B)
        srow = (slots.long() if slots is not None else torch.arange(B, device=dev)).repeat_interleave(S)   # cache row of every token
        delta, res = ew.embedding(ids.reshape(-1), m.embed_tokens), None
        for li, layer in enumerate(m.layers):
            x, res = rms_norm(delta, layer.input_layernorm, c.rms_norm_eps, res)
            qkv = ew.qknorm_rope(layer.qkv_proj(x), layer.q_norm, layer.k_norm, cos, sin, pos, H, Hkv, D, False, c.rms_norm_eps)
            t = qkv.view(B * S, H + 2 * Hkv, D)
            cache.store(li, srow, pos.long(), t[:, H:H + Hkv], t[:, H + Hkv:])
            ctx = attention(qkv, B, S, H, Hkv, D, causal=True)
            x, res = rms_norm(layer.o_proj(ctx), layer.post_attention_layernorm, c.rms_norm_eps, res)
            delta = layer.down_proj(ew.swiglu(layer.gate_up_proj(x)))
        h, _ = rms_norm(delta, m.norm, c.rms_norm_eps, res)
        cache.len = S
        last = h.view(B, S, -1)[:, -1]
        return gemm(last.contiguous(), m.head_weight(), out_dtype=torch.float32)

    @torch.no_grad()
    def copy_prefix(self, cache: KVCache, src: int, dst: int, n: int) -> None:
        """K/V of the first ``n`` positions of cache row ``src`` -> row ``dst`` (a prefix-cache hit on a row that is still
        in use). A method of the generator so that tensor-parallel followers repeat it on their slice of the cache."""
        if cache.paged:
            return            # paged cache: the scheduler lists the prefix's physical blocks in both table rows, nothing to copy
        for li in range(len(cache.k)):
            cache.k[li][dst, :, :n] = cache.k[li][src, :, :n]
            cache.v[li][dst, :, :n] = cache.v[li][src, :, :n]

    # -- host tier of the prefix cache (serve/engine.py, ``kv_offload_gb``): generator methods so that tensor- / pipeline-parallel
    #    followers keep their own slice of every off-loaded row under the same key
    @torch.no_grad()
    def offload_row(self, cache: KVCache, slot: int, n: int, key: int) -> None:
        """K/V of the first ``n`` positions of sequence ``slot`` (every local layer) -> pinned host memory under ``key``."""
        store = self.__dict__.setdefault("_host_kv", {})
        ks, vs = [], []
        for li in range(len(cache.k)):
            k, v = cache.seq_view(li, slot, n)
            for src, dst in ((k, ks), (v, vs)):
                h = torch.empty((src.shape[0], n, src.shape[2]), dtype=src.dtype, device="cpu", pin_memory=src.is_cuda)
                h.copy_(src[:, :n], non_blocking=src.is_cuda)
                dst.append(h)
        if cache.k[0].is_cuda:
            torch.cuda.current_stream().synchronize()          # the row is about to be overwritten by its next owner
        store[key] = (ks, vs)

    @torch.no_grad()
    def restore_row(self, cache: KVCache, key: int, slot: int, n: int) -> None:
        """The first ``n`` positions stored under ``key`` -> sequence ``slot`` (its blocks are already allocated)."""
        ks, vs = self._host_kv[key]
        dev = cache.k[0].device
        rows = torch.full((n,), slot, dtype=torch.long, device=dev)
        pos = torch.arange(n, dtype=torch.long, device=dev)
        for li in range(len(cache.k)):
            cache.store(li, rows, pos, ks[li][:, :n].to(dev, non_blocking=True).transpose(0, 1),
                        vs[li][:, :n].to(dev, non_blocking=True).transpose(0, 1))

    def drop_host_row(self, key: int) -> None:
        self.__dict__.get("_host_kv", {}).pop(key, None)

    @torch.no_grad()
    def set_block_row(self, cache: KVCache, slot: int, blocks) -> None:
        """Paged cache: logical -> physical block list of sequence ``slot``. A generator method so that tensor-parallel
        followers apply the same table update to their slice of the cache."""
        cache.set_row(slot, blocks)

    @torch.no_grad()
    def _prefill_chunk(self, ids: torch.Tensor, cache: KVCache, slot: int, pos0: int) -> torch.Tensor:
        """Prefill ``ids`` [T] of ONE sequence at positions pos0 .. pos0 + T - 1 of cache row ``slot``, attending to what
        the row already holds below ``pos0``: the unit of chunked prefill and of a prefix-cache hit (the shared prefix is
        in the row, only the suffix is computed). Returns the logits of the last token [1, V]."""
        m, c = self.model, self.cfg
        n_real = ids.shape[0]
        dev = ids.device
        # the score / probability rows of the tensor-core path need 16-byte strides: the chunk is padded with dummy tokens
        # until pos0 + T is a multiple of 8. Their K/V land beyond the sequence's valid length (rewritten before they can
        # ever be attended to) and their outputs are dropped.
        padn = (-(pos0 + n_real)) % 8
        if padn and pos0 + n_real + padn <= cache.max_len and ids.is_cuda:
            ids = torch.cat([ids, ids.new_zeros(padn)])
        T = ids.shape[0]
        H, Hkv, D = c.num_attention_heads, c.num_key_value_heads, c.head_dim
        cos, sin = m.rope(dev, self.max_len)
        pos = torch.arange(pos0, pos0 + T, device=dev, dtype=torch.int32)
        srow = torch.full((T,), slot, device=dev, dtype=torch.long)
        from ..ops.attention import attention_with_past
        delta, res = ew.embedding(ids, m.embed_tokens), None
        for li, layer in enumerate(m.layers):
            x, res = rms_norm(delta, layer.input_layernorm, c.rms_norm_eps, res)
            qkv = ew.qknorm_rope(layer.qkv_proj(x), layer.q_norm, layer.k_norm, cos, sin, pos, H, Hkv, D, False, c.rms_norm_eps)
            t = qkv.view(T, H + 2 * Hkv, D)
            cache.store(li, srow, pos.long(), t[:, H:H + Hkv], t[:, H + Hkv:])
            kq, vq = cache.seq_view(li, slot, pos0 + T)
            ctx = attention_with_past(qkv, kq, vq, pos0, H, Hkv, D)
            x, res = rms_norm(layer.o_proj(ctx), layer.post_attention_layernorm, c.rms_norm_eps, res)
            delta = layer.down_proj(ew.swiglu(layer.gate_up_proj(x)))
        h, _ = rms_norm(delta, m.norm, c.rms_norm_eps, res)
        return gemm(h[n_real - 1:n_real].contiguous(), m.head_weight(), out_dtype=torch.float32)

    @torch.no_grad()
    def _prefill_packed_core(self, ids, pos, row, cu, last, max_seqlen: int, cache: KVCache) -> torch.Tensor:
        """Tensor-level body of the packed prefill (everything it reads is a device tensor, so it can be captured):
        ids / pos / row [T] (token, position and cache row of every packed token), cu [n + 1] int32, last [n] (row index
        of each sequence's final token)."""
        m, c = self.model, self.cfg
        H, Hkv, D = c.num_attention_heads, c.num_key_value_heads, c.head_dim
        cos, sin = m.rope(ids.device, self.max_len)
        pl = pos.long()
        from ..ops.flash import flash_attention_varlen
        delta, res = ew.embedding(ids, m.embed_tokens), None
        for li, layer in enumerate(m.layers):
            x, res = rms_norm(delta, layer.input_layernorm, c.rms_norm_eps, res)
            qkv = ew.qknorm_rope(layer.qkv_proj(x), layer.q_norm, layer.k_norm, cos, sin, pos, H, Hkv, D, False, c.rms_norm_eps)
            t = qkv.view(-1, H + 2 * Hkv, D)
            cache.store(li, row, pl, t[:, H:H + Hkv], t[:, H + Hkv:])
            ctx = flash_attention_varlen(qkv, cu, max_seqlen, H, Hkv, D, D ** -0.5)
            x, res = rms_norm(layer.o_proj(ctx), layer.post_attention_layernorm, c.rms_norm_eps, res)
            delta = layer.down_proj(ew.swiglu(layer.gate_up_proj(x)))
        h, _ = rms_norm(delta, m.norm, c.rms_norm_eps, res)
        return gemm(h[last].contiguous(), m.head_weight(), out_dtype=torch.float32)

    _PREFILL_BUCKETS = tuple(range(64, 513, 64))   # graphs pay off while the pass is launch-bound; larger passes run eagerly

    @torch.no_grad()
    def _prefill_packed(self, prompts, cache: KVCache, slots, scratch_slot: Optional[int] = None) -> torch.Tensor:
        """Prefill several prompts of different lengths in ONE pass: the tokens are packed back to back ([T, hidden]
        through every linear, so the GEMMs see T rows instead of one short problem per prompt), attention runs per
        sequence through the packed flash kernel, K/V rows are scattered to cache row ``slots[i]``. Returns the
        last-token logits [n, vocab].

        With ``scratch_slot`` (a cache row no sequence owns) on one GPU, passes of up to 512 tokens are replayed from a
        CUDA graph: T is padded to a multiple of 64, the sequence count to 8, the padding tokens write their K/V to the
        scratch row and belong to no sequence. A 259-token prefill is ~500 launches, i.e. bound by the eager launch
        rate (15.2 ms eager, 9.7 ms replayed); from ~1000 tokens the GPU is the bound and padding only costs."""
        m = self.model
        dev = m.embed_tokens.device
        lens = [len(p) for p in prompts]
        n, T = len(lens), sum(lens)
        tp = getattr(m, "tp", None)
        graphed = (scratch_slot is not None and dev.type == "cuda" and n <= 8 and T <= self._PREFILL_BUCKETS[-1]
                   and not (tp is not None and tp.world > 1) and self._graph_prefill)
        Tb = next(b for b in self._PREFILL_BUCKETS if b >= T) if graphed else T
        ns = 8 if graphed else n
        pad = Tb - T
        cu_l = [0]
        for k in lens:
            cu_l.append(cu_l[-1] + k)
        cu_l += [T] * (ns - n)
        ids_h = torch.tensor([t for p in prompts for t in p] + [0] * pad)
        pos_h = torch.cat([torch.arange(k, dtype=torch.int32) for k in lens] + [torch.zeros(pad, dtype=torch.int32)])
        row_h = torch.cat([torch.repeat_interleave(torch.as_tensor(list(slots), dtype=torch.long), torch.tensor(lens)),
                           torch.full((pad,), scratch_slot if scratch_slot is not None else 0, dtype=torch.long)])
        cu_h = torch.tensor(cu_l, dtype=torch.int32)
        last_h = torch.tensor([c - 1 for c in cu_l[1:n + 1]] + [0] * (ns - n), dtype=torch.long)
        if not graphed:
            return self._prefill_packed_core(ids_h.to(dev), pos_h.to(dev), row_h.to(dev), cu_h.to(dev), last_h.to(dev),
                                             max(lens), cache)
        key = ("prefill", Tb, id(cache))
        g = self._graphs.get(key)
        if g is None:
            st = {"ids": ids_h.to(dev), "pos": pos_h.to(dev), "row": row_h.to(dev), "cu": cu_h.to(dev), "last": last_h.to(dev)}
            ensure_graph_workspace(dev)
            self._prefill_packed_core(st["ids"], st["pos"], st["row"], st["cu"], st["last"], Tb, cache)   # warm-up
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                st["logits"] = self._prefill_packed_core(st["ids"], st["pos"], st["row"], st["cu"], st["last"], Tb, cache)
            g = self._graphs[key] = (graph, st)
        graph, st = g
        st["ids"].copy_(ids_h, non_blocking=True)
        st["pos"].copy_(pos_h, non_blocking=True)
        st["row"].copy_(row_h, non_blocking=True)
        st["cu"].copy_(cu_h, non_blocking=True)
        st["last"].copy_(last_h, non_blocking=True)
        graph.replay()
        return st["logits"][:n]

    def _fused_tail(self, x: torch.Tensor):
        """The fused-tail context (staging buffer + device state) when every o_proj / down_proj of the model can take the
        fused launch for this batch, else None. ``B200_FUSED_DECODE=0`` turns it off (A/B timing)."""
        if os.environ.get("B200_FUSED_DECODE", "1") == "0" or not x.is_cuda:
            return None
        m = self.model
        hid = self.cfg.hidden_size
        ok = getattr(self, "_fused_ok", None)
        if ok is None:
            l0 = m.layers[0]
            ok = (m.norm.dtype == x.dtype and l0.input_layernorm.dtype == x.dtype and
                  all(fused_decode.supported(getattr(l, n), x[:1], hid) for l in m.layers for n in ("o_proj", "down_proj")))
            self._fused_ok = ok
        if not ok or x.shape[0] > fused_decode.MAX_TOKENS:
            return None
        if getattr(self, "_tail", None) is None:
            tp = getattr(m, "tp", None)
            if tp is not None and tp.world > 1 and not getattr(tp, "fused", False):
                self._fused_ok = False
                return None
            if tp is not None and tp.world > 1:
                # ONE tail per tensor-parallel group and width: its epoch lives on the device and the flags it is compared
                # with live in the group's symmetric memory — a second tail on the same flag channel would start again at
                # epoch 1, find the previous tail's (larger) epochs in the flags and stop waiting for its peers
                tails = tp.__dict__.setdefault("_fused_tails", {})
                key = (hid, x.dtype)
                if key not in tails:
                    if tails:                     # the flag channel is taken by a tail of another width / dtype
                        self._fused_ok = False
                        return None
                    tails[key] = fused_decode.FusedTail(hid, x.device, tp, x.dtype)
                self._tail = tails[key]
            else:
                self._tail = fused_decode.FusedTail(hid, x.device, None, x.dtype)
        return self._tail

    @torch.no_grad()
    def _decode(self, tok: torch.Tensor, cache: KVCache, pos: Optional[torch.Tensor] = None,
                slots: Optional[torch.Tensor] = None) -> torch.Tensor:
        """One token per sequence. ``pos`` [B] int32 = tokens already cached per sequence (default: cache.len for
        all); ``slots`` [B] int32 = cache row of each batch row (continuous batching)."""
        m, c = self.model, self.cfg
        B = tok.shape[0]
        dev = tok.device
        H, Hkv, D = c.num_attention_heads, c.num_key_value_heads, c.head_dim
        cos, sin = m.rope(dev, self.max_len)
        if pos is None:
            pos = torch.full((B,), cache.len, device=dev, dtype=torch.int32)
        delta, res = ew.embedding(tok.reshape(-1), m.embed_tokens), None
        tail = self._fused_tail(delta)
        if tail is not None:
            # o_proj / down_proj run as ONE launch each with the tensor-parallel all-reduce, the residual add and the next
            # RMSNorm in their tail (ops/fused_decode.py): 7 launches per layer instead of 10 (12 under TP)
            layers = m.layers
            x, res = rms_norm(delta, layers[0].input_layernorm, c.rms_norm_eps, None)
            for li, layer in enumerate(layers):
                qkv = ew.qknorm_rope(layer.qkv_proj(x), layer.q_norm, layer.k_norm, cos, sin, pos, H, Hkv, D, False, c.rms_norm_eps)
                ctx = _decode_attn(qkv, cache.k[li], cache.v[li], pos, H, Hkv, D, D ** -0.5, slots, getattr(cache, "table", None))
                x, res = fused_decode.gemv_fused(layer.o_proj, ctx, res, layer.post_attention_layernorm, c.rms_norm_eps, tail)
                nw = layers[li + 1].input_layernorm if li + 1 < len(layers) else m.norm
                x, res = fused_decode.gemv_fused(layer.down_proj, ew.swiglu(layer.gate_up_proj(x)), res, nw, c.rms_norm_eps, tail)
            h = x
        else:
            for li, layer in enumerate(m.layers):
                x, res = rms_norm(delta, layer.input_layernorm, c.rms_norm_eps, res)
                qkv = ew.qknorm_rope(layer.qkv_proj(x), layer.q_norm, layer.k_norm, cos, sin, pos, H, Hkv, D, False, c.rms_norm_eps)
                ctx = _decode_attn(qkv, cache.k[li], cache.v[li], pos, H, Hkv, D, D ** -0.5, slots, getattr(cache, "table", None))
                x, res = rms_norm(layer.o_proj(ctx), layer.post_attention_layernorm, c.rms_norm_eps, res)
                delta = layer.down_proj(ew.swiglu(layer.gate_up_proj(x)))
            h, _ = rms_norm(delta, m.norm, c.rms_norm_eps, res)
        cache.len += 1
        logits = gemv16(h, m.head_weight(), out_dtype=torch.float32)
        return logits if logits is not None else gemm(h, m.head_weight(), out_dtype=torch.float32)

    @torch.no_grad()
    def decode_graphed(self, tok: torch.Tensor, cache: KVCache, pos: torch.Tensor, slots: torch.Tensor) -> torch.Tensor:
        """``_decode`` replayed from a CUDA graph (one per batch size and cache): a decode step is ~14 launches per
        layer of a few microseconds each, so eager execution is bound by the Python launch rate, not by the GPU. The
        kernels read token ids, positions and cache slots from device tensors, so one graph serves any set of
        sequences. Falls back to the eager path on CPU and under tensor parallelism (the peer all-reduce takes a host
        epoch). Inside a capture the GEMMs run without stream-K and 4-bit weights are expanded in line."""
        B = tok.shape[0]
        L = cache.len                                   # ``cache.len`` is the caller's to advance, never this method's
        tp = getattr(self.model, "tp", None)
        # under tensor parallelism the step is replayable only when its all-reduce keeps its epoch on the device
        tp_eager = tp is not None and tp.world > 1 and not (getattr(tp, "fused", False) and getattr(tp, "oneshot", False)
                                                            and B * self.cfg.hidden_size <= getattr(tp, "half16", 0))
        if not tok.is_cuda or B > 512 or tp_eager:
            out = self._decode(tok, cache, pos, slots)
            cache.len = L
            return out
        key = (B, id(cache))
        g = self._graphs.get(key)
        if g is None:
            st = {"tok": tok.clone(), "pos": pos.clone(), "slots": slots.clone()}
            ensure_graph_workspace(tok.device)                              # stream-K scratch of captured GEMMs
            self._decode(st["tok"], cache, st["pos"], st["slots"])          # warm-up: allocates every lazily built buffer
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                st["logits"] = self._decode(st["tok"], cache, st["pos"], st["slots"])
            g = self._graphs[key] = (graph, st)
            cache.len = L
        graph, st = g
        st["tok"].copy_(tok)
        st["pos"].copy_(pos)
        st["slots"].copy_(slots)
        graph.replay()
        return st["logits"]

    @torch.no_grad()
    def generate(self, input_ids: torch.Tensor, config: Optional[GenerationConfig] = None, streamer=None,
                 seed: Optional[int] = None, return_logprobs: bool = False, **kw):
        try:
            return self._generate(input_ids, config, streamer, seed, return_logprobs, **kw)
        finally:
            if streamer is not None:
                streamer.end()            # also when generation raised: the thread reading the streamer must not wait for ever

    def _generate(self, input_ids: torch.Tensor, config: Optional[GenerationConfig], streamer, seed: Optional[int],
                  return_logprobs: bool, **kw):
        cfg = config or GenerationConfig(**{k: v for k, v in kw.items() if k in GenerationConfig.__dataclass_fields__})
        eos = cfg.eos_token_id
        eos = [] if cfg.ignore_eos else ([eos] if isinstance(eos, int) else list(eos or []))
        was = self.model.training
        self.model.eval()
        B, S = input_ids.shape
        dev = input_ids.device
        total = min(self.max_len, S + cfg.max_new_tokens)
        if input_ids.is_cuda and B <= 8:
            # one persistent full-length cache per batch size: its CUDA graph (which bakes the addresses in) is reused
            cache = self._caches.get(B)
            if cache is None:
                cache = self._caches[B] = KVCache(self.cfg.num_hidden_layers, B, self.cfg.num_key_value_heads, self.max_len,
                                                  self.cfg.head_dim, self.model.embed_tokens.dtype, dev)
        else:
            cache = KVCache(self.cfg.num_hidden_layers, B, self.cfg.num_key_value_heads, total, self.cfg.head_dim,
                            self.model.embed_tokens.dtype, dev)
        gen = torch.Generator(device=dev).manual_seed(seed) if seed is not None else None
        logits = self._prefill(input_ids, cache)
        use_graph = input_ids.is_cuda and B <= 8 and kw.get("cuda_graph", True)
        if use_graph:
            pos_t = torch.zeros(B, dtype=torch.int32, device=dev)
            slots_t = torch.arange(B, dtype=torch.int32, device=dev)
        out = input_ids
        done = torch.zeros(B, dtype=torch.bool, device=dev)
        logprobs: List[torch.Tensor] = []
        for _ in range(total - S):
            nxt = sample_next(logits, cfg, out, gen)
            if return_logprobs:
                logprobs.append(logits.float().log_softmax(-1).gather(1, nxt[:, None]).squeeze(1))
            nxt = torch.where(done, torch.full_like(nxt, cfg.pad_token_id), nxt)
            out = torch.cat([out, nxt[:, None]], 1)
            if streamer is not None:
                streamer.put(nxt)
            for e in eos:
                done |= nxt == e
            if bool(done.all()):
                break
            if use_graph:
                pos_t.fill_(cache.len)
                logits = self.decode_graphed(nxt, cache, pos_t, slots_t)
                cache.len += 1
            else:
                logits = self._decode(nxt, cache)
        self.model.train(was)
        return (out, torch.stack(logprobs, 1)) if return_logprobs else out


class TextIteratorStreamer:
    """Iterator over decoded text pieces, fed by ``generate`` from another thread."""

    def __init__(self, tokenizer, skip_prompt: bool = True, timeout: Optional[float] = None, **decode_kw):
        self.tok, self.q, self.timeout, self.kw = tokenizer, queue.Queue(), timeout, decode_kw
        self._end = object()
        self._ids: List[int] = []          # tokens of the current line, decoded together: a character may span tokens
        self._sent = 0                     # characters of that decoding already handed out
        self._ended = False

    def put(self, ids: torch.Tensor) -> None:
        """New token(s) of ONE sequence. Byte-level vocabularies split multi-byte characters (most CJK text) over several
        tokens, so the text is re-decoded from the start of the line and only complete characters are emitted."""
        self._ids.extend(int(i) for i in ids.reshape(-1).tolist())
        text = self.tok.decode(self._ids, **self.kw)
        if text.endswith("\ufffd"):
            return                         # an unfinished character: wait for its remaining bytes
        piece, self._sent = text[self._sent:], len(text)
        if piece:
            self.q.put(piece)
        if text.endswith("\n"):
            self._ids, self._sent = [], 0  # a new line cannot change how the earlier text decodes

    def end(self) -> None:
        if self._ended:
            return
        self._ended = True
        if self._ids:
            try:
                tail = self.tok.decode(self._ids, **self.kw)[self._sent:]
            except Exception:              # the failure that ended generation must not also block the reader
                tail = ""
            if tail:
                self.q.put(tail)
        self.q.put(self._end)

    def __iter__(self) -> Iterator[str]:
        return self

    def __next__(self) -> str:
        v = self.q.get(timeout=self.timeout)
        if v is self._end:
            th = getattr(self, "thread", None)
            if th is not None and th is not threading.current_thread():
                th.join()                  # the producer is past its last token: let it finish before the caller moves on
            raise StopIteration
        return v


def stream_generate(gen: Generator, input_ids: torch.Tensor, tokenizer, **kw) -> Iterator[str]:
    st = TextIteratorStreamer(tokenizer)
    th = threading.Thread(target=lambda: gen.generate(input_ids, streamer=st, **kw), daemon=True)
    st.thread = th                         # joined by the iterator at the end of the stream (a daemon thread still inside
    th.start()                             # native code at interpreter exit aborts the process)
    return st


def chat(gen: Generator, tokenizer, query: str, history=None, system_prompt: str = "You are a helpful assistant.",
         max_new_tokens: int = 256, do_sample: bool = True, temperature: float = 0.7, top_p: float = 0.9, **kw):
    """One chat turn with hand-rendered ChatML; returns (response, new_history)."""
    from ..data.chat import build_messages, render_chatml
    history = list(history or [])
    text = render_chatml(build_messages(query, None, system_prompt, history), add_generation_prompt=True, space_before_end=False)
    ids = torch.tensor([tokenizer.encode(text)], device=gen.model.embed_tokens.device)
    eos = [t for t in (getattr(tokenizer, "eos_token_id", None),) if t is not None]
    out = gen.generate(ids, max_new_tokens=max_new_tokens, do_sample=do_sample, temperature=temperature, top_p=top_p,
                       eos_token_id=eos, **kw)
    resp = tokenizer.decode(out[0, ids.shape[1]:].tolist())
    history.append((query, resp))
    return resp, history
