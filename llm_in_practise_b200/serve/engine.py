"""Continuous-batching inference engine — the in-process replacement for the vLLM server every serving recipe of the
reference points at (``vllm serve … --max-model-len … --max-num-seqs … --gpu-memory-utilization …``;
Deployment/Ray/serve_run_examples/deepseek.py:5-41, Deployment/Basics-02/docker-compose.yml:56-65,
Quantization/README.md:172-176).

One scheduler thread owns the GPU. The KV cache is a fixed pool of ``max_num_seqs`` rows per layer
([slots, Hkv, max_model_len, D], sized once — 180 GB of HBM3e hold e.g. 64 x 32k-token rows of Qwen3-14B in bf16
next to the NF4 weights); each iteration admits waiting requests into free rows (one prefill each: tcgen05 GEMMs +
causal attention, K/V stored straight into the row) and then advances *all* running sequences by one token in a
single batched decode step (``b200_decode_attn`` takes per-row positions and slot indices, so sequences of different
lengths share the launch). Finished rows are recycled immediately, so the batch never drains.

Counters follow the vLLM metric names the reference's autoscaling rules read
(``vllm:num_requests_waiting``, ``vllm:time_to_first_token_seconds``;
LLM_on_Kubernetes/Inference_Platfrom/05-KEDA-AutoScale/keda-scaledobject.yaml:12-53).
"""
from __future__ import annotations

import itertools
import queue
import threading
import time
from collections import OrderedDict, deque
from dataclasses import dataclass, field
from typing import Deque, Dict, Iterator, List, Optional, Sequence, Tuple

import torch

from ..inference.generate import GenerationConfig, Generator, KVCache, PagedKVCache, sample_next


@dataclass
class EngineArgs:
    max_num_seqs: int = 16
    max_model_len: int = 4096
    max_prefill_per_step: int = 8        # bound on admissions per iteration so decode latency stays flat
    max_prefill_tokens: int = 2048       # ... and on their total prompt tokens (one packed prefill pass per iteration)
    seed: Optional[int] = None
    warmup: bool = True                  # capture the prefill / decode CUDA graphs in start() instead of in the first requests
    # the three engine_kwargs the reference's serving recipes set (Deployment/Ray/serve_run_examples/deepseek.py:29-34)
    enable_prefix_caching: bool = True   # finished sequences keep their KV row; a new prompt reuses the longest cached prefix
    enable_chunked_prefill: bool = True  # prompts longer than max_prefill_tokens are prefilled in pieces between decode steps
    prefix_block: int = 16               # prefix matches are counted in blocks of this many tokens (vLLM's block_size)
    gpu_memory_utilization: float = 0.0  # > 0: derive max_num_seqs from the free HBM (rows of max_model_len tokens)
    # paged KV cache (vLLM's PagedAttention): block_size > 0 (a power of two, e.g. 16) replaces the per-sequence rows of
    # max_model_len tokens by a pool of num_kv_blocks blocks of block_size tokens, handed out as sequences grow and shared
    # between sequences with a common prompt prefix. num_kv_blocks 0 = the capacity the slot pool would have had.
    block_size: int = 0
    num_kv_blocks: int = 0
    # host tier of the prefix cache (the LMCache stage of the reference's platform, LLM_on_Kubernetes/Inference_Platfrom/
    # 07-L1-Cache/LMCache): > 0 keeps the K/V of cached prefixes that lose their HBM row in this many GB of pinned host
    # memory (LRU) and copies them back on a later prefix hit instead of recomputing the prompt
    kv_offload_gb: float = 0.0


@dataclass
class Request:
    rid: int
    prompt_ids: List[int]
    cfg: GenerationConfig
    arrival: float = field(default_factory=time.perf_counter)
    first_token_time: Optional[float] = None
    finish_time: Optional[float] = None
    output_ids: List[int] = field(default_factory=list)
    finish_reason: Optional[str] = None
    slot: int = -1
    done: threading.Event = field(default_factory=threading.Event)
    stream: "queue.Queue" = field(default_factory=queue.Queue)
    error: Optional[str] = None
    token_times: List[float] = field(default_factory=list)   # perf_counter at every emitted token (TTFT / ITL statistics)
    prefilled: int = 0                                       # prompt tokens whose K/V are already in the cache row
    cached_tokens: int = 0                                   # ... of which came from the prefix cache
    aborted: bool = False                                    # set by LLMEngine.abort(): dropped at the next scheduler iteration
    reserved_blocks: int = 0                                 # paged cache: blocks promised to this request but not handed out yet

    @property
    def ttft(self) -> Optional[float]:
        return None if self.first_token_time is None else self.first_token_time - self.arrival

    def tokens(self) -> Iterator[int]:
        """Blocking iterator over generated ids (used by the SSE endpoint)."""
        while True:
            t = self.stream.get()
            if t is None:
                return
            yield t


class EngineMetrics:
    TTFT_BUCKETS = (0.001, 0.005, 0.01, 0.02, 0.04, 0.06, 0.08, 0.1, 0.25, 0.5, 0.75, 1.0, 2.5, 5.0, 7.5, 10.0)

    def __init__(self):
        self.lock = threading.Lock()
        self.prompt_tokens = 0
        self.generation_tokens = 0
        self.requests_finished: Dict[str, int] = {}
        self.ttft_counts = [0] * (len(self.TTFT_BUCKETS) + 1)
        self.ttft_sum = 0.0
        self.e2e_sum = 0.0
        self.steps = 0
        self.prefix_queries = 0          # prompt tokens looked up in the prefix cache (vllm:gpu_prefix_cache_queries)
        self.prefix_hits = 0             # ... and found there                        (vllm:gpu_prefix_cache_hits)
        self.queue_sum = 0.0             # seconds requests spent waiting before admission (vllm:request_queue_time_seconds)
        self.queue_count = 0
        self.host_hits = 0               # prompt tokens restored from the host tier instead of being recomputed
        self.host_offloads = 0           # rows copied to the host tier

    def observe_ttft(self, v: float) -> None:
        with self.lock:
            self.ttft_sum += v
            for i, b in enumerate(self.TTFT_BUCKETS):
                if v <= b:
                    self.ttft_counts[i] += 1
                    return
            self.ttft_counts[-1] += 1


class LLMEngine:
    generator_cls = Generator            # serve/tp_engine.PipelineParallelEngine runs a pipeline stage instead

    def __init__(self, model, tokenizer=None, args: Optional[EngineArgs] = None, model_name: str = "b200-model"):
        self.args = args or EngineArgs()
        self.gen = self.generator_cls(model, tokenizer, self.args.max_model_len)
        self.model, self.tokenizer, self.model_name = self.gen.model, tokenizer, model_name
        c = self.model.config
        self.device = self.model.embed_tokens.device
        if self.args.gpu_memory_utilization > 0 and self.device.type == "cuda":
            # vLLM's knob: the KV pool takes what is left of `utilization x total HBM` after the weights
            free, total = torch.cuda.mem_get_info(self.device)
            budget = free - (1.0 - self.args.gpu_memory_utilization) * total
            row = 2 * c.num_hidden_layers * c.num_key_value_heads * self.args.max_model_len * c.head_dim * self.model.embed_tokens.element_size()
            self.args.max_num_seqs = int(max(1, min(4096, budget // row - 1)))
        # one extra row: the scratch slot that padded rows of a bucketed (CUDA-graph) decode batch write into
        self.paged = self.args.block_size > 0
        if self.paged:
            BS = self.args.block_size
            per_seq = -(-self.args.max_model_len // BS)
            nb = self.args.num_kv_blocks or (self.args.max_num_seqs * per_seq)
            # physical block 0 is never handed out: table entries that were never set (the scratch row, positions a padded
            # prefill piece writes past a sequence's end) point at it
            self.cache = PagedKVCache(c.num_hidden_layers, nb + 1, c.num_key_value_heads, BS, c.head_dim,
                                      self.model.embed_tokens.dtype, self.device, self.args.max_num_seqs + 1, self.args.max_model_len)
            self._free_blocks: List[int] = list(range(nb, 0, -1))
            self._block_ref: List[int] = [0] * (nb + 1)
            self._slot_blocks: Dict[int, List[int]] = {}
            self._reserved = 0
            if self.args.prefix_block % BS:
                self.args.prefix_block = -(-self.args.prefix_block // BS) * BS      # prefixes are shared in whole blocks
        else:
            self.cache = KVCache(c.num_hidden_layers, self.args.max_num_seqs + 1, c.num_key_value_heads, self.args.max_model_len,
                                 c.head_dim, self.model.embed_tokens.dtype, self.device)
        self.free: List[int] = list(range(self.args.max_num_seqs))[::-1]
        self.waiting: Deque[Request] = deque()
        self.running: List[Request] = []
        self.prefilling: List[Request] = []                 # admitted, prompt only partly in the cache (chunked prefill)
        # prefix cache: rows of finished sequences stay valid until their slot is needed (LRU), indexed by the hash of
        # every whole block-prefix of their tokens
        self._cached: "OrderedDict[int, bool]" = OrderedDict()            # slots of finished sequences, least recently used first
        self._prefix_index: Dict[int, int] = {}                           # hash(tokens[:k * block]) -> slot
        self._row_tokens: Dict[int, List[int]] = {}                       # slot -> tokens with valid K/V (cached or running rows)
        self.metrics = EngineMetrics()
        # host tier: key -> (tokens whose K/V the entry holds, bytes); block-prefix hash -> key
        self._host_rows: "OrderedDict[int, Tuple[List[int], int]]" = OrderedDict()
        self._host_index: Dict[int, int] = {}
        self._host_bytes, self._host_budget = 0, int(self.args.kv_offload_gb * 2 ** 30)
        self._host_keys = itertools.count()
        self._host_pin = -1
        self._kv_token_bytes = 2 * c.num_hidden_layers * c.num_key_value_heads * c.head_dim * self.model.embed_tokens.element_size()
        self._ids = itertools.count()
        self._cv = threading.Condition()
        self._stop = False
        self._thread: Optional[threading.Thread] = None
        self._rng = None
        if self.args.seed is not None:
            self._rng = torch.Generator(device=self.device).manual_seed(self.args.seed)
        self.model.eval()

    # ------------------------------------------------------------------ client side
    @torch.no_grad()
    def warmup(self) -> None:
        """Capture the CUDA graphs of every prefill bucket and decode bucket up front (dummy tokens in the scratch cache
        row), so no request pays a capture (~0.1 s each) in its time to first token."""
        if self.device.type != "cuda":
            return
        scratch = self.args.max_num_seqs
        for tb in self.gen._PREFILL_BUCKETS:
            if tb < self.args.max_model_len:
                self.gen._prefill_packed([[0] * tb], self.cache, [scratch], scratch_slot=scratch)
        nb = 1
        while True:
            z = torch.zeros(nb, dtype=torch.int32, device=self.device)
            self.gen.decode_graphed(torch.zeros(nb, dtype=torch.long, device=self.device), self.cache, z,
                                    torch.full((nb,), scratch, dtype=torch.int32, device=self.device))
            if nb >= self.args.max_num_seqs:
                break
            nb *= 2
        torch.cuda.synchronize()

    def abort(self, r: Request) -> None:
        """Stop generating for ``r`` (client went away, or a stop string matched in the server layer): a waiting request
        finishes immediately, a running one leaves the batch and frees its KV slot at the next iteration
        (``finish_reason = "abort"``)."""
        with self._cv:
            if r.done.is_set():
                return
            r.aborted = True
            if r in self.waiting:
                self.waiting.remove(r)
                self._finish(r, "abort")
            self._cv.notify_all()

    def submit(self, prompt_ids: Sequence[int], cfg: Optional[GenerationConfig] = None) -> Request:
        cfg = cfg or GenerationConfig()
        ids = list(prompt_ids)
        if len(ids) == 0:
            raise ValueError("empty prompt")
        if len(ids) >= self.args.max_model_len:
            raise ValueError(f"prompt of {len(ids)} tokens does not fit max_model_len={self.args.max_model_len}")
        if self.paged:
            pool = len(self._block_ref) - 1
            if self._nblocks(len(ids) + 1) > pool:
                raise ValueError(f"prompt of {len(ids)} tokens does not fit the KV block pool ({pool} blocks of {self.args.block_size})")
            if self._nblocks(min(len(ids) + cfg.max_new_tokens, self.args.max_model_len)) > pool:
                # the pool could never promise this request its worst case: cap the generation at what the pool holds
                import dataclasses
                room = pool * self.args.block_size - 8 - len(ids)
                cfg = dataclasses.replace(cfg, max_new_tokens=max(1, room))
        r = Request(next(self._ids), ids, cfg)
        with self._cv:
            self.waiting.append(r)
            self._cv.notify()
        return r

    def generate(self, prompts: Sequence[Sequence[int]], cfg: Optional[GenerationConfig] = None) -> List[Request]:
        """Synchronous helper: submit everything, drive the scheduler in this thread (or wait for the background
        thread) until all requests finished."""
        reqs = [self.submit(p, cfg) for p in prompts]
        if self._thread is None:
            while not all(r.done.is_set() for r in reqs):
                self.step()
        else:
            for r in reqs:
                r.done.wait()
        return reqs

    def start(self) -> "LLMEngine":
        if self._thread is None:
            if self.args.warmup:
                self.warmup()
            self._stop = False
            self._thread = threading.Thread(target=self._loop, name="b200-engine", daemon=True)
            self._thread.start()
        return self

    def shutdown(self) -> None:
        with self._cv:
            self._stop = True
            self._cv.notify_all()
        if self._thread is not None:
            self._thread.join(timeout=30)
            self._thread = None

    @property
    def num_waiting(self) -> int:
        return len(self.waiting)

    @property
    def num_running(self) -> int:
        return len(self.running)

    def kv_cache_usage(self) -> float:
        if self.paged:
            total = len(self._block_ref) - 1
            return (total - len(self._free_blocks)) / float(max(1, total))
        used = sum(len(r.prompt_ids) + len(r.output_ids) for r in self.running)
        return used / float(self.args.max_num_seqs * self.args.max_model_len)

    # ------------------------------------------------------------------ paged KV cache: block allocator
    def _nblocks(self, n_tokens: int) -> int:
        """Blocks that hold ``n_tokens`` plus the <= 7 dummy positions a padded prefill piece may write past them."""
        return -(-(min(n_tokens + 8, self.args.max_model_len)) // self.args.block_size)

    def _blocks_free(self) -> int:
        return len(self._free_blocks) - self._reserved

    def _drop_blocks(self, slot: int) -> None:
        for b in self._slot_blocks.pop(slot, []):
            self._block_ref[b] -= 1
            if self._block_ref[b] == 0:
                self._free_blocks.append(b)

    def _truncate_blocks(self, slot: int, n_blocks: int) -> None:
        blocks = self._slot_blocks.get(slot, [])
        for b in blocks[n_blocks:]:
            self._block_ref[b] -= 1
            if self._block_ref[b] == 0:
                self._free_blocks.append(b)
        del blocks[n_blocks:]

    def _evict_cached_rows(self, need: int, keep: int = -1) -> None:
        """Give finished sequences' rows back (least recently used first) until ``need`` blocks are free."""
        for slot in list(self._cached):
            if self._blocks_free() >= need:
                return
            if slot != keep:
                del self._cached[slot]
                self._offload(slot)
                self._unindex_row(slot)
                self._drop_blocks(slot)
                self.free.append(slot)

    def _admit_blocks(self, r: Request, shared_from: int = -1, shared_tokens: int = 0) -> bool:
        """Reserve every block ``r`` can need (prompt + max_new_tokens, less the prefix blocks it shares) so that decode
        never runs out; returns False when the pool cannot promise that yet."""
        total = self._nblocks(min(len(r.prompt_ids) + r.cfg.max_new_tokens, self.args.max_model_len))
        have = len(self._slot_blocks.get(r.slot, [])) if r.slot >= 0 else 0
        shared = shared_tokens // self.args.block_size if shared_from >= 0 else 0
        need = max(0, total - max(have, shared))
        if self._blocks_free() < need:
            self._evict_cached_rows(need, keep=shared_from if shared_from >= 0 else r.slot)
        if self._blocks_free() < need:
            return False
        r.reserved_blocks = need
        self._reserved += need
        return True

    def _share_prefix(self, src: int, dst: int, n_tokens: int) -> None:
        blocks = self._slot_blocks[src][:n_tokens // self.args.block_size]
        self._drop_blocks(dst)
        for b in blocks:
            self._block_ref[b] += 1
        self._slot_blocks[dst] = list(blocks)
        self.gen.set_block_row(self.cache, dst, blocks)

    def _grow(self, r: Request, n_tokens: int) -> None:
        """Make sure ``r``'s row has blocks for ``n_tokens`` positions (taken out of its reservation)."""
        blocks = self._slot_blocks.setdefault(r.slot, [])
        want = self._nblocks(n_tokens)
        if len(blocks) >= want:
            return
        first_new = len(blocks)
        while len(blocks) < want:
            if not self._free_blocks:
                raise RuntimeError("KV block pool exhausted (reservation accounting is broken)")
            b = self._free_blocks.pop()
            self._block_ref[b] = 1
            blocks.append(b)
            if r.reserved_blocks > 0:
                r.reserved_blocks -= 1
                self._reserved -= 1
        self.gen.set_block_row(self.cache, r.slot, blocks) if first_new == 0 else self._set_row_tail(r.slot, blocks, first_new)

    def _set_row_tail(self, slot: int, blocks: List[int], first_new: int) -> None:
        self.gen.set_block_row(self.cache, slot, blocks)          # rows are short (max_model_len / block_size ints)

    def _unreserve(self, r: Request) -> None:
        self._reserved -= r.reserved_blocks
        r.reserved_blocks = 0

    # ------------------------------------------------------------------ prefix cache (enable_prefix_caching)
    def _block_hashes(self, toks: Sequence[int], upto: int):
        """(n, hash of toks[:n]) for every whole block-prefix n = block, 2 * block, ... <= upto (chained hashes)."""
        B = self.args.prefix_block
        h = 0
        for k in range(1, upto // B + 1):
            h = hash((h, tuple(toks[(k - 1) * B:k * B])))
            yield k * B, h

    def _index_row(self, slot: int, toks: List[int]) -> None:
        """Row ``slot`` holds valid K/V for ``toks`` (a running sequence's prompt, or a finished sequence)."""
        self._unindex_row(slot)
        self._row_tokens[slot] = toks
        for _, h in self._block_hashes(toks, len(toks)):
            self._prefix_index[h] = slot

    def _unindex_row(self, slot: int) -> None:
        toks = self._row_tokens.pop(slot, None)
        if toks is not None:
            for _, h in self._block_hashes(toks, len(toks)):
                if self._prefix_index.get(h) == slot:
                    del self._prefix_index[h]

    def _lookup_prefix(self, prompt: List[int]):
        """Longest cached block-prefix of ``prompt`` that leaves at least one token to compute: (slot, n) or (-1, 0)."""
        best = (-1, 0)
        for n, h in self._block_hashes(prompt, len(prompt) - 1):
            slot = self._prefix_index.get(h)
            if slot is None:
                break
            if self._row_tokens.get(slot, [])[:n] == prompt[:n]:        # guards against hash collisions / stale entries
                best = (slot, n)
            else:
                break
        return best

    def _take_slot(self, keep: int = -1) -> int:
        """A free KV row, else the least recently used cached one (never ``keep``, the row a prefix is copied from)."""
        if self.free:
            return self.free.pop()
        for slot in self._cached:
            if slot != keep:
                del self._cached[slot]
                self._offload(slot)
                self._unindex_row(slot)
                if self.paged:
                    self._drop_blocks(slot)
                return slot
        return -1

    # ------------------------------------------------------------------ host tier of the prefix cache (kv_offload_gb)
    def _drop_host(self, key: int) -> None:
        toks, nbytes = self._host_rows.pop(key)
        self._host_bytes -= nbytes
        for _, h in self._block_hashes(toks, len(toks)):
            if self._host_index.get(h) == key:
                del self._host_index[h]
        self.gen.drop_host_row(key)

    def _offload(self, slot: int) -> None:
        """A cached row is about to lose its HBM: keep its whole-block prefix in host memory (least recently used out)."""
        if self._host_budget <= 0:
            return
        toks = self._row_tokens.get(slot)
        B = self.args.prefix_block
        n = (len(toks) // B * B) if toks else 0
        nbytes = n * self._kv_token_bytes
        if n < B or nbytes > self._host_budget:
            return
        hashes = list(self._block_hashes(toks, n))
        if hashes[-1][1] in self._host_index:                 # the host tier already holds this prefix (or a longer one)
            self._host_rows.move_to_end(self._host_index[hashes[-1][1]])
            return
        while self._host_bytes + nbytes > self._host_budget:
            victim = next((k for k in self._host_rows if k != self._host_pin), None)
            if victim is None:
                return
            self._drop_host(victim)
        key = next(self._host_keys)
        self.gen.offload_row(self.cache, slot, n, key)
        self._host_rows[key] = (list(toks[:n]), nbytes)
        self._host_bytes += nbytes
        for _, h in hashes:
            self._host_index[h] = key
        with self.metrics.lock:
            self.metrics.host_offloads += 1

    def _lookup_host(self, prompt: List[int]):
        """Longest block-prefix of ``prompt`` the host tier holds (leaving one token to compute): (key, n) or (-1, 0)."""
        best = (-1, 0)
        if not self._host_rows:
            return best
        for n, h in self._block_hashes(prompt, len(prompt) - 1):
            key = self._host_index.get(h)
            if key is None or self._host_rows[key][0][:n] != prompt[:n]:
                break
            best = (key, n)
        return best

    def _slots_available(self) -> int:
        return len(self.free) + len(self._cached)

    def _release_slot(self, r: Request) -> None:
        """The request is done with its row: keep it as a cached prefix (LRU) or hand it back."""
        slot, r.slot = r.slot, -1
        if slot < 0:
            return
        if self.paged:
            self._unreserve(r)
        valid = r.prompt_ids[:r.prefilled] if r.prefilled < len(r.prompt_ids) else r.prompt_ids + r.output_ids[:-1]
        if self.args.enable_prefix_caching and len(valid) >= self.args.prefix_block and r.finish_reason != "error":
            self._index_row(slot, list(valid))
            self._cached[slot] = True
            self._cached.move_to_end(slot)
        else:
            self._unindex_row(slot)
            if self.paged:
                self._drop_blocks(slot)
            self.free.append(slot)

    # ------------------------------------------------------------------ scheduler
    def _loop(self) -> None:
        if self.device.type == "cuda":
            torch.cuda.set_device(self.device)
        while True:
            with self._cv:
                while not self._stop and not self.waiting and not self.running and not self.prefilling:
                    self._cv.wait(timeout=0.5)
                if self._stop:
                    break
            try:
                self.step()
            except Exception as e:  # fail the in-flight requests, keep serving
                for r in self.running + self.prefilling + list(self.waiting):
                    r.error = f"{type(e).__name__}: {e}"
                    self._finish(r, "error")
                self.running.clear()
                self.prefilling.clear()
                self.waiting.clear()
                self.free = list(range(self.args.max_num_seqs))[::-1]
                self._cached.clear()
                self._prefix_index.clear()
                self._row_tokens.clear()

    def _eos(self, cfg: GenerationConfig) -> List[int]:
        if cfg.ignore_eos:
            return []
        e = cfg.eos_token_id
        if e is None:
            t = getattr(self.tokenizer, "eos_token_id", None)
            return [] if t is None else [t]
        return [e] if isinstance(e, int) else list(e)

    def _finish(self, r: Request, reason: str) -> None:
        r.finish_reason = reason
        r.finish_time = time.perf_counter()
        self._release_slot(r)
        with self.metrics.lock:
            self.metrics.requests_finished[reason] = self.metrics.requests_finished.get(reason, 0) + 1
            self.metrics.e2e_sum += r.finish_time - r.arrival
        r.stream.put(None)
        r.done.set()

    def _emit(self, r: Request, tok: int) -> bool:
        """Record one sampled token; returns True when the request is finished."""
        now = time.perf_counter()
        if r.first_token_time is None:
            r.first_token_time = now
            self.metrics.observe_ttft(now - r.arrival)
        if tok in self._eos(r.cfg):
            self._finish(r, "stop")
            return True
        r.output_ids.append(tok)
        r.token_times.append(now)
        r.stream.put(tok)
        with self.metrics.lock:
            self.metrics.generation_tokens += 1
        if len(r.output_ids) >= r.cfg.max_new_tokens or len(r.prompt_ids) + len(r.output_ids) >= self.args.max_model_len:
            self._finish(r, "length")
            return True
        return False

    def _sample(self, logits: torch.Tensor, reqs: List[Request]) -> List[int]:
        """Per-request sampling parameters: requests sharing a config are sampled together."""
        out = [0] * len(reqs)
        groups: Dict[tuple, List[int]] = {}
        for i, r in enumerate(reqs):
            c = r.cfg
            groups.setdefault((c.do_sample, c.temperature, c.top_p, c.top_k, c.repetition_penalty), []).append(i)
        for key, idx in groups.items():
            cfg = reqs[idx[0]].cfg
            prev = None
            if cfg.repetition_penalty != 1.0:
                seqs = [reqs[i].prompt_ids + reqs[i].output_ids for i in idx]
                L = max(len(s) for s in seqs)
                prev = torch.tensor([s + [s[-1]] * (L - len(s)) for s in seqs], device=logits.device)
            ids = sample_next(logits[idx], cfg, prev, self._rng)
            for i, t in zip(idx, ids.tolist()):
                out[i] = t
        return out

    def _start_running(self, r: Request) -> None:
        self.running.append(r)
        if self.args.enable_prefix_caching:
            self._index_row(r.slot, r.prompt_ids)          # later prompts may copy this (still running) row's prefix

    def _prefill_piece(self, r: Request, budget: int) -> int:
        """Prefill the next piece (at most ``budget`` tokens when chunking is on) of a partly cached prompt; on the last
        piece sample the first token and move the request to the decode batch. Returns the tokens processed."""
        L = len(r.prompt_ids)
        n = L - r.prefilled
        if self.args.enable_chunked_prefill:
            n = min(n, max(budget, 8))
            if r.prefilled + n < L:
                n = max(8, n // 8 * 8)                       # pieces end on 8-token boundaries (aligned score rows)
        ids = torch.tensor(r.prompt_ids[r.prefilled:r.prefilled + n], device=self.device)
        if self.paged:
            self._grow(r, r.prefilled + n)
        logits = self.gen._prefill_chunk(ids, self.cache, r.slot, r.prefilled)
        r.prefilled += n
        with self.metrics.lock:
            self.metrics.prompt_tokens += n
        if r.prefilled >= L:
            self.prefilling.remove(r)
            tok = self._sample(logits, [r])[0]
            if not self._emit(r, tok):
                self._start_running(r)
        return n

    @torch.no_grad()
    def step(self) -> None:
        """One scheduler iteration: admit + prefill, then one batched decode step."""
        batch: List[Request] = []
        budget = self.args.max_prefill_tokens
        # (1) sequences whose prompt is only partly in the cache (chunked prefill / prefix hits): one more piece each
        for r in list(self.prefilling):
            if r.aborted:
                self.prefilling.remove(r)
                self._finish(r, "abort")
            elif budget > 0:
                budget -= self._prefill_piece(r, budget)
        # (2) admissions
        while (self.waiting and self._slots_available() and budget > 0
               and len(batch) + len(self.prefilling) < self.args.max_prefill_per_step):
            with self._cv:
                r = self.waiting[0]
                L = len(r.prompt_ids)
                src, P = self._lookup_prefix(r.prompt_ids) if self.args.enable_prefix_caching else (-1, 0)
                hkey, Ph = self._lookup_host(r.prompt_ids) if self.args.enable_prefix_caching else (-1, 0)
                from_host = Ph > P                            # the host tier knows a longer prefix than any HBM row
                piecewise = P > 0 or from_host or (L > budget and self.args.enable_chunked_prefill and not batch)
                if not piecewise and batch and L > budget:
                    break
                self.waiting.popleft()
            with self.metrics.lock:
                if not getattr(r, "_queued_counted", False):      # a request put back for lack of blocks is counted once
                    r._queued_counted = True
                    self.metrics.queue_sum += time.perf_counter() - r.arrival
                    self.metrics.queue_count += 1
                self.metrics.prefix_queries += L
                self.metrics.prefix_hits += max(P, Ph) if from_host else P
                self.metrics.host_hits += Ph if from_host else 0
            if from_host:
                self._host_pin = hkey                         # taking a slot may off-load another row: not at this entry's cost
                self._host_rows.move_to_end(hkey)
                r.slot = self._take_slot()
                self._host_pin = -1
                if self.paged:
                    if not self._admit_blocks(r):
                        self.free.append(r.slot)
                        r.slot = -1
                        with self._cv:
                            self.waiting.appendleft(r)
                        break
                    self._grow(r, Ph)
                self.gen.restore_row(self.cache, hkey, r.slot, Ph)
                r.prefilled = r.cached_tokens = Ph
                self.prefilling.append(r)
                budget -= self._prefill_piece(r, budget)
                continue
            if not piecewise:
                r.slot = self._take_slot()
                if self.paged:
                    if not self._admit_blocks(r):
                        self.free.append(r.slot)
                        r.slot = -1
                        with self._cv:
                            self.waiting.appendleft(r)        # the pool cannot promise its blocks yet: retry next iteration
                        break
                    self._grow(r, L)
                r.prefilled = L
                budget -= L
                batch.append(r)
                continue
            if P > 0 and src in self._cached:
                # the prefix lives in the row of a finished sequence: continue in that very row, nothing is copied
                del self._cached[src]
                self._unindex_row(src)
                r.slot = src
                if self.paged:
                    # positions >= P of the row are rewritten: blocks past the prefix go back to the pool (some may be listed
                    # by other sequences that share a longer prefix of the old row - those keep them)
                    self._truncate_blocks(src, P // self.args.block_size)
                    if not self._admit_blocks(r):
                        self._drop_blocks(src)
                        self.free.append(src)
                        r.slot = -1
                        with self._cv:
                            self.waiting.appendleft(r)
                        break
            else:
                r.slot = self._take_slot(keep=src)
                if self.paged:
                    if not self._admit_blocks(r, src if P > 0 else -1, P):
                        self.free.append(r.slot)
                        r.slot = -1
                        with self._cv:
                            self.waiting.appendleft(r)
                        break
                    if P > 0:
                        self._share_prefix(src, r.slot, P)               # both table rows list the prefix's physical blocks
                elif P > 0:
                    self.gen.copy_prefix(self.cache, src, r.slot, P)     # the prefix belongs to a running sequence: copy its K/V
            r.prefilled = r.cached_tokens = P
            self.prefilling.append(r)
            budget -= self._prefill_piece(r, budget)
        if batch:
            # the admitted prompts run as ONE packed pass (token-level batching): GEMMs see the sum of the prompt lengths
            if self.device.type == "cuda":
                # ... replayed from a CUDA graph per token-count bucket; padding tokens land in the scratch cache row
                logits = self.gen._prefill_packed([r.prompt_ids for r in batch], self.cache, [r.slot for r in batch],
                                                  scratch_slot=self.args.max_num_seqs)
            elif len(batch) == 1:
                r = batch[0]
                logits = self.gen._prefill(torch.tensor([r.prompt_ids], device=self.device), self.cache,
                                           torch.tensor([r.slot], device=self.device))
            else:
                logits = self.gen._prefill_packed([r.prompt_ids for r in batch], self.cache, [r.slot for r in batch])
            with self.metrics.lock:
                self.metrics.prompt_tokens += sum(len(r.prompt_ids) for r in batch)
            for r, tok in zip(batch, self._sample(logits, batch)):
                if not self._emit(r, tok):
                    self._start_running(r)
        if any(r.aborted for r in self.running):
            for r in [r for r in self.running if r.aborted]:
                self._finish(r, "abort")
            self.running = [r for r in self.running if not r.aborted]
        if not self.running:
            return
        reqs = self.running
        n = len(reqs)
        if self.paged:
            for r in reqs:
                self._grow(r, len(r.prompt_ids) + len(r.output_ids))
        nb = n
        if self.device.type == "cuda":          # pad to a power-of-two bucket: one captured graph per bucket, not per batch size
            nb = 1
            while nb < n:
                nb *= 2
        pad = nb - n
        scratch = self.args.max_num_seqs
        last = torch.tensor([r.output_ids[-1] for r in reqs] + [0] * pad, device=self.device)
        pos = torch.tensor([len(r.prompt_ids) + len(r.output_ids) - 1 for r in reqs] + [0] * pad, device=self.device, dtype=torch.int32)
        slots = torch.tensor([r.slot for r in reqs] + [scratch] * pad, device=self.device, dtype=torch.int32)
        logits = self.gen.decode_graphed(last, self.cache, pos, slots)[:n]
        toks = self._sample(logits, reqs)
        self.running = [r for r, t in zip(reqs, toks) if not self._emit(r, t)]
        self.metrics.steps += 1
