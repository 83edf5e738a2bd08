"""DistributedDataParallel with bucketed gradient all-reduce overlapped with backward (K-C1).

Call shape of ``torch.nn.parallel.DistributedDataParallel`` as used in
LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py:271-279 (``device_ids``, ``output_device``,
``find_unused_parameters=False``), ``.module`` access (:322-326), ``no_sync()`` for gradient accumulation
(…/temp/ddp_gpt_bpe_tokenizer_02.py:399-419) and the constructor-time parameter broadcast (C2).

B200 design: trainable parameters and gradients are re-homed into one flat arena (symmetric peer memory on
>1 GPU). Buckets are contiguous arena ranges in reverse registration order (the order backward produces
gradients), ``bucket_cap_mb`` each. When the last gradient of a bucket lands — either through autograd's
accumulate hook or through the ``_grad_ready`` callback of the wgrad GEMM that wrote ``main_grad`` directly — the
bucket's all-reduce is launched on a side stream: one two-shot kernel over NVLink peer memory that sums, applies
1/world and writes the result into every replica (``b200_allreduce``), while backward keeps running on the main
stream. ``torch.distributed.all_reduce`` is used instead on CPU/gloo, without symmetric memory, or when
``B200_COMM=nccl`` (the baseline to compare against).
"""
from __future__ import annotations

import contextlib
import os
from typing import List, Optional

import torch
import torch.distributed as dist
from torch import nn

from ..utils.arena import ParamArena, flatten_module, grad_landed, reset_uses
from .dist import get_info
from .symm import PeerFlags, SymmBuffer, allreduce_, symm_available


class _Bucket:
    __slots__ = ("lo", "hi", "params", "pending", "launched", "ready", "work")

    def __init__(self, lo: int, hi: int):
        self.lo, self.hi = lo, hi
        self.params: List[nn.Parameter] = []
        self.pending = 0
        self.launched = False
        self.ready = False
        self.work = None


class DistributedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, device_ids=None, output_device=None, find_unused_parameters: bool = False,
                 bucket_cap_mb: float = 25.0, broadcast_buffers: bool = True, process_group=None,
                 gradient_as_bucket_view: bool = True, mixed_precision: Optional[torch.dtype] = None):
        super().__init__()
        self.mixed = mixed_precision
        self.module = module
        self.info = get_info()
        self.world = self.info.world_size if (dist.is_available() and dist.is_initialized()) else 1
        self.rank = self.info.rank
        self.find_unused = find_unused_parameters
        self.require_sync = True
        dev = next(module.parameters()).device
        self.device = dev
        self.use_symm = (self.world > 1 and dev.type == "cuda" and symm_available()
                         and os.environ.get("B200_COMM", "symm") == "symm")
        self._bufs: List[SymmBuffer] = []
        alloc = None
        if self.use_symm:
            def alloc(numel, dtype, device):
                b = SymmBuffer(numel, dtype, device)
                self._bufs.append(b)
                return b.local
        if mixed_precision is not None:
            # bf16/fp16 working weights, fp32 masters + fp32 main grads (the all-reduce runs on the fp32 arena)
            from ..utils.arena import flatten_module_mixed
            self.arena = flatten_module_mixed(module, mixed_precision, pad_to=4 * max(self.world, 1), alloc=alloc)
        else:
            self.arena: ParamArena = flatten_module(module, pad_to=4 * max(self.world, 1), alloc=alloc)
        if self.use_symm:
            self.grad_buf = self._bufs[1]
            self.flags = PeerFlags(dev)
            self.comm_stream = torch.cuda.Stream(device=dev)
            from ..ops.gemm import overlapped_collectives
            overlapped_collectives(True)        # collective kernels will overlap GEMMs: no stream-K (see ops/gemm.py)
        elif dev.type == "cuda":
            self.comm_stream = torch.cuda.Stream(device=dev)
        else:
            self.comm_stream = None
        if self.world > 1:
            dist.broadcast(self.arena.data, src=0)                      # C2: every replica starts from rank 0
            if mixed_precision is not None:
                self.arena.lowp.copy_(self.arena.data)
            if broadcast_buffers:
                for b in module.buffers():
                    dist.broadcast(b, src=0)
        self._build_buckets(int(bucket_cap_mb * 1024 * 1024))
        self._next_bucket = 0
        self._callback_queued = False
        self._exposed_events: List[tuple] = []

    # ------------------------------------------------------------------ buckets
    def _build_buckets(self, cap_bytes: int) -> None:
        es = self.arena.data.element_size()
        # module parameters are views of the working-weight buffer: the fp32 arena itself, or its low-precision twin
        lowp = getattr(self.arena, "lowp", None)
        pbase, pes = (lowp.data_ptr(), lowp.element_size()) if lowp is not None else (self.arena.data.data_ptr(), es)
        owners = {}
        for _, p in self.module.named_parameters():
            if p.requires_grad:
                owners.setdefault(p.data_ptr(), p)
        self.buckets: List[_Bucket] = []
        hi = self.arena.numel
        params: List[nn.Parameter] = []
        lo = hi
        for _, shape, off in reversed(list(self.arena.named_slots())):   # reverse registration order
            lo = off
            p = owners.get(pbase + off * pes)
            if p is not None:
                params.append(p)
            if (hi - lo) * es >= cap_bytes:
                bk = _Bucket(lo, hi)
                bk.params = params
                self.buckets.append(bk)
                hi, params = lo, []
        if params or hi > 0:
            bk = _Bucket(0, hi)
            bk.params = params
            self.buckets.append(bk)
        for bi, bk in enumerate(self.buckets):
            for p in bk.params:
                p._ddp_bucket = bi
                p._grad_ready = (lambda pp=p: self._on_grad_ready(pp, True))
                p.register_post_accumulate_grad_hook(lambda pp: self._on_grad_ready(pp, False))

    def _reset_buckets(self) -> None:
        for b in self.buckets:
            b.pending = set(id(p) for p in b.params)
            b.launched = False
            b.ready = not b.pending          # a bucket without parameters (padding) never blocks its successors
            b.work = None
        self._next_bucket = 0

    # ------------------------------------------------------------------ reduction
    def make_optimizer(self, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.01,
                       eight_bit: bool = False):
        """One-launch AdamW over the flat arena (updates fp32 masters and rewrites the low-precision copy)."""
        from ..optim.adamw import FusedAdamW, PagedAdamW8bit
        cls = PagedAdamW8bit if eight_bit else FusedAdamW
        return cls(self.arena, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay,
                   lowp_copy=getattr(self.arena, "lowp", None))

    def finish_local_grads(self) -> None:
        """World size 1 / no_sync boundary helper: fold low-precision gradients into the fp32 arena."""
        if self.mixed is not None:
            self.arena.fold()

    def _on_grad_ready(self, p: nn.Parameter, direct: bool = False) -> None:
        """One contribution to ``p``'s gradient has landed (``direct``: written into ``main_grad`` by an op's backward;
        otherwise accumulated by autograd). A parameter used several times in the forward — a tied embedding / lm_head —
        is complete only after its last contribution (``utils.arena.note_use`` counts the uses during forward)."""
        if not grad_landed(p, direct):
            return
        if self.world == 1 or not self.require_sync:
            return
        if not self._callback_queued:
            self._reset_buckets()
            self._callback_queued = True
            torch.autograd.Variable._execution_engine.queue_callback(self._finalize_backward)
        b = self.buckets[p._ddp_bucket]
        b.pending.discard(id(p))
        if not b.pending:
            b.ready = True
        # Buckets are reduced strictly in index order, like torch's Reducer: the collectives of one step share the
        # epoch-flag channels, so every rank has to issue the same sequence even when gradients become ready in a
        # different order on different ranks (data-dependent MoE routing, unused parameters). A ready bucket waits
        # for its predecessors; whatever is left is flushed, in order, when backward ends.
        while self._next_bucket < len(self.buckets) and self.buckets[self._next_bucket].ready:
            nb = self.buckets[self._next_bucket]
            self._next_bucket += 1
            if not nb.launched:
                self._launch(nb)

    def _launch(self, b: _Bucket) -> None:
        b.launched = True
        n = b.hi - b.lo
        if n <= 0:
            return
        if self.mixed is not None:
            self.arena.fold(b.lo, b.hi)
        if self.comm_stream is not None:
            self.comm_stream.wait_stream(torch.cuda.current_stream(self.device))
        if self.use_symm:
            with torch.cuda.stream(self.comm_stream):
                allreduce_(self.grad_buf, self.flags, b.lo, n, 1.0 / self.world)
        else:
            g = self.arena.grad[b.lo:b.hi]
            if self.comm_stream is not None:
                with torch.cuda.stream(self.comm_stream):
                    b.work = dist.all_reduce(g, op=dist.ReduceOp.SUM, async_op=True)
            else:
                b.work = dist.all_reduce(g, op=dist.ReduceOp.SUM, async_op=True)

    def _finalize_backward(self) -> None:
        self._callback_queued = False
        reset_uses(p for b in self.buckets for p in b.params)
        # buckets whose parameters never produced a gradient this step (unused / frozen paths) still need
        # their (zero) contribution reduced so every rank launches the same sequence of collectives
        for b in self.buckets:
            if not b.launched:
                self._launch(b)
        for b in self.buckets:
            if b.work is not None:
                b.work.wait()
                if self.comm_stream is not None:
                    with torch.cuda.stream(self.comm_stream):
                        self.arena.grad[b.lo:b.hi].div_(self.world)
                else:
                    self.arena.grad[b.lo:b.hi].div_(self.world)
        if self.comm_stream is not None:
            cur = torch.cuda.current_stream(self.device)
            # exposed communication = how long the compute stream sits in this wait (events are read lazily)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(cur)
            cur.wait_stream(self.comm_stream)
            e1.record(cur)
            self._exposed_events.append((e0, e1))
            if len(self._exposed_events) > 64:
                self._exposed_events.pop(0)

    @property
    def exposed_ms(self) -> List[float]:
        """Per-step time (ms) the compute stream waited for the gradient all-reduce after backward had finished."""
        out = []
        for e0, e1 in self._exposed_events:
            if e1.query():
                out.append(e0.elapsed_time(e1))
        return out

    # ------------------------------------------------------------------ module surface
    def forward(self, *args, **kwargs):
        reset_uses(p for b in self.buckets for p in b.params)      # use counts of a pass without backward (eval) are stale
        return self.module(*args, **kwargs)

    @contextlib.contextmanager
    def no_sync(self):
        old = self.require_sync
        self.require_sync = False
        try:
            yield
        finally:
            self.require_sync = old

    def zero_grad(self, set_to_none: bool = False) -> None:
        self.arena.zero_grad()

    def state_dict(self, *a, **k):
        return {f"module.{n}": v for n, v in self.module.state_dict(*a, **k).items()}

    def load_state_dict(self, sd, strict: bool = True):
        sd = {(k[7:] if k.startswith("module.") else k): v for k, v in sd.items()}
        return self.module.load_state_dict(sd, strict=strict)

    def parameters(self, recurse: bool = True):
        return self.module.parameters(recurse)

    def named_parameters(self, *a, **k):
        for n, p in self.module.named_parameters(*a, **k):
            yield f"module.{n}", p
