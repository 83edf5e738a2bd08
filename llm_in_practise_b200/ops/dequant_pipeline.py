"""Expansion of 4-bit weights one GEMM ahead, on a side stream.

The dequant kernels are HBM-bound (4.5 TB/s for ~0.1 ms), the tcgen05 GEMMs are bound by the tensor pipe and the
shared-memory port and leave most of the DRAM bandwidth idle — so the next weight is expanded *while* the current GEMM
runs. Dequant blocks (256 threads, 64 B of shared memory) co-reside with the persistent GEMM CTAs on the same SMs.

The order of requests in a training step is the same every step (q|k|v, o, gate|up, down per layer, then the transposed
ones in reverse), so the pipeline simply remembers which (weight, layout) followed which last time and expands that one
into the other buffer as soon as the current request has been served. A misprediction costs nothing but the wasted
expansion: every request checks that its buffer really holds what it asked for and otherwise expands in line.

All expansions — predicted or not — run on the side stream (the compute stream waits on their event), which also gives
the ZeRO-3 shard prefetcher (parallel/zero.py) a single consumer stream.
"""
from __future__ import annotations

import os
from typing import Dict, Optional, Tuple

import torch

ENABLED = os.environ.get("B200_DEQUANT_AHEAD", "1") != "0"


class _Slot:
    __slots__ = ("buf", "key", "ready", "used")

    def __init__(self):
        self.buf: Optional[torch.Tensor] = None
        self.key = None
        self.ready = torch.cuda.Event()
        self.used = torch.cuda.Event()      # recorded on the compute stream once the consumer GEMM is enqueued


class DequantPipeline:
    def __init__(self, device: torch.device, nslots: int = 2):
        self.device = device
        self.stream = torch.cuda.Stream(device)
        self.slots = [_Slot() for _ in range(nslots)]
        self.next_of: Dict[tuple, Tuple[object, torch.dtype, bool]] = {}
        self.prev_key = None
        self.cur = -1                       # slot handed out by the last request
        self.hits = self.misses = 0

    def _expand(self, slot: _Slot, q, dtype, transpose: bool, key) -> None:
        n = q.N * q.K
        if slot.buf is None or slot.buf.dtype != dtype or slot.buf.numel() < n:
            slot.buf = torch.empty(n, dtype=dtype, device=self.device)
        self.stream.wait_event(slot.used)                                # the GEMM that last read this buffer is done
        with torch.cuda.stream(self.stream):
            shape = (q.K, q.N) if transpose else (q.N, q.K)
            q.dequantize(dtype, out=slot.buf[:n].view(shape), transpose=transpose)
            slot.ready.record(self.stream)
        slot.key = key

    def get(self, q, dtype: torch.dtype, transpose: bool) -> torch.Tensor:
        key = (id(q), dtype, transpose)
        comp = torch.cuda.current_stream(self.device)
        if self.cur >= 0:
            self.slots[self.cur].used.record(comp)                       # previous consumer is enqueued before this point
        hit = next((i for i, s in enumerate(self.slots) if s.key == key), None)
        if hit is None:
            self.misses += 1
            hit = (self.cur + 1) % len(self.slots)
            self._expand(self.slots[hit], q, dtype, transpose, key)
        else:
            self.hits += 1
        slot = self.slots[hit]
        comp.wait_event(slot.ready)
        if self.prev_key is not None:
            self.next_of[self.prev_key] = (q, dtype, transpose)
        self.prev_key, self.cur = key, hit
        nxt = self.next_of.get(key)
        if nxt is not None:
            nq, ndt, ntr = nxt
            nkey = (id(nq), ndt, ntr)
            other = (hit + 1) % len(self.slots)
            if self.slots[other].key != nkey:
                self._expand(self.slots[other], nq, ndt, ntr, nkey)
        n = q.N * q.K
        return slot.buf[:n].view((q.K, q.N) if transpose else (q.N, q.K))


_pipes: Dict[int, DequantPipeline] = {}


def pipeline(device: torch.device) -> DequantPipeline:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    p = _pipes.get(idx)
    if p is None:
        p = _pipes[idx] = DequantPipeline(torch.device("cuda", idx))
    return p


def reset() -> None:
    _pipes.clear()
