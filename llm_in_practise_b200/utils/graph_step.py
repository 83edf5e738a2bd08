"""A whole training step (forward, backward, optimizer, zero_grad) captured in ONE CUDA graph.

The reference's small models are launch-bound on a B200: MiniGPT2 (llm-demo/minigpt2/model.py, 1.9 M parameters) spends
6 ms per step issuing ~400 launches whose kernels take a fraction of that. Capturing the step removes the Python / launch
path from the loop. Two things would go wrong in a naive capture and are handled here:

* **dropout** — Philox seeds are kernel arguments and would be frozen: every replay would drop the same elements. The
  library adds a per-step offset that lives in device memory (``b200_set_graph_state``); the graph starts with a copy of
  that offset from pinned host memory, which ``__call__`` advances before every replay. Forward and backward of one replay
  see the same offset, so the regenerated masks agree.
* **AdamW** — the learning rate and the bias corrections are kernel arguments too: ``FusedAdamW.enable_graph_mode`` moves them
  to device memory the same way.

Static shapes only; single process (the overlapped DDP / FSDP collectives take their epochs from the host); models whose
step does not read anything back to the host (the framework's losses normalise on the device).
"""
from __future__ import annotations

import ctypes
from typing import Callable, Optional, Sequence

import torch

from ..ops import _native as nv

_SEED_STRIDE = 0x9E3779B97F4A7C15 & 0x7FFFFFFFFFFFFFFF
_keep = []          # the registered seed-offset tensors must outlive every kernel that may read them


class GraphedTrainStep:
    def __init__(self, loss_fn: Callable[..., torch.Tensor], optimizer, example_inputs: Sequence[torch.Tensor],
                 zero_grad: Optional[Callable[[], None]] = None, warmup: int = 3):
        if not (torch.cuda.is_available() and example_inputs and example_inputs[0].is_cuda):
            raise RuntimeError("GraphedTrainStep needs CUDA tensors")
        dev = example_inputs[0].device
        self.loss_fn, self.opt = loss_fn, optimizer
        self.zero_grad = zero_grad or (lambda: optimizer.zero_grad(set_to_none=False))
        self.static_in = [t.clone() for t in example_inputs]
        self.seed_host = torch.zeros(1, dtype=torch.int64).pin_memory()
        self.seed_dev = torch.zeros(1, dtype=torch.int64, device=dev)
        _keep.append((self.seed_host, self.seed_dev))
        fn = nv.proto("b200_set_graph_state", [ctypes.c_void_p])
        nv.check(fn(self.seed_dev.data_ptr()), "set_graph_state")
        if hasattr(optimizer, "enable_graph_mode"):
            optimizer.enable_graph_mode()
        # torch's rule for whole-network capture: warm up on a side stream first (lazy initialisations, workspaces,
        # cudaFuncSetAttribute calls and the autograd graph's buffers all happen here, not under capture)
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(warmup):
                self._advance()
                self._body()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        step_before = getattr(optimizer, "_step", None)
        with torch.cuda.graph(self.graph):
            self.loss = self._body()
        if step_before is not None:
            optimizer._step = step_before          # capture records the step, it does not run it
        self.replays = 0

    def _advance(self) -> None:
        self.seed_host[0] = (int(self.seed_host[0]) + _SEED_STRIDE) & 0x7FFFFFFFFFFFFFFF

    def _body(self) -> torch.Tensor:
        self.seed_dev.copy_(self.seed_host, non_blocking=True)      # memcpy node: re-read from pinned memory at every replay
        loss = self.loss_fn(*self.static_in)
        loss.backward()
        self.opt.step()
        self.zero_grad()
        return loss.detach()

    def __call__(self, *inputs: torch.Tensor) -> torch.Tensor:
        for s, x in zip(self.static_in, inputs):
            if x is not s:
                s.copy_(x, non_blocking=True)
        self._advance()
        if hasattr(self.opt, "advance_graph_step"):
            self.opt.advance_graph_step()
        self.graph.replay()
        self.replays += 1
        return self.loss
