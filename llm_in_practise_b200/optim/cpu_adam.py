"""ZeRO-Offload optimizer: AdamW state in pinned host memory, stepped by the C++ kernel in
``csrc/host/cpu_adam.cpp`` (the role of ``DeepSpeedCPUAdam`` in
LLM_Distributed_Trainning/DeepSpeed/DeepSpeed-GPTLike-ZeRO-Offload/DeepSpeed-GPTLike-ZeRO-Offload.py:27,228 with
``offload_optimizer: {device: cpu, pin_memory: true}``). Per step: D2H of the owned gradient shard, host AdamW on
the fp32 master shard, H2D of the updated shard. On a 180 GB B200 this is a supported-but-discouraged mode."""
from __future__ import annotations

import ctypes
import math
from typing import Optional

import torch

from ..ops import _native as nv
from ..utils.arena import ParamArena


class CPUAdamW(torch.optim.Optimizer):
    eight_bit = False

    def __init__(self, arena: ParamArena, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.01,
                 shard: Optional[slice] = None, pin_memory: bool = True):
        self.arena = arena
        self.shard = shard or slice(0, arena.numel)
        anchor = torch.nn.Parameter(arena.data[:1].detach(), requires_grad=False)
        super().__init__([anchor], dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        n = self.shard.stop - self.shard.start
        pin = pin_memory and torch.cuda.is_available()
        mk = lambda: torch.zeros(n, dtype=torch.float32, pin_memory=pin)  # noqa: E731
        self.master, self.grad_host, self.m, self.v = mk(), mk(), mk(), mk()
        self.master.copy_(arena.data[self.shard].detach().float().cpu())
        self._step = 0
        self.grad_scale = 1.0
        self.dyn_scale = None
        self.found_inf = None
        self._ranges = None

    @torch.no_grad()
    def step(self, closure=None):
        if self.found_inf is not None and float(self.found_inf) != 0.0:
            return None
        self._step += 1
        g = self.param_groups[0]
        self.grad_host.copy_(self.arena.grad[self.shard], non_blocking=False)
        gs = self.grad_scale * (float(self.dyn_scale) if self.dyn_scale is not None else 1.0)
        b1, b2 = g["betas"]
        L = nv.host_lib()
        if L is not None and hasattr(L, "b200_cpu_adamw"):
            L.b200_cpu_adamw.restype = None
            L.b200_cpu_adamw.argtypes = [ctypes.c_void_p] * 5 + [ctypes.c_int64] + [ctypes.c_float] * 5 + \
                                        [ctypes.c_int64, ctypes.c_float]
            L.b200_cpu_adamw(self.master.data_ptr(), self.grad_host.data_ptr(), self.m.data_ptr(), self.v.data_ptr(), None,
                             self.master.numel(), g["lr"], b1, b2, g["eps"], g["weight_decay"], self._step, gs)
        else:
            gf = self.grad_host * gs
            self.m.mul_(b1).add_(gf, alpha=1 - b1)
            self.v.mul_(b2).addcmul_(gf, gf, value=1 - b2)
            bc1, bc2 = 1 - b1 ** self._step, 1 - b2 ** self._step
            self.master.mul_(1 - g["lr"] * g["weight_decay"])
            self.master.addcdiv_(self.m, self.v.sqrt() / math.sqrt(bc2) + g["eps"], value=-g["lr"] / bc1)
        self.arena.data[self.shard].copy_(self.master.to(self.arena.data.dtype), non_blocking=False)
        torch.autograd.graph.increment_version(self.arena.data)
        return None

    def zero_grad(self, set_to_none: bool = False):
        self.arena.zero_grad()

    def state_dict(self):
        return {"step": self._step, "m": self.m.clone(), "v": self.v.clone(), "master": self.master.clone(),
                "param_groups": [{k: v for k, v in g.items() if k != "params"} for g in self.param_groups]}

    def load_state_dict(self, sd):
        self._step = sd["step"]
        self.m.copy_(sd["m"]); self.v.copy_(sd["v"]); self.master.copy_(sd["master"])
        for g, s in zip(self.param_groups, sd["param_groups"]):
            g.update(s)
