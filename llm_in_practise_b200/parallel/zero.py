"""ZeRO engine (stages 0-3) with a DeepSpeed-shaped surface.

    engine = initialize(model=model, config="ds_config.json" | dict, lora_config=...)   # or ds_initialize(...) -> 4-tuple
    loss = engine(x, labels=y)["loss"];  engine.backward(loss);  engine.step()
    engine.save_checkpoint(dir, tag);  engine.module;  engine.train_micro_batch(ids, labels)

Reference behaviour reproduced: ``deepspeed.initialize`` + ``engine(x) / engine.backward / engine.step`` loops
of LLM_Distributed_Trainning/DeepSpeed/DeepSpeed-GPTLike-ZeRO-{1,2,3,Offload}/*.py, and the HF-Trainer +
``deepspeed="ds_zero3_config.json"`` path of Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:132-172.

How the stages map onto B200 (one process per GPU, NVLink peer memory):

* trainable parameters, gradients and optimizer state live in flat fp32 arenas (``utils/arena.py``);
  on >1 GPU the parameter and gradient arenas are symmetric memory.
* stage 0: gradients are all-reduced in place by the two-shot peer kernel (K-C1); every rank steps.
* stage 1/2/3: ONE kernel per optimizer step (K-C2, ``b200_zero_step``) does reduce-scatter -> global
  norm / clip -> AdamW (fp32 or 8-bit state) on the owned 1/W shard -> writes the new parameters into
  every replica. Optimizer state exists only for the owned shard (that is the ZeRO-1 memory saving);
  the full-size gradient arena is transient accumulation space (stage 2's saving is moot when the
  reduce happens once per step out of the accumulation buffer).
* stage 3: frozen 4-bit base weights are row-sharded across ranks in symmetric memory and never
  gathered into a packed copy: the dequantisation kernel that feeds each GEMM pulls the rows it
  does not own straight from the peers (K-C3; ``ShardedNF4``). Trainable tensors below
  ``stage3_param_persistence_threshold`` stay replicated — every LoRA matrix does, as in DeepSpeed.
* ``offload_optimizer: cpu`` keeps the AdamW state in pinned host memory and steps with the C++
  CPU AdamW (``csrc/host/cpu_adam.cpp``): supported for parity with the ZeRO-Offload lesson, not a
  mode a 180 GB B200 needs.

Without CUDA (gloo) or without symmetric memory the same engine runs on ``torch.distributed``
collectives — that path is the CPU test vehicle and the "NCCL baseline" of the comparisons.
"""
from __future__ import annotations

import contextlib
import ctypes
import json
import logging
import os
import time
from typing import Any, Dict, List, Optional, Tuple

import torch
import torch.distributed as dist
from torch import nn

from ..ops import _native as nv
from ..ops import elementwise as ew
from ..ops.linear import FusedLinear
from ..ops.w4 import NF4Weight, w4_dequant
from ..optim.adamw import FusedAdamW, PagedAdamW8bit, clip_coef, grad_sumsq
from ..utils.arena import ParamArena, flatten_module
from .config import DSConfig
from .dist import get_info
from .symm import MAX_RANKS, PeerFlags, SymmBuffer, symm_available

log = logging.getLogger("llm_in_practise_b200.zero")
vp, i32, i64, u32, f32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_uint, ctypes.c_float


class _ZeroStepArgs(ctypes.Structure):
    _fields_ = [("grad_ptrs", vp), ("param_ptrs", vp), ("flag_ptrs", vp), ("norm_ptrs", vp),
                ("gred", vp), ("m", vp), ("v", vp), ("s1", vp), ("s2", vp), ("absmax1", vp), ("absmax2", vp),
                ("qmap1", vp), ("qmap2", vp), ("stats_out", vp),
                ("n", i64), ("world", i32), ("rank", i32), ("eight_bit", i32), ("reduce_only", i32), ("epoch", u32),
                ("lr", f32), ("beta1", f32), ("beta2", f32), ("eps", f32), ("weight_decay", f32), ("bias_c1", f32),
                ("bias_c2", f32), ("grad_scale", f32), ("max_norm", f32),
                ("grid_ctr", vp), ("grid_target_base", u32), ("grid", i32)]


# --------------------------------------------------------------------------------------------------
# stage-3 sharding of the frozen NF4 base
# --------------------------------------------------------------------------------------------------
class ShardedNF4:
    """An NF4 matrix whose rows are partitioned across ranks; duck-types ``NF4Weight.dequantize``.
    Codes and block scales of a shard live in ONE symmetric byte buffer (``wbuf``), grouped by decoder layer so a whole
    layer's shard is one contiguous range per owner."""

    def __init__(self, N, K, wbuf: SymmBuffer, p_off: int, s_off: int, world: int, rank: int, group: int = 64):
        self.N, self.K, self.group = N, K, group
        self.world, self.rank = world, rank
        self.rows = N // world
        self._wbuf, self._poff, self._soff = wbuf, p_off, s_off          # byte offsets into wbuf
        self.packed = wbuf.local  # for .device
        self.shards = [(wbuf.ptrs[r] + p_off, wbuf.ptrs[r] + s_off, 0) for r in range(world)]
        self.prefetcher: Optional["ShardPrefetcher"] = None
        self.layer = -1
        self._stage_shards: List[list] = []

    def dequantize(self, dtype=torch.bfloat16, out=None, transpose: bool = False) -> torch.Tensor:
        shards = self.shards
        if self.prefetcher is not None:
            # the layer's shards were pulled into a local staging buffer while the previous layer computed
            shards = self._stage_shards[self.prefetcher.acquire(self.layer)]
        return w4_dequant(self._wbuf.local, self._wbuf.local, None, self.N, self.K, self.group, None, dtype, out,
                          shards=shards, rows_per_shard=self.rows, transpose=transpose)

    def can_transpose(self) -> bool:
        return self.N % 64 == 0 and self.K % 64 == 0 and self.rows % 64 == 0

    def local_shard(self) -> Tuple[torch.Tensor, torch.Tensor]:
        nb = self.rows * self.K // 2
        ns = self.rows * self.K // self.group
        return (self._wbuf.local[self._poff:self._poff + nb],
                self._wbuf.local[self._soff:self._soff + 4 * ns].view(torch.float32))


class ShardPrefetcher:
    """ZeRO-3 parameter prefetch for the 4-bit base (the ``stage3_prefetch_bucket_size`` / ``stage3_max_live_parameters``
    idea of Fine-Tuning/ds_zero3_config.json, done per decoder layer: one layer ahead, two layers live). While layer L
    computes, layer L+1's (backward: L-1's) shards — packed codes and scales, one contiguous range per owner — are
    copied out of the owners' memory into one of two local staging buffers by the COPY ENGINES (one ``cudaMemcpyAsync``
    per owner on a side stream), so the pull takes no SM away from the persistent GEMMs and the dequant kernels read
    local HBM instead of stalling on NVLink latency. ``B200_PREFETCH_SM=1`` restores the SM pull kernel (the round-1
    path: measured 29 ms of exposed time per step at 8 GPUs because it shares SMs and shared memory with the GEMMs).

    ``exposed_ms()``: time the consuming stream sat in the wait for a prefetch (CUDA event pairs around every wait)."""

    def __init__(self, wbuf: SymmBuffer, regions: Dict[int, Tuple[int, int]], world: int, device):
        from .symm import peer_gather, peer_gather_ce
        self._gather = peer_gather if os.environ.get("B200_PREFETCH_SM", "0") == "1" else peer_gather_ce
        self.wbuf, self.regions, self.world, self.rank = wbuf, regions, world, wbuf.rank
        rmax = max(hi - lo for lo, hi in regions.values())
        self.rmax = rmax
        self.stage = [torch.empty(world * rmax, dtype=torch.uint8, device=device) for _ in range(2)]
        self.resident: List[Optional[int]] = [None, None]
        self.ready = [torch.cuda.Event(), torch.cuda.Event()]
        self.free = [torch.cuda.Event(), torch.cuda.Event()]
        self.stream = torch.cuda.Stream(device)
        self.last: Optional[int] = None
        self.direction = 1
        self.issued = 0
        self.measure = True
        self._waits: List[tuple] = []

    def _issue(self, layer: int, b: int) -> None:
        lo, hi = self.regions[layer]
        cur = torch.cuda.current_stream()
        self.free[b].record(cur)               # every dequant that read this buffer is already queued on `cur`
        self.stream.wait_event(self.free[b])
        with torch.cuda.stream(self.stream):
            self._gather(self.wbuf, self.stage[b][: self.world * (hi - lo)], lo, hi - lo, self.rank)
            self.ready[b].record(self.stream)
        self.resident[b] = layer
        self.issued += 1

    def acquire(self, layer: int) -> int:
        if layer in self.resident:
            b = self.resident.index(layer)
        else:
            b = 1 - self.resident.index(self.last) if self.last in self.resident else 0
            self._issue(layer, b)
        cur = torch.cuda.current_stream()
        if self.measure and layer != self.last:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(cur)
            cur.wait_event(self.ready[b])
            e1.record(cur)
            self._waits.append((e0, e1))
            if len(self._waits) > 4096:
                del self._waits[:2048]
        else:
            cur.wait_event(self.ready[b])
        if self.last is not None and layer != self.last:
            self.direction = 1 if layer > self.last else -1
        nxt = layer + self.direction
        if nxt in self.regions and nxt not in self.resident:
            self._issue(nxt, 1 - b)
        self.last = layer
        return b

    def reset_exposed(self) -> None:
        self._waits.clear()

    def exposed_ms(self) -> float:
        """Sum over the recorded waits (since ``reset_exposed``) of the time between reaching the wait and passing it."""
        tot = 0.0
        for e0, e1 in self._waits:
            if e1.query():
                tot += e0.elapsed_time(e1)
        return tot


def shard_nf4_base(model: nn.Module, world: int, rank: int, device) -> Dict[str, int]:
    """Replace every ``NF4Weight`` of ``model`` by a ``ShardedNF4`` holding 1/world of the rows locally."""
    lins = [(n, m) for n, m in model.named_modules() if isinstance(m, FusedLinear) and isinstance(m.nf4, NF4Weight)]
    if not lins:
        return {"packed_bytes": 0, "scale_floats": 0}
    al = lambda x: (x + 255) // 256 * 256
    # group by decoder layer (layer_uid = 4 * layer + slot); linears without a layer id get a group of their own
    groups: Dict[int, list] = {}
    for i, (name, lin) in enumerate(lins):
        q = lin.nf4
        if q.N % world:
            raise ValueError(f"{name}: {q.N} rows do not divide over {world} ranks")
        L = lin.layer_uid // 4 if hasattr(lin, "layer_uid") and len(lins) > 1 else -(i + 1)
        groups.setdefault(L, []).append(lin)
    total = 0
    plan = []                                   # (lin, p_off, s_off, packed bytes, scale bytes, rows)
    regions: Dict[int, Tuple[int, int]] = {}
    p_bytes = s_floats = 0
    for L, members in groups.items():
        lo = total
        offs = []
        for lin in members:                     # codes of every linear of the layer, then their scales
            q = lin.nf4
            rows = q.N // world
            offs.append([lin, total, 0, rows * q.K // 2, rows * q.K // q.group * 4, rows])
            total += al(offs[-1][3])
        for o in offs:
            o[2] = total
            total += al(o[4])
            p_bytes += o[3]
            s_floats += o[4] // 4
        plan += [tuple(o) for o in offs]
        regions[L] = (lo, total)
    wbuf = SymmBuffer(total, torch.uint8, device, zero=False)
    for lin, po, so, pb, sb, rows in plan:
        q = lin.nf4
        wbuf.local[po:po + pb].copy_(q.packed.view(q.N, q.K // 2)[rank * rows:(rank + 1) * rows].reshape(-1))
        wbuf.local[so:so + sb].view(torch.float32).copy_(
            q.scales().view(q.N, q.K // q.group)[rank * rows:(rank + 1) * rows].reshape(-1))
        lin.nf4 = ShardedNF4(q.N, q.K, wbuf, po, so, world, rank, q.group)
    # ---- one-layer-ahead prefetch on the copy engines (default: on for every world size > 1) ----
    info_pf = 0
    pf_on = os.environ.get("B200_ZERO_PREFETCH", "1") != "0"
    if pf_on and len(regions) > 1:
        pf = ShardPrefetcher(wbuf, regions, world, device)
        layer_of = {id(lin): L for L, members in groups.items() for lin in members}
        for lin, po, so, pb, sb, rows in plan:
            L = layer_of[id(lin)]
            lo, hi = regions[L]
            q = lin.nf4
            q.prefetcher, q.layer = pf, L
            # remote shards come from the staging buffer, the local one is read where it lives
            q._stage_shards = [[q.shards[r] if r == rank else
                                (pf.stage[b].data_ptr() + r * (hi - lo) + (po - lo),
                                 pf.stage[b].data_ptr() + r * (hi - lo) + (so - lo), 0) for r in range(world)]
                               for b in range(2)]
        model._shard_prefetcher = pf
        info_pf = 1
    torch.cuda.synchronize()
    if dist.is_initialized():
        dist.barrier()
    torch.cuda.empty_cache()
    return {"packed_bytes": p_bytes, "scale_floats": s_floats, "prefetch": info_pf, "buffer_bytes": total}


# --------------------------------------------------------------------------------------------------
class LossScaler:
    """fp16 dynamic loss scaling with DeepSpeed's knobs (initial_scale_power, window, hysteresis, min)."""

    def __init__(self, cfg):
        self.dynamic = cfg.loss_scale == 0
        self.scale = float(2 ** cfg.initial_scale_power) if self.dynamic else float(cfg.loss_scale)
        self.window, self.min_scale = cfg.loss_scale_window, cfg.min_loss_scale
        self.hysteresis = self.cur_hysteresis = cfg.hysteresis
        self.good_steps = 0

    def update(self, overflow: bool) -> None:
        if not self.dynamic:
            return
        if overflow:
            if self.cur_hysteresis <= 1:
                self.scale = max(self.scale / 2, self.min_scale)
            else:
                self.cur_hysteresis -= 1
            self.good_steps = 0
        else:
            self.good_steps += 1
            if self.good_steps % self.window == 0:
                self.scale *= 2
                self.cur_hysteresis = self.hysteresis


class WarmupLR:
    """DeepSpeed ``WarmupLR`` (linear by default here) — LLM_Distributed_Trainning/DeepSpeed/*/ds_config.json."""

    def __init__(self, optimizer, warmup_min_lr=0.0, warmup_max_lr=1e-3, warmup_num_steps=1000, warmup_type="linear", **_):
        self.opt, self.lo, self.hi, self.n, self.kind = optimizer, warmup_min_lr, warmup_max_lr, max(1, warmup_num_steps), warmup_type
        self.t = 0
        self._apply()

    def _apply(self):
        import math
        if self.t >= self.n:
            f = 1.0
        elif self.kind == "log":
            f = math.log(self.t + 1) / math.log(self.n)
        else:
            f = self.t / self.n
        lr = self.lo + (self.hi - self.lo) * f
        for g in self.opt.param_groups:
            g["lr"] = lr

    def step(self):
        self.t += 1
        self._apply()

    def get_last_lr(self):
        return [g["lr"] for g in self.opt.param_groups]

    def state_dict(self):
        return {"t": self.t}

    def load_state_dict(self, sd):
        self.t = sd["t"]
        self._apply()


# --------------------------------------------------------------------------------------------------
class ZeroEngine(nn.Module):
    def __init__(self, model: nn.Module, config=None, optimizer=None, lora_config=None, lr_scheduler=None,
                 arena: Optional[ParamArena] = None, force_collectives: bool = False):
        super().__init__()
        self.info = get_info()
        W, R = self.info.world_size, self.info.rank
        self.world, self.rank = W, R
        self.module = model
        hidden = getattr(getattr(model, "config", None), "hidden_size", None)
        self.config = DSConfig.from_any(config).resolve_auto(hidden_size=hidden, world_size=W)
        self.config.check_batch(W)
        self.stage = int(self.config.zero.stage)
        self.ga = int(self.config.gradient_accumulation_steps)
        self.max_norm = float(self.config.gradient_clipping or 0.0)
        dev = next(model.parameters()).device
        self.device = dev
        self.use_symm = (W > 1 and dev.type == "cuda" and symm_available() and not force_collectives
                         and os.environ.get("B200_COMM", "symm") == "symm")
        self.unit = 256 * W  # arena length granularity: every rank's shard is a whole number of 8-bit blocks

        # ---- ZeRO-2 / ZeRO-3 of DENSE trainable parameters: really partition gradients (2) and parameters (3) ----
        self._fsdp = None
        self.base_shard_info = None
        dense = (lora_config is None and arena is None and getattr(model, "_lora_arena", None) is None
                 and not any(isinstance(m, FusedLinear) and m.nf4 is not None for m in model.modules()))
        want_param_offload = self.config.zero.offload_param_device == "cpu"
        if dense and W > 1 and (self.stage >= 2 or want_param_offload) and os.environ.get("B200_ZERO_DENSE_SHARD", "1") != "0":
            self._init_dense_sharded(model, optimizer, lr_scheduler, want_param_offload)
            return
        if want_param_offload and not (dense and W > 1):
            log.warning("offload_param: cpu is honoured for dense models on > 1 rank (parameter shards in pinned host "
                        "memory); this run keeps parameters on the GPU%s",
                        "" if dense else " (LoRA / 4-bit base: the trainable set is a few MB)")

        # ---- trainable parameters -> flat arena (symmetric when fused collectives are on) ----
        alloc = None
        if self.use_symm:
            self._symm_bufs: List[SymmBuffer] = []

            def alloc(numel, dtype, device):
                b = SymmBuffer(numel, dtype, device)
                self._symm_bufs.append(b)
                return b.local
        if lora_config is not None:
            from ..peft.lora import apply_lora
            arena = apply_lora(model, lora_config, pad_to=self.unit, alloc=alloc)
        elif arena is None:
            arena = getattr(model, "_lora_arena", None)
            if arena is not None and hasattr(arena, "_symm_bufs") and self.use_symm:
                self._symm_bufs = list(arena._symm_bufs)
            if arena is None or arena.numel % self.unit or (self.use_symm and not getattr(arena, "symmetric", False)):
                if arena is not None:
                    raise ValueError("existing adapter arena is not laid out for this world size; pass lora_config "
                                     "to the engine instead of calling get_peft_model first")
                arena = flatten_module(model, dtype=torch.float32, pad_to=self.unit, alloc=alloc)
        self.arena = arena
        if self.use_symm:
            self.param_buf, self.grad_buf = self._symm_bufs[0], self._symm_bufs[1]
            self.flags = PeerFlags(dev)
        self.shard_len = arena.numel // W
        self.shard = slice(R * self.shard_len, (R + 1) * self.shard_len) if (self.stage >= 1 and W > 1) else slice(0, arena.numel)

        # ---- stage 3: partition the frozen 4-bit base ----
        if self.stage >= 3 and W > 1 and self.use_symm:
            self.base_shard_info = shard_nf4_base(model, W, R, dev)

        # ---- make every rank start from rank 0's trainable values (DDP ctor broadcast / sync_module_states) ----
        if W > 1:
            dist.broadcast(arena.data, src=0)

        # ---- optimizer ----
        self.offload = self.config.zero.offload_optimizer_device == "cpu"
        self.optimizer = self._build_optimizer(optimizer)
        self.lr_scheduler = lr_scheduler or self._build_scheduler()
        self.scaler = LossScaler(self.config.fp16) if self.config.fp16.enabled else None
        self.micro_steps = 0
        self.global_steps = 0
        self.skipped_steps = 0
        self._stats = torch.zeros(4, dtype=torch.float32, device=dev)
        self._gred = torch.zeros(self.shard_len, dtype=torch.float32, device=dev) if self.use_symm else None
        self._timers: Dict[str, List] = {}
        self._breakdown = bool(self.config.wall_clock_breakdown) and self.device.type == "cuda"
        self._last_norm = None

    # ------------------------------------------------------------------ dense ZeRO-2 / ZeRO-3 backend
    def _init_dense_sharded(self, model, optimizer, lr_scheduler, offload_param: bool) -> None:
        """Stage 2 (``SHARD_GRAD_OP``: gradients and optimizer state partitioned, parameters kept gathered between
        forward and backward) and stage 3 (``FULL_SHARD``: parameters partitioned too, gathered per block one block
        ahead on the copy engines and freed after use) of the reference's hand-written GPTLike loops
        (LLM_Distributed_Trainning/DeepSpeed/DeepSpeed-GPTLike-ZeRO-{2,3,Offload}/ds_config.json). Every
        ``nn.ModuleList`` member is one unit, the rest of the model another. ``overlap_comm``: the reduce-scatter of
        a unit starts on a side stream as soon as backward has produced its last gradient; ``contiguous_gradients``:
        a unit's gradients are one flat buffer that exists only between its backward and its reduce-scatter;
        ``offload_param`` / ``offload_optimizer`` = cpu: fp32 shards, their gradients and the AdamW state live in
        pinned host memory. Peak memory per rank falls as 1 / world for parameters (3), gradients and state (2, 3)."""
        from .fsdp import CPUOffload, FullyShardedDataParallel, MixedPrecision, ShardingStrategy
        dev = self.device
        kinds = {type(m) for ml in model.modules() if isinstance(ml, nn.ModuleList) for m in ml
                 if any(p.requires_grad for p in m.parameters())}
        policy = (lambda m, recurse=False, n=0: type(m) in kinds) if kinds else None
        first = next(p for p in model.parameters() if p.requires_grad)
        rd = torch.bfloat16 if (self.config.bf16 and first.dtype == torch.bfloat16) else torch.float32
        offload = offload_param or self.config.zero.offload_optimizer_device == "cpu"
        strat = ShardingStrategy.FULL_SHARD if self.stage >= 3 else ShardingStrategy.SHARD_GRAD_OP
        self._fsdp = FullyShardedDataParallel(model, auto_wrap_policy=policy, sharding_strategy=strat,
                                              mixed_precision=MixedPrecision(param_dtype=None, reduce_dtype=rd),
                                              cpu_offload=CPUOffload(offload_params=True) if offload else None,
                                              sync_module_states=True)
        self.use_symm = self._fsdp.use_symm
        self.arena = None
        self.offload = offload
        kind, hp = self._optimizer_hparams(optimizer)
        self.opt_kind = kind
        if offload:
            self.optimizer = torch.optim.AdamW(list(self._fsdp.parameters()), **hp)       # host tensors: torch's CPU AdamW
        else:
            self.optimizer = (PagedAdamW8bit if kind == "adamw8bit" else FusedAdamW)(list(self._fsdp.parameters()), **hp)
        self.lr_scheduler = lr_scheduler or self._build_scheduler()
        self.scaler = LossScaler(self.config.fp16) if self.config.fp16.enabled else None
        self.micro_steps = self.global_steps = self.skipped_steps = 0
        self._timers = {}
        self._breakdown = bool(self.config.wall_clock_breakdown) and dev.type == "cuda"
        self._last_norm = None
        self.shard = None
        log.info("ZeRO stage %d (dense): %d units, strategy %s, reduce dtype %s, offload %s", self.stage,
                 len(self._fsdp.units), strat.name, rd, offload)

    def _step_dense(self, inv_scale: float) -> None:
        f = self._fsdp
        if inv_scale != 1.0:
            for u in f.units:
                u.master.grad.mul_(inv_scale)
        overflow = False
        if self.max_norm > 0 or self.scaler is not None:
            norm = f.clip_grad_norm_(self.max_norm if self.max_norm > 0 else float("inf"))
            self._last_norm = norm
            if self.scaler is not None:
                overflow = not bool(torch.isfinite(norm).item())
        if not overflow:
            self.optimizer.step()
            for u in f.units:        # the flat kernels write through .data: tell the units their shards changed
                torch.autograd.graph.increment_version(u.master)
        if self.scaler is not None:
            self.skipped_steps += int(overflow)
            self.scaler.update(overflow)
        f.zero_grad()

    # ------------------------------------------------------------------ construction helpers
    def _optimizer_hparams(self, client):
        """(kind, hyper-parameters) from a client optimizer or the DeepSpeed JSON ``optimizer`` block."""
        if client is not None and not isinstance(client, FusedAdamW):
            # a client torch optimizer (e.g. optim.AdamW built by the script, DeepSpeed-GPTLike-ZeRO-1.py:272):
            # adopt its hyper-parameters, run our flat implementation
            g = client.param_groups[0]
            kind = "adamw8bit" if "8bit" in type(client).__name__.lower() else "adamw"
            hp = dict(lr=g.get("lr", 1e-3), betas=tuple(g.get("betas", (0.9, 0.999))), eps=g.get("eps", 1e-8),
                      weight_decay=g.get("weight_decay", 0.0))
        elif isinstance(client, FusedAdamW):
            g = client.param_groups[0]
            kind = "adamw8bit" if client.eight_bit else "adamw"
            hp = dict(lr=g["lr"], betas=g["betas"], eps=g["eps"], weight_decay=g["weight_decay"])
        else:
            oc = self.config.optimizer or {"type": "AdamW", "params": {}}
            t = str(oc.get("type", "AdamW")).lower()
            kind = "adamw8bit" if "8bit" in t else "adamw"
            p = dict(oc.get("params", {}))
            hp = dict(lr=p.get("lr", 1e-3), betas=tuple(p.get("betas", (0.9, 0.999))), eps=p.get("eps", 1e-8),
                      weight_decay=p.get("weight_decay", 0.0 if "adamw" not in t else 0.01) if "weight_decay" not in p else p["weight_decay"])
        return kind, hp

    def _build_optimizer(self, client):
        kind, hp = self._optimizer_hparams(client)
        self.opt_kind = kind
        cls = PagedAdamW8bit if kind == "adamw8bit" else FusedAdamW
        if self.offload:
            from ..optim.cpu_adam import CPUAdamW
            return CPUAdamW(self.arena, shard=self.shard, **hp)
        return cls(self.arena, shard=self.shard, **hp)

    def _build_scheduler(self):
        sc = self.config.scheduler
        if not sc:
            return None
        t = str(sc.get("type", "")).lower()
        p = dict(sc.get("params", {}))
        if t == "warmuplr":
            return WarmupLR(self.optimizer, **p)
        if t == "warmupdecaylr":
            total = p.pop("total_num_steps", 10000)
            w = WarmupLR(self.optimizer, **p)
            w.total = total
            return w
        raise ValueError(f"unsupported scheduler type {sc.get('type')}")

    # ------------------------------------------------------------------ training surface
    @contextlib.contextmanager
    def _phase(self, name: str):
        """``wall_clock_breakdown``: a device timer (CUDA event pair, read lazily by ``comm_stats`` /
        ``wall_clock_report``) and an NVTX range around one phase; free when the option is off."""
        if not self._breakdown:
            yield
            return
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.nvtx.range_push(f"b200.{name}")
        e0.record()
        try:
            yield
        finally:
            e1.record()
            torch.cuda.nvtx.range_pop()
            lst = self._timers.setdefault(name, [])
            lst.append((e0, e1))
            if len(lst) > 256:
                del lst[:128]

    def wall_clock_report(self) -> Dict[str, float]:
        """Mean milliseconds per phase (forward / backward / step [/ zero_step]) over the recorded steps, max over ranks."""
        torch.cuda.synchronize()
        out = {}
        for k, evs in self._timers.items():
            ms = [a.elapsed_time(b) for a, b in evs]
            if ms:
                t = torch.tensor([sum(ms) / len(ms)], device=self.device)
                if self.world > 1:
                    dist.all_reduce(t, op=dist.ReduceOp.MAX)
                out[k + "_ms"] = float(t)
        return out

    def _loss_factor(self) -> float:
        s = 1.0 / self.ga
        if self.scaler is not None:
            s *= self.scaler.scale
        return s

    def forward(self, *args, **kwargs):
        ew.set_step(self.global_steps, self.micro_steps % self.ga)
        from ..ops.loss import grad_scale_context
        # the factor ``backward`` will apply is declared up front so a fused lm_head + loss can scale the gradients
        # it forms during the forward launches (tied / trainable heads write their weight gradient there)
        with self._phase("forward"), grad_scale_context(self._loss_factor() if torch.is_grad_enabled() else None):
            if self._fsdp is not None:
                return self._fsdp(*args, **kwargs)
            return self.module(*args, **kwargs)

    def backward(self, loss: torch.Tensor) -> None:
        """Scales by 1/accumulation (and the fp16 loss scale) like DeepSpeed, then backpropagates into the arena."""
        s = self._loss_factor()
        with self._phase("backward"):
            if self._fsdp is not None and (self.micro_steps + 1) % self.ga:
                with self._fsdp.no_sync():           # non-boundary micro-step: accumulate locally, no collective
                    (loss * s).backward()
            else:
                (loss * s).backward()
        self.micro_steps += 1

    def train_micro_batch(self, input_ids, labels=None, **kw) -> torch.Tensor:
        out = self(input_ids, labels=labels if labels is not None else input_ids, **kw)
        loss = out["loss"] if isinstance(out, dict) else out
        self.backward(loss)
        return loss.detach()

    def supports_fused_accumulation(self) -> bool:
        import inspect
        if self._fsdp is not None:
            return False
        try:
            return "micro_batches" in inspect.signature(self.module.forward).parameters
        except (TypeError, ValueError):
            return False

    def train_batch(self, input_ids, labels=None, **kw) -> torch.Tensor:
        """All ``gradient_accumulation_steps`` micro-batches of one optimizer step in ONE layer-major pass (the name is
        DeepSpeed's ``PipelineEngine.train_batch``, which also consumes a whole step's micro-batches).

        ``input_ids`` / ``labels``: a list of G micro-batches [B, S] (or already stacked [G * B, S]). Every layer
        processes the G * B sequences together, so each frozen 4-bit weight is gathered from its owners and expanded
        ONCE per direction per optimizer step instead of G times — the ZeRO-3 parameter traffic and the dequantisation
        work both drop by G — and the GEMMs see G times more rows. The loss is the mean over micro-batches of each
        one's mean token loss and the gradients equal those of G separate ``backward(loss / G)`` calls (attention is
        per sequence, the normalisation per micro-batch; LoRA dropout draws one mask per element either way).
        Falls back to G sequential micro-steps for modules without ``micro_batches`` support. Call ``step()`` after."""
        G = self.ga
        ids_l = list(input_ids) if isinstance(input_ids, (list, tuple)) else list(input_ids.chunk(G, 0))
        lab_l = ids_l if labels is None else (list(labels) if isinstance(labels, (list, tuple)) else list(labels.chunk(G, 0)))
        if len(ids_l) != G:
            raise ValueError(f"train_batch needs {G} micro-batches (gradient_accumulation_steps), got {len(ids_l)}")
        if G == 1 or not self.supports_fused_accumulation() or len({tuple(t.shape) for t in ids_l}) != 1:
            loss = None
            for a, b in zip(ids_l, lab_l):
                l = self.train_micro_batch(a, b, **kw)
                loss = l if loss is None else loss + l
            return loss / G
        ids = torch.cat(ids_l, 0)
        lab = ids if labels is None else torch.cat(lab_l, 0)
        ew.set_step(self.global_steps, 0)
        from ..ops.loss import grad_scale_context
        s = self.scaler.scale if self.scaler is not None else 1.0        # 1 / G is inside the per-row normalisation
        with self._phase("forward"), grad_scale_context(s):
            out = self.module(ids, labels=lab, micro_batches=G, **kw)
        loss = out["loss"] if isinstance(out, dict) else out
        with self._phase("backward"):
            (loss * s if s != 1.0 else loss).backward()
        self.micro_steps += G
        return loss.detach()

    def is_gradient_accumulation_boundary(self) -> bool:
        return self.micro_steps % self.ga == 0 and self.micro_steps > 0

    def step(self) -> None:
        if not self.is_gradient_accumulation_boundary():
            return  # non-boundary micro-steps issue no collective at all
        inv_scale = 1.0 / self.scaler.scale if self.scaler is not None else 1.0
        with self._phase("step"):
            if self._fsdp is not None:
                self._step_dense(inv_scale)
            elif self.world == 1:
                self._step_single(inv_scale)
            elif self.use_symm and not self.offload:
                self._step_fused(inv_scale)
            else:
                self._step_collectives(inv_scale)
        if self.arena is not None:
            self.arena.zero_grad()
        if self.lr_scheduler is not None:
            self.lr_scheduler.step()
        self.global_steps += 1

    # ---- single GPU ----
    def _step_single(self, inv_scale: float) -> None:
        opt = self.optimizer
        opt.grad_scale = inv_scale
        need_stats = self.max_norm > 0 or self.scaler is not None
        if need_stats:
            ss = grad_sumsq([self.arena.grad], pre_scale=inv_scale)
            coef, norm = clip_coef(ss, self.max_norm)
            opt.dyn_scale = coef
            self._last_norm = norm
            if self.scaler is not None:
                opt.found_inf = ss[1:2]
        opt.step()
        if self.scaler is not None:
            overflow = bool(ss[1].item() != 0)
            self.skipped_steps += int(overflow)
            self.scaler.update(overflow)

    # ---- multi GPU, fused peer-memory kernel (K-C2) ----
    def _step_fused(self, inv_scale: float) -> None:
        opt = self.optimizer
        if opt._ranges is None:
            opt._ranges = opt._build_ranges()
        r = opt._ranges[0]
        if not r.state:
            opt._init_state(r)
        opt._step += 1
        g = opt.param_groups[0]
        b1, b2 = g["betas"]
        a = _ZeroStepArgs()
        F = self.flags
        a.grad_ptrs = ctypes.cast(self.grad_buf.c_ptrs(), vp)
        a.param_ptrs = ctypes.cast(self.param_buf.c_ptrs(), vp)
        a.flag_ptrs = ctypes.cast(F.flags.c_ptrs(), vp)
        a.norm_ptrs = ctypes.cast(F.norms.c_ptrs(), vp)
        a.gred = self._gred.data_ptr()
        st = r.state
        eight = opt.eight_bit
        if eight:
            a.s1, a.s2 = st["state1"].data_ptr(), st["state2"].data_ptr()
            a.absmax1, a.absmax2 = st["absmax1"].data_ptr(), st["absmax2"].data_ptr()
            a.qmap1, a.qmap2 = st["qmap1"].data_ptr(), st["qmap2"].data_ptr()
        else:
            a.m, a.v = st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr()
        a.stats_out = self._stats.data_ptr()
        a.n, a.world, a.rank = self.arena.numel, self.world, self.rank
        a.eight_bit, a.reduce_only = int(eight), 0
        a.epoch = F.next_epoch()
        a.lr, a.beta1, a.beta2, a.eps, a.weight_decay = g["lr"], b1, b2, g["eps"], g["weight_decay"]
        a.bias_c1, a.bias_c2 = 1.0 - b1 ** opt._step, 1.0 - b2 ** opt._step
        a.grad_scale = inv_scale / self.world
        a.max_norm = self.max_norm
        grid = max(1, min(64, self.shard_len // 256))
        a.grid = grid
        a.grid_ctr = F.grid_ctr.data_ptr()
        a.grid_target_base = F.take_grid_target(grid, 2)
        if getattr(self, "time_skew", False):
            # measurement mode: a device-side barrier across ranks first, so the K-C2 timer below holds kernel time
            # only and the time spent waiting for the slowest rank shows up separately ("step_skew")
            from .symm import peer_barrier
            b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            b0.record()
            peer_barrier(self.flags, channel=6)
            b1.record()
            sk = self._timers.setdefault("step_skew", [])
            sk.append((b0, b1))
            if len(sk) > 256:
                del sk[:128]
        ev0 = torch.cuda.Event(enable_timing=True)
        ev1 = torch.cuda.Event(enable_timing=True)
        ev0.record()
        fn = nv.proto("b200_zero_step", [ctypes.POINTER(_ZeroStepArgs), vp])
        nv.check(fn(ctypes.byref(a), nv.stream()), "zero_step")
        nv.count_launch()
        ev1.record()
        lst = self._timers.setdefault("zero_step", [])
        lst.append((ev0, ev1))
        if len(lst) > 256:
            del lst[:128]
        torch.autograd.graph.increment_version(self.arena.data)
        if self.scaler is not None:
            overflow = bool(self._stats[2].item() != 0)
            if overflow:
                opt._step -= 1          # a skipped step does not advance the bias-correction counter (DeepSpeed semantics)
            self.skipped_steps += int(overflow)
            self.scaler.update(overflow)

    # ---- multi process, torch.distributed collectives (CPU tests / NCCL baseline / offload) ----
    def _step_collectives(self, inv_scale: float) -> None:
        opt = self.optimizer
        W = self.world
        grad = self.arena.grad
        if self.stage >= 2 and not self.offload:
            shard = torch.empty(self.shard_len, dtype=grad.dtype, device=grad.device)
            dist.reduce_scatter_tensor(shard, grad, op=dist.ReduceOp.SUM) if grad.is_cuda else \
                self._gloo_reduce_scatter(shard, grad)
            grad[self.shard].copy_(shard)
        else:
            dist.all_reduce(grad, op=dist.ReduceOp.SUM)
        gs = inv_scale / W
        local = grad[self.shard] if self.stage >= 1 else grad
        ss = grad_sumsq([local], pre_scale=gs)
        if self.stage >= 1:
            dist.all_reduce(ss[0:1], op=dist.ReduceOp.SUM)
            dist.all_reduce(ss[1:2], op=dist.ReduceOp.MAX)
        coef, norm = clip_coef(ss, self.max_norm)
        self._last_norm = norm
        opt.grad_scale = gs
        opt.dyn_scale = coef
        if self.scaler is not None:
            opt.found_inf = ss[1:2]
        opt.step()
        if self.stage >= 1:
            # updated shards -> every replica (allgather_partitions)
            dist.all_gather_into_tensor(self.arena.data, self.arena.data[self.shard].clone())
            torch.autograd.graph.increment_version(self.arena.data)
        if self.scaler is not None:
            overflow = bool(ss[1].item() != 0)
            self.skipped_steps += int(overflow)
            self.scaler.update(overflow)

    def _gloo_reduce_scatter(self, out, full):
        dist.all_reduce(full, op=dist.ReduceOp.SUM)
        out.copy_(full[self.shard])

    # ------------------------------------------------------------------ introspection
    def get_global_grad_norm(self) -> Optional[float]:
        if self.use_symm and self.world > 1:
            return float(self._stats[0].item())
        return None if self._last_norm is None else float(self._last_norm.item())

    def get_lr(self):
        return [g["lr"] for g in self.optimizer.param_groups]

    def comm_stats(self) -> Dict[str, Any]:
        out: Dict[str, Any] = {"backend": "nvlink-peer-fused" if self.use_symm else ("single" if self.world == 1 else "torch.distributed"),
                               "zero_stage": self.stage}
        if self._fsdp is not None:
            out["dense_sharding"] = self._fsdp.strategy.name
            out["units"] = len(self._fsdp.units)
        for k, evs in self._timers.items():
            ms = [a.elapsed_time(b) for a, b in evs[-20:]]
            if ms:
                out[k + "_ms"] = round(sum(ms) / len(ms), 4)
        if self.base_shard_info:
            out["nf4_shard_bytes_per_rank"] = self.base_shard_info["packed_bytes"] + 4 * self.base_shard_info["scale_floats"]
        pf = getattr(self.module, "_shard_prefetcher", None)
        if pf is not None:
            out["prefetch"] = "copy-engine" if pf._gather.__name__.endswith("_ce") else "sm-kernel"
            out["prefetch_issued"] = pf.issued
        return out

    def reset_comm_timers(self) -> None:
        """Start a fresh measurement window for ``exposed_comm_ms`` / ``zero_step_ms``."""
        self._timers.pop("zero_step", None)
        self._timers.pop("step_skew", None)
        pf = getattr(self.module, "_shard_prefetcher", None)
        if pf is not None:
            pf.reset_exposed()
        self._window_steps = self.global_steps

    def exposed_comm_ms(self) -> Optional[float]:
        """Exposed communication per optimizer step since ``reset_comm_timers`` (ms, this rank): the time the stream that
        consumes the prefetched parameter shards sat waiting for them, plus the fused reduce-scatter / AdamW / all-gather
        launch (K-C2), whose whole duration is serial with compute. ``None`` on one GPU."""
        if self.world == 1:
            return None
        torch.cuda.synchronize()
        steps = max(1, self.global_steps - getattr(self, "_window_steps", 0))
        tot = 0.0
        pf = getattr(self.module, "_shard_prefetcher", None)
        if pf is not None:
            tot += pf.exposed_ms()
        for a, b in self._timers.get("zero_step", []):
            tot += a.elapsed_time(b)
        return tot / steps

    # ------------------------------------------------------------------ checkpoints
    def consolidated_state_dict(self) -> Dict[str, torch.Tensor]:
        """Full trainable state dict on every rank (replicas are always current after ``step``) —
        what ``stage3_gather_16bit_weights_on_model_save`` / FSDP full_state_dict produce."""
        if self._fsdp is not None:
            return self._fsdp.full_state_dict(cpu=False)         # gathers every unit (C8), re-shards afterwards
        return {k: v.detach().clone() for k, v in self.module.state_dict().items()}

    def save_checkpoint(self, save_dir: str, tag: Optional[str] = None, client_state: Optional[dict] = None) -> None:
        """DeepSpeed directory layout: ``<dir>/<tag>/mp_rank_00_model_states.pt`` (rank 0) +
        ``zero_pp_rank_<r>_mp_rank_00_optim_states.pt`` (every rank) + ``<dir>/latest``."""
        tag = tag or f"global_step{self.global_steps}"
        d = os.path.join(save_dir, tag)
        os.makedirs(d, exist_ok=True)
        full = self.consolidated_state_dict() if (self._fsdp is not None or self.rank == 0) else None   # a collective when sharded
        if self.rank == 0:
            torch.save({"module": full, "global_steps": self.global_steps,
                        "micro_steps": self.micro_steps, "skipped_steps": self.skipped_steps,
                        "lr_scheduler": None if self.lr_scheduler is None else self.lr_scheduler.state_dict(),
                        "ds_config": self.config.extra, "client_state": client_state or {}},
                       os.path.join(d, "mp_rank_00_model_states.pt"))
            with open(os.path.join(save_dir, "latest"), "w") as f:
                f.write(tag)
        torch.save({"optimizer_state_dict": self.optimizer.state_dict(), "shard": (self.shard.start, self.shard.stop) if self.shard is not None else None,
                    "master_shard": (self.arena.data[self.shard].detach().cpu() if self.arena is not None else
                                     [u.master.detach().cpu() for u in self._fsdp.units]), "zero_stage": self.stage,
                    "world_size": self.world, "loss_scale": None if self.scaler is None else self.scaler.scale},
                   os.path.join(d, f"zero_pp_rank_{self.rank}_mp_rank_00_optim_states.pt"))
        if self.world > 1:
            dist.barrier()

    def load_checkpoint(self, load_dir: str, tag: Optional[str] = None):
        if tag is None:
            with open(os.path.join(load_dir, "latest")) as f:
                tag = f.read().strip()
        d = os.path.join(load_dir, tag)
        ms = torch.load(os.path.join(d, "mp_rank_00_model_states.pt"), map_location="cpu", weights_only=False)
        (self._fsdp if self._fsdp is not None else self.module).load_state_dict(ms["module"], strict=False)
        self.global_steps, self.micro_steps = ms["global_steps"], ms["micro_steps"]
        if self.lr_scheduler is not None and ms.get("lr_scheduler"):
            self.lr_scheduler.load_state_dict(ms["lr_scheduler"])
        os_path = os.path.join(d, f"zero_pp_rank_{self.rank}_mp_rank_00_optim_states.pt")
        if os.path.exists(os_path):
            st = torch.load(os_path, map_location="cpu", weights_only=False)
            if st["world_size"] == self.world:
                self.optimizer.load_state_dict(st["optimizer_state_dict"])
                if self.scaler is not None and st.get("loss_scale"):
                    self.scaler.scale = st["loss_scale"]
            else:
                log.warning("checkpoint %s was written by %d ranks, this job has %d: optimizer state NOT restored "
                            "(parameters are); the moments restart from zero", d, st["world_size"], self.world)
        if self.arena is not None:
            torch.autograd.graph.increment_version(self.arena.data)
        return d, ms.get("client_state", {})


def initialize(args=None, model: Optional[nn.Module] = None, optimizer=None, model_parameters=None, config=None,
               config_params=None, lora_config=None, lr_scheduler=None, **kw) -> ZeroEngine:
    """``deepspeed.initialize`` with our engine. ``config`` may be a path, a dict, or come from
    ``args.deepspeed_config`` / ``args.ds_config`` (the flag names the reference scripts use)."""
    if config is None:
        config = config_params
    if config is None and args is not None:
        config = getattr(args, "deepspeed_config", None) or getattr(args, "ds_config", None)
    return ZeroEngine(model, config=config, optimizer=optimizer, lora_config=lora_config, lr_scheduler=lr_scheduler, **kw)


def ds_initialize(*a, **k):
    """Drop-in for ``engine, optimizer, _, scheduler = deepspeed.initialize(...)``."""
    e = initialize(*a, **k)
    return e, e.optimizer, None, e.lr_scheduler


def init_distributed(dist_backend: Optional[str] = None, **_):
    """``deepspeed.init_distributed()``."""
    from .dist import setup_distributed
    return setup_distributed(force_backend=dist_backend)
