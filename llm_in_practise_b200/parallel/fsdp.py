"""Fully-sharded data parallel: the FSDP1 wrapper and the FSDP2 ``fully_shard`` call of the reference.

    FullyShardedDataParallel(model, auto_wrap_policy=transformer_auto_wrap_policy{TransformerBlock},
        sharding_strategy=ShardingStrategy.FULL_SHARD, mixed_precision=MixedPrecision(bf16, bf16, bf16),
        cpu_offload=CPUOffload(offload_params=False), device_id=..., sync_module_states=True)
                                    — LLM_Distributed_Trainning/PyTorch/fsdp_basics/fsdp_gpt_wikitext2.py:271-312
    mesh = DeviceMesh("cuda", list(range(world)));  fully_shard(model, mesh=mesh, mp_policy=..., offload_policy=...)
                                    — …/fsdp_basics/fsdp2_gpt_wikitext2.py:261-293
    get_state_dict(model, optimizers=(), options=StateDictOptions(full_state_dict=True))   — fsdp…:360-383

Mechanics (one process per GPU):
* every wrapped unit owns a flat parameter: an fp32 *master shard* (1/world of it, exposed as the unit's single
  ``nn.Parameter`` so any optimizer works on ``model.parameters()``) and a transient *full* buffer in the compute
  dtype whose storage is released between uses (``untyped_storage().resize_``), exactly when FULL_SHARD says so;
* the compute-dtype copy of every master shard is staged in symmetric memory ONCE per optimizer step (two device-side
  barriers per step, none per unit); unshard = a pull-based gather of the peers' staged shards over NVLink on the COPY
  ENGINES (``b200_peer_gather_ce``, K-C3), issued on a side stream one unit AHEAD of the compute (forward: the next
  unit in execution order, backward: the previous one), so the gather of unit i+1 overlaps the compute of unit i —
  or ``all_gather_into_tensor`` on CPU/gloo and without symmetric memory;
* gradients accumulate in a transient fp32 full buffer (``p.main_grad`` for the wgrad GEMMs, autograd grads are
  folded in), then ``b200_reduce_scatter`` (peer-memory reads, cast to ``MixedPrecision.reduce_dtype`` and 1/world
  fused, double-buffered staging, on a side stream) leaves each rank its 1/world slice in the master shard's ``.grad``;
* SHARD_GRAD_OP keeps the full buffer between forward and backward; NO_SHARD never frees it.
"""
from __future__ import annotations

import enum
import os
from dataclasses import dataclass
from typing import Any, Callable, Dict, Iterable, List, Optional, Sequence, Set, Type

import torch
import torch.distributed as dist
from torch import nn

from .dist import get_info
from .symm import SymmBuffer, peer_gather, symm_available


class ShardingStrategy(enum.Enum):
    FULL_SHARD = 1
    SHARD_GRAD_OP = 2
    NO_SHARD = 3


@dataclass
class MixedPrecision:
    param_dtype: Optional[torch.dtype] = None
    reduce_dtype: Optional[torch.dtype] = None
    buffer_dtype: Optional[torch.dtype] = None


@dataclass
class MixedPrecisionPolicy:  # FSDP2 spelling
    param_dtype: Optional[torch.dtype] = None
    reduce_dtype: Optional[torch.dtype] = None
    output_dtype: Optional[torch.dtype] = None
    cast_forward_inputs: bool = True


@dataclass
class CPUOffload:
    offload_params: bool = False


class OffloadPolicy:  # FSDP2 no-op base class, as the reference passes it
    pass


@dataclass
class CPUOffloadPolicy(OffloadPolicy):
    pin_memory: bool = True


@dataclass
class StateDictOptions:
    full_state_dict: bool = False
    cpu_offload: bool = False
    strict: bool = True


class DeviceMesh:
    """1-D mesh (the only kind the reference builds)."""

    def __init__(self, device_type: str, mesh: Sequence[int], mesh_dim_names: Optional[Sequence[str]] = None):
        self.device_type = device_type
        self.mesh = list(mesh)
        self.mesh_dim_names = tuple(mesh_dim_names or ("dp",))

    def size(self, dim: int = 0) -> int:
        return len(self.mesh)

    def get_rank(self) -> int:
        return get_info().rank


def init_device_mesh(device_type: str, mesh_shape: Sequence[int], mesh_dim_names=None) -> DeviceMesh:
    n = 1
    for s in mesh_shape:
        n *= s
    return DeviceMesh(device_type, list(range(n)), mesh_dim_names)


def transformer_auto_wrap_policy(module: Optional[nn.Module] = None, recurse: bool = True, nonwrapped_numel: int = 0,
                                 transformer_layer_cls: Optional[Set[Type[nn.Module]]] = None):
    """``functools.partial(transformer_auto_wrap_policy, transformer_layer_cls={Block})`` compatible."""
    if module is None:
        return lambda m, r=True, n=0: transformer_auto_wrap_policy(m, r, n, transformer_layer_cls)
    if recurse:
        return True
    return isinstance(module, tuple(transformer_layer_cls or ()))


def size_based_auto_wrap_policy(module: nn.Module = None, recurse: bool = True, nonwrapped_numel: int = 0,
                                min_num_params: int = int(1e8)):
    if recurse:
        return True
    return nonwrapped_numel >= min_num_params


class _Unit:
    """One flat parameter: the parameters of a wrapped module (minus those of nested units)."""

    def __init__(self, owner: "FullyShardedDataParallel", module: nn.Module, params: List[nn.Parameter], name: str):
        self.o, self.module, self.name = owner, module, name
        self.params = params
        W, R = owner.world, owner.rank
        dev = params[0].device
        self.pdtype = owner.param_dtype or params[0].dtype
        total = 0
        self.offs = []
        for p in params:
            self.offs.append(total)
            total += (p.numel() + 7) // 8 * 8
        self.n = (total + W * 64 - 1) // (W * 64) * (W * 64)
        self.shard_len = self.n // W
        flat = torch.zeros(self.n, dtype=torch.float32, device=dev)
        for p, off in zip(params, self.offs):
            flat[off:off + p.numel()].copy_(p.data.reshape(-1).float())
        if W > 1 and owner.sync_states:
            dist.broadcast(flat, src=0)
        # fp32 master shard = the Parameter the optimizer sees. CPUOffload(offload_params=True) / CPUOffloadPolicy /
        # DeepSpeed ``offload_param: {device: cpu}``: the master shard and its gradient live in (pinned) host memory,
        # the optimizer steps there, and the GPU only ever holds the compute-dtype copies.
        self.offload = bool(owner.cpu_offload is not None and getattr(owner.cpu_offload, "offload_params", False))
        shard = flat[R * self.shard_len:(R + 1) * self.shard_len].clone()
        if self.offload:
            shard = shard.cpu()
            if dev.type == "cuda":
                shard = shard.pin_memory()
        self.master = nn.Parameter(shard)
        self.master.grad = torch.zeros_like(self.master.data)
        if self.offload and dev.type == "cuda":
            self.master.grad = self.master.grad.pin_memory()
        self.gshard = torch.zeros(self.shard_len, dtype=torch.float32, device=dev) if self.offload else None
        self.master._fsdp_unit = self
        # compute-dtype shard staged for the gather (symmetric memory when peers can read it)
        self.symm = SymmBuffer(self.shard_len, self.pdtype, dev) if owner.use_symm else None
        self.shard_lowp = self.symm.local if self.symm is not None else torch.empty(self.shard_len, dtype=self.pdtype, device=dev)
        # transient full buffers
        self.full = torch.empty(self.n, dtype=self.pdtype, device=dev)
        self.full.copy_(flat)
        self.views_set = False
        self._point_params()
        self.gfull: Optional[torch.Tensor] = None
        self.sharded = False
        self.gather_ev = None
        self.pending: Set[int] = set()
        self.reduced = True
        self.master_version = -1
        del flat

    def _point_params(self) -> None:
        for p, off in zip(self.params, self.offs):
            p.data = self.full[off:off + p.numel()].view(p.shape)
            p._fsdp_unit = self

    # ---- parameters ----
    def stage(self) -> None:
        """fp32 master shard -> compute dtype, into the buffer the peers read."""
        self.shard_lowp.copy_(self.master.data)
        self.master_version = self.master._version

    def _alloc_full(self) -> None:
        es = self.full.element_size()
        if self.full.untyped_storage().size() == 0:
            self.full.untyped_storage().resize_(self.n * es)

    def issue_gather(self) -> None:
        """Start the gather of this unit's peers' shards on the gather stream (copy engines); ``gather_ev`` marks its end."""
        o = self.o
        self._alloc_full()
        cur = torch.cuda.current_stream()
        o.gather_stream.wait_stream(cur)            # the buffer exists (and its last reader is done) in stream order
        with torch.cuda.stream(o.gather_stream):
            from .symm import peer_gather_ce
            peer_gather_ce(self.symm, self.full, 0, self.shard_len * self.full.element_size())
            self.gather_ev = torch.cuda.Event()
            self.gather_ev.record(o.gather_stream)

    def unshard(self, direction: int = 1) -> None:
        o = self.o
        W = o.world
        if W > 1 and self.symm is not None:
            o.ensure_staged()
            if self.sharded:
                if self.gather_ev is None:
                    self.issue_gather()
                o.wait_gather(self)
                self.sharded = False
            o.prefetch_after(self, direction)
            return
        if not self.sharded and self.master_version == self.master._version:
            return
        self._alloc_full()
        self.stage()
        if W == 1:
            self.full.copy_(self.shard_lowp)
        else:
            dist.all_gather_into_tensor(self.full, self.shard_lowp)
        self.sharded = False

    def reshard(self) -> None:
        if self.o.strategy == ShardingStrategy.NO_SHARD or self.o.world == 1:
            return
        self.full.untyped_storage().resize_(0)
        self.sharded = True
        self.gather_ev = None

    # ---- gradients ----
    def grad_target(self) -> torch.Tensor:
        """Where the reduced gradient shard is accumulated on the device (the host copy follows in ``flush_offload``)."""
        return self.gshard if self.offload else self.master.grad

    def flush_offload(self) -> None:
        if self.offload:
            self.master.grad.add_(self.gshard.to(self.master.grad.device))
            self.gshard.zero_()

    def prepare_backward(self) -> None:
        self.unshard(direction=-1)
        if self.gfull is None:
            self.gfull = torch.zeros(self.n, dtype=torch.float32, device=self.full.device)
            for p, off in zip(self.params, self.offs):
                p.main_grad = self.gfull[off:off + p.numel()].view(p.shape)
        self.pending = set(id(p) for p in self.params if p.requires_grad)
        self.reduced = False

    def reduce_grads(self) -> None:
        if self.reduced or self.gfull is None:
            return
        self.reduced = True
        for p in self.params:                                    # fold what autograd produced in the compute dtype
            if p.grad is not None:
                p.main_grad += p.grad.float()
                p.grad = None
        W = self.o.world
        if W == 1:
            self.grad_target().add_(self.gfull)
        elif self.o.use_symm:
            self.o.peer_reduce_scatter(self)
        else:
            out = torch.empty(self.shard_len, dtype=torch.float32, device=self.gfull.device)
            if self.gfull.is_cuda:
                dist.reduce_scatter_tensor(out, self.gfull, op=dist.ReduceOp.SUM)
            else:
                dist.all_reduce(self.gfull, op=dist.ReduceOp.SUM)
                out.copy_(self.gfull[self.o.rank * self.shard_len:(self.o.rank + 1) * self.shard_len])
            self.grad_target().add_(out / W)
        if self.o.require_sync:
            for p in self.params:
                if hasattr(p, "main_grad"):
                    del p.main_grad
            self.gfull = None
            if self.o.strategy == ShardingStrategy.FULL_SHARD:
                self.reshard()


class FullyShardedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, process_group=None, sharding_strategy: ShardingStrategy = ShardingStrategy.FULL_SHARD,
                 cpu_offload: Optional[CPUOffload] = None, auto_wrap_policy: Optional[Callable] = None,
                 mixed_precision: Optional[MixedPrecision] = None, device_id=None, sync_module_states: bool = False,
                 use_orig_params: bool = False, limit_all_gathers: bool = True, **_):
        super().__init__()
        self.module = module
        info = get_info()
        self.world = info.world_size if (dist.is_available() and dist.is_initialized()) else 1
        self.rank = info.rank if self.world > 1 else 0
        self.strategy = sharding_strategy or ShardingStrategy.FULL_SHARD
        self.param_dtype = mixed_precision.param_dtype if mixed_precision else None
        self.reduce_dtype = (mixed_precision.reduce_dtype if mixed_precision else None) or torch.float32
        if self.reduce_dtype not in (torch.float32, torch.bfloat16):
            raise ValueError("reduce_dtype must be float32 or bfloat16")
        self.sync_states = sync_module_states
        self.cpu_offload = cpu_offload
        self.require_sync = True
        dev = next(module.parameters()).device
        self.use_symm = (self.world > 1 and dev.type == "cuda" and symm_available()
                         and os.environ.get("B200_COMM", "symm") == "symm")
        if self.use_symm:
            from .symm import PeerFlags
            self.flags = PeerFlags(dev)
            from ..ops.gemm import overlapped_collectives
            overlapped_collectives(True)        # collective kernels will overlap GEMMs: no stream-K (see ops/gemm.py)
        if mixed_precision and mixed_precision.buffer_dtype:
            for b in module.buffers():
                if b.is_floating_point():
                    b.data = b.data.to(mixed_precision.buffer_dtype)
        # ---- unit discovery: modules selected by the policy, innermost first; the root takes the rest ----
        chosen: List[nn.Module] = []
        if auto_wrap_policy is not None:
            pol = auto_wrap_policy
            for m in module.modules():
                if m is module:
                    continue
                try:
                    ok = pol(m, False, sum(p.numel() for p in m.parameters()))
                except TypeError:
                    ok = pol(m)
                if ok:
                    chosen.append(m)
        claimed: Set[int] = set()
        self.units: List[_Unit] = []
        for m in chosen:
            ps = [p for p in m.parameters() if p.requires_grad and id(p) not in claimed]
            if ps:
                claimed.update(id(p) for p in ps)
                self.units.append(_Unit(self, m, ps, type(m).__name__))
        rest = [p for p in module.parameters() if p.requires_grad and id(p) not in claimed]
        if rest:
            self.units.append(_Unit(self, module, rest, "root"))
        self._flat_params = nn.ParameterList([u.master for u in self.units])
        self._order: List[_Unit] = []           # forward execution order (recorded on the first forward)
        self._exposed: List[tuple] = []
        self._rs_epoch = 0
        if self.use_symm:
            self.gather_stream = torch.cuda.Stream(dev)
            self.rs_stream = torch.cuda.Stream(dev)
            nmax = max(u.n for u in self.units)
            self._wire_half = (nmax + 63) // 64 * 64
            self._wire = SymmBuffer(2 * self._wire_half, self.reduce_dtype, dev, zero=False)
        # ---- hooks ----
        for u in self.units:
            u.module.register_forward_pre_hook(lambda m, a, uu=u: uu.unshard())
            if u.module is not module:
                u.module.register_forward_hook(lambda m, a, o, uu=u: self._post_forward(uu))
                u.module.register_full_backward_pre_hook(lambda m, g, uu=u: uu.prepare_backward())
            for p in u.params:
                p._grad_ready = (lambda pp=p: self._on_grad(pp, True))
                p.register_post_accumulate_grad_hook(lambda pp: self._on_grad(pp, False))
        self._cb_queued = False

    # ------------------------------------------------------------------ peer-memory collectives
    def ensure_staged(self) -> None:
        """After an optimizer step the compute-dtype shards the peers read are stale. Restage ALL units at once between
        two device-side barriers (peers are done reading the old ones / every new one is visible) — two barriers per
        optimizer step instead of two per unit per pass."""
        if all(u.master_version == u.master._version for u in self.units):
            return
        from .symm import peer_barrier
        peer_barrier(self.flags, channel=7)
        for u in self.units:
            u.stage()
            u.sharded = True                # whatever is gathered is stale now
            u.gather_ev = None
        peer_barrier(self.flags, channel=7)

    def wait_gather(self, u: "_Unit") -> None:
        cur = torch.cuda.current_stream()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(cur)
        cur.wait_event(u.gather_ev)
        e1.record(cur)
        self._note_exposed(e0, e1)

    def _note_exposed(self, e0, e1) -> None:
        self._exposed.append((e0, e1))
        if len(self._exposed) > 4096:
            del self._exposed[:2048]

    def prefetch_after(self, u: "_Unit", direction: int) -> None:
        """Issue the gather of the unit that runs after ``u`` (forward order, or reverse order in backward)."""
        if u not in self._order:
            if direction > 0:
                self._order.append(u)
            return
        i = self._order.index(u) + direction
        if 0 <= i < len(self._order):
            nxt = self._order[i]
            if nxt.sharded and nxt.gather_ev is None and self.strategy != ShardingStrategy.NO_SHARD:
                nxt.issue_gather()

    def peer_reduce_scatter(self, u: "_Unit") -> None:
        """``master.grad += sum over ranks of gfull / world`` for this rank's slice: one kernel on the reduce stream."""
        import ctypes
        from ..ops import _native as nv
        vp, i32, i64, u32, f32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_uint, ctypes.c_float
        cur = torch.cuda.current_stream()
        self.rs_stream.wait_stream(cur)                          # the unit's gradients are complete in stream order
        u.gfull.record_stream(self.rs_stream)
        self._rs_epoch += 1
        grid = max(1, min(64, u.n // 4 // 256))
        base = self.flags.take_grid_target(grid, 2)
        fn = nv.proto("b200_reduce_scatter", [vp, vp, vp, vp, i64, i64, i32, i32, u32, f32, i32, i32, vp, u32, i32, vp])
        with torch.cuda.stream(self.rs_stream):
            nv.check(fn(nv.ptr(u.gfull), ctypes.cast(self._wire.c_ptrs(), vp), ctypes.cast(self.flags.flags.c_ptrs(), vp),
                        nv.ptr(u.grad_target()), u.n, self._wire_half, self.world, self.rank, self._rs_epoch, 1.0 / self.world, 1,
                        nv.dtype_code(self.reduce_dtype), nv.ptr(self.flags.grid_ctr), base, grid, nv.stream()),
                     "reduce_scatter")
            nv.count_launch()

    def exposed_ms(self, reset: bool = True) -> float:
        """Time (ms, this rank) the compute stream spent waiting for gathers and for the last reduce-scatters."""
        torch.cuda.synchronize()
        t = sum(a.elapsed_time(b) for a, b in self._exposed)
        if reset:
            self._exposed.clear()
        return t

    # ------------------------------------------------------------------ scheduling
    def _post_forward(self, u: _Unit) -> None:
        if self.strategy == ShardingStrategy.FULL_SHARD:
            u.reshard()

    def _on_grad(self, p: nn.Parameter, direct: bool = False) -> None:
        from ..utils.arena import grad_landed
        if not grad_landed(p, direct):          # a tied weight: wait for its last contribution
            return
        u: _Unit = p._fsdp_unit
        if not self._cb_queued:
            self._cb_queued = True
            torch.autograd.Variable._execution_engine.queue_callback(self._finalize)
        if u.gfull is None:
            u.prepare_backward()
        u.pending.discard(id(p))
        if not u.pending and self.require_sync:
            u.reduce_grads()

    def _finalize(self) -> None:
        self._cb_queued = False
        from ..utils.arena import reset_uses
        reset_uses(p for u in self.units for p in u.params)
        if not self.require_sync:
            for u in self.units:                 # no_sync: keep accumulating in the full fp32 buffers
                for p in u.params:
                    if p.grad is not None and u.gfull is not None:
                        p.main_grad += p.grad.float()
                        p.grad = None
            return
        for u in self.units:
            if not u.reduced:
                u.reduce_grads()
        if self.use_symm:
            cur = torch.cuda.current_stream()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(cur)
            cur.wait_stream(self.rs_stream)                      # the optimizer may read master.grad after this
            e1.record(cur)
            self._note_exposed(e0, e1)
        for u in self.units:
            u.flush_offload()

    def forward(self, *args, **kwargs):
        from ..utils.arena import reset_uses
        reset_uses(p for u in self.units for p in u.params)        # use counts of a pass without backward (eval) are stale
        root = self.units[-1] if self.units and self.units[-1].module is self.module else None
        if root is not None:
            root.unshard()
            if torch.is_grad_enabled():
                root.prepare_backward()
        return self.module(*args, **kwargs)

    # ------------------------------------------------------------------ optimizer-facing surface
    def parameters(self, recurse: bool = True):
        return iter(self._flat_params)

    def named_parameters(self, prefix: str = "", recurse: bool = True, remove_duplicate: bool = True):
        for i, u in enumerate(self.units):
            yield f"{prefix}_fsdp_flat_param.{i}.{u.name}", u.master

    def zero_grad(self, set_to_none: bool = False) -> None:
        for u in self.units:
            u.master.grad.zero_()

    def clip_grad_norm_(self, max_norm: float) -> torch.Tensor:
        dev = self.units[0].full.device
        ss = torch.zeros(1, device=dev)
        for u in self.units:
            ss += u.master.grad.float().pow(2).sum().to(dev)
        if self.world > 1:
            dist.all_reduce(ss)
        norm = ss.sqrt()
        coef = (max_norm / (norm + 1e-6)).clamp(max=1.0)
        for u in self.units:
            u.master.grad.mul_(coef.to(u.master.grad.device))
        return norm[0]

    class _NoSync:
        def __init__(self, o):
            self.o = o

        def __enter__(self):
            self.old = self.o.require_sync
            self.o.require_sync = False

        def __exit__(self, *a):
            self.o.require_sync = self.old

    def no_sync(self):
        return FullyShardedDataParallel._NoSync(self)

    # ------------------------------------------------------------------ state dicts
    def full_state_dict(self, cpu: bool = True, rank0_only: bool = False) -> Dict[str, torch.Tensor]:
        for u in self.units:
            u.unshard()
        sd = {}
        if not rank0_only or self.rank == 0:
            for k, v in self.module.state_dict().items():
                sd[k] = v.detach().float().cpu().clone() if cpu else v.detach().clone()
        if self.strategy == ShardingStrategy.FULL_SHARD:
            for u in self.units:
                u.reshard()
        return sd

    def state_dict(self, *a, **k):
        return self.full_state_dict(cpu=False)

    def load_state_dict(self, sd, strict: bool = True):
        for u in self.units:
            u.unshard()
        res = self.module.load_state_dict(sd, strict=strict)
        with torch.no_grad():
            for u in self.units:
                lo, hi = self.rank * u.shard_len, (self.rank + 1) * u.shard_len
                u.master.data.copy_(u.full[lo:hi].float())
        return res


FSDP = FullyShardedDataParallel


def fully_shard(module: nn.Module, *, mesh: Optional[DeviceMesh] = None, reshard_after_forward: bool = True,
                mp_policy: Optional[MixedPrecisionPolicy] = None, offload_policy: Optional[OffloadPolicy] = None,
                **_) -> FullyShardedDataParallel:
    """FSDP2 entry point. Applied to the root only (as in the reference) it yields ONE flat group: a single
    gather before forward and a single reduce-scatter after backward. Call it on each block first, then on the
    root, to get per-block groups — nested calls are honoured because already-sharded parameters are skipped."""
    mp = MixedPrecision(param_dtype=mp_policy.param_dtype, reduce_dtype=mp_policy.reduce_dtype) if mp_policy else None
    strat = ShardingStrategy.FULL_SHARD if reshard_after_forward else ShardingStrategy.SHARD_GRAD_OP
    # the base ``OffloadPolicy()`` the reference passes means "no offload"; ``CPUOffloadPolicy`` keeps the fp32 shards,
    # their gradients and the optimizer state in (pinned) host memory
    off = CPUOffload(offload_params=True) if isinstance(offload_policy, CPUOffloadPolicy) else None
    return FullyShardedDataParallel(module, sharding_strategy=strat, mixed_precision=mp, cpu_offload=off,
                                    sync_module_states=True)


def get_state_dict(model: nn.Module, optimizers: Iterable = (), options: Optional[StateDictOptions] = None):
    """``torch.distributed.checkpoint.state_dict.get_state_dict``: (model_state_dict, optim_state_dict)."""
    opts = options or StateDictOptions()
    if isinstance(model, FullyShardedDataParallel):
        msd = model.full_state_dict(cpu=opts.cpu_offload or opts.full_state_dict)
    else:
        msd = model.state_dict()
    osd = {}
    for i, o in enumerate(optimizers if isinstance(optimizers, (list, tuple)) else [optimizers]):
        if o is not None and o != ():
            osd[i] = o.state_dict()
    return msd, osd
