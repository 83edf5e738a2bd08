"""Symmetric peer memory: the substrate of the fused compute+collective kernels (``csrc/comm.cu``).

A ``SymmBuffer`` is an allocation that exists at the same size on every rank of the node and that
every rank has mapped: rank i holds a device pointer to rank j's copy, so kernels reach peer HBM
over NVLink / NVSwitch with plain loads and stores. Allocation and the pointer exchange use
``torch.distributed._symmetric_memory`` (CUDA VMM + fabric handles under the hood); nothing here
calls NCCL on the data path. ``PeerFlags`` holds the epoch flags of the signalling protocol and the
grid-barrier counter.

On CPU, or in a single-process run, the same objects degrade to ordinary tensors with a
one-element pointer table, so engine code has one shape. The reference has no counterpart: all
of its communication is implicit inside DDP / FSDP / DeepSpeed.
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops import _native as nv

MAX_RANKS = 8
_enabled_groups = set()


def symm_available() -> bool:
    if not (torch.cuda.is_available() and dist.is_available() and dist.is_initialized()):
        return False
    if dist.get_backend() != "nccl" or dist.get_world_size() > MAX_RANKS:
        return False
    try:
        import importlib
        importlib.import_module("torch.distributed._symmetric_memory")
        return True
    except Exception:
        return False


class SymmBuffer:
    """``.local``: this rank's tensor; ``.ptrs``: device pointers of every rank's copy (index = rank)."""

    def __init__(self, numel: int, dtype: torch.dtype, device, group=None, zero: bool = True):
        self.numel, self.dtype = numel, dtype
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.rank = dist.get_rank(group) if self.world > 1 else 0
        self.handle = None
        if self.world > 1 and symm_available():
            import torch.distributed._symmetric_memory as sm
            g = group or dist.group.WORLD
            if g.group_name not in _enabled_groups:
                try:
                    sm.enable_symm_mem_for_group(g.group_name)
                except Exception:
                    pass
                _enabled_groups.add(g.group_name)
            self.local = sm.empty(numel, dtype=dtype, device=device)
            self.handle = sm.rendezvous(self.local, g.group_name)
            self.ptrs: List[int] = [int(p) for p in self.handle.buffer_ptrs]
            self.multicast_ptr = int(getattr(self.handle, "multicast_ptr", 0) or 0)
        else:
            self.local = torch.empty(numel, dtype=dtype, device=device)
            self.ptrs = [self.local.data_ptr()]
            self.multicast_ptr = 0
        if zero:
            self.local.zero_()
        self._carr = (ctypes.c_void_p * MAX_RANKS)(*(self.ptrs + [0] * (MAX_RANKS - len(self.ptrs))))

    @property
    def is_symmetric(self) -> bool:
        return self.handle is not None

    def c_ptrs(self, byte_offset: int = 0):
        """ctypes array of the per-rank pointers (optionally offset), for the C ABI."""
        if byte_offset == 0:
            return self._carr
        return (ctypes.c_void_p * MAX_RANKS)(*([p + byte_offset for p in self.ptrs] + [0] * (MAX_RANKS - len(self.ptrs))))

    def barrier(self) -> None:
        if self.handle is not None:
            self.handle.barrier()


class PeerFlags:
    """Epoch flags (``uint32 [channels][MAX_RANKS]`` per rank, symmetric) + local grid-barrier counter
    + partial-norm slots. One instance serves one stream of fused collectives."""

    CHANNELS = 8

    def __init__(self, device, group=None):
        self.flags = SymmBuffer(self.CHANNELS * MAX_RANKS, torch.int32, device, group)
        self.norms = SymmBuffer(MAX_RANKS * 2, torch.float32, device, group)
        self.grid_ctr = torch.zeros(1, dtype=torch.int32, device=device)
        self.grid_count = 0        # value the counter has after all launches issued so far
        self.epoch = 0
        self.world, self.rank = self.flags.world, self.flags.rank
        if self.flags.is_symmetric:
            torch.cuda.synchronize()
            dist.barrier(group)

    def next_epoch(self) -> int:
        self.epoch += 1
        return self.epoch

    def take_grid_target(self, grid: int, barriers: int) -> int:
        base = self.grid_count
        self.grid_count += grid * barriers
        return base


def peer_barrier(flags: PeerFlags, channel: int = 7) -> None:
    """Device-side barrier across ranks on the current stream (no host involvement)."""
    vp, i32, u32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_uint
    fn = nv.proto("b200_peer_barrier", [vp, i32, i32, i32, u32, vp])
    nv.check(fn(ctypes.cast(flags.flags.c_ptrs(), vp), channel, flags.world, flags.rank, flags.next_epoch(), nv.stream()),
             "peer_barrier")
    nv.count_launch()


def peer_gather(src: SymmBuffer, out: torch.Tensor, src_off_bytes: int, shard_bytes: int, skip: int = -1) -> torch.Tensor:
    """out = concat over ranks of src[r][off : off + shard_bytes]  (pull-based all-gather); ``skip`` >= 0 leaves that
    rank's slot of ``out`` untouched."""
    vp, i32, i64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong
    fn = nv.proto("b200_peer_gather_skip", [vp, i64, vp, i64, i32, i32, vp])
    nv.check(fn(ctypes.cast(src.c_ptrs(), vp), src_off_bytes, nv.ptr(out), shard_bytes, src.world, skip, nv.stream()), "peer_gather")
    nv.count_launch()
    return out


def peer_gather_ce(src: SymmBuffer, out: torch.Tensor, src_off_bytes: int, shard_bytes: int, skip: int = -1) -> torch.Tensor:
    """``peer_gather`` on the copy engines: one ``cudaMemcpyAsync`` per owner on the current stream (no kernel)."""
    vp, i32, i64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong
    fn = nv.proto("b200_peer_gather_ce", [vp, i64, vp, i64, i32, i32, vp])
    nv.check(fn(ctypes.cast(src.c_ptrs(), vp), src_off_bytes, nv.ptr(out), shard_bytes, src.world, skip, nv.stream()), "peer_gather_ce")
    return out


def nvls_enabled(*bufs: Optional[SymmBuffer], min_world: int = 1) -> bool:
    """True when every given buffer has a multicast (NVLS) address, B200_NVLS is not 0 and the group has at least
    ``min_world`` ranks (B200_NVLS=1 forces it at any size)."""
    env = os.environ.get("B200_NVLS", "")
    if env == "0":
        return False
    real = [b for b in bufs if b is not None]
    if not all(getattr(b, "multicast_ptr", 0) for b in real):
        return False
    return env == "1" or all(b.world >= min_world for b in real)


def allreduce_(buf: SymmBuffer, flags: PeerFlags, offset: int, n: int, scale: float,
               lowp: Optional[SymmBuffer] = None, grid: int = 64) -> None:
    """In-place average (``scale`` = 1/world) of fp32 ``buf[offset : offset+n]`` across ranks — K-C1. With a multicast
    address (NVLS) the reduction runs inside NVSwitch (``b200_allreduce_nvls``), otherwise over peer pointers."""
    vp, i32, i64, u32, f32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_uint, ctypes.c_float
    grid = max(1, min(grid, (n // 4 + 255) // 256))
    base = flags.take_grid_target(grid, 1)
    # measured at 2 ranks (profiles/r2): the peer-pointer kernel moves the same bytes faster (0.44 vs 0.66 ms for 252 MiB);
    # the switch only saves traffic from 3 ranks up (2 n / W instead of 2 n (W - 1) / W per rank)
    if nvls_enabled(buf, lowp, min_world=3) and (offset * 4) % 16 == 0:
        fn = nv.proto("b200_allreduce_nvls", [vp, vp, vp, i64, i32, i32, u32, f32, i32, vp, u32, i32, vp])
        lp = (lowp.multicast_ptr + offset * 2) if lowp is not None else None
        nv.check(fn(buf.multicast_ptr + offset * 4, ctypes.cast(flags.flags.c_ptrs(), vp), lp, n, buf.world, buf.rank,
                    flags.next_epoch(), scale, nv.dtype_code(torch.float32), nv.ptr(flags.grid_ctr), base, grid, nv.stream()),
                 "allreduce_nvls")
        nv.count_launch()
        return
    fn = nv.proto("b200_allreduce", [vp, vp, vp, i64, i32, i32, u32, f32, vp, u32, i32, vp])
    lp = ctypes.cast(lowp.c_ptrs(offset * 2), vp) if lowp is not None else None
    nv.check(fn(ctypes.cast(buf.c_ptrs(offset * 4), vp), ctypes.cast(flags.flags.c_ptrs(), vp), lp, n, buf.world, buf.rank,
                flags.next_epoch(), scale, nv.ptr(flags.grid_ctr), base, grid, nv.stream()), "allreduce")
    nv.count_launch()


def allreduce_bf16_nvls_(buf: SymmBuffer, flags: PeerFlags, offset: int, n: int, scale: float = 1.0, grid: int = 64) -> None:
    """In-place sum (times ``scale``) of bf16 ``buf[offset : offset+n]`` across ranks, reduced by the switch with fp32
    accumulation. ``buf`` needs a multicast address; n % 8 == 0, offset % 8 == 0."""
    vp, i32, i64, u32, f32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_uint, ctypes.c_float
    assert buf.dtype == torch.bfloat16 and buf.multicast_ptr and n % 8 == 0 and offset % 8 == 0
    grid = max(1, min(grid, (n // 8 + 255) // 256))
    base = flags.take_grid_target(grid, 1)
    fn = nv.proto("b200_allreduce_nvls", [vp, vp, vp, i64, i32, i32, u32, f32, i32, vp, u32, i32, vp])
    nv.check(fn(buf.multicast_ptr + offset * 2, ctypes.cast(flags.flags.c_ptrs(), vp), None, n, buf.world, buf.rank,
                flags.next_epoch(), scale, nv.dtype_code(torch.bfloat16), nv.ptr(flags.grid_ctr), base, grid, nv.stream()),
             "allreduce_nvls")
    nv.count_launch()
