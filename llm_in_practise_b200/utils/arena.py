"""Flat parameter / gradient arenas.

Every trainable tensor of a model lives inside one contiguous buffer per dtype, with its gradient
at the same offset of a twin buffer (``p.grad`` is pre-assigned as a view, and aliased as
``p.main_grad`` for the backward kernels that accumulate into it directly). This is what lets the
optimizer, the gradient-norm scan, the DDP bucket all-reduce and the ZeRO reduce-scatter each be
ONE kernel launch over a flat range instead of a Python loop over parameters — the role played by
DeepSpeed's ``contiguous_gradients`` / flat fp32 partitions and by the c10d Reducer's buckets in the
reference stack (LLM_Distributed_Trainning/DeepSpeed/DeepSpeed-GPTLike-ZeRO-2/ds_config.json:4-12,
LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py:271-279).
"""
from __future__ import annotations

from typing import Dict, Iterable, List, Optional, Tuple

import torch
from torch import nn

_ALIGN = 64  # elements; 256 B for fp32 — keeps every tensor 16-byte aligned for vector / TMA access


def _round_up(n: int, a: int) -> int:
    return (n + a - 1) // a * a


class ParamArena:
    """Two-phase allocator: ``request`` shapes while building the model, then ``finalize`` once."""

    def __init__(self, dtype: torch.dtype = torch.float32, device=None, pad_to: int = 1):
        self.dtype = dtype
        self.device = device
        self.pad_to = pad_to  # total size is padded to a multiple of this (world size * block for sharding)
        self._specs: List[Tuple[str, Tuple[int, ...], int]] = []
        self._size = 0
        self.data: Optional[torch.Tensor] = None
        self.grad: Optional[torch.Tensor] = None
        self.symmetric = False

    def request(self, name: str, shape) -> int:
        if self.data is not None:
            raise RuntimeError("arena already finalised")
        shape = tuple(int(s) for s in shape)
        n = 1
        for s in shape:
            n *= s
        off = self._size
        self._specs.append((name, shape, off))
        self._size = _round_up(off + n, _ALIGN)
        return len(self._specs) - 1

    def finalize(self, with_grad: bool = True, alloc=None) -> "ParamArena":
        """``alloc(numel, dtype, device) -> zeroed 1-D tensor`` lets the caller place the arena in
        symmetric (peer-mapped) memory; the data buffer is requested first, then the gradient buffer."""
        total = _round_up(max(self._size, _ALIGN), max(self.pad_to, 1))
        if alloc is None:
            alloc = lambda n, dt, dev: torch.zeros(n, dtype=dt, device=dev)  # noqa: E731
        else:
            self.symmetric = True
        self.data = alloc(total, self.dtype, self.device)
        self.grad = alloc(total, self.dtype, self.device) if with_grad else None
        return self

    @property
    def numel(self) -> int:
        return 0 if self.data is None else self.data.numel()

    def view(self, idx: int) -> torch.Tensor:
        _, shape, off = self._specs[idx]
        n = 1
        for s in shape:
            n *= s
        return self.data[off:off + n].view(shape)

    def grad_view(self, idx: int) -> torch.Tensor:
        _, shape, off = self._specs[idx]
        n = 1
        for s in shape:
            n *= s
        return self.grad[off:off + n].view(shape)

    def parameter(self, idx: int, rows: Optional[Tuple[int, int]] = None, requires_grad: bool = True) -> nn.Parameter:
        """A Parameter aliasing slot ``idx`` (optionally only rows [lo, hi) of it)."""
        d = self.view(idx)
        g = self.grad_view(idx) if self.grad is not None else None
        if rows is not None:
            d = d[rows[0]:rows[1]]
            g = None if g is None else g[rows[0]:rows[1]]
        p = nn.Parameter(d, requires_grad=requires_grad)
        if g is not None and requires_grad:
            p.grad = g
            p.main_grad = g
        return p

    def zero_grad(self) -> None:
        if self.grad is not None:
            self.grad.zero_()

    def named_slots(self) -> Iterable[Tuple[str, Tuple[int, ...], int]]:
        return iter(self._specs)


def note_use(p, direct: bool) -> None:
    """Called in the FORWARD of an op that will contribute to ``p``'s gradient (only while gradients are recorded).
    A data-parallel wrapper must not reduce a parameter before its LAST contribution has landed, and a parameter can be
    used more than once in a forward (tied embedding / lm_head): ``direct`` contributions are written into ``p.main_grad``
    by the op's backward, which then calls ``p._grad_ready()``; the others arrive through autograd, whose accumulate hook
    fires once after all of them. The wrappers (DDP / FSDP) count both kinds down in ``grad_landed``."""
    if getattr(p, "_grad_ready", None) is None:
        return
    if direct:
        p._uses_direct = getattr(p, "_uses_direct", 0) + 1
    else:
        p._uses_auto = 1


def grad_landed(p, direct: bool) -> bool:
    """One contribution to ``p``'s gradient has landed; True when none is outstanding."""
    if direct:
        if getattr(p, "_uses_direct", 0) > 0:
            p._uses_direct -= 1
    else:
        p._uses_auto = 0
    return getattr(p, "_uses_direct", 0) == 0 and getattr(p, "_uses_auto", 0) == 0


def reset_uses(params) -> None:
    for p in params:
        p._uses_direct = 0
        p._uses_auto = 0


def flatten_module(module: nn.Module, dtype: Optional[torch.dtype] = None, device=None,
                   pad_to: int = 1, only_trainable: bool = True, alloc=None) -> ParamArena:
    """Re-home the (trainable) parameters of an arbitrary ``nn.Module`` into one arena, in
    registration order, preserving values; ``p.data`` / ``p.grad`` become views. Tied parameters
    (the same Parameter object reachable twice) are placed once."""
    params: List[Tuple[str, nn.Parameter]] = []
    seen: Dict[int, str] = {}
    for name, p in module.named_parameters(remove_duplicate=True):
        if only_trainable and not p.requires_grad:
            continue
        if id(p) in seen:
            continue
        seen[id(p)] = name
        params.append((name, p))
    if not params:
        raise ValueError("module has no parameters to flatten")
    dtype = dtype or params[0][1].dtype
    device = device or params[0][1].device
    arena = ParamArena(dtype=dtype, device=device, pad_to=pad_to)
    idxs = [arena.request(n, p.shape) for n, p in params]
    arena.finalize(alloc=alloc)
    with torch.no_grad():
        for (n, p), i in zip(params, idxs):
            v = arena.view(i)
            v.copy_(p.data.to(dtype))
            p.data = v
            g = arena.grad_view(i)
            p.grad = g
            p.main_grad = g
    return arena


class MixedArena(ParamArena):
    """fp32 master + fp32 main-gradient arena (``data`` / ``grad``) with a low-precision twin:
    ``lowp`` holds the working weights the modules compute with (``p.data`` views) and ``lowp_grad`` receives
    whatever autograd produces in that dtype. Kernels that know about ``p.main_grad`` (wgrad GEMM, fused CE)
    accumulate straight into the fp32 arena; ``fold()`` adds the rest in one launch. The optimizer updates the
    master and rewrites ``lowp`` in the same pass (``p_lowp`` of ``b200_adamw``)."""

    lowp: Optional[torch.Tensor] = None
    lowp_grad: Optional[torch.Tensor] = None

    def fold(self, lo: int = 0, hi: Optional[int] = None) -> None:
        hi = self.numel if hi is None else hi
        x, y = self.lowp_grad[lo:hi], self.grad[lo:hi]
        from ..ops import _native as nv
        if nv.use_native(x):
            import ctypes
            vp, i64, i32, f32 = ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_float
            fn = nv.proto("b200_fold_grad", [vp, vp, i64, f32, i32, i32, vp])
            nv.check(fn(nv.ptr(x), nv.ptr(y), hi - lo, 1.0, 1, nv.dtype_code(x.dtype), nv.stream()), "fold_grad")
            nv.count_launch()
        else:
            y += x.float()
            x.zero_()

    def zero_grad(self) -> None:
        super().zero_grad()
        if self.lowp_grad is not None:
            self.lowp_grad.zero_()


def flatten_module_mixed(module: nn.Module, lowp_dtype: torch.dtype = torch.bfloat16, pad_to: int = 1,
                         alloc=None) -> MixedArena:
    """Mixed-precision re-homing: parameters become ``lowp_dtype`` views, masters and main grads are fp32."""
    params: List[Tuple[str, nn.Parameter]] = []
    seen = set()
    for name, p in module.named_parameters(remove_duplicate=True):
        if p.requires_grad and id(p) not in seen:
            seen.add(id(p))
            params.append((name, p))
    if not params:
        raise ValueError("module has no trainable parameters")
    device = params[0][1].device
    arena = MixedArena(dtype=torch.float32, device=device, pad_to=pad_to)
    idxs = [arena.request(n, p.shape) for n, p in params]
    arena.finalize(alloc=alloc)
    arena.lowp = torch.zeros(arena.numel, dtype=lowp_dtype, device=device)
    arena.lowp_grad = torch.zeros(arena.numel, dtype=lowp_dtype, device=device)
    with torch.no_grad():
        for (n, p), i in zip(params, idxs):
            _, shape, off = arena._specs[i]
            k = p.numel()
            arena.view(i).copy_(p.data.float())
            lv = arena.lowp[off:off + k].view(shape)
            lv.copy_(p.data.to(lowp_dtype))
            p.data = lv
            p.grad = arena.lowp_grad[off:off + k].view(shape)
            p.main_grad = arena.grad_view(i)
    return arena
