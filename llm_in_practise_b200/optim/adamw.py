"""K11/K12/K13/K15 — optimizers over flat arenas (``csrc/optim.cu``).

``FusedAdamW``       torch.optim.AdamW semantics (LLM_Distributed_Trainning/PyTorch/ddp_basics/
                     ddp_gpt_wikitext2.py:281; DeepSpeed ``"optimizer": {"type": "AdamW"}``): one launch per
                     flat range, fp32 state, optional bf16 working copy written in the same pass.
``PagedAdamW8bit``   the reference's ``optim="paged_adamw_8bit"`` (Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:151):
                     block-wise (256) dynamically quantised 8-bit m and v. "Paged" in bitsandbytes means the
                     state may spill to host through CUDA managed memory; with 180 GB of HBM the state of a
                     LoRA run (~10 MB) is simply resident, so paging is a no-op here.
``clip_grad_norm_``  global L2 norm + scale; the coefficient stays on the device and is consumed by the
                     next optimizer launch, so a step has no host synchronisation.

All accept either a ``ParamArena`` (fast path: one kernel for every parameter) or a plain iterable
of parameters (one kernel per tensor). CPU tensors take a PyTorch implementation of the same maths.
"""
from __future__ import annotations

import ctypes
import math
from typing import Iterable, List, Optional, Union

import torch

from ..ops import _native as nv
from ..ops.w4 import create_dynamic_map, dequantize_blockwise_8bit, quantize_blockwise_8bit
from ..utils.arena import ParamArena

vp, i32, i64, f32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float


class _Hyper(ctypes.Structure):
    _fields_ = [(n, ctypes.c_float) for n in
                ("lr", "beta1", "beta2", "eps", "weight_decay", "bias_c1", "bias_c2", "grad_scale")]


def _hyper(group, step, grad_scale):
    b1, b2 = group["betas"]
    return _Hyper(group["lr"], b1, b2, group["eps"], group["weight_decay"], 1.0 - b1 ** step, 1.0 - b2 ** step,
                  grad_scale)


class _Range:
    """One flat (param, grad) range plus its optimizer state."""

    def __init__(self, p: torch.Tensor, g: torch.Tensor, lowp: Optional[torch.Tensor] = None):
        self.p, self.g, self.lowp = p, g, lowp
        self.state: dict = {}


class FusedAdamW(torch.optim.Optimizer):
    eight_bit = False

    def __init__(self, params: Union[ParamArena, Iterable], lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.01, lowp_copy: Optional[torch.Tensor] = None, block_size: int = 256,
                 shard: Optional[slice] = None):
        self.arena = params if isinstance(params, ParamArena) else None
        self.block_size = block_size
        self._shard = shard
        if self.arena is not None:
            anchor = torch.nn.Parameter(self.arena.data[:1].detach(), requires_grad=False)  # keeps Optimizer API happy
            plist = [anchor]
        else:
            plist = list(params)
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        super().__init__(plist, defaults)
        self._step = 0
        self._ranges: Optional[List[_Range]] = None
        self._lowp = lowp_copy
        self.dyn_scale: Optional[torch.Tensor] = None   # device scalar multiplied into every gradient
        self.found_inf: Optional[torch.Tensor] = None   # device flag: skip the step when non-zero
        self.grad_scale = 1.0
        self._graph_mode = False

    # ------------------------------------------------------------------ CUDA-graph capturable mode
    def enable_graph_mode(self) -> None:
        """Hyper-parameters (lr, bias corrections, ...) travel through device memory instead of kernel arguments, so a
        ``step()`` captured in a CUDA graph keeps following the schedule: ``step()`` writes them to pinned host memory and
        issues the host->device copy (a memcpy node under capture); before each replay ``advance_graph_step()`` moves the
        step counter on and rewrites the pinned values. fp32 / bf16 AdamW only (``utils/graph_step.GraphedTrainStep``)."""
        if self.eight_bit:
            raise NotImplementedError("the 8-bit AdamW keeps its hyper-parameters in kernel arguments")
        if self._ranges is None:
            self._ranges = self._build_ranges()
        dev = self._ranges[0].p.device
        self._hyper_host = torch.zeros(len(self._ranges), 8, dtype=torch.float32).pin_memory()
        self._hyper_dev = torch.zeros(len(self._ranges), 8, dtype=torch.float32, device=dev)
        self._graph_mode = True

    def _write_hyper(self) -> None:
        for i, r in enumerate(self._ranges):
            h = _hyper(getattr(r, "group", self.param_groups[0]), self._step, self.grad_scale)
            self._hyper_host[i] = torch.tensor([h.lr, h.beta1, h.beta2, h.eps, h.weight_decay, h.bias_c1, h.bias_c2, h.grad_scale])

    def advance_graph_step(self) -> None:
        """Call before replaying a graph that contains ``step()``."""
        self._step += 1
        self._write_hyper()

    # ------------------------------------------------------------------
    def _build_ranges(self) -> List[_Range]:
        if self.arena is not None:
            sl = self._shard or slice(0, self.arena.numel)
            lp = None if self._lowp is None else self._lowp[sl]
            return [_Range(self.arena.data[sl], self.arena.grad[sl], lp)]
        out = []
        for group in self.param_groups:
            for p in group["params"]:
                if p.requires_grad:
                    out.append(_Range(p.data.view(-1), None))
                    out[-1].param = p
                    out[-1].group = group
        return out

    def _init_state(self, r: _Range) -> None:
        n = r.p.numel()
        dev = r.p.device
        if self.eight_bit:
            nb = (n + self.block_size - 1) // self.block_size
            r.state = {
                "state1": torch.full((n,), 0, dtype=torch.uint8, device=dev),
                "state2": torch.zeros(n, dtype=torch.uint8, device=dev),
                "absmax1": torch.zeros(nb, dtype=torch.float32, device=dev),
                "absmax2": torch.zeros(nb, dtype=torch.float32, device=dev),
                "qmap1": create_dynamic_map(signed=True).to(dev),
                "qmap2": create_dynamic_map(signed=False).to(dev),
            }
            # code index of 0.0 so a fresh state dequantises to exactly zero
            z1 = int((r.state["qmap1"] == 0).nonzero()[0].item())
            z2 = int((r.state["qmap2"] == 0).nonzero()[0].item())
            r.state["state1"].fill_(z1)
            r.state["state2"].fill_(z2)
        else:
            r.state = {"exp_avg": torch.zeros(n, dtype=torch.float32, device=dev),
                       "exp_avg_sq": torch.zeros(n, dtype=torch.float32, device=dev)}

    @torch.no_grad()
    def step(self, closure=None):
        loss = closure() if closure is not None else None
        if self._ranges is None:
            self._ranges = self._build_ranges()
        self._step += 1
        if self._graph_mode:
            self._write_hyper()
            self._hyper_dev.copy_(self._hyper_host, non_blocking=True)      # under capture: a memcpy node, re-read every replay
        for ri, r in enumerate(self._ranges):
            group = getattr(r, "group", self.param_groups[0])
            g = r.g
            if g is None:
                pg = r.param.grad
                if pg is None:
                    continue
                g = pg.reshape(-1)
            if not r.state:
                self._init_state(r)
            h = _hyper(group, self._step, self.grad_scale)
            if self._graph_mode and nv.use_native(r.p):
                fn = nv.proto("b200_adamw_graph", [vp] * 5 + [i64, vp, vp, vp, i32, i32, vp])
                st = r.state
                nv.check(fn(nv.ptr(r.p), nv.ptr(g), nv.ptr(st["exp_avg"]), nv.ptr(st["exp_avg_sq"]), nv.ptr(r.lowp), r.p.numel(),
                            self._hyper_dev[ri].data_ptr(), nv.ptr(self.dyn_scale), nv.ptr(self.found_inf),
                            nv.dtype_code(r.p.dtype), nv.dtype_code(g.dtype), nv.stream()), "adamw_graph")
                nv.count_launch()
                torch.autograd.graph.increment_version(r.p)
            elif nv.use_native(r.p):
                self._native(r, g, h)
            else:
                self._reference(r, g, h)
        return loss

    def _native(self, r: _Range, g: torch.Tensor, h: _Hyper) -> None:
        n = r.p.numel()
        if self.eight_bit:
            if self.block_size != 256:
                raise ValueError("the CUDA 8-bit AdamW uses 256-value blocks")
            fn = nv.proto("b200_adamw8bit", [vp] * 9 + [i64, ctypes.POINTER(_Hyper), vp, vp, i32, i32, vp])
            s = r.state
            nv.check(fn(nv.ptr(r.p), nv.ptr(g), nv.ptr(s["state1"]), nv.ptr(s["state2"]), nv.ptr(s["absmax1"]),
                        nv.ptr(s["absmax2"]), nv.ptr(s["qmap1"]), nv.ptr(s["qmap2"]), nv.ptr(r.lowp), n, ctypes.byref(h),
                        nv.ptr(self.dyn_scale), nv.ptr(self.found_inf), nv.dtype_code(r.p.dtype),
                        nv.dtype_code(g.dtype), nv.stream()), "adamw8bit")
        else:
            fn = nv.proto("b200_adamw", [vp] * 5 + [i64, ctypes.POINTER(_Hyper), vp, vp, i32, i32, vp])
            s = r.state
            nv.check(fn(nv.ptr(r.p), nv.ptr(g), nv.ptr(s["exp_avg"]), nv.ptr(s["exp_avg_sq"]), nv.ptr(r.lowp), n,
                        ctypes.byref(h), nv.ptr(self.dyn_scale), nv.ptr(self.found_inf), nv.dtype_code(r.p.dtype),
                        nv.dtype_code(g.dtype), nv.stream()), "adamw")
        nv.count_launch()
        # the kernel wrote behind autograd's back: bump the (shared) version counter so cached bf16
        # operand copies derived from these parameters are rebuilt
        torch.autograd.graph.increment_version(r.p)

    def _reference(self, r: _Range, g: torch.Tensor, h: _Hyper) -> None:
        if self.found_inf is not None and float(self.found_inf) != 0.0:
            return
        gs = h.grad_scale * (float(self.dyn_scale) if self.dyn_scale is not None else 1.0)
        gf = g.float() * gs
        s = r.state
        if self.eight_bit:
            m = dequantize_blockwise_8bit(s["state1"], s["absmax1"], s["qmap1"], self.block_size)
            v = dequantize_blockwise_8bit(s["state2"], s["absmax2"], s["qmap2"], self.block_size)
        else:
            m, v = s["exp_avg"], s["exp_avg_sq"]
        m = h.beta1 * m + (1 - h.beta1) * gf
        v = h.beta2 * v + (1 - h.beta2) * gf * gf
        pf = r.p.float() * (1 - h.lr * h.weight_decay)
        pf = pf - (h.lr / h.bias_c1) * m / (v.sqrt() / math.sqrt(h.bias_c2) + h.eps)
        r.p.copy_(pf.to(r.p.dtype))
        if r.lowp is not None:
            r.lowp.copy_(pf.to(r.lowp.dtype))
        if self.eight_bit:
            s["state1"], s["absmax1"] = quantize_blockwise_8bit(m, s["qmap1"], self.block_size)
            s["state2"], s["absmax2"] = quantize_blockwise_8bit(v, s["qmap2"], self.block_size)
        else:
            s["exp_avg"], s["exp_avg_sq"] = m, v

    # ------------------------------------------------------------------
    def zero_grad(self, set_to_none: bool = False):
        if self.arena is not None:
            self.arena.zero_grad()
            return
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is not None:
                    if set_to_none and not hasattr(p, "main_grad"):
                        p.grad = None
                    else:
                        p.grad.zero_()

    def state_dict(self):
        ranges = self._ranges or []
        return {"step": self._step, "param_groups": [{k: v for k, v in g.items() if k != "params"} for g in self.param_groups],
                "ranges": [{k: v.clone() for k, v in r.state.items()} for r in ranges], "eight_bit": self.eight_bit}

    def load_state_dict(self, sd):
        self._step = sd["step"]
        for g, s in zip(self.param_groups, sd["param_groups"]):
            g.update(s)
        if self._ranges is None:
            self._ranges = self._build_ranges()
        for r, st in zip(self._ranges, sd["ranges"]):
            r.state = {k: v.to(r.p.device) for k, v in st.items()}


class PagedAdamW8bit(FusedAdamW):
    eight_bit = True


AdamW8bit = PagedAdamW8bit


# ------------------------------------------------------------------------------------------------
# gradient norm / clipping / overflow
# ------------------------------------------------------------------------------------------------
def grad_sumsq(grads: Iterable[torch.Tensor], out: Optional[torch.Tensor] = None, pre_scale: float = 1.0) -> torch.Tensor:
    """out[0] = sum of squares over all tensors, out[1] = 1.0 if any value is inf/nan (device tensor)."""
    grads = [g for g in grads if g is not None]
    dev = grads[0].device
    if out is None:
        out = torch.zeros(2, dtype=torch.float32, device=dev)
    for g in grads:
        if nv.use_native(g):
            gc = g if g.is_contiguous() else g.contiguous()
            fn = nv.proto("b200_sumsq", [vp, i64, vp, f32, i32, vp])
            nv.check(fn(nv.ptr(gc), gc.numel(), nv.ptr(out), pre_scale, nv.dtype_code(gc.dtype), nv.stream()), "sumsq")
            nv.count_launch()
        else:
            gf = g.float() * pre_scale
            out[0] += (gf * gf).sum()
            if not torch.isfinite(gf).all():
                out[1] = 1.0
    return out


def clip_coef(sumsq: torch.Tensor, max_norm: float):
    """-> (coef, norm) device scalars: coef = min(1, max_norm / (norm + 1e-6))."""
    coef = torch.empty(1, dtype=torch.float32, device=sumsq.device)
    norm = torch.empty(1, dtype=torch.float32, device=sumsq.device)
    if nv.use_native(sumsq):
        fn = nv.proto("b200_clip_coef", [vp, f32, vp, vp, vp])
        nv.check(fn(nv.ptr(sumsq), max_norm, nv.ptr(coef), nv.ptr(norm), nv.stream()), "clip_coef")
        nv.count_launch()
    else:
        norm[0] = sumsq[0].sqrt()
        coef[0] = min(1.0, max_norm / (float(norm[0]) + 1e-6)) if max_norm > 0 else 1.0
    return coef, norm


def scale_(t: torch.Tensor, scale: float = 1.0, dyn_scale: Optional[torch.Tensor] = None) -> torch.Tensor:
    if nv.use_native(t) and t.is_contiguous():
        fn = nv.proto("b200_scale", [vp, i64, f32, vp, i32, vp])
        nv.check(fn(nv.ptr(t), t.numel(), scale, nv.ptr(dyn_scale), nv.dtype_code(t.dtype), nv.stream()), "scale")
        nv.count_launch()
    else:
        t.mul_(scale * (float(dyn_scale) if dyn_scale is not None else 1.0))
    return t


def clip_grad_norm_(params: Union[ParamArena, Iterable], max_norm: float, optimizer: Optional[FusedAdamW] = None) -> torch.Tensor:
    """torch.nn.utils.clip_grad_norm_ semantics. With ``optimizer`` given, the coefficient is handed to
    the optimizer (applied inside its update kernel) instead of rescaling the gradients in memory."""
    if isinstance(params, ParamArena):
        grads = [params.grad]
    else:
        grads = [p.grad for p in params if p.grad is not None]
    ss = grad_sumsq(grads)
    coef, norm = clip_coef(ss, max_norm)
    if optimizer is not None:
        optimizer.dyn_scale = coef
    else:
        for g in grads:
            scale_(g, 1.0, coef)
    return norm[0]
