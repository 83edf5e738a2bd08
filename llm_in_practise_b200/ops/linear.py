"""K1 + K2 + K3 — the linear layer of this framework.

``FusedLinear`` is one projection ``y = x W^T (+ b)`` whose weight is the row-concatenation of
several *named* sub-projections (``q_proj|k_proj|v_proj``, ``gate_proj|up_proj``), so that a
transformer block issues one tcgen05 GEMM where HF issues three. The base weight is either

* ``dense`` — a bf16/fp16/fp32 Parameter (trainable, or frozen under LoRA), or
* ``nf4``   — an ``NF4Weight`` (QLoRA): expanded to bf16 into a reusable workspace right before the
  GEMM that consumes it, forward and backward (bitsandbytes does the same with
  ``dequantize_4bit`` + cuBLAS; Fine-Tuning/qwen3-14b-qlora-dist-deepspeed.py:101-106).

LoRA adapters (peft ``lora.Linear``; Fine-Tuning/qwen3-8b-lora-dist.py:118-125) attach to any
subset of the sub-projections. For the merged weight they become a *stacked* A ``[r_total, K]`` and a
*block-diagonal* B ``[N, r_total]``, so the whole adapter costs one skinny product
``T = drop(x) A^T`` (CUDA cores, dropout fused) and ONE extra k-block inside the main GEMM
(``y = x W^T + T B^T`` on the tensor cores). The backward mirrors it. Adapter parameters live in a
flat fp32 arena (``utils/arena.py``); their gradients are accumulated in place by the kernels.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import torch
from torch import nn

from . import _native as nv
from . import dequant_pipeline as dq
from . import elementwise as ew
from .gemm import gemm, gemm_w4, gemv16
from .lora import lora_down, lora_dx, lora_wgrad
from .w4 import NF4Weight
from ..utils.arena import note_use

_workspaces: Dict[Tuple[int, torch.dtype, int], torch.Tensor] = {}


def weight_workspace(device, dtype, numel: int, slot: int = 0) -> torch.Tensor:
    """Reusable scratch for expanded 4-bit weights (one per device, dtype and slot)."""
    key = (device.index if device.index is not None else -1, dtype, slot)
    buf = _workspaces.get(key)
    if buf is None or buf.numel() < numel:
        buf = torch.empty(numel, dtype=dtype, device=device)
        _workspaces[key] = buf
    return buf[:numel]


# largest token count served by the 4-bit GEMV instead of dequant + GEMM (measured, profiles/inference.md: at batch 48 the
# table-free GEMV gives 12.4 ms per token vs 14.3 ms for dequant + GEMM, at batch 64 14.1 vs 14.7 ms)
_W4_GEMV_MAX_M = int(__import__("os").environ.get("B200_W4_GEMV_MAX_M", "64"))   # integer-code weights (table-free kernel); NF4: 32

class _LinearFn(torch.autograd.Function):
    """Dense linear. If ``w`` has ``main_grad`` (flat-arena fp32 gradient) the weight gradient is
    accumulated there by the wgrad GEMM (beta = 1) and autograd receives None for it."""

    @staticmethod
    def forward(ctx, x, w, bias, compute_dtype):
        wc = w if w.dtype == compute_dtype else w.to(compute_dtype)
        y = gemm(x, wc, bias=bias)
        if w.requires_grad:
            note_use(w, hasattr(w, "main_grad"))
        if bias is not None and bias.requires_grad:
            note_use(bias, hasattr(bias, "main_grad"))
        ctx.save_for_backward(x, w, bias)
        ctx.compute_dtype = compute_dtype
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, bias = ctx.saved_tensors
        cd = ctx.compute_dtype
        dy = dy.contiguous()
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            wc = w if w.dtype == cd else w.to(cd)
            dx = gemm(dy, wc, b_mn=True)
        if ctx.needs_input_grad[1]:
            mg = getattr(w, "main_grad", None)
            if mg is not None:
                gemm(dy, x, a_mn=True, b_mn=True, out=mg, beta=1.0)
                hook = getattr(w, "_grad_ready", None)
                if hook is not None:
                    hook()
            else:
                dw = gemm(dy, x, a_mn=True, b_mn=True, out_dtype=w.dtype)
        if bias is not None and ctx.needs_input_grad[2]:
            mg = getattr(bias, "main_grad", None)
            s = dy.float().sum(0)
            if mg is not None:
                mg += s.to(mg.dtype)
                hook = getattr(bias, "_grad_ready", None)
                if hook is not None:
                    hook()
            else:
                db = s.to(bias.dtype)
        return dx, dw, db, None


class _FrozenLinearFn(torch.autograd.Function):
    """Linear with a frozen base (dense or NF4): nothing is saved for backward — the input is not
    needed without a weight gradient, and a 4-bit weight is re-expanded into the workspace."""

    @staticmethod
    def forward(ctx, x, layer):
        ctx.layer = layer
        return layer._fwd(x)

    @staticmethod
    def backward(ctx, dy):
        if not ctx.needs_input_grad[0]:
            return None, None
        return ctx.layer._dgrad(dy.contiguous()), None


class _LoraLinearFn(torch.autograd.Function):
    """x -> x W^T + (drop(x) A^T)(sB)^T with W frozen (dense or NF4). ``layer`` carries the buffers;
    the adapter Parameters are passed only so autograd routes the backward here."""

    @staticmethod
    def forward(ctx, x, layer, seed, *adapter_params):
        layer._refresh_adapter_buffers()
        p = layer.lora_dropout if layer.training else 0.0
        T = lora_down(x, layer._A_stack, 1.0, seed, p)                       # [M, r_total]
        y = layer._fwd(x, lora=(T, layer._B_block))
        ctx.save_for_backward(x, T)
        ctx.layer, ctx.seed, ctx.p = layer, seed, p
        return y

    @staticmethod
    def backward(ctx, dy):
        x, T = ctx.saved_tensors
        layer, seed, p = ctx.layer, ctx.seed, ctx.p
        dy = dy.contiguous()
        # dT = dY (sB): the block-diagonal B transposed is a [r_total, N] "A" for the same skinny kernel
        dT = lora_down(dy, layer._Bt_block, 1.0, 0, 0.0)                     # [M, r_total]
        # adapter gradients, accumulated straight into the flat fp32 gradient arena
        lora_wgrad(dT, x, layer._A_grad, 1.0, seed, p)                       # dA_stack [r_total, K]
        for (lo, hi, c0, c1, scale, gBt) in layer._b_slices:
            lora_wgrad(T[:, c0:c1], dy[:, lo:hi], gBt, scale, 0, 0.0)        # d(B_i^T) [r, N_i]
        dx = None
        if ctx.needs_input_grad[0]:
            if p > 0.0:
                dx = layer._dgrad(dy)
                lora_dx(dT, layer._A_stack, dx, 1.0, seed, p)
            else:
                dx = layer._dgrad(dy, lora=(dT, layer._At))
        return (dx, None, None) + (None,) * (len(ctx.needs_input_grad) - 3)


class FusedLinear(nn.Module):
    def __init__(self, in_features: int, parts: Sequence[Tuple[str, int]], bias: bool = False,
                 dtype: torch.dtype = torch.bfloat16, device=None, init: str = "normal", init_std: float = 0.02,
                 materialize: bool = True):
        super().__init__()
        self.in_features = in_features
        self.part_names = [n for n, _ in parts]
        self.part_sizes = [int(s) for _, s in parts]
        self.out_features = sum(self.part_sizes)
        self.compute_dtype = dtype if dtype != torch.float32 else torch.float32
        self.nf4: Optional[NF4Weight] = None
        self.lora_dropout = 0.0
        self._lora: Dict[str, dict] = {}
        self._b_slices: List[tuple] = []
        self._versions = None
        self.layer_uid = 0
        if materialize:
            w = torch.empty(self.out_features, in_features, dtype=dtype, device=device)
            if init == "normal":
                nn.init.normal_(w, std=init_std)
            elif init == "xavier":
                nn.init.xavier_uniform_(w)
            elif init == "kaiming":
                nn.init.kaiming_uniform_(w, a=math.sqrt(5))
            elif init == "zeros":
                w.zero_()
            self.weight = nn.Parameter(w)
        else:
            self.register_parameter("weight", None)
        if bias:
            self.bias = nn.Parameter(torch.zeros(self.out_features, dtype=dtype, device=device))
        else:
            self.register_parameter("bias", None)

    # ------------------------------------------------------------------ base weight handling
    def part_range(self, name: str) -> Tuple[int, int]:
        i = self.part_names.index(name)
        lo = sum(self.part_sizes[:i])
        return lo, lo + self.part_sizes[i]

    def quantize_nf4(self, double_quant: bool = True) -> None:
        """Freeze and replace the dense weight by its NF4 quantisation (QLoRA base)."""
        with torch.no_grad():
            self.nf4 = NF4Weight.quantize(self.weight.data, double_quant=double_quant)
        self.weight = None

    def set_nf4(self, q: NF4Weight) -> None:
        self.nf4 = q
        self.weight = None

    def _materialize_weight(self, dtype=None) -> torch.Tensor:
        if self.nf4 is not None:
            dtype = dtype or torch.bfloat16
            if dq.ENABLED and self.nf4.packed.is_cuda and nv.available() and not torch.cuda.is_current_stream_capturing():
                return dq.pipeline(self.nf4.packed.device).get(self.nf4, dtype, False)   # expanded one GEMM ahead
            ws = weight_workspace(self.nf4.packed.device, dtype, self.out_features * self.in_features)
            return self.nf4.dequantize(dtype, out=ws.view(self.out_features, self.in_features))
        return self.weight

    def _fwd(self, x: torch.Tensor, lora=None) -> torch.Tensor:
        """y = x W^T (+ low-rank term) (+ bias); a 4-bit base goes through the dequant-fused tcgen05 GEMM."""
        if self.nf4 is not None:
            y = gemm_w4(x, self.nf4, lora=lora, bias=self.bias)
            if y is not None:
                return y
        return gemm(x, self._materialize_weight(x.dtype), lora=lora, bias=self.bias)

    def _dgrad(self, dy: torch.Tensor, lora=None) -> torch.Tensor:
        """dX = dY x W (+ low-rank term). A 4-bit base is expanded straight into W^T so the GEMM reads two K-major
        operands (the MN-major B path of the tcgen05 kernel is ~25 % slower)."""
        q = self.nf4
        if q is not None:
            dx = gemm_w4(dy, q, b_mn=True, lora=lora)          # codes expanded inside the GEMM (peer shards included)
            if dx is not None:
                return dx
        if q is not None and dy.is_cuda and getattr(q, "can_transpose", lambda: False)() and dy.dtype != torch.float32:
            if dq.ENABLED and nv.available() and not torch.cuda.is_current_stream_capturing():
                return gemm(dy, dq.pipeline(q.packed.device).get(q, dy.dtype, True), lora=lora)
            ws = weight_workspace(q.packed.device, dy.dtype, self.out_features * self.in_features)
            wt = q.dequantize(dy.dtype, out=ws.view(self.in_features, self.out_features), transpose=True)
            return gemm(dy, wt, lora=lora)
        return gemm(dy, self._materialize_weight(dy.dtype), b_mn=True, lora=lora)

    def dense_weight(self) -> torch.Tensor:
        """A fresh dense copy of the base weight (for merging / export)."""
        if self.nf4 is not None:
            return self.nf4.dequantize(torch.bfloat16).clone()
        return self.weight.detach()

    # ------------------------------------------------------------------ LoRA
    def request_lora(self, arena, targets: Dict[str, Tuple[int, float]], dropout: float, prefix: str) -> None:
        """Phase 1: reserve arena space. ``targets``: part name -> (r, lora_alpha)."""
        names = [n for n in self.part_names if n in targets]
        if not names:
            return
        r_total = sum(targets[n][0] for n in names)
        if r_total > 64:
            raise ValueError("sum of LoRA ranks on one fused projection must be <= 64")
        self.lora_dropout = float(dropout)
        self._lora_names = names
        self._lora_cfg = {n: targets[n] for n in names}
        self._lora_rtotal = r_total
        self._arena = arena
        self._a_slot = arena.request(prefix + ".lora_A_stack", (r_total, self.in_features))
        self._b_slots = {n: arena.request(f"{prefix}.{n}.lora_Bt", (targets[n][0], self.part_sizes[self.part_names.index(n)]))
                         for n in names}
        if self.weight is not None:
            self.weight.requires_grad_(False)
        if self.bias is not None:
            self.bias.requires_grad_(False)

    def materialize_lora(self) -> None:
        """Phase 2 (after ``arena.finalize()``): create the Parameters and initialise like peft
        (A: kaiming-uniform(a=sqrt 5), B: zeros)."""
        if not getattr(self, "_lora_names", None):
            return
        arena = self._arena
        c = 0
        self._b_slices = []
        self.lora_A = nn.ParameterDict()
        self.lora_Bt = nn.ParameterDict()
        for n in self._lora_names:
            r, alpha = self._lora_cfg[n]
            pa = arena.parameter(self._a_slot, rows=(c, c + r))
            with torch.no_grad():
                nn.init.kaiming_uniform_(pa, a=math.sqrt(5))
            pb = arena.parameter(self._b_slots[n])
            self.lora_A[n] = pa
            self.lora_Bt[n] = pb
            lo, hi = self.part_range(n)
            self._b_slices.append((lo, hi, c, c + r, alpha / r, arena.grad_view(self._b_slots[n])))
            c += r
        self._A_grad = arena.grad_view(self._a_slot)
        self._A_master = arena.view(self._a_slot)
        dev = self._A_master.device
        rt, K, N = self._lora_rtotal, self.in_features, self.out_features
        cd = torch.bfloat16
        self._A_stack = torch.zeros(rt, K, dtype=cd, device=dev)
        self._At = torch.zeros(K, rt, dtype=cd, device=dev)
        self._B_block = torch.zeros(N, rt, dtype=cd, device=dev)
        self._Bt_block = torch.zeros(rt, N, dtype=cd, device=dev)
        self._versions = None

    @property
    def has_lora(self) -> bool:
        return bool(self._b_slices)

    def _adapter_params(self) -> List[nn.Parameter]:
        return [self.lora_A[n] for n in self._lora_names] + [self.lora_Bt[n] for n in self._lora_names]

    def _refresh_adapter_buffers(self) -> None:
        """Rebuild the bf16 operand buffers when an adapter master changed (optimizer step, load)."""
        vers = tuple(p._version for p in self._adapter_params()) + (self._A_master._version,)
        if vers == self._versions:
            return
        with torch.no_grad():
            self._A_stack.copy_(self._A_master)
            self._At.copy_(self._A_master.t())
            for n, (lo, hi, c0, c1, scale, _) in zip(self._lora_names, self._b_slices):
                bt = self.lora_Bt[n].detach() * scale        # [r, N_i], scaling folded in
                self._Bt_block[c0:c1, lo:hi].copy_(bt)
                self._B_block[lo:hi, c0:c1].copy_(bt.t())
        self._versions = tuple(p._version for p in self._adapter_params()) + (self._A_master._version,)

    def invalidate_adapter_buffers(self) -> None:
        self._versions = None

    def lora_delta(self, name: str) -> torch.Tensor:
        """(alpha / r) * B A for one sub-projection, fp32 [N_i, K] (for merge_and_unload)."""
        r, alpha = self._lora_cfg[name]
        return (alpha / r) * (self.lora_Bt[name].detach().float().t() @ self.lora_A[name].detach().float())

    # ------------------------------------------------------------------ forward
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        shape = x.shape
        x2 = x.reshape(-1, shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        if self.has_lora:
            seed = ew.layer_seed(self.layer_uid) if (self.training and self.lora_dropout > 0) else 0
            y = _LoraLinearFn.apply(x2, self, seed, *self._adapter_params())
        elif (self.nf4 is not None and not torch.is_grad_enabled() and x2.is_cuda
              and (x2.shape[0] <= 8 or (x2.shape[0] <= (_W4_GEMV_MAX_M if hasattr(self.nf4, "_lut") else 32)
                                        and self.in_features % 128 == 0))
              and hasattr(self.nf4, "group") and not hasattr(self.nf4, "shards")
              and x2.dtype in (torch.bfloat16, torch.float16)):
            # decode: fused 4-bit GEMV straight from the packed weights (memory-bound, 4.25 bits/weight read)
            from .w4 import w4_gemv
            q = self.nf4
            sc = q.scales() if hasattr(q, "scales") else q.scale
            lut = q._lut() if hasattr(q, "_lut") else None
            y = w4_gemv(x2, q.packed, sc, getattr(q, "zs", None), q.N, q.K, q.group, lut, self.bias)
        elif self.nf4 is not None or not (self.weight.requires_grad or
                                          (self.bias is not None and self.bias.requires_grad)):
            y = None
            # measured on Qwen3-8B bf16: 6.1 vs 8.5 ms per step at 8 rows, but 20.6 vs 15.1 ms at 64 (mma.sync-bound): <= 16 only
            if self.nf4 is None and x2.shape[0] <= 16 and not torch.is_grad_enabled():
                y = gemv16(x2, self.weight, self.bias)       # decode / serving step: weights read once, no 128-row tile padding
            if y is None:
                y = _FrozenLinearFn.apply(x2, self)
        else:
            if x2.dtype != self.compute_dtype:
                x2 = x2.to(self.compute_dtype)
            y = _LinearFn.apply(x2, self.weight, self.bias, self.compute_dtype)
        return y.view(*shape[:-1], self.out_features)

    def extra_repr(self) -> str:
        base = "nf4" if self.nf4 is not None else "dense"
        return (f"in={self.in_features}, parts={list(zip(self.part_names, self.part_sizes))}, base={base}, "
                f"lora={getattr(self, '_lora_names', [])}")
