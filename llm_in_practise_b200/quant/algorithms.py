"""Post-training W4A16 quantisation algorithms: GPTQ and AWQ (plus the calibration capture they share).

GPTQ (``GPTQModel.quantize`` / llm-compressor ``GPTQModifier``; Quantization/GPTQModel/quantize_qwen3_4b_gptq.py:16-42,
Quantization/LLM-Compressor/GPTQ/quantize_qwen3_4b_gptq.py:19-62): per linear layer, with H = 2 X^T X accumulated over
the calibration activations, quantise the weight column by column in blocks of 128, propagating each column's
rounding error onto the not-yet-quantised columns through the Cholesky factor of H^-1 (dampening 1 % of mean diag);
group 128, ``desc_act=False`` (natural column order), symmetric or asymmetric grids.

AWQ (``AWQModifier``; Quantization/LLM-Compressor/AWQ/quantize_qwen3_4b_awq.py:17-58): per group of linears sharing an
input, search a per-input-channel scale s = mean|x|^alpha (alpha on a 20-point grid) that minimises the output error
of RTN-quantising W·diag(s); the inverse scale is folded into the producer of that input (the preceding RMSNorm
weight, or the rows of up_proj for down_proj), so the quantised model needs no extra op.
"""
from __future__ import annotations

from typing import Callable, Dict, Iterable, List, Optional, Tuple

import torch
from torch import nn

from ..ops.linear import FusedLinear
from .formats import INT4_LUT, W4Weight, pack_codes, quantize_rtn


class ActivationStats:
    """Forward-pre hook accumulator: Hessian X^T X (GPTQ) and mean |x| per channel (AWQ) for one linear."""

    def __init__(self, K: int, device, need_hessian: bool):
        self.H = torch.zeros(K, K, dtype=torch.float32, device=device) if need_hessian else None
        self.abs_mean = torch.zeros(K, dtype=torch.float32, device=device)
        self.n = 0
        self.sample: Optional[torch.Tensor] = None

    def add(self, x: torch.Tensor) -> None:
        x = x.reshape(-1, x.shape[-1]).float()
        if self.H is not None:
            self.H.addmm_(x.t(), x)
        self.abs_mean += x.abs().sum(0)
        self.n += x.shape[0]
        if self.sample is None:
            self.sample = x[:512].clone()


def capture(model: nn.Module, batches: Iterable[torch.Tensor], need_hessian: bool) -> Dict[str, ActivationStats]:
    stats: Dict[str, ActivationStats] = {}
    hooks = []
    for name, m in model.named_modules():
        if isinstance(m, FusedLinear):
            st = ActivationStats(m.in_features, next(model.parameters()).device, need_hessian)
            stats[name] = st
            hooks.append(m.register_forward_pre_hook(lambda mod, a, s=st: s.add(a[0])))
    was = model.training
    model.eval()
    with torch.no_grad():
        for ids in batches:
            model(ids)
    model.train(was)
    for h in hooks:
        h.remove()
    return stats


def _grid_params(g: torch.Tensor, symmetric: bool):
    """scale / zero for rows of one group: g [N, group]."""
    if symmetric:
        scale = g.abs().amax(-1).clamp_min(1e-10) / 7.0
        zero = torch.full_like(scale, 8.0)
    else:
        lo, hi = g.amin(-1), g.amax(-1)
        scale = ((hi - lo) / 15.0).clamp_min(1e-10)
        zero = torch.clamp(torch.round(-lo / scale), 0, 15)
    return scale, zero


@torch.no_grad()
def gptq_quantize(w: torch.Tensor, H: torch.Tensor, group: int = 128, symmetric: bool = True, damp: float = 0.01,
                  block: int = 128) -> W4Weight:
    N, K = w.shape
    W = w.float().clone()
    H = H.clone()
    dead = torch.diag(H) == 0
    H[dead, dead] = 1.0
    W[:, dead] = 0.0
    H += torch.eye(K, device=H.device) * damp * torch.mean(torch.diag(H))
    L = torch.linalg.cholesky(H)
    Hinv = torch.cholesky_inverse(L)
    Hinv = torch.linalg.cholesky(Hinv, upper=True)
    Q = torch.zeros(N, K, dtype=torch.uint8, device=w.device)
    scales = torch.zeros(N, K // group, device=w.device)
    zeros = torch.zeros(N, K // group, device=w.device)
    for b0 in range(0, K, block):
        b1 = min(K, b0 + block)
        Wb = W[:, b0:b1].clone()
        Eb = torch.zeros_like(Wb)
        Hb = Hinv[b0:b1, b0:b1]
        for j in range(b1 - b0):
            col = b0 + j
            if col % group == 0:
                s, z = _grid_params(W[:, col:col + group], symmetric)
                scales[:, col // group], zeros[:, col // group] = s, z
            s, z = scales[:, col // group], zeros[:, col // group]
            wq = torch.clamp(torch.round(Wb[:, j] / s) + z, 0, 15)
            Q[:, col] = wq.to(torch.uint8)
            err = (Wb[:, j] - (wq - z) * s) / Hb[j, j]
            Wb[:, j:] -= err.unsqueeze(1) * Hb[j, j:].unsqueeze(0)
            Eb[:, j] = err
        W[:, b0:b1] = Wb
        W[:, b1:] -= Eb @ Hinv[b0:b1, b1:]
    return W4Weight(pack_codes(Q), scales.reshape(-1).contiguous(), (zeros * scales).reshape(-1).contiguous(), N, K, group,
                    INT4_LUT.to(w.device))


@torch.no_grad()
def awq_search_scale(weights: List[torch.Tensor], abs_mean: torch.Tensor, x: torch.Tensor, group: int = 128,
                     symmetric: bool = False, n_grid: int = 20) -> torch.Tensor:
    """Best per-input-channel scale for a set of weights [N_i, K] sharing the input x [T, K]."""
    Wcat = torch.cat([w.float() for w in weights], 0)
    ref = x @ Wcat.t()
    am = abs_mean.clamp_min(1e-8)
    best, best_s = None, torch.ones_like(am)
    for i in range(n_grid):
        alpha = i / n_grid
        s = am.pow(alpha)
        s = s / (s.max() * s.min()).sqrt()
        wq = quantize_rtn(Wcat * s, group, symmetric).dequantize(torch.float32) / s
        err = ((x @ wq.t()) - ref).pow(2).mean().item()
        if best is None or err < best:
            best, best_s = err, s
    return best_s
