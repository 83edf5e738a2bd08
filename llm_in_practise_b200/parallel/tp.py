"""Tensor parallelism for inference (Megatron layout) over NVLink peer memory.

The reference only ever sets ``--tensor-parallel-size 1`` / ``tensor_parallel_size: 1`` on its vLLM servers
(Deployment/Basics-02/docker-compose.yml:56-65, Deployment/Ray/serve_run_examples/deepseek.py:26, Quantization/README.md:175);
the task's config 5 asks for W4A16 inference at TP 2/4/8, so this is the framework's own design:

* q|k|v and gate|up are **column-parallel** (each rank keeps H/W query heads, Hkv/W kv heads, F/W ffn channels — for 4-bit
  weights that is a row slice of the packed codes and scales, nothing is requantised);
* o_proj and down_proj are **row-parallel** (each rank keeps a K slice aligned to the quantisation group) and their
  partial outputs are summed with ``b200_allreduce`` — one kernel that signals, pulls the W partials over NVLink,
  reduces in fp32, pushes the sum (and its bf16 copy) to every rank and releases the buffer (csrc/comm.cu, K-C1);
* embeddings, norms and lm_head are replicated; the KV cache holds only the local kv heads.

``shard_qwen3`` rewrites a loaded (dense, NF4 or GPTQ/AWQ) model in place; ``Generator`` / ``LLMEngine`` then run unchanged
because they read the head counts from ``model.config``.
"""
from __future__ import annotations

import copy
import ctypes
import os
from typing import List, Optional

import torch
import torch.distributed as dist
from torch import nn

from ..ops import _native as nv
from ..ops.linear import FusedLinear
from ..ops.w4 import NF4Weight
from .dist import get_info
from .symm import PeerFlags, SymmBuffer, allreduce_, allreduce_bf16_nvls_, nvls_enabled, symm_available


class TensorParallel:
    def __init__(self, hidden: int, max_tokens: int = 8192, device=None, group=None):
        info = get_info()
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.rank = dist.get_rank(group) if self.world > 1 else 0
        self.device = device or info.device
        self.group = group
        self.n_max = (max_tokens * hidden + 3) // 4 * 4
        self.fused = self.world > 1 and self.device.type == "cuda" and symm_available()
        if self.fused:
            self.buf = SymmBuffer(self.n_max, torch.float32, self.device, group)
            self.lowp = SymmBuffer(self.n_max, torch.bfloat16, self.device, group)
            self.flags = PeerFlags(self.device)
            # decode-sized messages: one-shot kernel whose epoch lives on the device (CUDA-graph replayable)
            self.half16 = min(self.n_max, 64 * hidden) // 8 * 8
            self.stage16 = SymmBuffer(2 * self.half16, torch.bfloat16, self.device, group)
            self.state16 = torch.zeros(4, dtype=torch.int32, device=self.device)
            self.oneshot = os.environ.get("B200_TP_ONESHOT", "1") != "0"
            # K-C5: all-reduce inside the row-parallel GEMM (prefill-sized products). Tile flags: [max_tiles][2][8] uint32.
            self.max_tiles = 16384
            self.tile_flags = SymmBuffer(self.max_tiles * 16, torch.int32, self.device, group)
            self.ticket = torch.zeros(1, dtype=torch.int32, device=self.device)
            self.fuse_gemm = os.environ.get("B200_TP_GEMM_AR", "1") != "0"
        self.calls = 0
        self.fused_gemm_calls = 0

    def can_fuse_gemm(self, x: torch.Tensor, n_out: int) -> bool:
        """Row-parallel product of ``x`` [tokens, k] with an [n_out, k] weight as ONE launch that also all-reduces."""
        return (self.world > 1 and self.fused and self.fuse_gemm and x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 2
                and x.shape[0] > 64 and n_out % 8 == 0 and x.shape[0] * n_out <= self.n_max and nvls_enabled(self.lowp)
                and nv.use_native(x) and not torch.cuda.is_current_stream_capturing())

    def gemm_allreduce(self, x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
        """sum over ranks of ``x_r @ w_r^T``: the tcgen05 GEMM writes its partial tile into the symmetric bf16 buffer and the
        same launch sums the tiles in place inside NVSwitch as they complete on every rank (``b200_gemm_allreduce``)."""
        from ..ops.gemm import gemm
        M, N = x.shape[0], w.shape[0]
        out = self.lowp.local[:M * N].view(M, N)
        ar = nv.GemmAllReduce()
        ar.d_mc = self.lowp.multicast_ptr
        self._ar_keep = (self.tile_flags.c_ptrs(), self.flags.flags.c_ptrs(), self.lowp.c_ptrs())
        ar.d_ptrs = ctypes.cast(self._ar_keep[2], ctypes.c_void_p)
        ar.tile_flag_ptrs = ctypes.cast(self._ar_keep[0], ctypes.c_void_p)
        ar.sync_flag_ptrs = ctypes.cast(self._ar_keep[1], ctypes.c_void_p)
        ar.ticket = self.ticket.data_ptr()
        ar.world, ar.rank, ar.channel, ar.epoch = self.world, self.rank, 7, self.flags.next_epoch()
        ar.max_tiles = self.max_tiles
        gemm(x, w, out=out, ar=ar)
        self.calls += 1
        self.fused_gemm_calls += 1
        return out.clone()

    def all_reduce(self, y: torch.Tensor) -> torch.Tensor:
        """Sum of the ranks' partial activations ``y`` [tokens, hidden]; returns a tensor of y's dtype."""
        if self.world == 1:
            return y
        self.calls += 1
        n = y.numel()
        if not self.fused or n > self.n_max or n % 4:
            out = y.float()
            dist.all_reduce(out, group=self.group)
            return out.to(y.dtype)
        if self.oneshot and y.dtype in (torch.bfloat16, torch.float16) and n <= self.half16 and n % 8 == 0:
            y = y.contiguous()
            out = torch.empty_like(y)
            vp, i32, i64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong
            fn = nv.proto("b200_allreduce_oneshot", [vp, vp, vp, vp, i64, i64, i32, i32, i32, vp, i32, vp])
            nv.check(fn(nv.ptr(y), ctypes.cast(self.stage16.c_ptrs(), vp), ctypes.cast(self.flags.flags.c_ptrs(), vp), nv.ptr(out),
                        n, self.half16, self.world, self.rank, 6, nv.ptr(self.state16), nv.dtype_code(y.dtype), nv.stream()),
                     "allreduce_oneshot")
            nv.count_launch()
            return out
        if y.dtype == torch.bfloat16 and n % 8 == 0 and nvls_enabled(self.lowp):
            # prefill-sized bf16 activations: summed inside NVSwitch (multimem.ld_reduce, fp32 accumulation), 2 n / W
            # bytes per rank on the wire
            self.lowp.local[:n].view(y.shape).copy_(y)
            allreduce_bf16_nvls_(self.lowp, self.flags, 0, n)
            return self.lowp.local[:n].view(y.shape).clone()
        self.buf.local[:n].view(y.shape).copy_(y)                       # fp32 partial into the symmetric buffer
        allreduce_(self.buf, self.flags, 0, n, 1.0, lowp=self.lowp if y.dtype == torch.bfloat16 else None)
        src = self.lowp.local if y.dtype == torch.bfloat16 else self.buf.local
        return src[:n].view(y.shape).to(y.dtype, copy=True)


# ------------------------------------------------------------------------------------------------
# weight slicing
# ------------------------------------------------------------------------------------------------
def _slice_rows_q(q, lo: int, hi: int):
    """Rows [lo, hi) of a 4-bit matrix (NF4Weight or W4Weight)."""
    kp, kg = q.K // 2, q.K // q.group
    packed = q.packed.view(q.N, kp)[lo:hi].reshape(-1).contiguous()
    if isinstance(q, NF4Weight):
        return NF4Weight(packed=packed, N=hi - lo, K=q.K, absmax=q.scales().view(q.N, kg)[lo:hi].reshape(-1).contiguous(), group=q.group)
    from ..quant.formats import W4Weight
    return W4Weight(packed, q.scale.view(q.N, kg)[lo:hi].reshape(-1).contiguous(),
                    None if q.zs is None else q.zs.view(q.N, kg)[lo:hi].reshape(-1).contiguous(), hi - lo, q.K, q.group, q.lut)


def _slice_cols_q(q, lo: int, hi: int):
    """Columns [lo, hi) (a multiple of the group) of a 4-bit matrix."""
    if lo % q.group or hi % q.group:
        raise ValueError(f"TP slice [{lo}, {hi}) is not aligned to the quantisation group {q.group}")
    kp, kg = q.K // 2, q.K // q.group
    packed = q.packed.view(q.N, kp)[:, lo // 2:hi // 2].reshape(-1).contiguous()
    g0, g1 = lo // q.group, hi // q.group
    if isinstance(q, NF4Weight):
        return NF4Weight(packed=packed, N=q.N, K=hi - lo, absmax=q.scales().view(q.N, kg)[:, g0:g1].reshape(-1).contiguous(), group=q.group)
    from ..quant.formats import W4Weight
    return W4Weight(packed, q.scale.view(q.N, kg)[:, g0:g1].reshape(-1).contiguous(),
                    None if q.zs is None else q.zs.view(q.N, kg)[:, g0:g1].reshape(-1).contiguous(), q.N, hi - lo, q.group, q.lut)


def column_parallel(lin: FusedLinear, world: int, rank: int) -> FusedLinear:
    """Every named part keeps rows [rank, rank+1) / world of its own range."""
    parts, ranges, off = [], [], 0
    for name, size in zip(lin.part_names, lin.part_sizes):
        if size % world:
            raise ValueError(f"{name}: {size} output features do not divide over {world} ranks")
        s = size // world
        parts.append((name, s))
        ranges.append((off + rank * s, off + (rank + 1) * s))
        off += size
    dev = lin.nf4.packed.device if lin.nf4 is not None else lin.weight.device
    dt = lin.bias.dtype if lin.bias is not None else (lin.weight.dtype if lin.weight is not None else torch.bfloat16)
    new = FusedLinear(lin.in_features, parts, bias=lin.bias is not None, dtype=dt, device=dev, materialize=False)
    new.layer_uid, new.compute_dtype = lin.layer_uid, lin.compute_dtype
    with torch.no_grad():
        if lin.nf4 is not None:
            subs = [_slice_rows_q(lin.nf4, lo, hi) for lo, hi in ranges]
            q0 = subs[0]
            cat = lambda xs: None if xs[0] is None else torch.cat(xs)  # noqa: E731
            if isinstance(q0, NF4Weight):
                new.set_nf4(NF4Weight(packed=torch.cat([s.packed for s in subs]), N=sum(s.N for s in subs), K=q0.K,
                                      absmax=torch.cat([s.absmax for s in subs]), group=q0.group))
            else:
                from ..quant.formats import W4Weight
                new.set_nf4(W4Weight(torch.cat([s.packed for s in subs]), torch.cat([s.scale for s in subs]), cat([s.zs for s in subs]),
                                     sum(s.N for s in subs), q0.K, q0.group, q0.lut))
        else:
            new.weight = nn.Parameter(torch.cat([lin.weight.data[lo:hi] for lo, hi in ranges]).contiguous(), requires_grad=False)
        if lin.bias is not None:
            new.bias.copy_(torch.cat([lin.bias.data[lo:hi] for lo, hi in ranges]))
    return new


class RowParallelLinear(nn.Module):
    """y = all_reduce(x_local W_local^T) (+ bias once)."""

    def __init__(self, lin: FusedLinear, tp: TensorParallel):
        super().__init__()
        W, r = tp.world, tp.rank
        if lin.in_features % W:
            raise ValueError(f"{lin.in_features} input features do not divide over {W} ranks")
        k = lin.in_features // W
        dev = lin.nf4.packed.device if lin.nf4 is not None else lin.weight.device
        dt = lin.weight.dtype if lin.weight is not None else torch.bfloat16
        self.inner = FusedLinear(k, list(zip(lin.part_names, lin.part_sizes)), bias=False, dtype=dt, device=dev, materialize=False)
        self.inner.layer_uid, self.inner.compute_dtype = lin.layer_uid, lin.compute_dtype
        with torch.no_grad():
            if lin.nf4 is not None:
                self.inner.set_nf4(_slice_cols_q(lin.nf4, r * k, (r + 1) * k))
            else:
                self.inner.weight = nn.Parameter(lin.weight.data[:, r * k:(r + 1) * k].contiguous(), requires_grad=False)
        self.bias = None if lin.bias is None else nn.Parameter(lin.bias.data.clone(), requires_grad=False)
        self.tp = tp
        self.in_features, self.out_features = k, lin.out_features

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.dim() == 2 and not getattr(self.inner, "has_lora", False) and self.tp.can_fuse_gemm(x, self.out_features):
            y = self.tp.gemm_allreduce(x.contiguous(), self.inner._materialize_weight(x.dtype))
        else:
            y = self.tp.all_reduce(self.inner(x))
        return y if self.bias is None else y + self.bias


def shard_qwen3(model, tp: Optional[TensorParallel] = None, max_tokens: int = 8192):
    """In-place Megatron sharding of a Qwen3/Qwen2-style model for inference. Returns the ``TensorParallel`` handle."""
    cfg = model.config
    tp = tp or TensorParallel(cfg.hidden_size, max_tokens, model.embed_tokens.device)
    W, r = tp.world, tp.rank
    if W == 1:
        return tp
    if cfg.num_attention_heads % W or cfg.num_key_value_heads % W or cfg.intermediate_size % W:
        raise ValueError(f"heads ({cfg.num_attention_heads}/{cfg.num_key_value_heads}) or ffn ({cfg.intermediate_size}) do not divide over TP {W}")
    for layer in model.layers:
        layer.qkv_proj = column_parallel(layer.qkv_proj, W, r)
        layer.gate_up_proj = column_parallel(layer.gate_up_proj, W, r)
        layer.o_proj = RowParallelLinear(layer.o_proj, tp)
        layer.down_proj = RowParallelLinear(layer.down_proj, tp)
        local = copy.copy(layer.cfg)
        local.num_attention_heads, local.num_key_value_heads = cfg.num_attention_heads // W, cfg.num_key_value_heads // W
        local.intermediate_size = cfg.intermediate_size // W
        layer.cfg = local
        if hasattr(layer, "fused_linears"):
            layer.fused_linears = (lambda l=layer: {"qkv_proj": l.qkv_proj, "o_proj": l.o_proj.inner, "gate_up_proj": l.gate_up_proj,
                                                    "down_proj": l.down_proj.inner})
    model.config = copy.copy(cfg)
    model.config.num_attention_heads, model.config.num_key_value_heads = cfg.num_attention_heads // W, cfg.num_key_value_heads // W
    model.config.intermediate_size = cfg.intermediate_size // W
    model.tp = tp
    if tp.device.type == "cuda":
        torch.cuda.empty_cache()
    return tp
