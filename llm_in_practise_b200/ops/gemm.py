"""K1/K3 — GEMM on the tcgen05 tensor cores (``csrc/gemm_tcgen05.cu``), with a PyTorch oracle.

``gemm`` is the single entry point: ``D = alpha * A @ B^T (+ A2 @ B2^T) (+ bias) + beta * D`` where A is
logically ``[.., M, K]`` and B logically ``[.., N, K]``; either may be handed over *stored transposed*
(``a_mn`` / ``b_mn``) so that the three products of a linear layer read their operands in place:

    forward  y  = x  W^T   -> gemm(x,  W)
    dgrad    dx = dy W     -> gemm(dy, W, b_mn=True)
    wgrad    dW = dy^T x   -> gemm(dy, x, a_mn=True, b_mn=True)

Replaces the cuBLAS calls behind ``nn.Linear`` / ``torch.matmul`` in every reference script
(e.g. LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py:86-165) and the three small
GEMMs of peft's LoRA layer (Fine-Tuning/qwen3-8b-lora-dist.py:118-125).
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional, Tuple

import torch

from . import _native as nv


_sk_workspaces = {}


def ensure_graph_workspace(device: torch.device) -> None:
    """Allocate (outside any capture) the stream-K scratch that GEMMs captured into CUDA graphs use. The stream-K epoch
    lives in the workspace itself, so captured launches are replayable; all graphs of a device share this one scratch and
    must therefore be replayed on one stream at a time (the serving engine and ``Generator`` do)."""
    if os.environ.get("B200_GRAPH_STREAMK", "0") != "1":
        # measured on the serving sweep (Qwen3-8B bf16, decode GEMMs at M = 32..128 are weight-bandwidth-bound): stream-K
        # inside the decode graph costs 10-14 % of the throughput (fp32 partial tiles + fix-up traffic), so captured
        # launches run one CTA pair per tile unless asked otherwise
        return
    key = (device.index, "graph")
    if key not in _sk_workspaces and not torch.cuda.is_current_stream_capturing():
        fn = nv.lib().b200_gemm_workspace_bytes
        fn.restype = ctypes.c_longlong
        _sk_workspaces[key] = torch.zeros(int(fn()), dtype=torch.uint8, device=device)


_overlapped_collectives = 0


def overlapped_collectives(active: bool) -> None:
    """Called by the wrappers that run a peer-memory collective kernel on a side stream WHILE compute continues (DDP
    bucket all-reduce, FSDP reduce-scatter). Such a kernel spins on flags and has a grid barrier, i.e. it needs all of
    its CTAs resident; a stream-K GEMM needs all of ITS CTAs resident too (tile owners wait for the partial sums of
    other CTAs). Two kernels of that kind pending on one GPU can each get only part of their CTAs placed and then wait
    for each other forever (seen as "stream-K wait timed out" in the 2-GPU DDP suite). While any such wrapper is alive,
    GEMMs therefore run one CTA pair per tile (no cross-CTA dependency: their CTAs always drain)."""
    global _overlapped_collectives
    _overlapped_collectives = max(0, _overlapped_collectives + (1 if active else -1))


def _stream_k_workspace(device: torch.device) -> Optional[torch.Tensor]:
    """Per (device, stream) scratch for stream-K GEMMs: zeroed once, then owned by that stream."""
    if _overlapped_collectives:
        return None
    if torch.cuda.is_current_stream_capturing():
        return _sk_workspaces.get((device.index, "graph"))       # None: captured without stream-K
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    ws = _sk_workspaces.get(key)
    if ws is None:
        fn = nv.lib().b200_gemm_workspace_bytes
        fn.restype = ctypes.c_longlong
        ws = torch.zeros(int(fn()), dtype=torch.uint8, device=device)
        _sk_workspaces[key] = ws
    return ws


def _as4d(t: torch.Tensor) -> torch.Tensor:
    while t.dim() < 4:
        t = t.unsqueeze(0)
    if t.dim() != 4:
        raise ValueError("gemm operands must have at most 4 dims")
    return t


def _ref_gemm(a, b, a_mn, b_mn, out, out_dtype, alpha, beta, bias, lora):
    A = a.transpose(-1, -2) if a_mn else a
    B = b.transpose(-1, -2) if b_mn else b
    if A.dim() > 2 and B.dim() > 2 and A.shape[-3] != B.shape[-3]:  # grouped (GQA) batch
        rep = A.shape[-3] // B.shape[-3]
        B = B.repeat_interleave(rep, dim=-3)
    acc = torch.matmul(A.float(), B.float().transpose(-1, -2))
    if lora is not None:
        acc = acc + torch.matmul(lora[0].float(), lora[1].float().t())
    acc = acc * alpha
    if bias is not None:
        acc = acc + bias.float()
    if beta != 0.0:
        acc = acc + beta * out.float()
    res = acc.to(out_dtype)
    if out is not None:
        out.copy_(res)
        return out
    return res


def gemm(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False,
         out: Optional[torch.Tensor] = None, out_dtype: Optional[torch.dtype] = None,
         alpha: float = 1.0, beta: float = 0.0, bias: Optional[torch.Tensor] = None,
         lora: Optional[Tuple[torch.Tensor, torch.Tensor]] = None,
         beta_rows: Optional[torch.Tensor] = None, ar=None) -> torch.Tensor:
    """``beta_rows`` (float32 [M], 2-D problems): per-row beta, ``out[m] = alpha * acc[m] + beta_rows[m] * out[m]``.
    ``ar`` (``nv.GemmAllReduce``): ``out`` is this rank's copy of a symmetric bf16 buffer and is summed over the
    tensor-parallel ranks inside the launch (K-C5, see ``parallel/tp.TensorParallel.gemm_allreduce``).
    See module docstring. ``a``/``b`` are the *stored* tensors (last dim contiguous); leading dims
    (up to two) are batch dims with arbitrary strides; ``b`` may have fewer batch entries in its
    inner batch dim (an integer divisor: grouped-query attention)."""
    if a.dtype != b.dtype:
        raise TypeError("gemm operands must share a dtype")
    if out_dtype is None:
        out_dtype = out.dtype if out is not None else a.dtype
    if beta != 0.0 and out is None:
        raise ValueError("beta != 0 needs `out`")
    if not nv.use_native(a):
        if beta_rows is not None:
            prev = out.float() * beta_rows.float().unsqueeze(-1)
            res = _ref_gemm(a, b, a_mn, b_mn, None, torch.float32, alpha, 0.0, bias, lora) + prev
            out.copy_(res.to(out.dtype))
            return out
        return _ref_gemm(a, b, a_mn, b_mn, out, out_dtype, alpha, beta, bias, lora)

    if a.dtype == torch.float32 and (a_mn or b_mn):
        # the tf32 path only takes K-major operands: materialise the transposes (cold path —
        # fp32 training is the reference's teaching configuration, bf16 is the fast one)
        if a_mn:
            a, a_mn = a.transpose(-1, -2).contiguous(), False
        if b_mn:
            b, b_mn = b.transpose(-1, -2).contiguous(), False
    if a.stride(-1) != 1:
        a = a.contiguous()
    if b.stride(-1) != 1:
        b = b.contiguous()
    a4, b4 = _as4d(a), _as4d(b)
    if a_mn:
        K, M = a4.shape[-2], a4.shape[-1]
    else:
        M, K = a4.shape[-2], a4.shape[-1]
    if b_mn:
        Kb, N = b4.shape[-2], b4.shape[-1]
    else:
        N, Kb = b4.shape[-2], b4.shape[-1]
    if K != Kb:
        raise ValueError(f"contraction mismatch: {K} vs {Kb}")
    nb1, nb0 = a4.shape[0], a4.shape[1]
    if b4.shape[0] not in (nb1, 1) and nb1 != 1:
        raise ValueError("outer batch dims must match")
    if nb0 % b4.shape[1] != 0:
        raise ValueError("inner batch of b must divide inner batch of a")
    b_div = nb0 // b4.shape[1]
    if b4.shape[0] == 1 and nb1 > 1:
        b4 = b4.expand(nb1, *b4.shape[1:])
    batch = nb0 * nb1

    out_shape = (*a.shape[:-2], M, N)
    if out is None:
        out = torch.empty(out_shape, dtype=out_dtype, device=a.device)
    elif tuple(out.shape) != tuple(out_shape) or out.stride(-1) != 1:
        raise ValueError(f"`out` must have shape {out_shape} with a contiguous last dim")
    d4 = _as4d(out)

    args = nv.GemmArgs()
    args.A, args.B, args.D = a4.data_ptr(), b4.data_ptr(), d4.data_ptr()
    args.bias = 0 if bias is None else bias.data_ptr()
    args.bias_dtype = nv.dtype_code(bias.dtype) if bias is not None else 0
    args.M, args.N, args.K = M, N, K
    args.batch, args.nb0, args.b_div = batch, nb0, b_div
    args.a_rs, args.a_bs0, args.a_bs1 = a4.stride(2), a4.stride(1) if nb0 > 1 else 0, a4.stride(0) if nb1 > 1 else 0
    args.b_rs = b4.stride(2)
    args.b_bs0 = b4.stride(1) if b4.shape[1] > 1 else 0
    args.b_bs1 = b4.stride(0) if nb1 > 1 else 0
    args.d_rs, args.d_bs0, args.d_bs1 = d4.stride(2), d4.stride(1) if nb0 > 1 else 0, d4.stride(0) if nb1 > 1 else 0
    args.a_mn, args.b_mn = int(a_mn), int(b_mn)
    args.in_dtype, args.out_dtype = nv.dtype_code(a.dtype), nv.dtype_code(out.dtype)
    args.alpha, args.beta = alpha, beta
    args.beta_rows = 0 if beta_rows is None else beta_rows.data_ptr()
    args.ar = ctypes.addressof(ar) if ar is not None else 0
    keep = None
    if lora is not None:
        a2, b2 = lora
        if a2.dtype != a.dtype or b2.dtype != a.dtype:
            raise TypeError("lora operands must match the main dtype")
        a2 = a2 if a2.stride(-1) == 1 else a2.contiguous()
        b2 = b2 if b2.stride(-1) == 1 else b2.contiguous()
        keep = (a2, b2)
        args.A2, args.B2, args.K2 = a2.data_ptr(), b2.data_ptr(), a2.shape[-1]
        args.a2_rs, args.b2_rs = a2.stride(0), b2.stride(0)
    else:
        args.A2 = args.B2 = 0
        args.K2 = 0
    ws = _stream_k_workspace(a.device)
    if ws is None:
        args.workspace, args.workspace_bytes = 0, 0
    else:
        args.workspace, args.workspace_bytes = ws.data_ptr(), ws.numel()
    rc = nv.lib().b200_gemm(ctypes.byref(args), nv.stream())
    nv.check(rc, f"gemm M={M} N={N} K={K} batch={batch}")
    nv.count_launch()
    del keep
    return out


def linear_fwd(x, w, bias=None, lora=None, out=None, out_dtype=None):
    """y[M,N] = x[M,K] @ w[N,K]^T (+ bias) (+ lora)."""
    return gemm(x, w, bias=bias, lora=lora, out=out, out_dtype=out_dtype)


def linear_dgrad(dy, w, out=None, beta=0.0, out_dtype=None):
    """dx[M,K] = dy[M,N] @ w[N,K]."""
    return gemm(dy, w, b_mn=True, out=out, beta=beta, out_dtype=out_dtype)


def linear_wgrad(dy, x, out=None, beta=0.0, out_dtype=None):
    """dw[N,K] = dy[M,N]^T @ x[M,K]  (beta=1 accumulates into an existing gradient)."""
    return gemm(dy, x, a_mn=True, b_mn=True, out=out, beta=beta, out_dtype=out_dtype)


def gemv16(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None,
           out_dtype: Optional[torch.dtype] = None) -> Optional[torch.Tensor]:
    """``x @ w^T`` for 1..64 rows of 16-bit ``x`` [M, K] and a dense 16-bit ``w`` [N, K], reading ``w`` once
    (csrc/gemv.cu). Returns None when the shape is outside the kernel's domain (the caller then uses ``gemm``)."""
    if not nv.use_native(x) or x.dim() != 2 or x.dtype not in (torch.bfloat16, torch.float16) or w.dtype != x.dtype:
        return None
    M, K = x.shape
    N = w.shape[0]
    out_dtype = out_dtype or x.dtype
    if M > 64 or K % 128 or w.stride(-1) != 1 or w.stride(0) % 8 or out_dtype not in (x.dtype, torch.float32) \
            or (bias is not None and bias.dtype != x.dtype) or w.data_ptr() % 16:
        return None
    if x.stride(-1) != 1 or x.stride(0) % 8 or x.data_ptr() % 16:
        x = x.contiguous()
    y = torch.empty(M, N, dtype=out_dtype, device=x.device)
    vp, i32, i64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong
    fn = nv.proto("b200_gemv16", [vp, vp, vp, vp, i32, i64, i32, i64, i64, i64, i32, i32, vp])
    nv.check(fn(nv.ptr(x), nv.ptr(w), nv.ptr(bias), nv.ptr(y), M, N, K, x.stride(0), w.stride(0), y.stride(0),
                nv.dtype_code(x.dtype), int(out_dtype == torch.float32), nv.stream()), f"gemv16 M={M} N={N} K={K}")
    nv.count_launch()
    return y


# ------------------------------------------------------------------------------------------------
# 4-bit B operand expanded inside the GEMM (csrc/gemm_tcgen05.cu: gemm_w4_tcgen05_kernel)
# ------------------------------------------------------------------------------------------------
class W4Desc(ctypes.Structure):
    _fields_ = [("packed", ctypes.c_void_p * 8), ("scale", ctypes.c_void_p * 8), ("zs", ctypes.c_void_p * 8),
                ("lut", ctypes.c_void_p), ("nshards", ctypes.c_int32), ("rows", ctypes.c_int32), ("kw", ctypes.c_int32),
                ("group", ctypes.c_int32), ("rows_per_shard", ctypes.c_int64)]


# Off by default: on B200 the dense kernel is already bound by the shared-memory port (TMA fill + UMMA operand reads),
# so the extra smem traffic and barrier hand-offs of in-kernel expansion cost more than the separate bandwidth-bound
# dequant pass saves at M = 1024 (gate_up fwd: 0.61 ms fused vs 0.27 + 0.10 ms; profiles/w4_fused_gemm.md).
# B200_W4_FUSED=1 turns it on (it wins when the 16-bit copy would not fit or M is large enough to amortise nothing).
W4_FUSED = os.environ.get("B200_W4_FUSED", "0") == "1"


def w4_desc(q) -> Optional[W4Desc]:
    """Build (and cache on ``q``) the C descriptor of a 4-bit weight: NF4Weight, ShardedNF4 (peer shards) or W4Weight."""
    d = getattr(q, "_w4desc", None)
    if d is not None:
        return d
    d = W4Desc()
    shards = getattr(q, "shards", None)
    if shards is not None:
        if len(shards) > 8:
            return None
        for i, (pp, sp, zp) in enumerate(shards):
            d.packed[i], d.scale[i], d.zs[i] = pp, sp, zp or None
        d.nshards, d.rows_per_shard = len(shards), q.rows
        keep = None
    else:
        sc = q.scales() if hasattr(q, "scales") else q.scale
        zs = getattr(q, "zs", None)
        d.packed[0], d.scale[0], d.zs[0] = q.packed.data_ptr(), sc.data_ptr(), None if zs is None else zs.data_ptr()
        d.nshards, d.rows_per_shard = 1, q.N
        keep = (sc, zs)
    lut = q._lut() if hasattr(q, "_lut") else None
    d.lut = None if lut is None else lut.data_ptr()
    d.rows, d.kw, d.group = q.N, q.K, q.group
    q._w4desc, q._w4keep = d, (keep, lut)
    return d


def gemm_w4(a: torch.Tensor, q, *, b_mn: bool = False, out: Optional[torch.Tensor] = None, bias=None, lora=None,
            alpha: float = 1.0, beta: float = 0.0, force: bool = False) -> Optional[torch.Tensor]:
    """``a @ W^T`` (b_mn False) or ``a @ W`` (b_mn True) with W read as 4-bit codes. Returns None when the fused
    kernel does not cover the shape, so the caller can take the dequantise-then-GEMM path."""
    if not ((W4_FUSED or force) and nv.use_native(a)) or a.dtype not in (torch.bfloat16, torch.float16) or a.dim() != 2:
        return None
    M, K = a.shape
    N = q.K if b_mn else q.N
    if M <= 128 or N % 256 or K % 64 or q.group % 64 or K != (q.N if b_mn else q.K):
        return None
    d = w4_desc(q)
    if d is None or (d.nshards > 1 and d.rows_per_shard % (64 if b_mn else 128)):
        return None
    if a.stride(-1) != 1:
        a = a.contiguous()
    if out is None:
        out = torch.empty(M, N, dtype=a.dtype, device=a.device)
    args = nv.GemmArgs()
    args.A, args.B, args.D = a.data_ptr(), 0, out.data_ptr()
    args.bias = 0 if bias is None else bias.data_ptr()
    args.bias_dtype = nv.dtype_code(bias.dtype) if bias is not None else 0
    args.M, args.N, args.K = M, N, K
    args.batch, args.nb0, args.b_div = 1, 1, 1
    args.a_rs, args.d_rs = a.stride(0), out.stride(0)
    args.a_mn, args.b_mn = 0, int(b_mn)
    args.in_dtype, args.out_dtype = nv.dtype_code(a.dtype), nv.dtype_code(out.dtype)
    args.alpha, args.beta = alpha, beta
    keep = None
    if lora is not None:
        a2, b2 = lora
        a2 = a2 if a2.stride(-1) == 1 else a2.contiguous()
        b2 = b2 if b2.stride(-1) == 1 else b2.contiguous()
        keep = (a2, b2)
        args.A2, args.B2, args.K2 = a2.data_ptr(), b2.data_ptr(), a2.shape[-1]
        args.a2_rs, args.b2_rs = a2.stride(0), b2.stride(0)
    ws = _stream_k_workspace(a.device)
    args.workspace, args.workspace_bytes = (ws.data_ptr(), ws.numel()) if ws is not None else (0, 0)
    fn = nv.proto("b200_gemm_w4", [ctypes.POINTER(nv.GemmArgs), ctypes.POINTER(W4Desc), ctypes.c_void_p])
    rc = fn(ctypes.byref(args), ctypes.byref(d), nv.stream())
    if rc == 10006:      # B200_ERR_UNSUPPORTED_: shape outside the fused kernel's domain
        return None
    nv.check(rc, f"gemm_w4 M={M} N={N} K={K}")
    nv.count_launch()
    del keep
    return out
