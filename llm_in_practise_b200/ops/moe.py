"""K16: mixture-of-experts FFN as GROUPED GEMMs with the token permutation computed on the device.

The reference routes with a python loop over experts (``index_select`` / ``index_add_`` per expert, one cuBLAS GEMM pair per
expert and a host read of every expert's token count — LLM_Distributed_Trainning/PyTorch/transformer_basics/
DeepSeekLike_spare_MoE_wikitext2.py:300-329). Here:

1. ``dispatch``: (token, slot) assignments are counting-sorted by expert with device ops only (bincount, cumsum, stable
   argsort); every expert's segment is padded to a multiple of the 128-row GEMM tile, so a row tile belongs to exactly one
   expert. ``tile_map[t]`` = expert of row tile t or -1 (unused tail of the statically sized buffer). No ``.item()`` /
   ``.tolist()``: the host never learns the counts on the forward path.
2. ``grouped_linear``: ONE tcgen05 GEMM launch for all experts — batch entry = row tile, the B matrix (and bias row) of a
   tile is picked on the device through ``b_map`` (``b200_gemm_args.b_map``); unused tiles are skipped by all three warp
   roles. Forward (x W_e^T + b_e) and dgrad (dy W_e) are grouped launches; wgrad runs one GEMM per expert over its
   segment, with the segment bounds taken from a count copy that was started asynchronously during the forward pass.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional

import torch

from . import _native as nv
from .gemm import gemm

TILE = 128


@dataclass
class Dispatch:
    dest: torch.Tensor          # [R] row of assignment r = (token r // k, slot r % k) in the padded, expert-sorted buffer
    tok: torch.Tensor           # [R] token of assignment r
    tile_map: torch.Tensor      # [NT] int32: expert of row tile t, -1 = unused
    counts: torch.Tensor        # [E] tokens per expert (device)
    rows: int                   # NT * TILE (static upper bound: R + E * (TILE - 1), rounded up)
    num_experts: int
    counts_host: Optional[torch.Tensor] = None     # pinned copy of counts, valid after `ready`
    ready: Optional[torch.cuda.Event] = None

    def host_counts(self):
        if self.counts_host is None:
            return self.counts.tolist()
        if self.ready is not None:
            self.ready.synchronize()
        return self.counts_host.tolist()


def dispatch(top_i: torch.Tensor, num_experts: int, tile: int = TILE) -> Dispatch:
    """top_i [T, k] expert indices -> expert-sorted, tile-padded row layout. Device ops only."""
    T, k = top_i.shape
    R, E = T * k, num_experts
    dev = top_i.device
    flat_e = top_i.reshape(-1)
    counts = torch.bincount(flat_e, minlength=E)
    padc = (counts + (tile - 1)) // tile * tile
    seg_end = torch.cumsum(padc, 0)
    seg_start = seg_end - padc
    start_unp = torch.cumsum(counts, 0) - counts
    order = torch.argsort(flat_e, stable=True)
    sorted_e = flat_e.index_select(0, order)
    rank_in = torch.arange(R, device=dev) - start_unp.index_select(0, sorted_e)
    dest = torch.empty(R, dtype=torch.long, device=dev)
    dest.index_copy_(0, order, seg_start.index_select(0, sorted_e) + rank_in)
    nt = (R + E * (tile - 1) + tile - 1) // tile
    tile_start = torch.arange(nt, device=dev) * tile
    e_of_tile = torch.searchsorted(seg_end, tile_start, right=True)
    tile_map = torch.where(tile_start < seg_end[-1], e_of_tile, torch.full_like(e_of_tile, -1)).to(torch.int32)
    d = Dispatch(dest, torch.arange(R, device=dev) // k, tile_map, counts, nt * tile, E)
    if dev.type == "cuda" and not torch.cuda.is_current_stream_capturing():
        d.counts_host = torch.empty(E, dtype=counts.dtype, pin_memory=True)
        d.counts_host.copy_(counts, non_blocking=True)
        d.ready = torch.cuda.Event()
        d.ready.record()
    return d


def _grouped_ref(xs, tile_map, w, b, tile):
    nt = tile_map.numel()
    sel = tile_map.clamp_min(0).long()
    y = torch.bmm(xs.view(nt, tile, -1), w.index_select(0, sel).transpose(1, 2))
    if b is not None:
        y = y + b.index_select(0, sel).unsqueeze(1)
    y = y * (tile_map >= 0).to(y.dtype).view(nt, 1, 1)
    return y.reshape(nt * tile, -1)


def _grouped_gemm(a: torch.Tensor, w: torch.Tensor, tile_map: torch.Tensor, bias: Optional[torch.Tensor], b_mn: bool,
                  tile: int) -> torch.Tensor:
    """a [NT*tile, Ka]; w [E, N, K]; forward (b_mn False): out[t] = a[t] w[e]^T (+ bias[e]); dgrad (b_mn True): a[t] w[e]."""
    nt = tile_map.numel()
    E, N, K = w.shape
    n_out = K if b_mn else N
    a3 = a.view(nt, tile, a.shape[-1])
    out = torch.zeros(nt, tile, n_out, dtype=a.dtype, device=a.device)
    args = nv.GemmArgs()
    args.A, args.B, args.D = a3.data_ptr(), w.data_ptr(), out.data_ptr()
    args.bias = 0 if bias is None else bias.data_ptr()
    args.bias_dtype = nv.dtype_code(bias.dtype) if bias is not None else 0
    args.M, args.N, args.K = tile, n_out, a3.shape[-1]
    args.batch, args.nb0, args.b_div = nt, nt, 1
    args.a_rs, args.a_bs0, args.a_bs1 = a3.stride(1), a3.stride(0), 0
    args.b_rs, args.b_bs0, args.b_bs1 = w.stride(1), w.stride(0), 0
    args.d_rs, args.d_bs0, args.d_bs1 = out.stride(1), out.stride(0), 0
    args.a_mn, args.b_mn = 0, int(b_mn)
    args.in_dtype, args.out_dtype = nv.dtype_code(a.dtype), nv.dtype_code(out.dtype)
    args.alpha, args.beta = 1.0, 0.0
    args.b_map, args.b_count = tile_map.data_ptr(), E
    args.bias_bs = 0 if bias is None else bias.stride(0)
    nv.check(nv.lib().b200_gemm(ctypes.byref(args), nv.stream()), f"grouped gemm tiles={nt} N={n_out} K={a3.shape[-1]} E={E}")
    nv.count_launch()
    return out.view(nt * tile, n_out)


class _GroupedLinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, xs, w, b, d: Dispatch, tile: int):
        ctx.save_for_backward(xs, w)
        ctx.d, ctx.tile, ctx.has_bias = d, tile, b is not None
        return _grouped_gemm(xs, w, d.tile_map, b, False, tile)

    @staticmethod
    def backward(ctx, dy):
        xs, w = ctx.saved_tensors
        d, tile = ctx.d, ctx.tile
        dy = dy.contiguous()
        dxs = _grouped_gemm(dy, w, d.tile_map, None, True, tile) if ctx.needs_input_grad[0] else None
        dw = db = None
        if ctx.needs_input_grad[1] or ctx.has_bias:
            dw = torch.zeros_like(w)
            db = torch.zeros(w.shape[0], w.shape[1], dtype=torch.float32, device=w.device) if ctx.has_bias else None
            start = 0
            for e, n in enumerate(d.host_counts()):       # the copy was started in the forward pass: no stall here
                n_pad = (n + tile - 1) // tile * tile
                if n:
                    seg_y, seg_x = dy[start:start + n_pad], xs[start:start + n_pad]     # padding rows are zero in dy
                    gemm(seg_y, seg_x, a_mn=True, b_mn=True, out=dw[e])
                    if db is not None:
                        db[e] = seg_y[:n].float().sum(0)
                start += n_pad
            if db is not None:
                db = db.to(w.dtype)
        return dxs, dw, db, None, None


def grouped_linear(xs: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor], d: Dispatch, tile: int = TILE) -> torch.Tensor:
    """xs [d.rows, K] (expert-sorted, tile-padded) -> [d.rows, N]: row tile t uses w[tile_map[t]] / b[tile_map[t]]."""
    if (nv.use_native(xs) and xs.dtype in (torch.bfloat16, torch.float16) and w.dtype == xs.dtype and tile == TILE
            and xs.shape[-1] % 8 == 0 and w.shape[1] % 8 == 0 and xs.is_contiguous() and w.is_contiguous()):
        return _GroupedLinearFn.apply(xs, w, b, d, tile)
    return _grouped_ref(xs, d.tile_map, w, b, tile)


def moe_ffn(xf: torch.Tensor, top_i: torch.Tensor, top_p: torch.Tensor, w1, b1, w2, b2, act, post=None, tile: int = TILE):
    """sum over the k selected experts of top_p * post(act(x W1_e^T + b1_e) W2_e^T + b2_e) -> ([T, D], Dispatch).
    xf [T, D]; w1 [E, H, D]; w2 [E, D, H]; ``post`` (e.g. the expert's output dropout) acts on each expert's own rows."""
    E = w1.shape[0]
    d = dispatch(top_i, E, tile)
    xs = torch.zeros(d.rows, xf.shape[1], dtype=xf.dtype, device=xf.device).index_copy(0, d.dest, xf.index_select(0, d.tok))
    h = act(grouped_linear(xs, w1, b1, d, tile))
    y = grouped_linear(h, w2, b2, d, tile)
    if post is not None:
        y = post(y)
    ys = y.index_select(0, d.dest) * top_p.reshape(-1, 1).to(y.dtype)
    return torch.zeros_like(xf).index_add(0, d.tok, ys), d
