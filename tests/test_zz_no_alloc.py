"""b200_norm_bwd_noalloc — the norm backward for callers that are not allowed to allocate.

Some users of libb200ops.so come in through the C ABI with no allocator they may call at that
moment: a serving process that already gave all free HBM to its KV-cache pool, a host application
that bans cudaMalloc after start-up because it stalls the whole device, or a plain C/C++ harness
with no caching allocator at all. The regular ``b200_norm_bwd`` wants a scratch buffer from them;
this variant keeps its scratch in a fixed 8 MiB array inside the library instead.

The test drives the new entry point exactly as such a caller would — through ctypes, raw pointers
only — and checks dx / dw / db against the existing entry point (which gets its scratch from
torch), on two calls in a row with different inputs: the second call sees the array as the first
one left it, so stale partial rows would show up here. It also checks that an over-sized request
is refused rather than run.

The file name sorts last on purpose: a ``pytest -x`` run reaches it only after every other test.
It must also pass alone in a fresh process (the first call then performs the one-time address
look-up of the built-in array).
"""
import ctypes

import pytest
import torch

from llm_in_practise_b200.ops import _native as nv

pytestmark = pytest.mark.gpu

vp, i32 = ctypes.c_void_p, ctypes.c_int
_ARGS_NOALLOC = [i32, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp]
_ARGS_REGULAR = [i32, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp]


def _ptr(t):
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def _run_pair(layer_norm, rows, cols, dtype, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    dev = "cuda"
    h = torch.randn(rows, cols, device=dev, generator=g).to(dtype)
    dy = torch.randn(rows, cols, device=dev, generator=g).to(dtype)
    dres = torch.randn(rows, cols, device=dev, generator=g).to(dtype)
    w = (1 + 0.1 * torch.randn(cols, device=dev, generator=g)).to(dtype)
    hf = h.float()
    mean = hf.mean(-1) if layer_norm else torch.zeros(rows, device=dev)
    var = hf.var(-1, unbiased=False) if layer_norm else hf.pow(2).mean(-1)
    stats = torch.stack([mean, torch.rsqrt(var + 1e-6)], dim=1).contiguous()
    dt = nv.dtype_code(dtype)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    lib = nv.lib()

    # existing entry point: scratch supplied by the caller (torch allocates it here)
    dx_a, dw_a, db_a = torch.empty_like(h), torch.zeros_like(w), torch.zeros_like(w)
    parts = lib.b200_norm_bwd_parts(rows)
    scratch = torch.empty(2 * parts * cols, dtype=torch.float32, device=dev)
    f_reg = lib.b200_norm_bwd
    f_reg.argtypes, f_reg.restype = _ARGS_REGULAR, ctypes.c_int
    rc = f_reg(int(layer_norm), _ptr(dy), _ptr(h), _ptr(w), _ptr(stats), _ptr(dres), _ptr(dx_a), _ptr(dw_a),
               _ptr(db_a) if layer_norm else None, _ptr(scratch), 0, rows, cols, dt, dt, stream)
    assert rc == 0, rc

    # new entry point: inputs and outputs only
    dx_b, dw_b, db_b = torch.empty_like(h), torch.zeros_like(w), torch.zeros_like(w)
    f_na = lib.b200_norm_bwd_noalloc
    f_na.argtypes, f_na.restype = _ARGS_NOALLOC, ctypes.c_int
    rc = f_na(int(layer_norm), _ptr(dy), _ptr(h), _ptr(w), _ptr(stats), _ptr(dres), _ptr(dx_b), _ptr(dw_b),
              _ptr(db_b) if layer_norm else None, 0, rows, cols, dt, dt, stream)
    assert rc == 0, rc
    torch.cuda.synchronize()
    return (dx_a, dw_a, db_a), (dx_b, dw_b, db_b)


def _rel(a, b):
    return ((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-6)).item()


def test_noalloc_matches_regular_entry_point_twice_in_a_row():
    # two back-to-back calls with different shapes and data: the second starts from the first's leftovers
    for k, (layer_norm, rows, cols, dtype) in enumerate([(True, 700, 5120, torch.bfloat16),
                                                         (False, 333, 768, torch.float32)]):
        ref, got = _run_pair(layer_norm, rows, cols, dtype, seed=100 + k)
        tol = 2e-2 if dtype == torch.bfloat16 else 1e-4   # same tolerances as tests/test_ops_gpu.py
        assert torch.equal(got[0], ref[0]), "dx must not depend on where the partial rows live"
        assert _rel(got[1], ref[1]) < tol
        if layer_norm:
            assert _rel(got[2], ref[2]) < tol


def test_noalloc_refuses_sizes_beyond_its_scratch():
    lib = nv.lib()
    lib.b200_norm_bwd_noalloc_max_cols.restype = ctypes.c_int
    limit = lib.b200_norm_bwd_noalloc_max_cols()
    assert limit == 8192
    f_na = lib.b200_norm_bwd_noalloc
    f_na.argtypes, f_na.restype = _ARGS_NOALLOC, ctypes.c_int
    cols = limit + 8
    t = torch.zeros(2, cols, device="cuda", dtype=torch.bfloat16)
    w = torch.ones(cols, device="cuda", dtype=torch.bfloat16)
    stats = torch.ones(2, 2, device="cuda")
    rc = f_na(0, _ptr(t), _ptr(t), _ptr(w), _ptr(stats), None, _ptr(t.clone()), _ptr(w.clone()), None, 0, 2, cols,
              nv.DT_BF16, nv.DT_BF16, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 10005  # B200_ERR_TOO_LARGE: refused, nothing launched


def test_noalloc_fp32_limit_is_half_the_16bit_one():
    """fp32 activations are held 4 columns per vector, so the limit is 4096 columns; 4104 is refused (with a message on
    stderr), 4096 runs."""
    lib = nv.lib()
    f_na = lib.b200_norm_bwd_noalloc
    f_na.argtypes, f_na.restype = _ARGS_NOALLOC, ctypes.c_int
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    for cols, want in ((4104, 10005), (4096, 0)):
        t = torch.randn(4, cols, device="cuda", dtype=torch.float32)
        w = torch.ones(cols, device="cuda", dtype=torch.float32)
        stats = torch.ones(4, 2, device="cuda")
        rc = f_na(0, _ptr(t), _ptr(t), _ptr(w), _ptr(stats), None, _ptr(torch.empty_like(t)), _ptr(torch.zeros_like(w)), None,
                  0, 4, cols, nv.DT_FP32, nv.DT_FP32, stream)
        assert rc == want, (cols, rc)
    torch.cuda.synchronize()
