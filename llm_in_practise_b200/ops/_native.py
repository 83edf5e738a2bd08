"""ctypes binding to ``libb200ops.so`` (C ABI declared in ``include/b200_ops.h``).

The library is built in-tree by ``llm_in_practise_b200._build`` and loaded lazily. On a machine
with a GPU a missing library is an error (the CUDA path must be the one that runs); on a CPU-only
machine the ops fall back to their plain-PyTorch reference implementations, which are also the
numerical oracles of the GPU tests.
"""
from __future__ import annotations

import ctypes
import os
from pathlib import Path

import torch

_LIBDIR = Path(__file__).resolve().parent.parent / "_lib"
_lib = None
_host_lib = None
launch_count = 0  # number of kernel launches issued through this binding (bench.py reports it)

DT_BF16, DT_FP16, DT_FP32 = 0, 1, 2
_DT = {torch.bfloat16: DT_BF16, torch.float16: DT_FP16, torch.float32: DT_FP32}

_ERRORS = {
    10001: "bad shape", 10002: "bad alignment (pointers and strides must be 16-byte multiples)",
    10003: "CUDA driver entry point unavailable", 10004: "cuTensorMapEncodeTiled failed",
    10005: "size beyond the supported maximum", 10006: "unsupported argument combination",
}


class NativeError(RuntimeError):
    pass


def dtype_code(dt: torch.dtype) -> int:
    try:
        return _DT[dt]
    except KeyError:
        raise TypeError(f"unsupported dtype {dt}") from None


def lib_path() -> Path:
    return _LIBDIR / "libb200ops.so"


def available() -> bool:
    """True when the CUDA kernels can run: library built and a CUDA device visible."""
    return torch.cuda.is_available() and lib_path().exists()


def use_native(*tensors) -> bool:
    """Dispatch rule shared by all ops: CUDA tensors go to the native kernels (and a missing
    library is fatal there); CPU tensors go to the PyTorch reference path."""
    if not tensors or not tensors[0].is_cuda:
        return False
    if os.environ.get("B200_FORCE_TORCH") == "1":
        return False
    lib()  # raises if missing
    return True


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        p = lib_path()
        if not p.exists():
            raise NativeError(
                f"{p} is missing: run `python -m llm_in_practise_b200._build` (or __graft_entry__.build()) first")
        _lib = ctypes.CDLL(str(p))
        _declare(_lib)
    return _lib


def host_lib() -> ctypes.CDLL | None:
    global _host_lib
    if _host_lib is None:
        p = _LIBDIR / "libb200host.so"
        if not p.exists():
            return None
        _host_lib = ctypes.CDLL(str(p))
    return _host_lib


def check(rc: int, what: str = "") -> None:
    if rc == 0:
        return
    if rc >= 10000:
        msg = _ERRORS.get(rc, "unknown error")
    else:
        msg = f"CUDA error {rc}"
        try:
            msg += " (" + torch.cuda.cudart().cudaGetErrorString(rc) + ")"  # type: ignore[attr-defined]
        except Exception:
            pass
    raise NativeError(f"b200 {what}: {msg}")


def stream() -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t: torch.Tensor | None) -> ctypes.c_void_p:
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def count_launch(n: int = 1) -> None:
    global launch_count
    launch_count += n


class GemmArgs(ctypes.Structure):
    _fields_ = [
        ("A", ctypes.c_void_p), ("B", ctypes.c_void_p), ("D", ctypes.c_void_p), ("bias", ctypes.c_void_p),
        ("A2", ctypes.c_void_p), ("B2", ctypes.c_void_p),
        ("M", ctypes.c_int32), ("N", ctypes.c_int32), ("K", ctypes.c_int32), ("K2", ctypes.c_int32),
        ("batch", ctypes.c_int32), ("nb0", ctypes.c_int32), ("b_div", ctypes.c_int32), ("bias_dtype", ctypes.c_int32),
        ("a_rs", ctypes.c_int64), ("a_bs0", ctypes.c_int64), ("a_bs1", ctypes.c_int64),
        ("b_rs", ctypes.c_int64), ("b_bs0", ctypes.c_int64), ("b_bs1", ctypes.c_int64),
        ("d_rs", ctypes.c_int64), ("d_bs0", ctypes.c_int64), ("d_bs1", ctypes.c_int64),
        ("a2_rs", ctypes.c_int64), ("b2_rs", ctypes.c_int64),
        ("a_mn", ctypes.c_int32), ("b_mn", ctypes.c_int32), ("in_dtype", ctypes.c_int32), ("out_dtype", ctypes.c_int32),
        ("alpha", ctypes.c_float), ("beta", ctypes.c_float),
        ("workspace", ctypes.c_void_p), ("workspace_bytes", ctypes.c_int64),
        ("beta_rows", ctypes.c_void_p),
        ("ar", ctypes.c_void_p),
        ("b_map", ctypes.c_void_p), ("b_count", ctypes.c_int32), ("bias_bs", ctypes.c_int64),
    ]


class GemmAllReduce(ctypes.Structure):
    """``b200_gemm_allreduce`` (include/b200_ops.h): all-reduce of D over tensor-parallel ranks inside the GEMM launch."""
    _fields_ = [
        ("d_mc", ctypes.c_void_p), ("d_ptrs", ctypes.c_void_p), ("tile_flag_ptrs", ctypes.c_void_p), ("sync_flag_ptrs", ctypes.c_void_p),
        ("ticket", ctypes.c_void_p),
        ("world", ctypes.c_int32), ("rank", ctypes.c_int32), ("channel", ctypes.c_int32), ("epoch", ctypes.c_uint32),
        ("max_tiles", ctypes.c_int64),
    ]


def _declare(L: ctypes.CDLL) -> None:
    vp, i32, i64, f32 = ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64, ctypes.c_float
    L.b200_gemm.argtypes = [ctypes.POINTER(GemmArgs), vp]
    L.b200_gemm.restype = ctypes.c_int
    # remaining prototypes are declared next to their Python wrappers via `proto`


_declared = set()


def proto(name: str, argtypes) -> ctypes._CFuncPtr:
    """Fetch an entry point and (once) attach its argument types."""
    fn = getattr(lib(), name)
    if name not in _declared:
        fn.argtypes = argtypes
        fn.restype = ctypes.c_int
        _declared.add(name)
    return fn
