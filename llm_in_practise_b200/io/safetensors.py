"""safetensors reader / writer (format: 8-byte little-endian header length, JSON header, raw
little-endian tensor bytes) — byte-compatible with the ``safetensors`` wheel that HF
``save_pretrained`` / ``from_pretrained`` use in the reference, implemented here so
checkpoints can be memory-mapped shard by shard without that (Rust) dependency.

``load_file`` memory-maps the file and returns zero-copy CPU tensors; ``safe_open`` gives lazy
per-tensor access (used by the sharded HF loader); ``save_file`` writes tensors sorted by name with
64-byte aligned offsets, as the reference implementation does.
"""
from __future__ import annotations

import json
import mmap
import os
import struct
from typing import Dict, Iterable, Optional

import numpy as np
import torch

_DTYPES = {
    "F64": torch.float64, "F32": torch.float32, "F16": torch.float16, "BF16": torch.bfloat16,
    "I64": torch.int64, "I32": torch.int32, "I16": torch.int16, "I8": torch.int8, "U8": torch.uint8,
    "BOOL": torch.bool, "F8_E4M3": getattr(torch, "float8_e4m3fn", None), "F8_E5M2": getattr(torch, "float8_e5m2", None),
}
_NAMES = {v: k for k, v in _DTYPES.items() if v is not None}


def _nbytes(dtype: torch.dtype, shape) -> int:
    n = 1
    for s in shape:
        n *= int(s)
    return n * torch.empty((), dtype=dtype).element_size()


class safe_open:
    """Lazy reader: ``with safe_open(path) as f: f.keys(); f.get_tensor(name); f.metadata()``."""

    def __init__(self, path: str, framework: str = "pt", device: str = "cpu"):
        self.path = path
        self.device = device
        self._fh = open(path, "rb")
        (hlen,) = struct.unpack("<Q", self._fh.read(8))
        if hlen > 100 * 1024 * 1024:
            raise ValueError("safetensors header too large")
        self._header = json.loads(self._fh.read(hlen).decode("utf-8"))
        self._meta = self._header.pop("__metadata__", None)
        self._base = 8 + hlen
        self._mm = mmap.mmap(self._fh.fileno(), 0, access=mmap.ACCESS_READ)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def close(self):
        # tensors returned by get_tensor hold their own reference to the mapping
        self._fh.close()

    def keys(self) -> Iterable[str]:
        return list(self._header.keys())

    def metadata(self) -> Optional[dict]:
        return self._meta

    def info(self, name: str):
        e = self._header[name]
        return _DTYPES[e["dtype"]], tuple(e["shape"])

    def get_tensor(self, name: str) -> torch.Tensor:
        e = self._header[name]
        dtype = _DTYPES[e["dtype"]]
        if dtype is None:
            raise TypeError(f"dtype {e['dtype']} unsupported by this torch build")
        lo, hi = e["data_offsets"]
        shape = tuple(e["shape"])
        if hi == lo:
            return torch.empty(shape, dtype=dtype)
        buf = np.frombuffer(self._mm, dtype=np.uint8, count=hi - lo, offset=self._base + lo)
        t = torch.from_numpy(buf).view(dtype).view(shape)  # zero-copy view of the mapping
        if self.device != "cpu":
            t = t.to(self.device, non_blocking=True)
        return t


def load_file(path: str, device: str = "cpu") -> Dict[str, torch.Tensor]:
    with safe_open(path, device=device) as f:
        return {k: f.get_tensor(k) for k in f.keys()}


def save_file(tensors: Dict[str, torch.Tensor], path: str, metadata: Optional[Dict[str, str]] = None) -> None:
    header = {}
    if metadata:
        header["__metadata__"] = {str(k): str(v) for k, v in metadata.items()}
    off = 0
    order = sorted(tensors.keys())
    blobs = []
    for k in order:
        t = tensors[k].detach()
        if t.device.type != "cpu":
            t = t.cpu()
        t = t.contiguous()
        if t.dim() and t.stride(-1) != 1:          # size-1 dims keep odd strides through contiguous()
            t = torch.empty(t.shape, dtype=t.dtype).copy_(t)
        if t.dtype not in _NAMES:
            raise TypeError(f"cannot store dtype {t.dtype}")
        n = t.numel() * t.element_size()
        header[k] = {"dtype": _NAMES[t.dtype], "shape": list(t.shape), "data_offsets": [off, off + n]}
        blobs.append(t)
        off += n
    hjson = json.dumps(header, separators=(",", ":")).encode("utf-8")
    pad = (-(8 + len(hjson))) % 8
    hjson += b" " * pad
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        f.write(struct.pack("<Q", len(hjson)))
        f.write(hjson)
        for t in blobs:
            if t.numel():
                f.write(t.view(torch.uint8).numpy().tobytes() if t.dim() else t.reshape(1).view(torch.uint8).numpy().tobytes())
    os.replace(tmp, path)
