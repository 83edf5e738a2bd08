"""In-tree build of the native libraries.

``libb200ops.so``  — every CUDA kernel, compiled by nvcc for sm_100a only
                     (``-gencode arch=compute_100a,code=sm_100a -lineinfo``), C ABI in ``include/b200_ops.h``.
``libb200host.so`` — the host-side C++ runtime (safetensors reader, BPE tokenizer, CPU AdamW, ...).

The reference has no build at all (it is pure Python); this replaces the
pip-installed native libraries it leans on. Nothing is JIT-compiled at run time: the .so files live
under ``llm_in_practise_b200/_lib`` so they travel with the source tree.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
LIBDIR = Path(__file__).resolve().parent / "_lib"
OBJDIR = ROOT / "build" / "obj"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "-Xcompiler", "-fvisibility=hidden",
    "--expt-relaxed-constexpr",
    "-I", str(ROOT / "include"), "-I", str(CSRC),
]
CXX_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-march=x86-64-v3", "-fopenmp", "-fvisibility=hidden",
             "-I", str(ROOT / "include"), "-I", str(CSRC)]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found")


def _stale(target: Path, deps) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(Path(d).stat().st_mtime > t for d in deps)


def _run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("build failed: %s\n%s\n%s" % (" ".join(map(str, cmd)), r.stdout, r.stderr))
    return r.stdout + r.stderr


def build_cuda(force: bool = False, verbose: bool = False) -> Path:
    LIBDIR.mkdir(parents=True, exist_ok=True)
    OBJDIR.mkdir(parents=True, exist_ok=True)
    srcs = sorted(CSRC.glob("*.cu"))
    hdrs = list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h")) + list((ROOT / "include").glob("*.h"))
    out = LIBDIR / "libb200ops.so"
    nvcc = _nvcc()
    jobs = []
    for s in srcs:
        o = OBJDIR / (s.stem + ".o")
        if force or _stale(o, [s] + hdrs):
            jobs.append([nvcc, *NVCC_FLAGS, *(["-Xptxas", "-v"] if verbose else []), "-c", str(s), "-o", str(o)])
    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for log in ex.map(_run, jobs):
                if verbose and log.strip():
                    print(log)
    objs = [OBJDIR / (s.stem + ".o") for s in srcs]
    if force or jobs or _stale(out, objs):
        _run([nvcc, "-shared", "-o", str(out), *map(str, objs), "-gencode", "arch=compute_100a,code=sm_100a"])
    return out


def build_host(force: bool = False) -> Path:
    LIBDIR.mkdir(parents=True, exist_ok=True)
    srcs = sorted((CSRC / "host").glob("*.cpp"))
    out = LIBDIR / "libb200host.so"
    if not srcs:
        return out
    hdrs = list((CSRC / "host").glob("*.h")) + list((ROOT / "include").glob("*.h"))
    if force or _stale(out, srcs + hdrs):
        _run(["g++", *CXX_FLAGS, "-shared", "-o", str(out), *map(str, srcs)])
    return out


def build_all(force: bool = False, verbose: bool = False):
    return build_cuda(force, verbose), build_host(force)


if __name__ == "__main__":
    f = "--force" in sys.argv
    v = "-v" in sys.argv
    print(*build_all(f, v), sep="\n")
