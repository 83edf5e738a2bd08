"""Spawn helper for multi-process CPU tests (gloo, 127.0.0.1 rendez-vous)."""
import os
import socket
import sys
import traceback

import torch.multiprocessing as mp


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _entry(rank, world, port, fn, args, errq):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    try:
        fn(rank, world, *args)
    except Exception:
        errq.put((rank, traceback.format_exc()))
        raise
    finally:
        from llm_in_practise_b200.parallel.dist import cleanup_distributed
        cleanup_distributed()


def run_distributed(fn, world: int = 2, args=()):
    ctx = mp.get_context("spawn")
    errq = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_entry, args=(r, world, port, fn, args, errq)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
    errs = []
    while not errq.empty():
        errs.append(errq.get())
    for p in procs:
        if p.is_alive():
            p.terminate()
            errs.append((-1, "timeout"))
    assert not errs and all(p.exitcode == 0 for p in procs), "\n".join(f"[rank {r}] {t}" for r, t in errs)
