"""Process-group bootstrap: the launcher environment contract of the reference scripts.

``setup_distributed()`` reads ``RANK / WORLD_SIZE / LOCAL_RANK / MASTER_ADDR / MASTER_PORT`` (torchrun,
``accelerate launch`` and ``deepspeed`` all export them; a ``--local_rank`` CLI argument is tolerated),
initialises ``torch.distributed`` with NCCL on GPUs and gloo on CPU only when the world is larger than
one, and degrades to single-process on rendez-vous failure when asked to
(LLM_Distributed_Trainning/PyTorch/ddp_basics/ddp_gpt_wikitext2.py:170-186,
…/temp/ddp_gpt_bpe_tokenizer_02.py:94-102). NCCL is the plumbing (rendez-vous, barriers, cold-path
collectives, and the baseline path); the hot collectives are the fused peer-memory kernels in
``csrc/comm.cu``. Unlike the reference's consumer-GPU setup this never exports
``NCCL_P2P_DISABLE`` / ``NCCL_IB_DISABLE`` — NVLink peer access is what the fused kernels use.
"""
from __future__ import annotations

import datetime
import logging
import os
from dataclasses import dataclass
from typing import Optional

import torch
import torch.distributed as dist

log = logging.getLogger("llm_in_practise_b200.dist")


@dataclass
class DistInfo:
    rank: int = 0
    world_size: int = 1
    local_rank: int = 0
    device: torch.device = torch.device("cpu")
    backend: Optional[str] = None

    @property
    def is_main(self) -> bool:
        return self.rank == 0

    @property
    def distributed(self) -> bool:
        return self.world_size > 1


_info: Optional[DistInfo] = None


def env_int(name: str, default: int) -> int:
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def setup_distributed(local_rank_arg: Optional[int] = None, timeout_s: Optional[float] = None,
                      fallback_single: bool = False, force_backend: Optional[str] = None) -> DistInfo:
    """Initialise (if WORLD_SIZE > 1) and return rank/world/device. Idempotent."""
    global _info
    if _info is not None:
        return _info
    world = env_int("WORLD_SIZE", 1)
    rank = env_int("RANK", 0)
    local_rank = env_int("LOCAL_RANK", local_rank_arg if (local_rank_arg is not None and local_rank_arg >= 0) else 0)
    cuda = torch.cuda.is_available()
    if cuda:
        torch.cuda.set_device(local_rank % torch.cuda.device_count())
        device = torch.device("cuda", local_rank % torch.cuda.device_count())
    else:
        device = torch.device("cpu")
    backend = None
    if world > 1:
        backend = force_backend or ("nccl" if cuda else "gloo")
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        kw = {}
        if timeout_s is not None:
            kw["timeout"] = datetime.timedelta(seconds=timeout_s)
        if backend == "nccl":
            kw["device_id"] = device
        try:
            if not dist.is_initialized():
                dist.init_process_group(backend=backend, init_method="env://", rank=rank, world_size=world, **kw)
        except Exception as e:  # rendez-vous failure
            if not fallback_single:
                raise
            log.warning("process-group init failed (%s); continuing single-process", e)
            world, rank, backend = 1, 0, None
    _info = DistInfo(rank, world, local_rank, device, backend)
    return _info


def get_info() -> DistInfo:
    return _info if _info is not None else setup_distributed()


def cleanup_distributed() -> None:
    """barrier + destroy_process_group, swallowing errors at exit (ZeRO-1 script's ``finally:`` block,
    LLM_Distributed_Trainning/DeepSpeed/DeepSpeed-GPTLike-ZeRO-1/DeepSpeed-GPTLike-ZeRO-1.py:332-363)."""
    global _info
    if dist.is_available() and dist.is_initialized():
        try:
            dist.barrier()
        except Exception:
            pass
        try:
            dist.destroy_process_group()
        except Exception:
            pass
    _info = None


def barrier() -> None:
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()


def is_main_process() -> bool:
    return (not (dist.is_available() and dist.is_initialized())) or dist.get_rank() == 0


def world_size() -> int:
    return dist.get_world_size() if (dist.is_available() and dist.is_initialized()) else 1


def get_rank() -> int:
    return dist.get_rank() if (dist.is_available() and dist.is_initialized()) else 0


def all_reduce_mean_scalar(value: float, count: float, device) -> float:
    """Distributed validation loss: reduce(SUM) of value and count to rank 0, then broadcast of the
    mean (…/temp/ddp_gpt_bpe_tokenizer_02.py:325-340)."""
    if world_size() == 1:
        return value / max(count, 1e-12)
    t = torch.tensor([value, count], dtype=torch.float64, device=device)
    dist.reduce(t, dst=0, op=dist.ReduceOp.SUM)
    mean = torch.zeros(1, dtype=torch.float64, device=device)
    if get_rank() == 0:
        mean[0] = t[0] / t[1].clamp_min(1e-12)
    dist.broadcast(mean, src=0)
    return float(mean.item())
