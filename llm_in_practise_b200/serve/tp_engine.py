"""Tensor-parallel serving: one process per GPU, the scheduler on rank 0.

``vllm serve … --tensor-parallel-size T`` (Quantization/README.md:175, Fine-Tuning/README.md:345-348) and the
``tensor_parallel_size`` engine kwarg of the Ray Serve recipes (Deployment/Ray/serve_deploy_examples/*.yaml) start one
worker per GPU behind a single API server. Here the same shape is SPMD under ``torchrun``: every rank holds its shard of
the model (``parallel/tp.shard_qwen3``) and its slice of the KV cache (local kv heads only); rank 0 runs the
``LLMEngine`` scheduler and the HTTP server, and every model call it makes (packed prefill, prefill piece, graph-replayed
decode step) is first announced to the other ranks, which run the same call on their shards — the collectives inside the
layers (fused GEMV + all-reduce tails, peer all-reduce of the prefill GEMMs) need all ranks in the same launch sequence.
Logits are identical on every rank (the lm_head is replicated), so only rank 0 samples.

    torchrun --nproc-per-node T -m llm_in_practise_b200.recipes.infer serve --model … --tensor-parallel-size T
"""
from __future__ import annotations

from typing import Any, List, Optional

import torch
import torch.distributed as dist

from ..parallel.pp import PipelineGenerator
from .engine import EngineArgs, LLMEngine

_CALLS = ("_prefill_packed", "_prefill", "_prefill_chunk", "decode_graphed", "copy_prefix", "set_block_row",
          "offload_row", "restore_row", "drop_host_row")


def _to_wire(x: Any) -> Any:
    if isinstance(x, torch.Tensor):
        return ("__tensor__", x.detach().cpu())
    if isinstance(x, (list, tuple)):
        return type(x)(_to_wire(v) for v in x)
    return x


def _from_wire(x: Any, device) -> Any:
    if isinstance(x, tuple) and len(x) == 2 and x[0] == "__tensor__":
        return x[1].to(device)
    if isinstance(x, (list, tuple)):
        return type(x)(_from_wire(v, device) for v in x)
    return x


class _AnnouncingGenerator:
    """Rank 0's view of the ``Generator``: announce, then run. The KV cache argument is replaced by a marker — every rank
    passes its own cache object."""

    def __init__(self, gen, cache_of):
        self._gen, self._cache_of = gen, cache_of

    def __getattr__(self, name):
        attr = getattr(self._gen, name)
        if name not in _CALLS:
            return attr

        def call(*args, **kwargs):
            wire_args = tuple("__cache__" if a is self._cache_of() else _to_wire(a) for a in args)
            wire_kw = {k: _to_wire(v) for k, v in kwargs.items()}
            dist.broadcast_object_list([(name, wire_args, wire_kw)], src=0)
            return attr(*args, **kwargs)
        return call


class TensorParallelEngine(LLMEngine):
    """``LLMEngine`` over a model sharded with ``shard_qwen3``. Rank 0 uses it like any engine; the other ranks call
    ``serve_forever()``, which returns when rank 0 shuts the engine down."""

    def __init__(self, model, tokenizer=None, args: Optional[EngineArgs] = None, model_name: str = "b200-model"):
        if not (dist.is_available() and dist.is_initialized()):
            raise RuntimeError("TensorParallelEngine needs an initialised process group (launch with torchrun)")
        super().__init__(model, tokenizer, args, model_name)
        self.rank = dist.get_rank()
        if self.rank == 0:
            self.gen = _AnnouncingGenerator(self.gen, lambda: self.cache)

    def shutdown(self) -> None:
        super().shutdown()
        if self.rank == 0:
            dist.broadcast_object_list([("__stop__", (), {})], src=0)

    def warmup(self) -> None:
        if self.rank == 0:
            super().warmup()           # every captured call is announced: the followers capture the same graphs

    @torch.no_grad()
    def serve_forever(self) -> None:
        """Follower loop (ranks > 0): run what rank 0 announces."""
        assert self.rank != 0
        if self.device.type == "cuda":
            torch.cuda.set_device(self.device)
        while True:
            box: List[Any] = [None]
            dist.broadcast_object_list(box, src=0)
            name, args, kw = box[0]
            if name == "__stop__":
                return
            args = tuple(self.cache if (isinstance(a, str) and a == "__cache__") else _from_wire(a, self.device) for a in args)
            kw = {k: _from_wire(v, self.device) for k, v in kw.items()}
            getattr(self.gen, name)(*args, **kw)


def build_tp_engine(model, tokenizer=None, args: Optional[EngineArgs] = None, model_name: str = "b200-model",
                    max_tokens: int = 8192) -> TensorParallelEngine:
    """Shard ``model`` over the ranks of the default process group and wrap it in a ``TensorParallelEngine``."""
    from ..parallel.tp import shard_qwen3
    shard_qwen3(model, None, max_tokens=max_tokens)
    return TensorParallelEngine(model, tokenizer, args, model_name)


class PipelineParallelEngine(TensorParallelEngine):
    """The same SPMD engine over a ``stages x tp`` grid (``parallel/pp.py``): every rank runs its pipeline stage of each
    announced call, the last stage's logits reach rank 0, which schedules and samples. This is what
    ``engine_kwargs.pipeline_parallel_size`` of Deployment/Ray/serve_deploy_examples/qwen3_app_pipeline_parallel.yaml:21-31
    builds."""
    generator_cls = PipelineGenerator


def build_pp_engine(model, tokenizer=None, args: Optional[EngineArgs] = None, model_name: str = "b200-model",
                    pipeline_parallel_size: Optional[int] = None, tensor_parallel_size: int = 1,
                    max_tokens: int = 8192) -> PipelineParallelEngine:
    """Split ``model`` into pipeline stages (each optionally tensor-parallel) over the default process group."""
    from ..parallel.pp import shard_qwen3_pipeline
    shard_qwen3_pipeline(model, pipeline_parallel_size, tensor_parallel_size, max_tokens=max_tokens)
    return PipelineParallelEngine(model, tokenizer, args, model_name)
