"""Deployment descriptions in the shape of the reference's Ray Serve LLM apps, served by this framework's own engine.

Ray is not in the image, so ``LLMConfig`` / ``build_openai_app`` / ``load_serve_yaml`` accept the same dictionaries and
YAML as ``ray.serve.llm`` (Deployment/Ray/serve_run_examples/deepseek.py:5-41,
Deployment/Ray/serve_deploy_examples/*.yaml) and map them onto ``LLMEngine`` + the OpenAI-compatible FastAPI app:

* ``model_loading_config``: ``model_id`` (served name) and ``model_source`` (checkpoint directory, GPTQ / AWQ
  checkpoints are auto-detected — the "minimal AWQ" example);
* ``engine_kwargs``: ``max_num_seqs``, ``max_model_len``, ``dtype``; ``tensor_parallel_size`` x ``pipeline_parallel_size``
  ranks under torchrun form one replica (``serve/tp_engine.py``: Megatron sharding inside a stage, ``parallel/pp.py`` stages
  across; more GPUs otherwise = more replicas behind ``ReplicaRouter``), ``gpu_memory_utilization`` /
  ``enable_prefix_caching`` / ``enable_chunked_prefill`` / ``enforce_eager`` are accepted and recorded;
* ``deployment_config.autoscaling_config``: ``min_replicas`` / ``max_replicas`` / ``target_ongoing_requests`` drive
  ``ReplicaRouter.desired_replicas`` (the decision rule; starting processes is the launcher's job);
* several ``llm_configs`` / applications with different ``route_prefix`` give the "two models, two routes" layout.

``cluster_healthcheck`` is the single-node analogue of Deployment/Ray/scripts/ray_cluster_healthcheck.py:31-176: CPU task
fan-out, one ``nvidia-smi``-style probe per GPU, and a 16 MiB host<->device round trip per GPU.
"""
from __future__ import annotations

import math
import os
import time
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional

import torch


@dataclass
class LLMConfig:
    model_loading_config: Dict[str, Any]
    deployment_config: Dict[str, Any] = field(default_factory=dict)
    engine_kwargs: Dict[str, Any] = field(default_factory=dict)
    runtime_env: Dict[str, Any] = field(default_factory=dict)

    @property
    def model_id(self) -> str:
        return self.model_loading_config["model_id"]

    @property
    def model_source(self) -> str:
        return self.model_loading_config.get("model_source", self.model_id)

    def autoscaling(self) -> Dict[str, Any]:
        a = dict(self.deployment_config.get("autoscaling_config", {}))
        a.setdefault("min_replicas", 1)
        a.setdefault("max_replicas", a["min_replicas"])
        a.setdefault("target_ongoing_requests", 5)
        return a

    def parallel_size(self) -> int:
        """GPUs one replica spans: ``tensor_parallel_size x pipeline_parallel_size`` ranks under torchrun. With
        ``pipeline_parallel_size`` > 1 (serve_deploy_examples/qwen3_app_pipeline_parallel.yaml:21-31) the decoder layers are
        split into that many stages (``parallel/pp.py``), each stage tensor-parallel over ``tensor_parallel_size`` ranks."""
        ek = self.engine_kwargs
        return int(ek.get("tensor_parallel_size", 1)) * int(ek.get("pipeline_parallel_size", 1))

    def validate(self) -> None:
        import torch.distributed as dist
        want = self.parallel_size()
        have = dist.get_world_size() if (dist.is_available() and dist.is_initialized()) else 1
        if want != have:
            raise RuntimeError(f"engine_kwargs ask for {want} GPUs per replica (tensor_parallel_size x pipeline_parallel_size) but "
                               f"this process group has {have} rank(s): launch with `torchrun --nproc-per-node {want}`")


class ReplicaRouter:
    """Least-outstanding-requests choice among replicas + the Ray Serve autoscaling rule
    ``ceil(total_ongoing / target_ongoing_requests)`` clamped to [min_replicas, max_replicas]."""

    def __init__(self, engines: List[Any], autoscaling: Optional[Dict[str, Any]] = None):
        self.engines = list(engines)
        self.auto = autoscaling or {"min_replicas": 1, "max_replicas": len(self.engines), "target_ongoing_requests": 5}

    def ongoing(self) -> List[int]:
        return [e.num_running + e.num_waiting for e in self.engines]

    def pick(self):
        loads = self.ongoing()
        return self.engines[loads.index(min(loads))]

    def submit(self, prompt_ids, cfg=None):
        return self.pick().submit(prompt_ids, cfg)

    def desired_replicas(self) -> int:
        want = math.ceil(sum(self.ongoing()) / max(1, self.auto["target_ongoing_requests"]))
        return max(self.auto["min_replicas"], min(self.auto["max_replicas"], max(want, 1)))


class AutoscalingReplicaSet(ReplicaRouter):
    """A replica set that ACTS on the autoscaling block of the Ray Serve YAMLs
    (Deployment/Ray/serve_deploy_examples/qwen3_app_autoscaling.yaml: ``min_replicas`` / ``max_replicas`` /
    ``initial_replicas`` / ``target_ongoing_requests`` / ``upscale_delay_s`` / ``downscale_delay_s``): ``tick()`` compares the
    desired count (``ceil(ongoing / target)``) with the live one; a wish to grow that has lasted ``upscale_delay_s`` starts
    replicas through ``factory()`` (one engine per call, e.g. one per GPU), a wish to shrink that has lasted
    ``downscale_delay_s`` takes the least-loaded replica out of rotation, lets it drain and shuts it down. ``submit`` routes
    to the least-loaded live replica (and ticks)."""

    def __init__(self, factory, autoscaling: Optional[Dict[str, Any]] = None, clock=time.monotonic,
                 max_ongoing_requests: Optional[int] = None):
        a = dict(autoscaling or {})
        for short in ("upscale_delay", "downscale_delay"):          # the reference's YAML spells them without the _s suffix
            if short in a and short + "_s" not in a:
                a[short + "_s"] = a[short]
        self.max_ongoing = max_ongoing_requests                     # deployment_config.max_ongoing_requests: per-replica cap
        a.setdefault("min_replicas", 1)
        a.setdefault("max_replicas", max(1, a["min_replicas"]))
        a.setdefault("target_ongoing_requests", 5)
        a.setdefault("upscale_delay_s", 30.0)
        a.setdefault("downscale_delay_s", 600.0)
        self.factory, self._clock = factory, clock
        n0 = int(a.get("initial_replicas", a["min_replicas"]))
        n0 = max(a["min_replicas"], min(a["max_replicas"], n0))
        super().__init__([factory() for _ in range(n0)], a)
        self.draining: List[Any] = []
        self._up_since: Optional[float] = None
        self._down_since: Optional[float] = None
        self.events: List[tuple] = []          # (time, "up" | "down" | "stopped", live replicas afterwards)

    def tick(self) -> int:
        now = self._clock()
        for e in list(self.draining):                       # replicas taken out of rotation: stop them once they are idle
            if e.num_running + e.num_waiting == 0:
                self.draining.remove(e)
                if hasattr(e, "shutdown"):
                    e.shutdown()
                self.events.append((now, "stopped", len(self.engines)))
        want, have = self.desired_replicas(), len(self.engines)
        if want > have:
            self._down_since = None
            self._up_since = now if self._up_since is None else self._up_since
            if now - self._up_since >= float(self.auto["upscale_delay_s"]):
                while len(self.engines) < want:
                    self.engines.append(self.factory())
                self._up_since = None
                self.events.append((now, "up", len(self.engines)))
        elif want < have:
            self._up_since = None
            self._down_since = now if self._down_since is None else self._down_since
            if now - self._down_since >= float(self.auto["downscale_delay_s"]):
                loads = self.ongoing()
                victim = self.engines.pop(loads.index(min(loads)))        # one replica per decision, like Ray Serve's smoothing
                self.draining.append(victim)
                self._down_since = None
                self.events.append((now, "down", len(self.engines)))
        else:
            self._up_since = self._down_since = None
        return len(self.engines)

    def submit(self, prompt_ids, cfg=None):
        self.tick()
        if self.max_ongoing is not None and min(self.ongoing()) >= self.max_ongoing:
            raise RuntimeError(f"every replica holds max_ongoing_requests={self.max_ongoing} requests (back-pressure)")
        return super().submit(prompt_ids, cfg)


def build_engine(cfg: LLMConfig, device=None):
    from ..recipes.infer import _load
    from .engine import EngineArgs, LLMEngine
    cfg.validate()
    ek = cfg.engine_kwargs
    dt = {"auto": torch.bfloat16, "bfloat16": torch.bfloat16, "half": torch.float16, "float16": torch.float16}[str(ek.get("dtype", "auto"))]
    model, tok = _load(cfg.model_source, None, dt)
    args = EngineArgs(max_num_seqs=int(ek.get("max_num_seqs", 16)), max_model_len=int(ek.get("max_model_len", 4096)),
                      enable_prefix_caching=bool(ek.get("enable_prefix_caching", True)),
                      enable_chunked_prefill=bool(ek.get("enable_chunked_prefill", True)),
                      gpu_memory_utilization=float(ek.get("gpu_memory_utilization", 0.0)) if "max_num_seqs" not in ek else 0.0,
                      block_size=int(ek.get("block_size", 0)), kv_offload_gb=float(ek.get("kv_offloading_size", 0.0)))
    pp, tp = int(ek.get("pipeline_parallel_size", 1)), int(ek.get("tensor_parallel_size", 1))
    if pp > 1:
        from .tp_engine import build_pp_engine
        return build_pp_engine(model, tok, args, cfg.model_id, pp, tp)         # ranks > 0: call engine.serve_forever()
    if tp > 1:
        from .tp_engine import build_tp_engine
        return build_tp_engine(model, tok, args, model_name=cfg.model_id)     # ranks > 0: call engine.serve_forever()
    return LLMEngine(model, tok, args, model_name=cfg.model_id)


def build_openai_app(args: Dict[str, Any], engines: Optional[Dict[str, Any]] = None):
    """``{"llm_configs": [LLMConfig | dict, ...]}`` -> FastAPI app. With one config the API lives at ``/v1``; with several,
    each model is also mounted under ``/<model_id>/v1`` and ``/v1/chat/completions`` dispatches on the ``model`` field."""
    from fastapi import FastAPI, Request
    from fastapi.responses import JSONResponse
    from .openai_server import build_app
    cfgs = [c if isinstance(c, LLMConfig) else LLMConfig(**c) for c in args["llm_configs"]]
    engines = engines or {}
    apps = {}
    for c in cfgs:
        eng = engines.get(c.model_id) or build_engine(c)
        eng.start()
        apps[c.model_id] = build_app(eng)
    if len(apps) == 1:
        return next(iter(apps.values()))
    root = FastAPI(title="llm_in_practise_b200 multi-model")
    for mid, sub in apps.items():
        root.mount(f"/{mid}", sub)

    @root.get("/v1/models")
    def models():
        return {"object": "list", "data": [{"id": m, "object": "model"} for m in apps]}

    @root.middleware("http")
    async def by_model(request: Request, call_next):
        if request.url.path in ("/v1/chat/completions", "/v1/completions"):
            body = await request.body()
            import json
            try:
                mid = json.loads(body).get("model")
            except Exception:
                mid = None
            if mid not in apps:
                return JSONResponse(status_code=404, content={"error": f"unknown model {mid!r}; served: {list(apps)}"})
            request.scope["path"] = f"/{mid}{request.url.path}"
            request._body = body
        return await call_next(request)
    return root


def run(app, host: str = "0.0.0.0", port: int = 8000, route_prefix: str = "/", blocking: bool = True, **uvicorn_kw):
    """``serve.run(llm_app, route_prefix="/")`` (Deployment/Ray/serve_run_examples/deepseek.py:44-47): serve the app with
    uvicorn, under ``route_prefix`` when it is not the root."""
    import uvicorn
    from fastapi import FastAPI
    if route_prefix not in ("", "/"):
        root = FastAPI(title="llm_in_practise_b200 serve")
        root.mount(route_prefix.rstrip("/"), app)
        app = root
    if not blocking:
        import threading
        server = uvicorn.Server(uvicorn.Config(app, host=host, port=port, **uvicorn_kw))
        threading.Thread(target=server.run, name="b200-serve", daemon=True).start()
        return server
    return uvicorn.run(app, host=host, port=port, **uvicorn_kw)


def deploy_yaml(path: str, engines: Optional[Dict[str, Any]] = None):
    """``serve deploy <file>.yaml``: one FastAPI app for the whole file — every application of the YAML is built with
    ``build_openai_app`` and mounted at its ``route_prefix`` (the two-models-two-routes layout of
    Deployment/Ray/serve_deploy_examples/serve_app_two_models.yaml). Under torchrun (tensor- / pipeline-parallel replicas)
    the ranks other than 0 never return from here: they follow rank 0's engine until it shuts down."""
    from fastapi import FastAPI
    apps = load_serve_yaml(path)
    if not apps:
        raise ValueError(f"{path} lists no applications")
    built = []
    for a in apps:
        cfgs = a["llm_configs"]
        engs = dict(engines or {})
        for c in cfgs:
            if c.model_id not in engs:
                engs[c.model_id] = build_engine(c)
            if getattr(engs[c.model_id], "rank", 0) != 0:
                engs[c.model_id].serve_forever()               # follower rank of a parallel replica
                return None
        built.append((a["route_prefix"] or "/", build_openai_app({"llm_configs": cfgs}, engines=engs)))
    if len(built) == 1 and built[0][0] in ("", "/"):
        return built[0][1]
    root = FastAPI(title="llm_in_practise_b200 serve deploy")
    for prefix, app in sorted(built, key=lambda x: -len(x[0])):          # longest prefix first; "/" last
        if prefix in ("", "/"):
            root.mount("", app)
        else:
            root.mount(prefix.rstrip("/"), app)
    return root


def load_serve_yaml(path: str) -> List[Dict[str, Any]]:
    """Parse a ``serve deploy`` YAML: returns [{name, route_prefix, llm_configs: [LLMConfig]}]."""
    import yaml
    with open(path) as f:
        doc = yaml.safe_load(f)
    out = []
    for app in doc.get("applications", []):
        cfgs = [LLMConfig(**{k: v for k, v in c.items() if k in ("model_loading_config", "deployment_config", "engine_kwargs", "runtime_env")})
                for c in app.get("args", {}).get("llm_configs", [])]
        out.append({"name": app.get("name"), "route_prefix": app.get("route_prefix", "/"), "llm_configs": cfgs})
    return out


def cluster_healthcheck(cpu_tasks: int = 8, object_mib: int = 16) -> Dict[str, Any]:
    from concurrent.futures import ThreadPoolExecutor
    rep: Dict[str, Any] = {"host": os.uname().nodename, "cpus": os.cpu_count()}
    t0 = time.perf_counter()
    with ThreadPoolExecutor(cpu_tasks) as ex:
        rep["cpu_probe"] = sum(ex.map(lambda i: sum(j * j for j in range(20000)) > 0, range(cpu_tasks)))
    rep["cpu_probe_s"] = time.perf_counter() - t0
    gpus = []
    if torch.cuda.is_available():
        for i in range(torch.cuda.device_count()):
            p = torch.cuda.get_device_properties(i)
            free, total = torch.cuda.mem_get_info(i)
            buf = torch.empty(object_mib << 20, dtype=torch.uint8).pin_memory()
            torch.cuda.synchronize(i)
            t0 = time.perf_counter()
            back = buf.to(f"cuda:{i}", non_blocking=True).cpu()
            dt = time.perf_counter() - t0
            gpus.append({"index": i, "name": p.name, "sm": f"{p.major}.{p.minor}", "memory_free_gib": free / 2 ** 30, "memory_total_gib": total / 2 ** 30,
                         "roundtrip_ok": bool(back.numel() == buf.numel()), "roundtrip_gbs": 2 * buf.numel() / dt / 1e9})
    rep["gpus"] = gpus
    rep["healthy"] = rep["cpu_probe"] == cpu_tasks and all(g["roundtrip_ok"] for g in gpus)
    return rep


def cluster_healthcheck_distributed(object_mib: int = 16) -> Optional[Dict[str, Any]]:
    """The reference's Ray cluster check (Deployment/Ray/scripts/ray_cluster_healthcheck.py: driver reaches the cluster,
    worker processes start on every node, CPU tasks are scheduled, GPUs are visible and usable, the object store moves
    data between nodes) for a ``torchrun`` job: every rank is a worker process. Each rank reports host / pid / its GPU,
    runs a CPU probe and a small GPU kernel, then the data path is exercised the way the reference does it — an
    ``object_mib`` MiB object made on rank 0 is sent to every rank (broadcast), checksummed there, and the checksums
    travel back (all-gather). Rank 0 returns the aggregated report (other ranks return None). Launch under
    ``torchrun --nnodes N --nproc-per-node G -m llm_in_practise_b200.recipes.tools healthcheck``."""
    import socket
    import torch.distributed as dist
    from ..parallel.dist import get_info, setup_distributed
    if not (dist.is_available() and dist.is_initialized()):
        setup_distributed()
    info = get_info()
    rank, world, dev = info.rank, info.world_size, info.device
    me: Dict[str, Any] = {"rank": rank, "hostname": socket.gethostname(), "pid": os.getpid(), "device": str(dev)}
    t0 = time.perf_counter()
    me["cpu_probe_ok"] = sum(j * j for j in range(200000)) > 0
    me["cpu_probe_s"] = time.perf_counter() - t0
    if dev.type == "cuda":
        p = torch.cuda.get_device_properties(dev)
        free, total = torch.cuda.mem_get_info(dev)
        x = torch.ones(1 << 20, device=dev)
        me.update(gpu_name=p.name, gpu_sm=f"{p.major}.{p.minor}", gpu_memory_free_gib=free / 2 ** 30, gpu_memory_total_gib=total / 2 ** 30,
                  gpu_kernel_ok=bool(float((x * 2).sum()) == float(2 << 20)))
    # object path: rank 0 -> everyone, checksums back
    n = object_mib << 20
    obj = torch.arange(n, dtype=torch.int64, device=dev).to(torch.uint8) if rank == 0 else torch.empty(n, dtype=torch.uint8, device=dev)
    if dev.type == "cuda":
        torch.cuda.synchronize(dev)
    dist.barrier()
    t0 = time.perf_counter()
    dist.broadcast(obj, src=0)
    if dev.type == "cuda":
        torch.cuda.synchronize(dev)
    me["object_recv_s"] = time.perf_counter() - t0
    me["object_checksum"] = int(obj.to(torch.int64).sum())
    reports: list = [None] * world
    dist.all_gather_object(reports, me)
    if rank != 0:
        return None
    hosts: Dict[str, list] = {}
    for r in reports:
        hosts.setdefault(r["hostname"], []).append(r["rank"])
    want = reports[0]["object_checksum"]
    rep = {"world_size": world, "nodes": len(hosts), "ranks_per_host": hosts, "workers": reports,
           "distinct_pids": len({(r["hostname"], r["pid"]) for r in reports}) == world,
           "cpu_ok": all(r["cpu_probe_ok"] for r in reports),
           "gpu_ok": all(r.get("gpu_kernel_ok", True) for r in reports),
           "gpus_visible": sum(1 for r in reports if "gpu_name" in r),
           "object_path_ok": all(r["object_checksum"] == want for r in reports),
           "object_gbs_slowest": (n / max(r["object_recv_s"] for r in reports) / 1e9) if world > 1 else None}
    rep["healthy"] = rep["distinct_pids"] and rep["cpu_ok"] and rep["gpu_ok"] and rep["object_path_ok"]
    return rep
