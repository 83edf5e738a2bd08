"""Classic-ML projects and the Ray-Serve-shaped deployment layer, on CPU."""
import textwrap

import numpy as np
import torch
from fastapi.testclient import TestClient


def test_fault_prediction_pipeline_and_service():
    from llm_in_practise_b200.extras import classic_ml as M
    df = M.generate_system_metrics(1500)
    assert abs(df["label"].mean() - 0.05) < 0.01 and set(df["device_id"]) == {"server_001", "server_002"}
    d, feats = M.extract_features(df)
    assert feats == M.FAULT_FEATURES and d["cpu_usage_mean"].notna().all()
    Xb, yb = M.smote(np.random.rand(100, 3), np.array([0] * 90 + [1] * 10))
    assert (yb == 0).sum() == (yb == 1).sum() == 90
    out = M.train_fault_model(df, n_iter=2, cv=2)
    assert out["auc"] > 0.95
    c = TestClient(M.build_fault_app(out["model"], out["scaler"]))
    assert c.get("/health").json() == {"status": "healthy"}
    bad = {"cpu_usage": 97, "ram_usage": 92, "disk_io": 20, "temperature": 48, "error_count": 15, "timestamp": "2025-07-12 08:00:00", "device_id": "server_001"}
    r = c.post("/predict_fault", json=bad).json()
    assert r["is_fault_predicted"] and r["fault_probability"] > 0.5 and set(r["feature_importance"]) == set(feats)
    ok = dict(bad, cpu_usage=50, ram_usage=60, temperature=40, error_count=2)
    assert not c.post("/predict_fault", json=ok).json()["is_fault_predicted"]


def test_rca_pipeline_and_service(tmp_path):
    from llm_in_practise_b200.extras import classic_ml as M
    cfgp = tmp_path / "config.yaml"
    cfgp.write_text(textwrap.dedent("""
        data: {features: [cpu_usage, memory_usage, disk_io, network_latency], target: failure_cause}
        model: {type: random_forest, params: {n_estimators: 30, random_state: 42}}
        anomaly_detection: {contamination: 0.1}
    """))
    out = M.run_rca_pipeline(str(cfgp))
    assert out["accuracy"] > 0.9 and 0.05 < out["data"]["anomaly"].mean() < 0.15 and len(out["importance"]) == 4
    c = TestClient(M.build_rca_app(out["model"], out["scaler"], out["features"]))
    r = c.post("/predict", json={"cpu_usage": 90, "memory_usage": 50, "disk_io": 30, "network_latency": 20}).json()
    assert r["failure_cause"] == "cpu_overload"
    b = c.post("/batch_predict", json={"metrics": [{"cpu_usage": 40, "memory_usage": 95, "disk_io": 30, "network_latency": 20}] * 2}).json()
    assert [p["failure_cause"] for p in b["predictions"]] == ["memory_leak"] * 2
    assert c.get("/health").json()["status"] == "healthy"


def test_deploy_layer(tmp_path):
    from llm_in_practise_b200.data.tokenizers import ByteTokenizer
    from llm_in_practise_b200.models.qwen3 import QWEN3_PRESETS, Qwen3ForCausalLM
    from llm_in_practise_b200.serve import EngineArgs, LLMEngine
    from llm_in_practise_b200.serve.deploy import LLMConfig, ReplicaRouter, build_openai_app, cluster_healthcheck, load_serve_yaml
    y = tmp_path / "app.yaml"
    y.write_text(textwrap.dedent("""
        applications:
          - name: qwen3_app
            route_prefix: /
            import_path: ray.serve.llm:build_openai_app
            args:
              llm_configs:
                - model_loading_config: {model_id: qwen3, model_source: /models/q}
                  deployment_config: {autoscaling_config: {min_replicas: 1, max_replicas: 4, target_ongoing_requests: 5}}
                  engine_kwargs: {tensor_parallel_size: 1, pipeline_parallel_size: 2, max_model_len: 4096}
    """))
    apps = load_serve_yaml(str(y))
    cfg = apps[0]["llm_configs"][0]
    assert cfg.model_id == "qwen3" and cfg.autoscaling()["max_replicas"] == 4
    assert cfg.parallel_size() == 2                         # TP 1 x PP 2 -> two ranks, one pipeline stage each
    try:
        cfg.validate()
        raise AssertionError("a 2-GPU replica must not start in a single process")
    except RuntimeError as e:
        assert "torchrun --nproc-per-node 2" in str(e)
    torch.manual_seed(0)
    m = Qwen3ForCausalLM(QWEN3_PRESETS["qwen3-tiny"], dtype=torch.float32, device="cpu")
    e1, e2 = (LLMEngine(m, ByteTokenizer(), EngineArgs(max_num_seqs=2, max_model_len=64), model_name=n) for n in ("a", "b"))
    router = ReplicaRouter([e1, e2], {"min_replicas": 1, "max_replicas": 4, "target_ongoing_requests": 5})
    for _ in range(12):
        router.submit([1, 2, 3])
    assert router.ongoing() == [6, 6] and router.desired_replicas() == 3
    for e in (e1, e2):
        e.waiting.clear()
    app = build_openai_app({"llm_configs": [LLMConfig({"model_id": "a"}), LLMConfig({"model_id": "b"})]}, engines={"a": e1, "b": e2})
    c = TestClient(app)
    try:
        body = {"model": "b", "messages": [{"role": "user", "content": "hi"}], "max_tokens": 3, "temperature": 0}
        r = c.post("/v1/chat/completions", json=body)
        assert r.status_code == 200 and r.json()["model"] == "b"
        assert c.post("/a/v1/chat/completions", json=dict(body, model="a")).status_code == 200
        assert c.post("/v1/chat/completions", json=dict(body, model="zzz")).status_code == 404
        assert {d["id"] for d in c.get("/v1/models").json()["data"]} == {"a", "b"}
    finally:
        e1.shutdown()
        e2.shutdown()
    h = cluster_healthcheck(4, 1)
    assert h["healthy"] and h["cpu_probe"] == 4


def _hc_worker(rank, world, tmp):
    import json
    from llm_in_practise_b200.parallel.dist import setup_distributed
    from llm_in_practise_b200.serve.deploy import cluster_healthcheck_distributed
    setup_distributed()
    rep = cluster_healthcheck_distributed(object_mib=1)
    assert (rep is None) == (rank != 0)
    if rank == 0:
        json.dump(rep, open(f"{tmp}/hc.json", "w"), default=str)


def test_distributed_cluster_healthcheck(tmp_path):
    """The torchrun counterpart of the reference's Ray cluster health check: every rank is a worker process; rank 0 gets the
    aggregated report (nodes, workers, CPU / GPU probes, the rank-0 -> everyone object path verified by checksum)."""
    import json
    from dist_utils import run_distributed
    run_distributed(_hc_worker, 2, (str(tmp_path),))
    rep = json.load(open(tmp_path / "hc.json"))
    assert rep["healthy"] and rep["world_size"] == 2 and rep["nodes"] == 1 and rep["distinct_pids"]
    assert rep["object_path_ok"] and len(rep["workers"]) == 2 and rep["workers"][1]["rank"] == 1
    assert rep["object_gbs_slowest"] > 0


def test_autoscaling_replica_set_follows_the_ray_serve_rule():
    """deploy.AutoscalingReplicaSet acts on min / max / initial replicas, target_ongoing_requests and the up / down delays of
    the reference's Ray Serve autoscaling YAML: grows after upscale_delay_s of sustained load, drains and stops one
    replica per downscale decision, never leaves [min, max]."""
    from llm_in_practise_b200.serve.deploy import AutoscalingReplicaSet

    class FakeEngine:
        def __init__(self):
            self.num_running, self.num_waiting, self.stopped = 0, 0, False

        def submit(self, ids, cfg=None):
            self.num_running += 1
            return self

        def shutdown(self):
            self.stopped = True

    made = []
    t = [0.0]

    def factory():
        made.append(FakeEngine())
        return made[-1]

    rs = AutoscalingReplicaSet(factory, {"min_replicas": 1, "max_replicas": 3, "initial_replicas": 1, "target_ongoing_requests": 2,
                                         "upscale_delay_s": 10, "downscale_delay_s": 60}, clock=lambda: t[0])
    assert len(rs.engines) == 1
    for _ in range(7):
        rs.submit([1, 2, 3])
    assert rs.desired_replicas() == 3 and len(rs.engines) == 1          # the wish is younger than upscale_delay_s
    t[0] = 11
    assert rs.tick() == 3 and len(made) == 3
    for _ in range(4):                                                   # new work goes to the empty replicas
        rs.submit([1])
    assert sorted(rs.ongoing()) == [2, 2, 7]
    for e in made:                                                       # all work finishes
        e.num_running = 0
    t[0] = 20
    assert rs.tick() == 3                                                # wish to shrink, but downscale_delay_s has not passed
    t[0] = 85
    assert rs.tick() == 2 and len(rs.draining) == 1
    t[0] = 86
    rs.tick()
    assert not rs.draining and sum(e.stopped for e in made) == 1
    t[0] = 150
    assert rs.tick() == 1
    t[0] = 300
    assert rs.tick() == 1                                                # min_replicas holds
    assert [e[1] for e in rs.events].count("up") == 1 and [e[1] for e in rs.events].count("down") == 2
    # the reference's own YAML: keys without the _s suffix, per-replica max_ongoing_requests as back-pressure
    import os
    import yaml
    ref = "/root/reference/Deployment/Ray/serve_deploy_examples/qwen3_app_autoscaling.yaml"
    if os.path.exists(ref):
        dc = yaml.safe_load(open(ref))["applications"][0]["args"]["llm_configs"][0]["deployment_config"]
        rs2 = AutoscalingReplicaSet(factory, dc["autoscaling_config"], clock=lambda: t[0], max_ongoing_requests=2)
        assert rs2.auto["upscale_delay_s"] == 5.0 and rs2.auto["downscale_delay_s"] == 30.0 and rs2.auto["max_replicas"] == 2
        rs2.submit([1]); rs2.submit([1])
        try:
            rs2.submit([1])
            raise AssertionError("expected back-pressure")
        except RuntimeError as e:
            assert "back-pressure" in str(e)


def test_serve_deploy_yaml_mounts_every_application_at_its_route(tmp_path):
    """``serve deploy`` of the two-models-two-routes YAML: each application answers under its own ``route_prefix``."""
    import textwrap
    import torch
    from fastapi.testclient import TestClient
    from llm_in_practise_b200 import compat
    from llm_in_practise_b200.data.tokenizers import ByteTokenizer
    from llm_in_practise_b200.models.qwen3 import QWEN3_PRESETS, Qwen3ForCausalLM
    from llm_in_practise_b200.serve import EngineArgs, LLMEngine
    from llm_in_practise_b200.serve.deploy import deploy_yaml
    y = tmp_path / "two.yaml"
    y.write_text(textwrap.dedent("""
        applications:
          - name: qwen3_8b_app
            route_prefix: /qwen3-8b
            import_path: ray.serve.llm:build_openai_app
            args:
              llm_configs:
                - model_loading_config: {model_id: qwen3-8b, model_source: /models/a}
                  engine_kwargs: {tensor_parallel_size: 1, max_model_len: 64}
          - name: qwen3_4b_awq_app
            route_prefix: /qwen3-4b-awq
            import_path: ray.serve.llm:build_openai_app
            args:
              llm_configs:
                - model_loading_config: {model_id: qwen3-4b-awq, model_source: /models/b}
                  engine_kwargs: {quantization: awq, max_model_len: 64}
    """))
    torch.manual_seed(0)
    m = Qwen3ForCausalLM(QWEN3_PRESETS["qwen3-tiny"], dtype=torch.float32, device="cpu")
    engines = {n: LLMEngine(m, ByteTokenizer(), EngineArgs(max_num_seqs=2, max_model_len=64, warmup=False), model_name=n)
               for n in ("qwen3-8b", "qwen3-4b-awq")}
    app = deploy_yaml(str(y), engines=engines)
    try:
        c = TestClient(app)
        for name in engines:
            body = {"model": name, "messages": [{"role": "user", "content": "hi"}], "max_tokens": 3, "temperature": 0}
            r = c.post(f"/{name}/v1/chat/completions", json=body)
            assert r.status_code == 200 and r.json()["model"] == name
            assert c.get(f"/{name}/v1/models").json()["data"][0]["id"] == name
        assert c.post("/v1/chat/completions", json={"model": "qwen3-8b", "messages": []}).status_code == 404      # no app at the root
        # serve.run(app, blocking=False) returns a server object whose thread can be stopped
        assert callable(compat.serve.run) and compat.build_openai_app is not None
    finally:
        for e in engines.values():
            e.shutdown()
