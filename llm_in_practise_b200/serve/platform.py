"""The decision logic of the reference's Kubernetes inference platform, runnable without Kubernetes.

``LLM_on_Kubernetes/Inference_Platfrom`` is YAML: every stage configures a controller that watches the serving engine's
``/metrics`` (vLLM's metric names, which ``serve/openai_server.py`` exports) and acts on it. The controllers themselves are
outside the GPU build; what a user of the reference relies on is their *rules*, and those are small enough to state as code
that works on the same metrics text:

* ``parse_prometheus_text`` / ``MetricHistory`` / ``histogram_quantile`` — the three PromQL forms the YAMLs use
  (``sum(gauge)``, ``sum(rate(counter[2m]))``, ``histogram_quantile(0.99, sum(rate(..._bucket[2m])) by (le))``).
* ``KedaScaler`` — 05-KEDA-AutoScale/keda-scaledobject.yaml:12-53 (and 09-Canary-Deployment/keda-rollout.yaml): two
  Prometheus triggers (queue depth > 2 per replica, P99 TTFT > 600 ms), min / max replica counts, the HPA behaviour block
  (stabilisation windows, one pod per period), cool-down, and the ``fallback`` replica count after three failed scrapes.
* ``CanaryAnalysis`` / ``CanaryRollout`` — 09-Canary-Deployment/analysis-template.yaml:8-53 and vllm-rollout.yaml:13-33:
  P99 TTFT < 2 s (hard failure at 5 s), error rate < 1 %, GPU memory < 90 %, five measurements each; steps
  ``setWeight 25 -> pause -> analysis -> setWeight 50 -> pause -> analysis -> setWeight 100``, abort back to the stable
  version when an analysis fails, weighted traffic split in between.
* ``preload_model`` — 02-Preloader/model-preloader-daemonset.yaml:44-78: mirror a model directory into the node-local
  cache once, guarded by a ``.ready`` marker, then check the files an engine cannot start without.
* ``PrefixIndex`` — what 08-LLM-Router's ``--routing-logic prefixaware`` (vllm-router-deployment.yaml:33-34) and llm-d's
  ``load_aware_prefix`` (llm-d-config.yaml:13-25) need: which replica has most of this prompt in its prefix cache. The two
  strategies themselves are routing strategies of ``serve/router.Router``.
"""
from __future__ import annotations

import hashlib
import math
import os
import re
import shutil
import time
from collections import OrderedDict, deque
from dataclasses import dataclass
from typing import Any, Callable, Deque, Dict, Iterable, List, Optional, Sequence, Tuple

Labels = Tuple[Tuple[str, str], ...]
Samples = Dict[str, List[Tuple[Labels, float]]]

_LINE = re.compile(r'^([A-Za-z_:][A-Za-z0-9_:]*)(\{(.*)\})?\s+([^\s]+)(\s+\d+)?$')
_LABEL = re.compile(r'([A-Za-z_][A-Za-z0-9_]*)="((?:[^"\\]|\\.)*)"')


def parse_prometheus_text(text: str) -> Samples:
    """Prometheus exposition text -> {metric name: [(sorted label pairs, value), ...]}."""
    out: Samples = {}
    for raw in text.splitlines():
        line = raw.strip()
        if not line or line.startswith("#"):
            continue
        m = _LINE.match(line)
        if not m:
            continue
        try:
            v = float(m.group(4).replace("+Inf", "inf").replace("-Inf", "-inf"))
        except ValueError:
            continue
        labels = tuple(sorted((k, val.replace('\\"', '"')) for k, val in _LABEL.findall(m.group(3) or "")))
        out.setdefault(m.group(1), []).append((labels, v))
    return out


def metric_sum(samples: Samples, name: str, **match: str) -> float:
    """``sum(name{match})``; 0 when the series does not exist."""
    want = set(match.items())
    return sum(v for labels, v in samples.get(name, []) if want <= set(labels))


def histogram_quantile(q: float, buckets: Sequence[Tuple[float, float]]) -> float:
    """PromQL ``histogram_quantile``: ``buckets`` are (upper bound, cumulative count) pairs including ``+Inf``; linear
    interpolation inside the bucket the rank falls into, the highest finite bound when it falls into ``+Inf``, NaN when
    the histogram is empty."""
    b = sorted(buckets)
    if not b or not math.isinf(b[-1][0]) or b[-1][1] <= 0:
        return float("nan")
    rank = q * b[-1][1]
    for i, (le, cum) in enumerate(b):
        if cum >= rank:
            if math.isinf(le):
                return b[-2][0] if len(b) > 1 else float("nan")
            lo, lo_cum = (0.0, 0.0) if i == 0 else b[i - 1]
            if cum == lo_cum:
                return le
            return lo + (le - lo) * (rank - lo_cum) / (cum - lo_cum)
    return float("nan")


class MetricHistory:
    """Scrapes of one or more ``/metrics`` endpoints over time, enough for ``rate(...[window])``."""

    def __init__(self, window_s: float = 120.0, clock: Callable[[], float] = time.monotonic):
        self.window_s, self._clock = window_s, clock
        self._scrapes: Deque[Tuple[float, Samples]] = deque()

    def add(self, texts: Iterable[str]) -> Samples:
        """One scrape of every replica (their series are summed by the queries below)."""
        merged: Samples = {}
        for t in texts:
            for name, rows in parse_prometheus_text(t).items():
                merged.setdefault(name, []).extend(rows)
        now = self._clock()
        self._scrapes.append((now, merged))
        while len(self._scrapes) > 2 and now - self._scrapes[1][0] >= self.window_s:
            self._scrapes.popleft()                  # keep one sample at or beyond the window's far edge
        return merged

    @property
    def latest(self) -> Samples:
        return self._scrapes[-1][1] if self._scrapes else {}

    def _ends(self):
        if len(self._scrapes) < 2:
            return None
        (t0, a), (t1, b) = self._scrapes[0], self._scrapes[-1]
        return (t0, a, t1, b) if t1 > t0 else None

    def rate(self, name: str, **match: str) -> float:
        """``sum(rate(name{match}[window]))``: increase per second between the window's ends (a counter that went
        backwards — a restarted replica — counts from zero)."""
        e = self._ends()
        if e is None:
            return 0.0
        t0, a, t1, b = e
        inc = metric_sum(b, name, **match) - metric_sum(a, name, **match)
        if inc < 0:
            inc = metric_sum(b, name, **match)
        return inc / (t1 - t0)

    def quantile(self, q: float, bucket_metric: str) -> float:
        """``histogram_quantile(q, sum(rate(bucket_metric[window])) by (le))`` in the metric's own unit; NaN without
        observations inside the window."""
        e = self._ends()
        if e is None:
            return float("nan")
        _, a, _, b = e

        def by_le(s: Samples) -> Dict[float, float]:
            acc: Dict[float, float] = {}
            for labels, v in s.get(bucket_metric, []):
                le = dict(labels).get("le")
                if le is not None:
                    k = float(le.replace("+Inf", "inf"))
                    acc[k] = acc.get(k, 0.0) + v
            return acc

        old, new = by_le(a), by_le(b)
        inc = [(le, max(0.0, c - old.get(le, 0.0))) for le, c in new.items()]
        return histogram_quantile(q, inc)


# ---------------------------------------------------------------------------------------------------------------- KEDA
@dataclass
class Trigger:
    """One ``type: prometheus`` trigger: ``value(history)`` is the query, ``threshold`` the per-replica target."""
    name: str
    value: Callable[[MetricHistory], float]
    threshold: float


def queue_depth_trigger(threshold: float = 2.0) -> Trigger:
    """``sum(vllm:num_requests_waiting)`` against 2 (keda-scaledobject.yaml:38-45)."""
    return Trigger("vllm-num-requests-waiting", lambda h: metric_sum(h.latest, "vllm:num_requests_waiting"), threshold)


def p99_ttft_trigger(threshold_ms: float = 600.0) -> Trigger:
    """P99 time to first token over the last two minutes, in ms, against 600 (keda-scaledobject.yaml:46-53)."""
    def value(h: MetricHistory) -> float:
        v = h.quantile(0.99, "vllm:time_to_first_token_seconds_bucket")
        return 0.0 if math.isnan(v) else v * 1000.0
    return Trigger("vllm-p99-ttft-ms", value, threshold_ms)


@dataclass
class ScalePolicy:
    """One direction of ``horizontalPodAutoscalerConfig.behavior``: a stabilisation window and ``type: Pods`` policies."""
    stabilization_window_s: float = 0.0
    pods: int = 1
    period_s: float = 15.0


class KedaScaler:
    """``tick(current, scrape)`` -> the replica count the ScaledObject would set now.

    * every trigger proposes ``ceil(value / threshold)`` (the HPA ``AverageValue`` rule); the largest proposal wins and is
      clamped to ``[min_replicas, max_replicas]``;
    * scale-up takes the SMALLEST proposal seen inside its stabilisation window, scale-down the LARGEST inside its own
      (so a spike must last, and a lull must last longer, before the count moves);
    * a direction's policy caps how many pods may be added / removed inside one period;
    * ``scrape()`` raising counts as a failed poll: after ``failure_threshold`` consecutive failures the count goes to
      ``fallback_replicas`` until a poll succeeds again (keda-scaledobject.yaml:17-19)."""

    def __init__(self, triggers: Optional[Sequence[Trigger]] = None, min_replicas: int = 1, max_replicas: int = 2,
                 scale_up: Optional[ScalePolicy] = None, scale_down: Optional[ScalePolicy] = None,
                 failure_threshold: int = 3, fallback_replicas: Optional[int] = 1, polling_interval_s: float = 15.0,
                 cooldown_s: float = 300.0, window_s: float = 120.0, clock: Callable[[], float] = time.monotonic):
        self.triggers = list(triggers) if triggers is not None else [queue_depth_trigger(), p99_ttft_trigger()]
        self.min_replicas, self.max_replicas = min_replicas, max_replicas
        self.up = scale_up or ScalePolicy(30.0, 1, 10.0)
        self.down = scale_down or ScalePolicy(120.0, 1, 30.0)
        self.failure_threshold, self.fallback_replicas = failure_threshold, fallback_replicas
        self.polling_interval_s, self.cooldown_s = polling_interval_s, cooldown_s
        self._clock = clock
        self.history = MetricHistory(window_s, clock)
        self._proposals: Deque[Tuple[float, int]] = deque()
        self._changes: Deque[Tuple[float, int]] = deque()          # (time, signed pods) actually applied
        self.failures = 0
        self.in_fallback = False
        self.last_values: Dict[str, float] = {}

    @classmethod
    def from_scaled_object(cls, spec: Dict[str, Any], **kw) -> "KedaScaler":
        """Build from the ``spec`` block of a ScaledObject YAML (the reference's field names)."""
        s = spec.get("spec", spec)
        beh = s.get("advanced", {}).get("horizontalPodAutoscalerConfig", {}).get("behavior", {})

        def pol(d, default):
            if not d:
                return default
            p = (d.get("policies") or [{}])[0]
            return ScalePolicy(float(d.get("stabilizationWindowSeconds", default.stabilization_window_s)),
                               int(p.get("value", default.pods)), float(p.get("periodSeconds", default.period_s)))

        trig: List[Trigger] = []
        for t in s.get("triggers", []):
            md = t.get("metadata", {})
            q, th = md.get("query", ""), float(md.get("threshold", 1))
            if "time_to_first_token_seconds_bucket" in q:
                trig.append(p99_ttft_trigger(th))
            elif "num_requests_waiting" in q:
                trig.append(queue_depth_trigger(th))
            else:
                raise ValueError(f"trigger query not understood: {q.strip()!r}")
        fb = s.get("fallback") or {}
        return cls(trig or None, int(s.get("minReplicaCount", 1)), int(s.get("maxReplicaCount", 1)),
                   pol(beh.get("scaleUp"), ScalePolicy(0.0, 4, 15.0)), pol(beh.get("scaleDown"), ScalePolicy(300.0, 100, 15.0)),
                   int(fb.get("failureThreshold", 3)), (int(fb["replicas"]) if "replicas" in fb else None),
                   float(s.get("pollingInterval", 30)), float(s.get("cooldownPeriod", 300)), **kw)

    def _moved(self, now: float, period: float, sign: int) -> int:
        while self._changes and now - self._changes[0][0] > max(self.up.period_s, self.down.period_s):
            self._changes.popleft()
        return sum(abs(n) for t, n in self._changes if now - t < period and (n > 0) == (sign > 0))

    def tick(self, current: int, scrape: Callable[[], Iterable[str]]) -> int:
        now = self._clock()
        try:
            self.history.add(list(scrape()))
        except Exception:
            self.failures += 1
            if self.failures >= self.failure_threshold and self.fallback_replicas is not None:
                self.in_fallback = True
                return self.fallback_replicas
            return current
        self.failures, self.in_fallback = 0, False
        want = self.min_replicas
        for t in self.triggers:
            v = t.value(self.history)
            self.last_values[t.name] = v
            want = max(want, math.ceil(v / t.threshold) if v > 0 else 0)
        want = max(self.min_replicas, min(self.max_replicas, want))
        self._proposals.append((now, want))
        horizon = max(self.up.stabilization_window_s, self.down.stabilization_window_s)
        while self._proposals and now - self._proposals[0][0] > horizon:
            self._proposals.popleft()
        up_floor = min(w for t, w in self._proposals if now - t <= self.up.stabilization_window_s)
        down_ceil = max(w for t, w in self._proposals if now - t <= self.down.stabilization_window_s)
        target = current
        if up_floor > current:
            room = max(0, self.up.pods - self._moved(now, self.up.period_s, +1))
            target = min(up_floor, current + room)
        elif down_ceil < current:
            room = max(0, self.down.pods - self._moved(now, self.down.period_s, -1))
            target = max(down_ceil, current - room)
        target = max(self.min_replicas, min(self.max_replicas, target))
        if target != current:
            self._changes.append((now, target - current))
        return target


# -------------------------------------------------------------------------------------------------------------- canary
@dataclass
class AnalysisMetric:
    """One entry of an Argo ``AnalysisTemplate``: ``count`` measurements; a measurement is Successful when ``success(v)``,
    Failed when ``failure(v)`` (default: not successful), else Inconclusive. NaN (no data) is Inconclusive."""
    name: str
    value: Callable[[MetricHistory], float]
    success: Callable[[float], bool]
    failure: Optional[Callable[[float], bool]] = None
    count: int = 5
    interval_s: float = 60.0
    failure_limit: int = 0

    def judge(self, v: float) -> str:
        if math.isnan(v):
            return "Inconclusive"
        if self.success(v):
            return "Successful"
        if self.failure is None or self.failure(v):
            return "Failed"
        return "Inconclusive"


def _p99_ttft_ms(h: MetricHistory) -> float:
    return h.quantile(0.99, "vllm:time_to_first_token_seconds_bucket") * 1000.0


def _error_rate(h: MetricHistory) -> float:
    ok = h.rate("vllm:request_success_total")
    bad = h.rate("vllm:request_failure_total")
    if ok <= 0:
        return float("nan") if bad <= 0 else float("inf")
    return bad / ok


def _gpu_memory(h: MetricHistory) -> float:
    rows_u, rows_t = h.latest.get("nvidia_gpu_memory_used_bytes", []), h.latest.get("nvidia_gpu_memory_total_bytes", [])
    total = {labels: v for labels, v in rows_t}
    fr = [v / total[labels] for labels, v in rows_u if total.get(labels)]
    return sum(fr) / len(fr) if fr else float("nan")


def smoke_template() -> List[AnalysisMetric]:
    """``vllm-prometheus-smoke`` (analysis-template.yaml:8-53)."""
    return [AnalysisMetric("p99-ttft", _p99_ttft_ms, lambda v: v < 2000, lambda v: v >= 5000),
            AnalysisMetric("error-rate", _error_rate, lambda v: v < 0.01),
            AnalysisMetric("gpu-memory-saturation", _gpu_memory, lambda v: v < 0.9)]


def gpu_memory_metrics_text() -> str:
    """The two ``nvidia_gpu_memory_*_bytes`` gauges of the GPU exporter the canary gate reads, from this process's view of
    its GPUs (empty without CUDA)."""
    try:
        import torch
        if not torch.cuda.is_available():
            return ""
        out = []
        for i in range(torch.cuda.device_count()):
            free, total = torch.cuda.mem_get_info(i)
            out.append(f'nvidia_gpu_memory_used_bytes{{gpu="{i}"}} {total - free}')
            out.append(f'nvidia_gpu_memory_total_bytes{{gpu="{i}"}} {total}')
        return "\n".join(out) + "\n"
    except Exception:
        return ""


class CanaryAnalysis:
    """An ``AnalysisRun``: call ``measure(history)`` once per interval until ``phase`` is no longer ``Running``. A metric
    fails the run when more than ``failure_limit`` of its measurements Failed; the run is Successful when every metric
    finished its ``count`` measurements without that, Inconclusive when some measurement could not be judged."""

    def __init__(self, metrics: Optional[Sequence[AnalysisMetric]] = None):
        self.metrics = list(metrics) if metrics is not None else smoke_template()
        self.results: Dict[str, List[Tuple[float, str]]] = {m.name: [] for m in self.metrics}
        self.phase = "Running"

    def measure(self, history: MetricHistory) -> str:
        if self.phase != "Running":
            return self.phase
        for m in self.metrics:
            if len(self.results[m.name]) < m.count:
                v = m.value(history)
                self.results[m.name].append((v, m.judge(v)))
        for m in self.metrics:
            if sum(1 for _, r in self.results[m.name] if r == "Failed") > m.failure_limit:
                self.phase = "Failed"
                return self.phase
        if all(len(self.results[m.name]) >= m.count for m in self.metrics):
            inconclusive = any(r == "Inconclusive" for m in self.metrics for _, r in self.results[m.name])
            self.phase = "Inconclusive" if inconclusive else "Successful"
        return self.phase


class CanaryRollout:
    """The ``strategy.canary.steps`` of vllm-rollout.yaml:22-33 over two sets of replicas. ``advance()`` is called by the
    operator loop (once per analysis interval is enough): it applies ``setWeight`` steps at once, sits out ``pause`` steps
    on the clock, runs ``analysis`` steps against ``scrape()`` (the canary replicas' ``/metrics``) and either promotes
    (weight 100, phase ``Healthy``) or aborts (weight 0, phase ``Degraded``) — an Inconclusive analysis pauses the rollout
    for a human, as Argo does. ``pick(rng)`` splits traffic by the current weight, which is what the nginx canary ingress
    does with ``canary-weight``."""

    DEFAULT_STEPS = ({"setWeight": 25}, {"pause": {"duration": "3m"}}, {"analysis": {}}, {"setWeight": 50},
                     {"pause": {"duration": "5m"}}, {"analysis": {}}, {"setWeight": 100})

    def __init__(self, steps: Optional[Sequence[Dict[str, Any]]] = None, template: Optional[Callable[[], Sequence[AnalysisMetric]]] = None,
                 window_s: float = 120.0, clock: Callable[[], float] = time.monotonic):
        self.steps = list(steps if steps is not None else self.DEFAULT_STEPS)
        self.template = template or smoke_template
        self._clock, self._window = clock, window_s
        self.index, self.weight, self.phase = 0, 0, "Progressing"
        self._pause_until: Optional[float] = None
        self._run: Optional[CanaryAnalysis] = None
        self._hist: Optional[MetricHistory] = None
        self._next_measure = 0.0
        self.runs: List[CanaryAnalysis] = []

    @staticmethod
    def _seconds(d: Any) -> float:
        if isinstance(d, (int, float)):
            return float(d)
        m = re.fullmatch(r"(\d+(?:\.\d+)?)([smh]?)", str(d).strip())
        if not m:
            raise ValueError(f"duration not understood: {d!r}")
        return float(m.group(1)) * {"": 1, "s": 1, "m": 60, "h": 3600}[m.group(2)]

    def advance(self, scrape: Optional[Callable[[], Iterable[str]]] = None) -> str:
        now = self._clock()
        while self.phase == "Progressing" and self.index < len(self.steps):
            step = self.steps[self.index]
            if "setWeight" in step:
                self.weight = int(step["setWeight"])
                self.index += 1
            elif "pause" in step:
                dur = (step["pause"] or {}).get("duration")
                if dur is None:                                   # an indefinite pause waits for promote()
                    self.phase = "Paused"
                    break
                if self._pause_until is None:
                    self._pause_until = now + self._seconds(dur)
                if now < self._pause_until:
                    break
                self._pause_until = None
                self.index += 1
            elif "analysis" in step:
                if self._run is None:
                    self._run, self._hist = CanaryAnalysis(self.template()), MetricHistory(self._window, self._clock)
                    self.runs.append(self._run)
                    self._next_measure = now
                    if scrape is not None:
                        self._hist.add(list(scrape()))            # the far end of the first rate() window
                    break
                if scrape is None or now < self._next_measure:
                    break
                self._hist.add(list(scrape()))
                interval = min(m.interval_s for m in self._run.metrics)
                self._next_measure = now + interval
                verdict = self._run.measure(self._hist)
                if verdict == "Running":
                    break
                self._run = None
                if verdict == "Successful":
                    self.index += 1
                elif verdict == "Failed":
                    self.abort()
                else:
                    self.phase = "Paused"
            else:
                raise ValueError(f"rollout step not understood: {step}")
        if self.phase == "Progressing" and self.index >= len(self.steps):
            self.weight, self.phase = 100, "Healthy"
        return self.phase

    def promote(self) -> None:
        """Resume after an indefinite pause or an Inconclusive analysis (``kubectl argo rollouts promote``)."""
        if self.phase == "Paused":
            self.phase, self.index, self._run = "Progressing", self.index + 1, None

    def abort(self) -> None:
        self.weight, self.phase, self._run = 0, "Degraded", None

    def pick(self, stable: Sequence[Any], canary: Sequence[Any], rng=None) -> Any:
        """One replica for the next request: a canary one with probability ``weight`` %."""
        import random
        r = (rng or random).random() * 100.0
        pool = canary if (canary and (r < self.weight or not stable)) else stable
        return (rng or random).choice(list(pool))


# ----------------------------------------------------------------------------------------------------------- preloader
def preload_model(source: str, target: str, required: Sequence[str] = ("config.json", "tokenizer.json"),
                  ready_marker: str = ".ready") -> Dict[str, Any]:
    """Mirror ``source`` (the model store) into ``target`` (the node-local cache) unless ``target/.ready`` exists.

    The mirror has ``mc mirror --overwrite --remove`` semantics: files that differ in size or are older are replaced, files
    the source no longer has are removed. The marker is written after the copy and before the check of ``required`` files,
    as in the reference's init container — a cache with a marker but a missing key file fails loudly instead of serving."""
    marker = os.path.join(target, ready_marker)
    rep: Dict[str, Any] = {"target": target, "copied": 0, "removed": 0, "skipped": False}
    if os.path.isfile(marker):
        rep["skipped"] = True
    else:
        if not os.path.isdir(source):
            raise FileNotFoundError(f"model source {source!r} is not a directory")
        os.makedirs(target, exist_ok=True)
        seen = set()
        for root, _, files in os.walk(source):
            rel = os.path.relpath(root, source)
            os.makedirs(os.path.join(target, rel), exist_ok=True)
            for f in files:
                s, d = os.path.join(root, f), os.path.normpath(os.path.join(target, rel, f))
                seen.add(d)
                ss = os.stat(s)
                if not os.path.isfile(d) or os.path.getsize(d) != ss.st_size or os.path.getmtime(d) < ss.st_mtime:
                    shutil.copy2(s, d)
                    rep["copied"] += 1
        for root, _, files in os.walk(target):
            for f in files:
                d = os.path.normpath(os.path.join(root, f))
                if d not in seen:
                    os.remove(d)
                    rep["removed"] += 1
        with open(marker, "w"):
            pass
    missing = [f for f in required if not os.path.isfile(os.path.join(target, f))]
    if missing:
        raise FileNotFoundError(f"model cache {target!r} lacks {missing}")
    rep["bytes"] = sum(os.path.getsize(os.path.join(r, f)) for r, _, fs in os.walk(target) for f in fs)
    return rep


def wait_until_ready(target: str, timeout_s: float = 600.0, poll_s: float = 5.0, ready_marker: str = ".ready",
                     clock: Callable[[], float] = time.monotonic, sleep: Callable[[float], None] = time.sleep) -> bool:
    """The serving pod's ``model-ready-check`` init container (vllm-rollout.yaml:55-70): block until the preloader's marker
    exists."""
    end = clock() + timeout_s
    while not os.path.isfile(os.path.join(target, ready_marker)):
        if clock() >= end:
            return False
        sleep(poll_s)
    return True


# -------------------------------------------------------------------------------------------------------- prefix index
class PrefixIndex:
    """Which replica has seen which prompt prefixes. A prompt is cut into ``chunk``-character pieces; piece i is stored as
    the hash of everything up to and including it, so a lookup walks the prompt from the front and stops at the first piece
    a replica has not seen — the number of pieces matched is the length of the prefix that replica's prefix cache
    (``enable_prefix_caching``) can still hold. Per replica the index is an LRU of ``capacity`` pieces, so it forgets at
    about the rate the engine's cache does."""

    def __init__(self, chunk: int = 128, capacity: int = 65536):
        self.chunk, self.capacity = chunk, capacity
        self._seen: Dict[Any, "OrderedDict[bytes, None]"] = {}

    def _hashes(self, prompt: str) -> List[bytes]:
        out, h = [], hashlib.blake2b(digest_size=12)
        for i in range(0, len(prompt) - len(prompt) % self.chunk, self.chunk):
            h.update(prompt[i:i + self.chunk].encode("utf-8", "replace"))
            out.append(h.copy().digest())
        return out

    def match(self, replica: Any, prompt: str) -> int:
        """Characters of ``prompt`` that ``replica`` has already processed as a prefix."""
        seen = self._seen.get(replica)
        if not seen:
            return 0
        n = 0
        for d in self._hashes(prompt):
            if d not in seen:
                break
            seen.move_to_end(d)
            n += 1
        return n * self.chunk

    def insert(self, replica: Any, prompt: str) -> None:
        seen = self._seen.setdefault(replica, OrderedDict())
        for d in self._hashes(prompt):
            seen[d] = None
            seen.move_to_end(d)
        while len(seen) > self.capacity:
            seen.popitem(last=False)

    def forget(self, replica: Any) -> None:
        """A replica that restarted has an empty cache."""
        self._seen.pop(replica, None)


def prompt_text(body: Dict[str, Any]) -> str:
    """The text a serving engine will tokenise for this OpenAI request body, close enough for prefix matching."""
    if "messages" in body:
        return "\n".join(f"{m.get('role', '')}:{m.get('content', '')}" for m in body.get("messages") or [] if isinstance(m, dict))
    p = body.get("prompt", "")
    return p if isinstance(p, str) else "\n".join(str(x) for x in p)


# ------------------------------------------------------------------------------------------- KEDA acting on replicas
def keda_replica_set(factory: Callable[[], Any], scaler: KedaScaler, clock: Callable[[], float] = time.monotonic):
    """A replica set of in-process engines whose size follows ``scaler`` (the ScaledObject's rule) instead of the Ray Serve
    rule: every ``polling_interval_s`` the live engines' own ``/metrics`` text is scraped, ``KedaScaler.tick`` decides, new
    replicas come from ``factory()``, surplus ones leave the rotation, drain and shut down
    (``serve/deploy.AutoscalingReplicaSet`` does the starting / draining)."""
    from .deploy import AutoscalingReplicaSet
    from .openai_server import render_metrics

    class KedaReplicaSet(AutoscalingReplicaSet):
        def __init__(self):
            super().__init__(factory, {"min_replicas": scaler.min_replicas, "max_replicas": scaler.max_replicas}, clock)
            self.scaler, self._next_poll = scaler, clock()

        def tick(self) -> int:
            now = self._clock()
            for e in list(self.draining):
                if e.num_running + e.num_waiting == 0:
                    self.draining.remove(e)
                    if hasattr(e, "shutdown"):
                        e.shutdown()
                    self.events.append((now, "stopped", len(self.engines)))
            if now < self._next_poll:
                return len(self.engines)
            self._next_poll = now + self.scaler.polling_interval_s
            want = self.scaler.tick(len(self.engines), lambda: [render_metrics(e) for e in self.engines])
            while len(self.engines) < want:
                self.engines.append(self.factory())
                self.events.append((now, "up", len(self.engines)))
            while len(self.engines) > max(want, 1):
                loads = self.ongoing()
                self.draining.append(self.engines.pop(loads.index(min(loads))))
                self.events.append((now, "down", len(self.engines)))
            return len(self.engines)

    return KedaReplicaSet()


# ------------------------------------------------------------------------------------------ canary acting on replicas
class CanaryReplicaSet:
    """Two sets of in-process engines behind one ``submit``: requests go to the canary set with the rollout's current
    weight, ``tick()`` advances the rollout on the canary engines' own ``/metrics`` text (plus this process's GPU memory
    gauges), and the outcome is applied — ``Healthy``: the canary engines become the stable set and the old ones drain and
    shut down; ``Degraded``: the canary engines drain and shut down, all traffic is back on the stable set."""

    def __init__(self, stable: Sequence[Any], canary: Sequence[Any], rollout: Optional[CanaryRollout] = None, seed: Optional[int] = None):
        import random
        self.stable, self.canary = list(stable), list(canary)
        self.rollout = rollout or CanaryRollout()
        self.draining: List[Any] = []
        self._rng = random.Random(seed)
        self.sent = {"stable": 0, "canary": 0}

    def submit(self, prompt_ids, cfg=None):
        eng = self.rollout.pick(self.stable, self.canary, self._rng)
        self.sent["canary" if any(eng is c for c in self.canary) else "stable"] += 1
        return eng.submit(prompt_ids, cfg)

    def _scrape(self) -> List[str]:
        from .openai_server import render_metrics
        return [render_metrics(e) for e in self.canary] + [gpu_memory_metrics_text()]

    def tick(self) -> str:
        for e in list(self.draining):
            if e.num_running + e.num_waiting == 0:
                self.draining.remove(e)
                if hasattr(e, "shutdown"):
                    e.shutdown()
        before = self.rollout.phase
        phase = self.rollout.advance(self._scrape if self.canary else None)
        if phase != before and phase == "Healthy" and self.canary:
            self.draining += self.stable
            self.stable, self.canary = self.canary, []
        elif phase != before and phase == "Degraded":
            self.draining += self.canary
            self.canary = []
        return phase
