"""The two classic-ML ops projects of ``ML_Basics/`` (scikit-learn on the host CPU; no GPU work here).

* fault prediction (``ML_Basics/fault_prediction_project/src/*.py``): synthetic per-minute server metrics with ~5 % labelled
  faults (90 % high-load, 10 % low-load signatures; ``data_generation.py:6-91``), 60-minute rolling mean / std and a
  60-minute lag per device plus hour / weekday (``feature_engineering.py``), minority over-sampling (SMOTE — implemented
  here, imbalanced-learn is not in the image) + ``StandardScaler`` + ``GradientBoostingClassifier`` tuned by
  ``RandomizedSearchCV`` (``model_training.py:28-65``), and a service with ``POST /predict_fault`` and ``GET /health``
  (``model_service.py:17-51``; the reference uses Flask, this one FastAPI).
* root-cause analysis (``ML_Basics/server_failure_rca/{src,scripts}/*.py``): YAML-driven pipeline preprocess ->
  ``RandomForestClassifier`` on ``failure_cause`` -> ``IsolationForest`` anomaly flags -> feature-importance / anomaly
  tables (``scripts/run_pipeline.py:15-30``), and a service with ``/predict``, ``/batch_predict``, ``/health``
  (``scripts/api_server.py:69-129``).
"""
from __future__ import annotations

import os
from datetime import datetime, timedelta
from typing import Any, Dict, List, Optional, Tuple

import numpy as np
import pandas as pd

FAULT_FEATURES = ["cpu_usage", "ram_usage", "disk_io", "temperature", "error_count", "cpu_usage_mean", "cpu_usage_std",
                  "ram_usage_mean", "ram_usage_std", "hour", "day_of_week", "cpu_usage_lag1"]


# ------------------------------------------------------------------------------------------------ fault prediction
def generate_system_metrics(n_samples: int = 5000, start_time: str = "2025-07-12 06:00:00", fault_ratio: float = 0.05,
                            seed: int = 42, path: Optional[str] = None) -> pd.DataFrame:
    rng = np.random.default_rng(seed)
    t0 = datetime.strptime(start_time, "%Y-%m-%d %H:%M:%S")
    ts = [t0 + timedelta(minutes=i) for i in range(n_samples)]
    hours = np.array([t.hour for t in ts])
    dev = rng.choice(["server_001", "server_002"], size=n_samples)
    day = np.sin(2 * np.pi * hours / 24)
    cpu = np.clip(50 + 20 * day + rng.normal(0, 5, n_samples), 0, 100)
    ram = np.clip(60 + 15 * day + rng.normal(0, 5, n_samples), 0, 100)
    disk = np.clip(rng.normal(20, 5, n_samples), 0, 100)
    temp = 35 + 0.2 * cpu + rng.normal(0, 2, n_samples)
    err = np.clip(rng.poisson(2, n_samples), 0, 20).astype(float)
    label = np.zeros(n_samples, dtype=int)
    idx = rng.choice(n_samples, size=int(n_samples * fault_ratio), replace=False)
    label[idx] = 1
    hi = rng.random(len(idx)) < 0.9
    for i, high in zip(idx, hi):
        if high:       # overload signature
            cpu[i], ram[i], temp[i], err[i] = rng.uniform(90, 100), rng.uniform(85, 95), rng.uniform(45, 50), rng.integers(10, 20)
        else:          # hung / idle signature
            cpu[i], ram[i], disk[i], err[i] = rng.uniform(0, 10), rng.uniform(0, 15), rng.uniform(0, 5), rng.integers(5, 15)
    temp = np.clip(temp + 2 * (dev == "server_002"), 30, 50)
    df = pd.DataFrame({"timestamp": ts, "device_id": dev, "cpu_usage": cpu, "ram_usage": ram, "disk_io": disk,
                       "temperature": temp, "error_count": err, "label": label})
    if path:
        os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
        df.to_csv(path, index=False)
    return df


def extract_features(data: pd.DataFrame) -> Tuple[pd.DataFrame, List[str]]:
    d = data.copy()
    d["timestamp"] = pd.to_datetime(d["timestamp"])
    d["hour"], d["day_of_week"] = d["timestamp"].dt.hour, d["timestamp"].dt.dayofweek
    g = d.groupby("device_id")
    for col in ("cpu_usage", "ram_usage"):
        r = g[col].rolling(window=60, min_periods=1)
        d[f"{col}_mean"] = r.mean().reset_index(level=0, drop=True)
        d[f"{col}_std"] = r.std().reset_index(level=0, drop=True)
    d["cpu_usage_lag1"] = g["cpu_usage"].shift(60)
    return d, list(FAULT_FEATURES)


def smote(X: np.ndarray, y: np.ndarray, k: int = 5, seed: int = 42) -> Tuple[np.ndarray, np.ndarray]:
    """Synthetic minority over-sampling to a balanced set: new = x + u * (neighbour - x), neighbour among the k nearest
    minority samples."""
    from sklearn.neighbors import NearestNeighbors
    rng = np.random.default_rng(seed)
    cls, cnt = np.unique(y, return_counts=True)
    major = cnt.max()
    Xs, ys = [X], [y]
    for c, n in zip(cls, cnt):
        if n == major or n < 2:
            continue
        Xm = X[y == c]
        nn_ = NearestNeighbors(n_neighbors=min(k + 1, len(Xm))).fit(Xm)
        nbr = nn_.kneighbors(Xm, return_distance=False)[:, 1:]
        base = rng.integers(0, len(Xm), major - n)
        pick = nbr[base, rng.integers(0, nbr.shape[1], major - n)]
        u = rng.random((major - n, 1))
        Xs.append(Xm[base] + u * (Xm[pick] - Xm[base]))
        ys.append(np.full(major - n, c))
    return np.concatenate(Xs), np.concatenate(ys)


def train_fault_model(data: Optional[pd.DataFrame] = None, n_iter: int = 6, cv: int = 3, seed: int = 42, out_dir: Optional[str] = None) -> Dict[str, Any]:
    from sklearn.ensemble import GradientBoostingClassifier
    from sklearn.metrics import classification_report, roc_auc_score
    from sklearn.model_selection import RandomizedSearchCV, train_test_split
    from sklearn.preprocessing import StandardScaler
    data = generate_system_metrics() if data is None else data
    d, feats = extract_features(data)
    X, y = d[feats].fillna(0).to_numpy(), d["label"].to_numpy()
    Xtr, Xte, ytr, yte = train_test_split(X, y, test_size=0.2, stratify=y, random_state=seed)
    scaler = StandardScaler().fit(Xtr)
    Xb, yb = smote(scaler.transform(Xtr), ytr, seed=seed)
    search = RandomizedSearchCV(GradientBoostingClassifier(random_state=seed),
                                {"n_estimators": [50, 100, 200], "learning_rate": [0.01, 0.05, 0.1, 0.2], "max_depth": [2, 3, 5],
                                 "subsample": [0.8, 1.0]}, n_iter=n_iter, cv=cv, scoring="f1", random_state=seed, n_jobs=1)
    search.fit(Xb, yb)
    model = search.best_estimator_
    proba = model.predict_proba(scaler.transform(Xte))[:, 1]
    rep = classification_report(yte, proba > 0.5, output_dict=True, zero_division=0)
    out = {"model": model, "scaler": scaler, "features": feats, "best_params": search.best_params_, "auc": float(roc_auc_score(yte, proba)),
           "f1_fault": rep.get("True", rep.get("1", {})).get("f1-score", 0.0), "report": rep}
    if out_dir:
        import joblib
        os.makedirs(out_dir, exist_ok=True)
        joblib.dump(model, os.path.join(out_dir, "gbc_fault_prediction_model.pkl"))
        joblib.dump(scaler, os.path.join(out_dir, "scaler.pkl"))
    return out


def build_fault_app(model, scaler):
    from fastapi import FastAPI, HTTPException
    app = FastAPI(title="fault prediction")

    @app.get("/health")
    def health():
        return {"status": "healthy"}

    @app.post("/predict_fault")
    def predict_fault(row: Dict[str, Any]):
        try:
            df = pd.DataFrame([row], columns=["cpu_usage", "ram_usage", "disk_io", "temperature", "error_count", "timestamp", "device_id"])
            d, feats = extract_features(df)
            X = scaler.transform(d[feats].fillna(0).to_numpy())
            p = float(model.predict_proba(X)[0, 1])
            return {"is_fault_predicted": bool(p > 0.5), "fault_probability": p,
                    "feature_importance": dict(zip(feats, map(float, model.feature_importances_)))}
        except Exception as e:
            raise HTTPException(status_code=500, detail=str(e))
    return app


# ------------------------------------------------------------------------------------------------ root-cause analysis
RCA_DEFAULT_CONFIG = {"data": {"features": ["cpu_usage", "memory_usage", "disk_io", "network_latency"], "target": "failure_cause"},
                      "model": {"type": "random_forest", "params": {"n_estimators": 100, "random_state": 42}},
                      "anomaly_detection": {"contamination": 0.1}}
RCA_CAUSES = ["normal", "cpu_overload", "memory_leak", "disk_bottleneck", "network_issue"]


def generate_monitoring_data(n: int = 2000, seed: int = 0) -> pd.DataFrame:
    rng = np.random.default_rng(seed)
    cause = rng.choice(RCA_CAUSES, size=n, p=[0.6, 0.1, 0.1, 0.1, 0.1])
    X = np.stack([rng.normal(40, 10, n), rng.normal(50, 10, n), rng.normal(30, 8, n), rng.normal(20, 5, n)], 1)
    for j, c in enumerate(RCA_CAUSES[1:]):
        m = cause == c
        X[m, j] += rng.normal(45, 6, m.sum())
    df = pd.DataFrame(np.clip(X, 0, None), columns=RCA_DEFAULT_CONFIG["data"]["features"])
    df["failure_cause"] = cause
    return df


def load_rca_config(path: Optional[str]) -> Dict[str, Any]:
    if not path:
        return RCA_DEFAULT_CONFIG
    import yaml
    with open(path) as f:
        return yaml.safe_load(f)


def run_rca_pipeline(config_path: Optional[str] = None, data: Optional[pd.DataFrame] = None) -> Dict[str, Any]:
    """preprocess -> train classifier -> isolation forest -> importance / anomaly tables."""
    from sklearn.ensemble import IsolationForest, RandomForestClassifier
    from sklearn.model_selection import train_test_split
    from sklearn.preprocessing import StandardScaler
    cfg = load_rca_config(config_path)
    feats, target = cfg["data"]["features"], cfg["data"]["target"]
    if data is None:
        raw = cfg["data"].get("raw_data_path")
        data = pd.read_csv(raw) if raw and os.path.exists(raw) else generate_monitoring_data()
    data = data.dropna(subset=feats + [target])
    scaler = StandardScaler().fit(data[feats])
    X, y = scaler.transform(data[feats]), data[target].to_numpy()
    Xtr, Xte, ytr, yte = train_test_split(X, y, test_size=0.2, random_state=42, stratify=y)
    clf = RandomForestClassifier(**cfg["model"]["params"]).fit(Xtr, ytr)
    iso = IsolationForest(contamination=cfg["anomaly_detection"]["contamination"], random_state=42).fit(X)
    out = data.copy()
    out["anomaly"] = iso.predict(X) == -1
    imp = pd.DataFrame({"feature": feats, "importance": clf.feature_importances_}).sort_values("importance", ascending=False)
    return {"model": clf, "scaler": scaler, "iso": iso, "accuracy": float(clf.score(Xte, yte)), "importance": imp, "data": out,
            "features": feats}


def build_rca_app(model, scaler, features: List[str]):
    from fastapi import FastAPI, HTTPException
    app = FastAPI(title="server failure RCA")

    def _predict(rows: List[Dict[str, float]]):
        try:
            X = scaler.transform(pd.DataFrame(rows)[features])
        except Exception as e:
            raise HTTPException(status_code=400, detail=str(e))
        pred, proba = model.predict(X), model.predict_proba(X)
        return [{"failure_cause": str(p), "probabilities": dict(zip(map(str, model.classes_), map(float, pr)))} for p, pr in zip(pred, proba)]

    @app.post("/predict")
    def predict(metric: Dict[str, float]):
        return _predict([metric])[0]

    @app.post("/batch_predict")
    def batch_predict(batch: Dict[str, List[Dict[str, float]]]):
        return {"predictions": _predict(batch.get("metrics", []))}

    @app.get("/health")
    def health():
        return {"status": "healthy", "model_loaded": model is not None}
    return app
