"""Seeded synthetic run logs and power traces for the five BASELINE.json configs.

No request log ships with the reference, so every config is
synthesised with the load generator's own format rules:
  - requests.csv header / `%.3f` formatting / empty cells on failures:
    scripts/loadtest.py:572-608
  - steady arrival schedule (`target_rps = 2*concurrency`, arrivals i/target_rps):
    scripts/loadtest.py:182-185,244-248
  - power.json layout {"samples":[{"ts":..,"watts":..|null}],"interval_s":..}:
    energy/collector.py:249

Small configs are written as real files (cfg1); large ones are produced as columnar
numpy arrays (host) or torch tensors (device, for sizes that do not fit host RAM).
"""

from __future__ import annotations

import csv
import json
import math
import os
import random
from typing import Dict, List, Optional, Tuple

import numpy as np

CSV_HEADER = ["id", "scheduled_ms", "start_ms", "ttfb_ms", "tllt_ms", "latency_ms", "status",
              "prompt_tokens", "completion_tokens", "total_tokens", "error", "trace_id"]

F64_COLS = ("start_ms", "latency_ms", "ttfb_ms", "tllt_ms")
I64_COLS = ("status", "completion_tokens", "total_tokens")


# --------------------------------------------------------------------------- config 1
def cfg1_rows(n: int = 500, concurrency: int = 20, seed: int = 42,
              t_epoch_ms: float = 1.7e12) -> List[List[str]]:
    """Config 1: 500 requests, concurrency 20, steady pattern (README.md:67-68)."""
    rng = random.Random(seed)
    rps = concurrency * 2.0
    rows: List[List[str]] = []
    for i in range(n):
        arrival = i * (1.0 / rps)
        sched = t_epoch_ms + arrival * 1000.0
        start = sched + rng.uniform(0.0, 5.0)
        lat = rng.lognormvariate(5.5, 0.5)
        failed = rng.random() < 0.03
        ctok = rng.randint(20, 64)
        if failed:
            rows.append([str(i + 1), f"{sched:.3f}", f"{start:.3f}", "", "", f"{lat:.3f}", "500",
                         "12", "", "", "upstream error", f"{rng.getrandbits(128):032x}"])
        else:
            rows.append([str(i + 1), f"{sched:.3f}", f"{start:.3f}", f"{0.2 * lat:.3f}",
                         f"{0.98 * lat:.3f}", f"{lat:.3f}", "200", "12", str(ctok), str(12 + ctok),
                         "", f"{rng.getrandbits(128):032x}"])
    return rows


def write_requests_csv(path: str, rows: List[List[str]], header: Optional[List[str]] = None) -> None:
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with open(path, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(header or CSV_HEADER)
        w.writerows(rows)


def cfg1_power(rows: List[List[str]], seed: int = 43, hz: float = 2.0, pad_s: float = 10.0
               ) -> Dict[str, object]:
    """A power.json for config 1: samples from pad_s before the first request to pad_s
    after the last, a few missing (`null`) readings."""
    rng = random.Random(seed)
    starts = [float(r[2]) for r in rows]
    ends = [float(r[2]) + float(r[5]) for r in rows]
    t0, t1 = min(starts) / 1000.0 - pad_s, max(ends) / 1000.0 + pad_s
    samples = []
    k = 0
    while True:
        ts = t0 + k / hz + rng.uniform(0.0, 0.01)
        if ts > t1:
            break
        inside = (min(starts) / 1000.0) <= ts <= (max(ends) / 1000.0)
        w: Optional[float] = (420.0 if inside else 95.0) + rng.gauss(0.0, 12.0)
        if rng.random() < 0.02:
            w = None
        samples.append({"ts": ts, "watts": w})
        k += 1
    return {"samples": samples, "interval_s": 1.0 / hz}


def cfg1_pods(rows: List[List[str]], cold_offset_s: float = 2.0) -> Dict[str, object]:
    """Pod list a fake `kubectl get pods -o json` returns: one predictor pod whose
    container started cold_offset_s after the window start (-> cold requests)."""
    import datetime as dt
    t0 = min(float(r[2]) for r in rows) / 1000.0
    started = dt.datetime.fromtimestamp(math.floor(t0) + cold_offset_s, tz=dt.timezone.utc)
    return {"items": [{
        "metadata": {"name": "demo-llm-predictor-0", "namespace": "ml-prod"},
        "spec": {"nodeName": "gpu-node-0", "containers": [
            {"name": "kserve-container", "resources": {
                "requests": {"cpu": "4", "memory": "16Gi"},
                "limits": {"cpu": "8", "memory": "32Gi", "nvidia.com/gpu": "1"}}},
            {"name": "queue-proxy", "resources": {"requests": {"cpu": "100m", "memory": "128Mi"}}}]},
        "status": {"startTime": started.strftime("%Y-%m-%dT%H:%M:%SZ"), "containerStatuses": [
            {"name": "kserve-container",
             "state": {"running": {"startedAt": started.strftime("%Y-%m-%dT%H:%M:%SZ")}}},
            {"name": "queue-proxy",
             "state": {"running": {"startedAt": started.strftime("%Y-%m-%dT%H:%M:%SZ")}}}]},
    }]}


# --------------------------------------------------------------------------- config 2 / 5
def request_columns(n: int, seed: int = 1234, span_ms: float = 3.6e6, mu: float = 5.5,
                    sigma: float = 0.6, fail_p: float = 0.02, t_epoch_ms: float = 1.7e12
                    ) -> Dict[str, np.ndarray]:
    """Config 2 (n=1e7) / config 5 distribution: time-ordered columnar log."""
    rng = np.random.default_rng(seed)
    start = np.sort(t_epoch_ms + rng.uniform(0.0, span_ms, n))
    lat = np.round(rng.lognormal(mu, sigma, n), 3)
    ok = rng.random(n) >= fail_p
    ttfb = np.where(ok, np.round(0.15 * lat, 3), np.nan)
    tllt = np.where(ok, np.round(0.97 * lat, 3), np.nan)
    ctok = np.where(ok, rng.integers(16, 129, n), 0).astype(np.int64)
    ptok = rng.integers(8, 65, n).astype(np.int64)
    return {
        "start_ms": np.round(start, 3), "latency_ms": lat, "ttfb_ms": ttfb, "tllt_ms": tllt,
        "status": np.where(ok, 200, 500).astype(np.int64),
        "completion_tokens": ctok,
        "total_tokens": np.where(ok, ptok + ctok, 0).astype(np.int64),
    }


# --------------------------------------------------------------------------- config 3
def sweep_columns(runs: int, rows_per_run: int, seed0: int = 1000
                  ) -> Tuple[Dict[str, np.ndarray], np.ndarray, List[np.ndarray]]:
    """Config 3: `runs` independent logs concatenated, CSR offsets, per-run cold starts.
    Per-run mu~U(5,6.5), sigma~U(0.3,0.8), error rate~U(0,0.05), 0-2 cold starts in the
    first 10% of the window; rng seed = seed0 + run id."""
    parts: List[Dict[str, np.ndarray]] = []
    colds: List[np.ndarray] = []
    for r in range(runs):
        rng = np.random.default_rng(seed0 + r)
        mu, sigma, err = rng.uniform(5.0, 6.5), rng.uniform(0.3, 0.8), rng.uniform(0.0, 0.05)
        span = rows_per_run * 25.0  # 40 rps
        c = request_columns(rows_per_run, seed=seed0 * 7919 + r, span_ms=span, mu=mu, sigma=sigma,
                            fail_p=err, t_epoch_ms=1.7e12 + r * 1.0e7)
        nc = int(rng.integers(0, 3))
        t0 = float(c["start_ms"][0]) / 1000.0
        colds.append(np.sort(t0 + rng.uniform(0.0, 0.1 * span / 1000.0, nc)))
        parts.append(c)
    cols = {k: np.concatenate([p[k] for p in parts]) for k in parts[0]}
    offsets = np.arange(runs + 1, dtype=np.int64) * rows_per_run
    return cols, offsets, colds


# --------------------------------------------------------------------------- config 4
def power_trace(m: int, gpu: int = 0, hz: float = 1000.0, t0_s: float = 1.7e9,
                null_p: float = 1e-4) -> Tuple[np.ndarray, np.ndarray]:
    """Config 4: one DCGM-style trace: ts = T0 + k/hz + U(0,5e-5); watts = 400+300*util(t)
    + N(0,15) clipped to [60,1000]; null_p of the readings missing (nan).  seed = 77+gpu."""
    rng = np.random.default_rng(77 + gpu)
    k = np.arange(m, dtype=np.float64)
    ts = t0_s + k / hz + rng.uniform(0.0, 5e-5, m)
    util = 0.5 + 0.5 * np.sin(2.0 * np.pi * (k / hz) / 3600.0 + gpu)
    watts = np.clip(400.0 + 300.0 * util + rng.normal(0.0, 15.0, m), 60.0, 1000.0)
    watts[rng.random(m) < null_p] = np.nan
    return ts, watts


def power_json(ts: np.ndarray, watts: np.ndarray, interval_s: float) -> Dict[str, object]:
    return {"samples": [{"ts": float(t), "watts": (None if math.isnan(w) else float(w))}
                        for t, w in zip(ts.tolist(), watts.tolist())], "interval_s": interval_s}


def write_json(path: str, obj: object) -> None:
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    with open(path, "w") as f:
        json.dump(obj, f, indent=2)


# --------------------------------------------------------------------------- device-side (large n)
def request_columns_device(n: int, device, seed: int = 1234, span_ms: float = 3.6e6,
                           mu: float = 5.5, sigma: float = 0.6, fail_p: float = 0.02,
                           t_epoch_ms: float = 1.7e12, t_offset_ms: float = 0.0):
    """Same distribution as request_columns, generated on `device` with torch's RNG (a
    different stream than numpy's, so use only where no host copy is needed)."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    f64 = dict(dtype=torch.float64, device=device)
    start = torch.rand(n, generator=g, **f64).mul_(span_ms).add_(t_epoch_ms + t_offset_ms)
    start = torch.sort(start).values
    start = torch.round(start * 1000.0) / 1000.0
    lat = torch.randn(n, generator=g, **f64).mul_(sigma).add_(mu).exp_()
    lat = torch.round(lat * 1000.0) / 1000.0
    ok = torch.rand(n, generator=g, device=device) >= fail_p
    nan = torch.full((), float("nan"), **f64)
    ttfb = torch.where(ok, torch.round(0.15 * lat * 1000.0) / 1000.0, nan)
    tllt = torch.where(ok, torch.round(0.97 * lat * 1000.0) / 1000.0, nan)
    ctok = torch.randint(16, 129, (n,), generator=g, device=device, dtype=torch.int64)
    ptok = torch.randint(8, 65, (n,), generator=g, device=device, dtype=torch.int64)
    zero = torch.zeros((), dtype=torch.int64, device=device)
    return {
        "start_ms": start, "latency_ms": lat, "ttfb_ms": ttfb, "tllt_ms": tllt,
        "status": torch.where(ok, 200, 500).to(torch.int64),
        "completion_tokens": torch.where(ok, ctok, zero),
        "total_tokens": torch.where(ok, ptok + ctok, zero),
    }
