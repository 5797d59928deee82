"""Drop-in for the reference's analyze.py: same importable functions, same CLI, same
results.json / requests_classified.csv, with the per-request arithmetic on the GPU.

Reference interface mirrored here (file:line):
  read_requests_csv(path) -> List[dict]          analyze.py:34-56
  percentile(arr, p) -> float                    analyze.py:59-81
  compute_histograms(data, num_bins=20) -> dict  analyze.py:84-122
  window_bounds(rows) -> (start_s, end_s)        analyze.py:183-189
  main(): --run-dir --namespace --service [--prom-url]   analyze.py:463-618
"""

from __future__ import annotations

import argparse
import csv
import json
import math
import os
import sys
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _native as N
from . import cluster, columns, engine, report


def read_requests_csv(path: str) -> List[dict]:
    """Load requests.csv, coercing the known numeric cells to float (analyze.py:34-56)."""
    return columns.read_rows(path)


# --------------------------------------------------------------------------- list-level API
def _to_device_f64(arr: Sequence[float]) -> torch.Tensor:
    device = engine.require_cuda()
    host = np.asarray(list(arr) if not isinstance(arr, np.ndarray) else arr, dtype=np.float64)
    return torch.from_numpy(np.ascontiguousarray(host)).to(device)


def percentile(arr: Sequence[float], p: float) -> float:
    """p-quantile (0<=p<=1) with linear interpolation, NaNs ignored (analyze.py:59-81).

    The two neighbouring order statistics come from the device's exact radix select;
    the interpolation is the reference's expression in Python floats."""
    if len(arr) == 0:
        return float("nan")
    stat = engine.ColumnOps().quantile_stat(_to_device_f64(arr), float(p))
    return stat.interpolate(float(p))


def compute_histograms(data: Sequence[float], num_bins: int = 20) -> Dict[str, Any]:
    """Histogram dict of analyze.py:84-122; counts come from the device."""
    if len(data) == 0:
        return {"bins": [], "counts": [], "bin_edges": []}
    ops = engine.ColumnOps()
    v = _to_device_f64(data)
    count, lo, hi = ops.minmax_count(v)
    counts: List[int] = []
    if count and lo != hi:
        counts = ops.histogram_counts(v, lo, hi, num_bins)
    return report.histogram_dict(count, (lo, hi) if count else None, counts, num_bins)


def window_bounds(rows: List[dict]) -> Tuple[float, float]:
    """analyze.py:183-189 on row dicts (host; used by callers that hold rows)."""
    start = min(r.get("start_ms", 0.0) for r in rows) / 1000.0
    end = max((r.get("start_ms", 0.0) + r.get("latency_ms", 0.0)) for r in rows) / 1000.0
    return start, end


# --------------------------------------------------------------------------- run-level API
def analyze_log(log: columns.RequestLog, cold_times: Sequence[float],
                analyzer: Optional[engine.LogAnalyzer] = None, timeseries: bool = True
                ) -> Tuple[Dict[str, Any], engine.RawAnalysis]:
    """Columnar log + cold-start times -> (numeric core of results.json, raw device output)."""
    columns.check_analyzable(log)
    analyzer = analyzer or engine.LogAnalyzer()
    dlog = engine.DeviceLog.from_host(log, analyzer.device)
    raw = analyzer.analyze(dlog, cold_times, series_window_s=1.0 if timeseries else None)
    return report.core_results(raw, len(cold_times)), raw


def exemplar_trace_id(log: columns.RequestLog, target_ms: float,
                      ops: Optional[engine.ColumnOps] = None) -> Optional[str]:
    """trace_id of the successful request whose latency is closest to `target_ms` (the p95
    exemplar behind the report's trace deep-link, report_generator.py:437-454: first row of
    `sorted((abs(latency - p95), trace_id))` over rows with status 200 and a latency).  One
    16 B/row pass on the GPU; the id itself is read from the CSV (native handle or kept rows).
    The reference compares the raw status cell with "200"; here the coerced status == 200."""
    if log.n == 0:
        return None
    ops = ops or engine.ColumnOps()
    lat = torch.from_numpy(log.cols["latency_ms"]).to(ops.device)
    status = torch.from_numpy(log.cols["status"]).to(ops.device)
    idx, _ = ops.nearest(lat, target_ms, member_mode=N.MEMBER_STATUS, member=status)
    if idx < 0:
        return None
    if log.native is not None and "trace_id" in log.native.header:
        return log.native.cell(idx, "trace_id")
    if log.raw_rows is not None:
        return log.raw_rows[idx].get("trace_id")
    return None


def _write_classified_csv(path: str, log: columns.RequestLog, is_cold: np.ndarray) -> None:
    """analyze.py:597-604: every row dict + is_cold_start, floats re-serialised by csv."""
    if log.native is not None:
        if log.native.write_classified(path, is_cold):
            return
        log.raw_rows = columns.read_rows(log.source_path)   # a cell only Python's float() can judge
    rows = log.raw_rows or []
    with open(path, "w", newline="") as f:
        if rows:
            for r, c in zip(rows, is_cold.tolist()):
                r["is_cold_start"] = bool(c)
            writer = csv.DictWriter(f, fieldnames=list(rows[0].keys()))
            writer.writeheader()
            writer.writerows(rows)


def main(argv: Optional[List[str]] = None) -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--run-dir", required=True)
    ap.add_argument("--namespace", required=True)
    ap.add_argument("--service", required=True)
    ap.add_argument("--prom-url", default=None)
    args = ap.parse_args(argv)

    log = columns.load_requests_csv(os.path.join(args.run_dir, "requests.csv"), keep_rows=True)
    if log.n == 0:
        print("No request rows found", file=sys.stderr)
        sys.exit(1)
    columns.check_analyzable(log)

    analyzer = engine.LogAnalyzer()
    dlog = engine.DeviceLog.from_host(log, analyzer.device)
    # window first (K1 without cold starts), because kubectl is asked about that window
    pre = analyzer.scan(dlog, ())
    start, end = pre.min_start_ms / 1000.0, pre.max_end_ms / 1000.0
    cold_times = cluster.cold_start_times(args.namespace, args.service, start, end)
    raw = analyzer.analyze(dlog, cold_times)
    core = report.core_results(raw, len(cold_times))
    duration = core["window"]["seconds"]
    success = core["requests"]["success"]
    tokens = report.total_tokens(raw)

    util = {k: None for k in report.UTIL_KEYS}
    if args.prom_url:
        try:
            util = cluster.utilization(args.prom_url, args.namespace, args.service, start, end)
        except Exception as e:
            print(f"Prometheus query failed: {e}", file=sys.stderr)

    energy: Dict[str, Any] = {}
    if util.get("gpu_power_watts_avg") and duration:   # analyze.py:539-552
        energy_wh = util["gpu_power_watts_avg"] * (duration / 3600.0)
        energy = {
            "energy_wh": energy_wh,
            "energy_wh_per_request": (energy_wh / success) if success > 0 else None,
            "energy_wh_per_1k_tokens": (energy_wh / tokens * 1000.0) if tokens > 0 else None,
        }

    io_probe: Dict[str, Any] = {}
    io_path = os.path.join(args.run_dir, "io_probe.json")
    if os.path.exists(io_path):
        try:
            with open(io_path) as f:
                io_probe = json.load(f)
        except Exception:
            io_probe = {}

    try:
        chr_val = cluster.cache_hit_ratio(args.prom_url, args.namespace, args.service, start, end)
    except Exception:
        chr_val = None

    # splice host-side keys into the reference's order (analyze.py:573-595)
    results: Dict[str, Any] = {}
    for k, v in core.items():
        if k in report.UTIL_KEYS:
            results[k] = util.get(k)
            if k == report.UTIL_KEYS[-1]:
                results.update(energy)
        else:
            results[k] = v
    results["network_rtt_p50_ms"] = (io_probe.get("rtt", {}) or {}).get("p50_ms")
    results["network_rtt_p95_ms"] = (io_probe.get("rtt", {}) or {}).get("p95_ms")
    results["s3_avg_MBps"] = (io_probe.get("s3", {}) or {}).get("avg_MBps")
    results["cache_hit_ratio"] = chr_val
    series = report.trim_series(raw)
    if series is not None:
        results["timeseries"] = series   # additive key (extension A12)

    is_cold = (raw.flags.cpu().numpy() & N.F_COLD) != 0
    _write_classified_csv(os.path.join(args.run_dir, "requests_classified.csv"), log, is_cold)

    out_path = os.path.join(args.run_dir, "results.json")
    if os.path.exists(out_path):
        try:
            with open(out_path) as f:
                prev = json.load(f)
            prev.update(results)
            results = prev
        except Exception:
            pass
    with open(out_path, "w") as f:
        json.dump(results, f, indent=2)
    print(json.dumps(results, indent=2))
