"""Drop-ins for cost_estimator.py and cost_calculator.py.

The per-request part of both - success counts, token sums, the cold/warm split, the mean
success latency - comes out of the fused device pass (kvm_log_scan); the rest is a few
scalar expressions evaluated here in the reference's order.

Reference interface mirrored (file:line):
  parse_k8s_quantity(q)                       cost_estimator.py:48-83
  load_pricing(path) -> UnitPricing           cost_estimator.py:128-156
  calculate_cold_warm_costs(...)              cost_estimator.py:289-340
  main(): --run-dir --namespace --service [--cost-file]      cost_estimator.py:399-501
  cost_calculator main(): RESULTS_FILE GPU_HOURLY [--requests-per-1k-tokens N]
                                              cost_calculator.py:11-75
"""

from __future__ import annotations

import argparse
import json
import os
import sys
from dataclasses import dataclass
from fractions import Fraction
from typing import Any, Dict, List, Optional

import numpy as np
import torch

from . import cluster, columns, engine
from .cluster import parse_k8s_quantity  # noqa: F401  (re-exported: lint-test.yml:143)


@dataclass
class UnitPricing:
    """Unit prices and knobs of cost.yaml (cost_estimator.py:128-139)."""
    gpu_default: float
    gpu_map: Dict[str, float]
    cpu_per_core_hr: float
    mem_per_gib_hr: float
    overhead_fraction: float
    use_requests: bool
    include_sidecars: bool


def load_pricing(path: str) -> UnitPricing:
    try:
        import yaml  # type: ignore
    except Exception:
        print("ERROR: Missing 'pyyaml'. Install with: pip install pyyaml", file=sys.stderr)
        sys.exit(2)
    with open(path) as f:
        data = yaml.safe_load(f)
    gpu = data.get("gpu", {}) or {}
    calc = data.get("calculation", {})
    return UnitPricing(
        gpu_default=float(gpu.get("default", 1.50)),
        gpu_map={k: float(v) for k, v in gpu.items() if k != "default"},
        cpu_per_core_hr=float(data.get("cpu", {}).get("hourly_per_core", 0.04)),
        mem_per_gib_hr=float(data.get("memory", {}).get("hourly_per_gib", 0.005)),
        overhead_fraction=float(data.get("overhead", {}).get("fraction", 0.10)),
        use_requests=bool(calc.get("use_requests", True)),
        include_sidecars=bool(calc.get("include_sidecars", False)),
    )


def _float_sum(total: int, count: int):
    """builtin sum() over `count` integer-valued floats: int 0 when empty (cost_estimator.py:307)."""
    return float(total) if count > 0 else 0


def calculate_cold_warm_costs(cold_count: int, warm_count: int, cold_tokens, warm_tokens,
                              total_cost: float, success: int) -> Dict[str, Any]:
    """cost_estimator.py:289-340 from the four device-side reductions."""
    cold_frac = cold_count / success if success > 0 else 0
    warm_frac = warm_count / success if success > 0 else 0
    cold_cost = total_cost * cold_frac
    warm_cost = total_cost * warm_frac
    return {
        "cold_cost_per_request": (cold_cost / cold_count) if cold_count > 0 else None,
        "warm_cost_per_request": (warm_cost / warm_count) if warm_count > 0 else None,
        "cold_cost_per_1k_tokens": (cold_cost / cold_tokens * 1000.0) if cold_tokens > 0 else None,
        "warm_cost_per_1k_tokens": (warm_cost / warm_tokens * 1000.0) if warm_tokens > 0 else None,
        "cold_total_cost": cold_cost,
        "warm_total_cost": warm_cost,
        "cold_requests": cold_count,
        "warm_requests": warm_count,
        "cold_tokens": cold_tokens,
        "warm_tokens": warm_tokens,
    }


def rollup_from_stats(stats, pricing: UnitPricing, rsecs: Dict[str, float], gpu_hr: float,
                      product_label: Optional[str]) -> Dict[str, Any]:
    """The keys cost_estimator.py merges into results.json (:441-482)."""
    success = int(stats.n_success)
    total_tokens = _float_sum(int(stats.sum_total_tokens), success)
    cpu_hr, mem_hr = pricing.cpu_per_core_hr, pricing.mem_per_gib_hr
    total_cost = ((rsecs["gpu_seconds"] * (gpu_hr / 3600.0))
                  + (rsecs["cpu_core_seconds"] * (cpu_hr / 3600.0))
                  + (rsecs["mem_gib_seconds"] * (mem_hr / 3600.0)))
    total_cost *= 1.0 + pricing.overhead_fraction
    cold_n = int(stats.n_cold_success)
    warm_n = success - cold_n
    cold_tok = _float_sum(int(stats.sum_total_tokens_cold), cold_n)
    warm_tok = _float_sum(int(stats.sum_total_tokens) - int(stats.sum_total_tokens_cold), warm_n)
    out: Dict[str, Any] = {
        "cost_per_request": (total_cost / success) if success > 0 else None,
        "cost_per_1k_tokens": (total_cost / total_tokens * 1000.0) if total_tokens > 0 else None,
    }
    out.update(calculate_cold_warm_costs(cold_n, warm_n, cold_tok, warm_tok, total_cost, success))
    out["cost_breakdown"] = {
        "total_cost": total_cost,
        "gpu_seconds": rsecs["gpu_seconds"],
        "cpu_core_seconds": rsecs["cpu_core_seconds"],
        "mem_gib_seconds": rsecs["mem_gib_seconds"],
        "gpu_hourly": gpu_hr,
        "cpu_hourly_per_core": cpu_hr,
        "mem_hourly_per_gib": mem_hr,
        "overhead_fraction": pricing.overhead_fraction,
        "gpu_product_label": product_label,
    }
    return out


def rollup(raw: engine.RawAnalysis, pricing: UnitPricing, rsecs: Dict[str, float], gpu_hr: float,
           product_label: Optional[str]) -> Dict[str, Any]:
    return rollup_from_stats(raw.stats, pricing, rsecs, gpu_hr, product_label)


def estimator_main(argv: Optional[List[str]] = None) -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--run-dir", required=True)
    ap.add_argument("--namespace", required=True)
    ap.add_argument("--service", required=True)
    ap.add_argument("--cost-file", default="cost.yaml")
    args = ap.parse_args(argv)

    run_dir = args.run_dir
    req_csv = os.path.join(run_dir, "requests_classified.csv")
    if not os.path.exists(req_csv):
        req_csv = os.path.join(run_dir, "requests.csv")
    if not os.path.exists(req_csv):
        print(f"ERROR: {req_csv} not found", file=sys.stderr)
        sys.exit(1)

    log = columns.load_requests_csv(req_csv)
    if log.n == 0:
        raise SystemExit("No rows in requests.csv")
    for k in ("start_ms", "latency_ms", "total_tokens"):
        if k in log.unparsed:
            raise TypeError(f"non-numeric {k} cell in row {int(log.unparsed[k][0])}")

    analyzer = engine.LogAnalyzer()
    dlog = engine.DeviceLog.from_host(log, analyzer.device)
    cold_csv = log.is_cold_csv if log.is_cold_csv is not None else np.zeros(log.n, dtype=bool)
    mask = torch.from_numpy(np.ascontiguousarray(cold_csv.astype(np.uint8))).to(analyzer.device)
    stats = analyzer.scan(dlog, (), cold_mask=mask)
    start_ts, end_ts = stats.min_start_ms / 1000.0, stats.max_end_ms / 1000.0

    pricing = load_pricing(args.cost_file)
    pods = cluster.kubectl_pods(args.namespace, args.service)   # unguarded, like the reference
    rsecs = cluster.resource_seconds(pods, start_ts, end_ts)
    label = None
    items = pods.get("items", [])
    if items:
        label = cluster.node_gpu_label(args.namespace, items[0]["metadata"]["name"])
    gpu_hr = cluster.pick_gpu_cost(pricing.gpu_default, pricing.gpu_map, label)
    fields = rollup_from_stats(stats, pricing, rsecs, gpu_hr, label)

    results_path = os.path.join(run_dir, "results.json")
    results: Dict[str, Any] = {}
    if os.path.exists(results_path):
        with open(results_path) as f:
            try:
                results = json.load(f)
            except Exception:
                results = {}
    results.update(fields)
    with open(results_path, "w") as f:
        json.dump(results, f, indent=2)

    cold_cpr, warm_cpr = fields.get("cold_cost_per_request"), fields.get("warm_cost_per_request")
    print(json.dumps({
        "cost_per_request": fields["cost_per_request"],
        "cost_per_1k_tokens": fields["cost_per_1k_tokens"],
        "cold_cost_per_request": cold_cpr,
        "warm_cost_per_request": warm_cpr,
        "cold_cost_per_1k_tokens": fields.get("cold_cost_per_1k_tokens"),
        "warm_cost_per_1k_tokens": fields.get("warm_cost_per_1k_tokens"),
        "cold_vs_warm_multiplier": (cold_cpr / warm_cpr) if cold_cpr and warm_cpr else None,
    }, indent=2))


# --------------------------------------------------------------------------- cost_calculator.py
def mean_success_latency(latency_ms: np.ndarray, ok: np.ndarray):
    """(count, mean) of latency over rows whose status token == "200" - masked double-double
    sum on the device (cost_calculator.py:42-50)."""
    n = int(latency_ms.shape[0])
    analyzer = engine.LogAnalyzer()
    dev = analyzer.device
    nan = torch.full((n,), float("nan"), dtype=torch.float64, device=dev)
    zero = torch.zeros(n, dtype=torch.int64, device=dev)
    cols = {"start_ms": torch.zeros(n, dtype=torch.float64, device=dev),
            "latency_ms": torch.from_numpy(np.ascontiguousarray(latency_ms, dtype=np.float64)).to(dev),
            "ttfb_ms": nan, "tllt_ms": nan,
            "status": torch.from_numpy(np.where(ok, 200, 0).astype(np.int64)).to(dev),
            "completion_tokens": zero, "total_tokens": zero}
    stats = analyzer.scan(engine.DeviceLog.from_tensors(cols), ())
    count = int(stats.n_success)
    if count == 0:
        return 0, None
    return count, float((Fraction(stats.lat_sum_hi) + Fraction(stats.lat_sum_lo)) / count)


def calculator_main(argv: Optional[List[str]] = None) -> None:
    parser = argparse.ArgumentParser(description="Calculate the cost of running a model on a GPU.")
    parser.add_argument("results_file", type=str, help="Path to the load test results file.")
    parser.add_argument("gpu_hourly_cost", type=float, help="GPU hourly cost in USD.")
    parser.add_argument("--requests-per-1k-tokens", type=int, default=10,
                        help="Number of requests to produce 1K tokens.")
    args = parser.parse_args(argv)
    try:
        with open(args.results_file) as f:
            lines = f.readlines()
    except FileNotFoundError:
        print(f"Error: Results file '{args.results_file}' not found!", file=sys.stderr)
        sys.exit(1)

    fields = [line.split() for line in lines]
    ok = np.fromiter((fl[1] == "200" for fl in fields), dtype=bool, count=len(fields))
    lat = np.full(len(fields), np.nan)
    for i, fl in enumerate(fields):
        if ok[i]:
            lat[i] = float(fl[0])
    if not ok.any():
        print("No successful requests found. Cannot calculate cost.")
        sys.exit(1)
    count, avg_ms = mean_success_latency(lat, ok)
    avg_s = avg_ms / 1000
    per_s = args.gpu_hourly_cost / 3600
    cost = per_s * avg_s * args.requests_per_1k_tokens

    print("=== COST CALCULATION ===")
    print(f"GPU hourly cost: ${args.gpu_hourly_cost:.2f}")
    print(f"Requests per 1K tokens: {args.requests_per_1k_tokens}")
    print(f"Results file: {args.results_file}")
    print()
    print("=== METRICS ===")
    print(f"Successful requests: {count}")
    print(f"Average latency: {avg_ms:.2f}ms ({avg_s:.4f}s)")
    print()
    print("=== COST ESTIMATION ===")
    print(f"GPU price per second: ${per_s:.6f}")
    print(f"Average latency (seconds): {avg_s:.4f}")
    print(f"Requests per 1K tokens: {args.requests_per_1k_tokens}")
    print()
    print(f"Cost per 1K tokens: ${cost:.6f}")
