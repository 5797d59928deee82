"""Pure-Python restatement of the reference's run-analysis arithmetic (TEST ORACLE).

Inputs are columnar (one Python list per field) instead of the reference's list of
row dicts; every value is what the reference holds after its CSV coercion:

  start_ms, latency_ms, ttfb_ms, tllt_ms : float (nan where the cell is `nan`)
  status, completion_tokens, total_tokens : int  (0 where the reference's `or 0` applies)

Each function cites the reference lines it follows.  The evaluation ORDER of every
floating-point expression is the reference's, because bit-exactness of the
interpolated percentiles, histogram edges and throughput ratios depends on it.
"""

from __future__ import annotations

import math
import statistics
from typing import Any, Dict, List, Optional, Sequence, Tuple

SUCCESS = 200
NAN = float("nan")


# ----------------------------------------------------------------------------- A2
def percentile(values: Sequence[float], p: float) -> float:
    """Linear-interpolated p-quantile, NaN-ignoring.  reference: analyze.py:59-81.

    pos = clamp(p,0,1)*(n-1); lo/hi = floor/ceil; v[lo] + (pos-lo)*(v[hi]-v[lo]).
    """
    kept = sorted(v for v in values if not math.isnan(v))
    n = len(kept)
    if n == 0:
        return NAN
    if n == 1:
        return kept[0]
    pos = max(0.0, min(p, 1.0)) * (n - 1)
    lo = int(math.floor(pos))
    hi = int(math.ceil(pos))
    if lo == hi:
        return kept[lo]
    return kept[lo] + (pos - lo) * (kept[hi] - kept[lo])


def percentile_ranks(n: int, p: float) -> Tuple[int, int, float]:
    """(lo, hi, pos) exactly as analyze.py:75-77 computes them for n kept samples."""
    pos = max(0.0, min(p, 1.0)) * (n - 1)
    return int(math.floor(pos)), int(math.ceil(pos)), pos


def interpolate(v_lo: float, v_hi: float, lo: int, hi: int, pos: float) -> float:
    """analyze.py:78-81 given the two order statistics."""
    if lo == hi:
        return v_lo
    return v_lo + (pos - lo) * (v_hi - v_lo)


# ----------------------------------------------------------------------------- A3
def histogram(values: Sequence[float], num_bins: int = 20) -> Dict[str, Any]:
    """Uniform-bin histogram.  reference: analyze.py:84-122."""
    kept = [v for v in values if not math.isnan(v)]
    if not kept:
        return {"bins": [], "counts": [], "bin_edges": []}
    lo, hi = min(kept), max(kept)
    if lo == hi:
        return {"bins": [lo], "counts": [len(kept)], "bin_edges": [lo, hi]}
    width = (hi - lo) / num_bins
    edges = [lo + i * width for i in range(num_bins + 1)]
    counts = [0] * num_bins
    for v in kept:
        counts[min(num_bins - 1, int((v - lo) / width))] += 1
    centers = [(edges[i] + edges[i + 1]) / 2 for i in range(num_bins)]
    return {
        "bins": centers,
        "counts": counts,
        "bin_edges": edges,
        "min": lo,
        "max": hi,
        "total_samples": len(kept),
    }


# ----------------------------------------------------------------------------- A5 / A14
def window_bounds(start_ms: Sequence[float], latency_ms: Sequence[float]) -> Tuple[float, float]:
    """(min start, max start+latency) in seconds over ALL rows.
    reference: analyze.py:183-189, energy/collector.py:114-122 ((0,0) when empty)."""
    if len(start_ms) == 0:
        return (0.0, 0.0)
    t0 = min(start_ms) / 1000.0
    t1 = max(s + l for s, l in zip(start_ms, latency_ms)) / 1000.0
    return t0, t1


# ----------------------------------------------------------------------------- A7
def classify_cold(start_ms: Sequence[float], cold_times: Sequence[float],
                  cold_window_sec: float = 30.0) -> List[bool]:
    """Row is cold iff some cold start c has c <= start_ms/1000 <= c+window.
    reference: analyze.py:402-419."""
    out = []
    for s in start_ms:
        t = s / 1000.0
        out.append(any(c <= t <= c + cold_window_sec for c in cold_times))
    return out


# ----------------------------------------------------------------------------- A8
def _subset_metrics(prefix: str, idx: List[int], cols: Dict[str, list]) -> Dict[str, Any]:
    """reference: analyze.py:427-455."""
    if not idx:
        return {f"{prefix}_count": 0, f"{prefix}_p50_ms": None,
                f"{prefix}_p95_ms": None, f"{prefix}_p99_ms": None}
    ok = [i for i in idx if cols["status"][i] == SUCCESS]
    lats = [cols["latency_ms"][i] for i in ok]
    ttfs = [cols["ttfb_ms"][i] for i in ok if not math.isnan(cols["ttfb_ms"][i])]
    return {
        f"{prefix}_count": len(idx),
        f"{prefix}_success_count": len(ok),
        f"{prefix}_p50_ms": percentile(lats, 0.50) if lats else None,
        f"{prefix}_p95_ms": percentile(lats, 0.95) if lats else None,
        f"{prefix}_p99_ms": percentile(lats, 0.99) if lats else None,
        f"{prefix}_ttft_p50_ms": percentile(ttfs, 0.50) if ttfs else None,
        f"{prefix}_ttft_p95_ms": percentile(ttfs, 0.95) if ttfs else None,
        f"{prefix}_error_rate": (len(idx) - len(ok)) / len(idx),
    }


def cold_warm_metrics(cols: Dict[str, list], is_cold: Sequence[bool]) -> Dict[str, Any]:
    """reference: analyze.py:422-460."""
    cold = [i for i, c in enumerate(is_cold) if c]
    warm = [i for i, c in enumerate(is_cold) if not c]
    out = _subset_metrics("cold", cold, cols)
    out.update(_subset_metrics("warm", warm, cols))
    return out


# ----------------------------------------------------------------------------- A4
def token_timing(cols: Dict[str, list]) -> Dict[str, Any]:
    """reference: analyze.py:125-180."""
    ok = [i for i, s in enumerate(cols["status"]) if s == SUCCESS]
    ttfb = [cols["ttfb_ms"][i] for i in ok]
    tllt = [cols["tllt_ms"][i] for i in ok]
    toks = [cols["completion_tokens"][i] for i in ok]
    gen: List[float] = []
    per_tok: List[float] = []
    for a, b, k in zip(ttfb, tllt, toks):
        if not math.isnan(a) and not math.isnan(b) and k > 0:
            g = b - a
            gen.append(g)
            per_tok.append(g / k)

    def pct(xs, p):
        return percentile(xs, p) if xs else None

    return {
        "ttfb_histogram": histogram([x for x in ttfb if not math.isnan(x)]),
        "tllt_histogram": histogram([x for x in tllt if not math.isnan(x)]),
        "generation_time_histogram": histogram(gen),
        "per_token_time_histogram": histogram(per_tok),
        "ttfb_p50_ms": pct(ttfb, 0.50), "ttfb_p95_ms": pct(ttfb, 0.95), "ttfb_p99_ms": pct(ttfb, 0.99),
        "tllt_p50_ms": pct(tllt, 0.50), "tllt_p95_ms": pct(tllt, 0.95), "tllt_p99_ms": pct(tllt, 0.99),
        "generation_p50_ms": pct(gen, 0.50), "generation_p95_ms": pct(gen, 0.95),
        "per_token_p50_ms": pct(per_tok, 0.50), "per_token_p95_ms": pct(per_tok, 0.95),
        # statistics.mean of integer-valued floats == exact rational, correctly rounded
        "avg_tokens_per_response": statistics.mean(float(k) for k in toks) if toks else None,
    }


# ----------------------------------------------------------------------------- A9 (+A5,A7,A8,A4)
UTIL_KEYS = ("gpu_util_avg", "gpu_mem_used_avg", "gpu_power_watts_avg", "cpu_util_avg", "mem_used_avg")


def analyze(cols: Dict[str, list], cold_times: Sequence[float]) -> Dict[str, Any]:
    """Numeric core of analyze.py main(), keys in the reference's order
    (analyze.py:477-595) with the Prometheus / io-probe / cache keys left at None."""
    n = len(cols["status"])
    t0, t1 = window_bounds(cols["start_ms"], cols["latency_ms"])
    cold_times = sorted(cold_times)
    is_cold = classify_cold(cols["start_ms"], cold_times)

    ok = [i for i in range(n) if cols["status"][i] == SUCCESS]
    lats = [cols["latency_ms"][i] for i in ok]
    ttfs = [cols["ttfb_ms"][i] for i in ok if not math.isnan(cols["ttfb_ms"][i])]
    success = len(ok)
    error_rate = (n - success) / n if n else None
    duration = t1 - t0 if t1 > t0 else None
    throughput = (success / duration) if duration and duration > 0 else None
    completion = [float(cols["completion_tokens"][i]) for i in ok]
    tokens_per_sec = (sum(completion) / duration) if duration and duration > 0 else None

    res: Dict[str, Any] = {
        "p50_ms": percentile(lats, 0.50),
        "p95_ms": percentile(lats, 0.95),
        "p99_ms": percentile(lats, 0.99),
        "throughput_rps": throughput,
        "tokens_per_sec": tokens_per_sec,
        "error_rate": error_rate,
        "cold_start_count": len(cold_times),
        "time_to_first_token_ms": statistics.mean(ttfs) if ttfs else None,
        "ttft_p50_ms": percentile(ttfs, 0.50) if ttfs else None,
        "ttft_p95_ms": percentile(ttfs, 0.95) if ttfs else None,
    }
    for k in UTIL_KEYS:
        res[k] = None
    res.update(cold_warm_metrics(cols, is_cold))
    res.update(token_timing(cols))
    res["window"] = {"start": t0, "end": t1, "seconds": duration}
    res["requests"] = {"total": n, "success": success}
    for k in ("network_rtt_p50_ms", "network_rtt_p95_ms", "s3_avg_MBps", "cache_hit_ratio"):
        res[k] = None
    return res


def total_tokens_success(cols: Dict[str, list]) -> float:
    """analyze.py:507-511 / cost_estimator.py:418-422 / collector.py:351-355."""
    return sum(float(t) for t, s in zip(cols["total_tokens"], cols["status"]) if s == SUCCESS)


# ----------------------------------------------------------------------------- A12 (extension, UNPINNED)
def window_series(cols: Dict[str, list], window_s: float = 1.0) -> Dict[str, Any]:
    """1-second series (no reference function - parity unpinned).

    bin = floor((start_ms - min(start_ms)) / (window_s*1000)); number of bins =
    ceil((max(start+lat) - min(start))/1000 / window_s), at least 1, and at least
    max(bin)+1.  Per bin: requests, successes, sum completion_tokens of successes.
    """
    start = cols["start_ms"]
    if not start:
        return {"window_s": window_s, "requests": [], "success": [], "completion_tokens": [],
                "rps": [], "tokens_per_sec": []}
    s0 = min(start)
    t0, t1 = window_bounds(start, cols["latency_ms"])
    span_ms = window_s * 1000.0
    bins = [int(math.floor((s - s0) / span_ms)) for s in start]
    nb = max(1, int(math.ceil((t1 - t0) / window_s)), max(bins) + 1)
    req = [0] * nb
    suc = [0] * nb
    tok = [0] * nb
    for b, st, k in zip(bins, cols["status"], cols["completion_tokens"]):
        req[b] += 1
        if st == SUCCESS:
            suc[b] += 1
            tok[b] += int(k)
    return {
        "window_s": window_s,
        "requests": req, "success": suc, "completion_tokens": tok,
        "rps": [c / window_s for c in suc],
        "tokens_per_sec": [c / window_s for c in tok],
    }


# ----------------------------------------------------------------------------- A15
def trapezoid_wh(ts: Sequence[float], watts: Sequence[Optional[float]], t0: float, t1: float) -> float:
    """reference: energy/collector.py:133-149.  `watts[i] is None` == missing sample.

    keep non-missing samples with t0<=ts<=t1; <2 kept -> 0; STABLE sort by ts;
    sequential `wh += ((p0+p1)/2.0) * ((ts1-ts0)/3600.0)` for dt>0."""
    if len(ts) == 0 or t1 <= t0:
        return 0.0
    kept = [(t, w) for t, w in zip(ts, watts) if w is not None and t0 <= t <= t1]
    if len(kept) < 2:
        return 0.0
    kept.sort(key=lambda tw: tw[0])
    wh = 0.0
    for (ta, pa), (tb, pb) in zip(kept, kept[1:]):
        dt_h = (tb - ta) / 3600.0
        if dt_h > 0:
            wh += ((pa + pb) / 2.0) * dt_h
    return wh


# ----------------------------------------------------------------------------- A15b
def integrate_energy(cols: Dict[str, list], ts: Sequence[float], watts: Sequence[Optional[float]],
                     include_warmup: bool = False, idle_tax: Optional[str] = None
                     ) -> Optional[Dict[str, Any]]:
    """energy.json body.  reference: energy/collector.py:289-368.  None == `return 2`."""
    t0, t1 = window_bounds(cols["start_ms"], cols["latency_ms"])
    if include_warmup and len(ts):
        t0, t1 = min(ts), max(ts)
    if t1 <= t0:
        return None
    wh_active = trapezoid_wh(ts, watts, t0, t1)
    wh_idle = None
    if idle_tax == "series":
        if len(ts):
            smin, smax = min(ts), max(ts)
            before = trapezoid_wh(ts, watts, smin, min(t0, smax)) if smin < t0 else 0.0
            after = trapezoid_wh(ts, watts, max(smin, t1), smax) if smax > t1 else 0.0
            wh_idle = before + after
    elif idle_tax == "baseline":
        outside = [w for t, w in zip(ts, watts) if w is not None and (t < t0 or t > t1)]
        if outside:
            p_idle = sorted(outside)[len(outside) // 2]
            if len(ts) > 1:
                steps = [ts[i + 1] - ts[i] for i in range(len(ts) - 1)]
                avg_step = sum(steps) / len(steps)
            else:
                avg_step = t1 - t0
            n_out = sum(1 for t in ts if (t < t0 or t > t1))
            wh_idle = p_idle * ((n_out * avg_step) / 3600.0)
    success = sum(1 for s in cols["status"] if s == SUCCESS)
    tokens = total_tokens_success(cols)
    return {
        "Wh_active": wh_active,
        "Wh_idle_tax": wh_idle,
        "Wh_per_1k_tokens_active": ((wh_active / tokens) * 1000.0) if tokens > 0 else None,
        "Wh_per_request_active": (wh_active / success) if success > 0 else None,
        "window": {"start": t0, "end": t1},
    }


# ----------------------------------------------------------------------------- A16 (extension, UNPINNED)
def interval_energy(ts: Sequence[float], watts: Sequence[Optional[float]],
                    start_ms: Sequence[float], latency_ms: Sequence[float]) -> List[float]:
    """E_i = trapezoid_wh(samples, s_i, e_i) with [s_i,e_i]=[start, start+lat]/1000.
    (the one-interval case IS collector.py:133-149.)"""
    return [trapezoid_wh(ts, watts, s / 1000.0, (s + l) / 1000.0)
            for s, l in zip(start_ms, latency_ms)]


# ----------------------------------------------------------------------------- A17
def cost_calculator(latency_ms: Sequence[float], status: Sequence[str], gpu_hourly: float,
                    requests_per_1k: int = 10) -> Optional[Dict[str, float]]:
    """reference: cost_calculator.py:42-57.  None == 'no successful requests' (exit 1)."""
    lat = [l for l, s in zip(latency_ms, status) if s == "200"]
    if not lat:
        return None
    avg_ms = sum(lat) / len(lat)
    avg_s = avg_ms / 1000
    per_s = gpu_hourly / 3600
    return {"successful": len(lat), "avg_latency_ms": avg_ms, "avg_latency_s": avg_s,
            "gpu_price_per_second": per_s, "cost_per_1k_tokens": per_s * avg_s * requests_per_1k}


# ----------------------------------------------------------------------------- A19 / A20
def cold_warm_costs(cols: Dict[str, list], is_cold: Sequence[bool], total_cost: float,
                    success: int) -> Dict[str, Any]:
    """reference: cost_estimator.py:289-340."""
    ok = [i for i, s in enumerate(cols["status"]) if s == SUCCESS]
    cold_n = sum(1 for i in ok if is_cold[i])
    warm_n = len(ok) - cold_n
    # builtin sum: int 0 for an empty subset, float otherwise (cost_estimator.py:307-308)
    cold_tok = sum(float(cols["total_tokens"][i]) for i in ok if is_cold[i])
    warm_tok = sum(float(cols["total_tokens"][i]) for i in ok if not is_cold[i])
    cf = cold_n / success if success > 0 else 0
    wf = warm_n / success if success > 0 else 0
    cold_cost = total_cost * cf
    warm_cost = total_cost * wf
    return {
        "cold_cost_per_request": (cold_cost / cold_n) if cold_n > 0 else None,
        "warm_cost_per_request": (warm_cost / warm_n) if warm_n > 0 else None,
        "cold_cost_per_1k_tokens": (cold_cost / cold_tok * 1000.0) if cold_tok > 0 else None,
        "warm_cost_per_1k_tokens": (warm_cost / warm_tok * 1000.0) if warm_tok > 0 else None,
        "cold_total_cost": cold_cost, "warm_total_cost": warm_cost,
        "cold_requests": cold_n, "warm_requests": warm_n,
        "cold_tokens": cold_tok, "warm_tokens": warm_tok,
    }


def cost_rollup(gpu_seconds: float, cpu_core_seconds: float, mem_gib_seconds: float,
                gpu_hr: float, cpu_hr: float, mem_hr: float, overhead: float,
                success: int, total_tokens: float) -> Dict[str, Any]:
    """reference: cost_estimator.py:441-451 (operation order as written there)."""
    total = (gpu_seconds * (gpu_hr / 3600.0)) + (cpu_core_seconds * (cpu_hr / 3600.0)) \
        + (mem_gib_seconds * (mem_hr / 3600.0))
    total *= 1.0 + overhead
    return {
        "total_cost": total,
        "cost_per_request": (total / success) if success > 0 else None,
        "cost_per_1k_tokens": (total / total_tokens * 1000.0) if total_tokens > 0 else None,
    }
