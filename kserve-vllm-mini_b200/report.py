"""RawAnalysis -> the reference's results.json fields (host-side scalar arithmetic).

Everything here is O(1) per run: interpolation of two order statistics, ratios of exact
integer counts, histogram edges.  The expressions and their ORDER follow the reference
so the JSON is bit-identical on integer and order-statistic fields:

  analyze.py:573-595   key order of results.json
  analyze.py:427-455   cold/warm subset fields (keys present only for non-empty subsets)
  analyze.py:84-122    histogram dict forms (empty / min==max / regular)
  analyze.py:154-180   token timing fields
"""

from __future__ import annotations

import math
from fractions import Fraction
from typing import Any, Dict, List, Optional

from . import _native as N
from .engine import OrderStat, RawAnalysis

UTIL_KEYS = ("gpu_util_avg", "gpu_mem_used_avg", "gpu_power_watts_avg", "cpu_util_avg", "mem_used_avg")


def dd_mean(hi: float, lo: float, n: int) -> Optional[float]:
    """Mean from a double-double sum: the exact rational (hi+lo)/n rounded once, which is
    what statistics.mean returns for the exact sum (analyze.py:581)."""
    if n <= 0:
        return None
    return float((Fraction(hi) + Fraction(lo)) / n)


def histogram_dict(count: int, rng, counts: List[int], num_bins: int) -> Dict[str, Any]:
    """analyze.py:84-122 from (count of non-NaN samples, (min,max), per-bin counts)."""
    if count == 0 or rng is None:
        return {"bins": [], "counts": [], "bin_edges": []}
    lo, hi = rng
    if lo == hi:
        return {"bins": [lo], "counts": [count], "bin_edges": [lo, hi]}
    width = (hi - lo) / num_bins
    edges = [lo + i * width for i in range(num_bins + 1)]
    centers = [(edges[i] + edges[i + 1]) / 2 for i in range(num_bins)]
    return {"bins": centers, "counts": [int(c) for c in counts], "bin_edges": edges,
            "min": lo, "max": hi, "total_samples": count}


def _pct(stat: OrderStat, p: float, list_len: int) -> Optional[float]:
    """`percentile(xs, p) if xs else None` where xs has list_len entries (NaNs included)."""
    if list_len <= 0:
        return None
    return stat.interpolate(p)


def trim_series(raw: RawAnalysis) -> Optional[Dict[str, Any]]:
    """Extension A12 (additive key `timeseries`; consumers use .get - SURVEY.md 8b)."""
    if raw.series is None:
        return None
    s = raw.stats
    w = raw.series_window_s
    t0, t1 = s.min_start_ms / 1000.0, s.max_end_ms / 1000.0
    req = raw.series[0]
    nz = req.nonzero()[0]
    last = int(nz[-1]) + 1 if nz.size else 0
    nb = max(1, int(math.ceil((t1 - t0) / w)), last)
    req_l = raw.series[0][:nb].tolist()
    suc_l = raw.series[1][:nb].tolist()
    tok_l = raw.series[2][:nb].tolist()
    return {"window_s": w, "requests": req_l, "success": suc_l, "completion_tokens": tok_l,
            "rps": [c / w for c in suc_l], "tokens_per_sec": [c / w for c in tok_l]}


def core_results(raw: RawAnalysis, cold_start_count: int) -> Dict[str, Any]:
    """The numeric core of analyze.py main() in the reference's key order; utilisation,
    energy-from-Prometheus, io-probe and cache keys are spliced in by the CLI."""
    s = raw.stats
    o = raw.order
    cnt = list(s.sets.count)
    total, success = int(s.n_total), int(s.n_success)
    start, end = s.min_start_ms / 1000.0, s.max_end_ms / 1000.0   # analyze.py:185-188
    duration = end - start if end > start else None
    throughput = (success / duration) if duration and duration > 0 else None
    tokens_per_sec = (float(int(s.sum_completion_tokens)) / duration) if duration and duration > 0 else None
    n_ttfb = cnt[N.SET_TTFB_COLD] + cnt[N.SET_TTFB_WARM]

    res: Dict[str, Any] = {
        "p50_ms": o["lat_p50"].interpolate(0.50),
        "p95_ms": o["lat_p95"].interpolate(0.95),
        "p99_ms": o["lat_p99"].interpolate(0.99),
        "throughput_rps": throughput,
        "tokens_per_sec": tokens_per_sec,
        "error_rate": (total - success) / total if total else None,
        "cold_start_count": cold_start_count,
        "time_to_first_token_ms": dd_mean(s.ttfb_sum_hi, s.ttfb_sum_lo, n_ttfb),
        "ttft_p50_ms": _pct(o["ttfb_p50"], 0.50, n_ttfb),
        "ttft_p95_ms": _pct(o["ttfb_p95"], 0.95, n_ttfb),
    }
    for k in UTIL_KEYS:
        res[k] = None

    # cold / warm (analyze.py:427-455)
    n_cold = int(s.n_cold)
    subsets = (("cold", n_cold, int(s.n_cold_success), cnt[N.SET_TTFB_COLD]),
               ("warm", total - n_cold, success - int(s.n_cold_success), cnt[N.SET_TTFB_WARM]))
    for prefix, count, ok, n_tt in subsets:
        if count == 0:
            res.update({f"{prefix}_count": 0, f"{prefix}_p50_ms": None,
                        f"{prefix}_p95_ms": None, f"{prefix}_p99_ms": None})
            continue
        res.update({
            f"{prefix}_count": count,
            f"{prefix}_success_count": ok,
            f"{prefix}_p50_ms": _pct(o[f"{prefix}_lat_p50"], 0.50, ok),
            f"{prefix}_p95_ms": _pct(o[f"{prefix}_lat_p95"], 0.95, ok),
            f"{prefix}_p99_ms": _pct(o[f"{prefix}_lat_p99"], 0.99, ok),
            f"{prefix}_ttft_p50_ms": _pct(o[f"{prefix}_ttfb_p50"], 0.50, n_tt),
            f"{prefix}_ttft_p95_ms": _pct(o[f"{prefix}_ttfb_p95"], 0.95, n_tt),
            f"{prefix}_error_rate": (count - ok) / count,
        })

    # token timing (analyze.py:125-180)
    nb = raw.num_bins
    hc = raw.hist_counts
    n_gen_rows = int(s.n_gen_rows)
    res.update({
        "ttfb_histogram": histogram_dict(n_ttfb, raw.col_range(N.COL_TTFB), hc[0].tolist(), nb),
        "tllt_histogram": histogram_dict(cnt[N.SET_TLLT], raw.col_range(N.COL_TLLT), hc[1].tolist(), nb),
        "generation_time_histogram": histogram_dict(cnt[N.SET_GEN], raw.col_range(N.COL_GEN), hc[2].tolist(), nb),
        "per_token_time_histogram": histogram_dict(cnt[N.SET_PERTOK], raw.col_range(N.COL_PERTOK), hc[3].tolist(), nb),
        "ttfb_p50_ms": _pct(o["ttfb_p50"], 0.50, success),
        "ttfb_p95_ms": _pct(o["ttfb_p95"], 0.95, success),
        "ttfb_p99_ms": _pct(o["ttfb_p99"], 0.99, success),
        "tllt_p50_ms": _pct(o["tllt_p50"], 0.50, success),
        "tllt_p95_ms": _pct(o["tllt_p95"], 0.95, success),
        "tllt_p99_ms": _pct(o["tllt_p99"], 0.99, success),
        "generation_p50_ms": _pct(o["gen_p50"], 0.50, n_gen_rows),
        "generation_p95_ms": _pct(o["gen_p95"], 0.95, n_gen_rows),
        "per_token_p50_ms": _pct(o["pertok_p50"], 0.50, n_gen_rows),
        "per_token_p95_ms": _pct(o["pertok_p95"], 0.95, n_gen_rows),
        # statistics.mean of integer-valued floats == exact ratio rounded once
        "avg_tokens_per_response": (int(s.sum_completion_tokens) / success) if success else None,
    })
    res["window"] = {"start": start, "end": end, "seconds": duration}
    res["requests"] = {"total": total, "success": success}
    return res


def total_tokens(raw: RawAnalysis) -> float:
    """analyze.py:507-511: float sum of integer-valued token counts."""
    return float(int(raw.stats.sum_total_tokens))
