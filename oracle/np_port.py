"""numpy-vectorised restatement of oracle/py_port.py (TEST ORACLE, large sizes).

Same contract as py_port, with numpy arrays as columns:
  start_ms, latency_ms, ttfb_ms, tllt_ms : float64 (nan == missing)
  status, completion_tokens, total_tokens : int64

Order statistics come from np.sort and are then interpolated with the REFERENCE's
expression in Python floats (analyze.py:75-81) - not numpy.percentile, whose lerp
differs by 1 ULP in some cases.  Counts, ranks, order
statistics and histogram counts are bit-equal to py_port (tests/test_oracle.py);
fp sums (mean TTFT, energy) agree to <=1e-12 relative.
"""

from __future__ import annotations

import math
from typing import Any, Dict, Optional, Sequence, Tuple

import numpy as np

from . import py_port

SUCCESS = 200


def percentile_sorted(sorted_vals: np.ndarray, p: float) -> float:
    """analyze.py:72-81 on an already NaN-free ascending array."""
    n = int(sorted_vals.shape[0])
    if n == 0:
        return float("nan")
    if n == 1:
        return float(sorted_vals[0])
    lo, hi, pos = py_port.percentile_ranks(n, p)
    return py_port.interpolate(float(sorted_vals[lo]), float(sorted_vals[hi]), lo, hi, pos)


def percentile(values: np.ndarray, p: float) -> float:
    v = np.asarray(values, dtype=np.float64)
    return percentile_sorted(np.sort(v[~np.isnan(v)]), p)


def order_stats(values: np.ndarray, ps: Sequence[float]) -> Dict[float, Tuple[int, float, float]]:
    """For each p: (n, v[lo], v[hi]) - what the device select returns."""
    v = np.asarray(values, dtype=np.float64)
    s = np.sort(v[~np.isnan(v)])
    n = int(s.shape[0])
    out = {}
    for p in ps:
        if n == 0:
            out[p] = (0, float("nan"), float("nan"))
            continue
        lo, hi, _ = py_port.percentile_ranks(n, p) if n > 1 else (0, 0, 0.0)
        out[p] = (n, float(s[lo]), float(s[hi]))
    return out


def histogram(values: np.ndarray, num_bins: int = 20) -> Dict[str, Any]:
    """analyze.py:84-122, vectorised; edges/centers evaluated in Python floats."""
    v = np.asarray(values, dtype=np.float64)
    v = v[~np.isnan(v)]
    if v.size == 0:
        return {"bins": [], "counts": [], "bin_edges": []}
    lo, hi = float(v.min()), float(v.max())
    if lo == hi:
        return {"bins": [lo], "counts": [int(v.size)], "bin_edges": [lo, hi]}
    width = (hi - lo) / num_bins
    idx = np.minimum(num_bins - 1, ((v - lo) / width).astype(np.int64))
    counts = np.bincount(idx, minlength=num_bins).tolist()
    edges = [lo + i * width for i in range(num_bins + 1)]
    centers = [(edges[i] + edges[i + 1]) / 2 for i in range(num_bins)]
    return {"bins": centers, "counts": counts, "bin_edges": edges,
            "min": lo, "max": hi, "total_samples": int(v.size)}


def window_bounds(start_ms: np.ndarray, latency_ms: np.ndarray) -> Tuple[float, float]:
    """NaN cells are skipped: Python's min()/max() over a list holding NaN is order-dependent
    (analyze.py:185-188), so that case has no well-defined reference answer."""
    if start_ms.size == 0:
        return (0.0, 0.0)
    return float(np.nanmin(start_ms)) / 1000.0, float(np.nanmax(start_ms + latency_ms)) / 1000.0


def classify_cold(start_ms: np.ndarray, cold_times: Sequence[float], window: float = 30.0) -> np.ndarray:
    """analyze.py:402-419.  With sorted cold times the largest c <= t decides."""
    c = np.sort(np.asarray(cold_times, dtype=np.float64))
    t = start_ms / 1000.0
    if c.size == 0:
        return np.zeros(t.shape, dtype=bool)
    j = np.searchsorted(c, t, side="right") - 1
    ok = j >= 0
    return ok & (t <= (c[np.maximum(j, 0)] + window))


def _mean_exact(v: np.ndarray) -> Optional[float]:
    """statistics.mean equivalent to <=1 ULP: correctly rounded sum / n."""
    if v.size == 0:
        return None
    return math.fsum(v.tolist()) / int(v.size)


def analyze(cols: Dict[str, np.ndarray], cold_times: Sequence[float]) -> Dict[str, Any]:
    """py_port.analyze, vectorised.  Same keys, same order."""
    start, lat = cols["start_ms"], cols["latency_ms"]
    ttfb, tllt = cols["ttfb_ms"], cols["tllt_ms"]
    status, ctok = cols["status"], cols["completion_tokens"]
    n = int(status.shape[0])
    t0, t1 = window_bounds(start, lat)
    cold_times = sorted(cold_times)
    is_cold = classify_cold(start, cold_times)
    ok = status == SUCCESS
    success = int(ok.sum())
    duration = t1 - t0 if t1 > t0 else None
    lat_ok = np.sort(lat[ok & ~np.isnan(lat)])
    ttf_mask = ok & ~np.isnan(ttfb)
    ttf_sorted = np.sort(ttfb[ttf_mask])
    sum_ctok = int(ctok[ok].sum())

    def pct_or_none(sorted_vals, p, nonempty):
        return percentile_sorted(sorted_vals, p) if nonempty else None

    res: Dict[str, Any] = {
        "p50_ms": percentile_sorted(lat_ok, 0.50),
        "p95_ms": percentile_sorted(lat_ok, 0.95),
        "p99_ms": percentile_sorted(lat_ok, 0.99),
        "throughput_rps": (success / duration) if duration and duration > 0 else None,
        "tokens_per_sec": (float(sum_ctok) / duration) if duration and duration > 0 else None,
        "error_rate": (n - success) / n if n else None,
        "cold_start_count": len(cold_times),
        "time_to_first_token_ms": _mean_exact(ttfb[ttf_mask]),
        "ttft_p50_ms": pct_or_none(ttf_sorted, 0.50, ttf_sorted.size > 0),
        "ttft_p95_ms": pct_or_none(ttf_sorted, 0.95, ttf_sorted.size > 0),
    }
    for k in py_port.UTIL_KEYS:
        res[k] = None

    for prefix, sub in (("cold", is_cold), ("warm", ~is_cold)):
        cnt = int(sub.sum())
        if cnt == 0:
            res.update({f"{prefix}_count": 0, f"{prefix}_p50_ms": None,
                        f"{prefix}_p95_ms": None, f"{prefix}_p99_ms": None})
            continue
        sok = sub & ok
        nok = int(sok.sum())
        ls = np.sort(lat[sok & ~np.isnan(lat)])
        ts_ = np.sort(ttfb[sok & ~np.isnan(ttfb)])
        res.update({
            f"{prefix}_count": cnt,
            f"{prefix}_success_count": nok,
            f"{prefix}_p50_ms": percentile_sorted(ls, 0.50) if nok else None,
            f"{prefix}_p95_ms": percentile_sorted(ls, 0.95) if nok else None,
            f"{prefix}_p99_ms": percentile_sorted(ls, 0.99) if nok else None,
            f"{prefix}_ttft_p50_ms": percentile_sorted(ts_, 0.50) if ts_.size else None,
            f"{prefix}_ttft_p95_ms": percentile_sorted(ts_, 0.95) if ts_.size else None,
            f"{prefix}_error_rate": (cnt - nok) / cnt,
        })

    # token timing (analyze.py:125-180)
    a, b, k = ttfb[ok], tllt[ok], ctok[ok]
    gmask = ~np.isnan(a) & ~np.isnan(b) & (k > 0)
    gen = b[gmask] - a[gmask]
    per = gen / k[gmask].astype(np.float64)
    a_s, b_s = np.sort(a[~np.isnan(a)]), np.sort(b[~np.isnan(b)])
    g_s, p_s = np.sort(gen[~np.isnan(gen)]), np.sort(per[~np.isnan(per)])
    has = success > 0
    res.update({
        "ttfb_histogram": histogram(a), "tllt_histogram": histogram(b),
        "generation_time_histogram": histogram(gen), "per_token_time_histogram": histogram(per),
        "ttfb_p50_ms": pct_or_none(a_s, 0.50, has), "ttfb_p95_ms": pct_or_none(a_s, 0.95, has),
        "ttfb_p99_ms": pct_or_none(a_s, 0.99, has),
        "tllt_p50_ms": pct_or_none(b_s, 0.50, has), "tllt_p95_ms": pct_or_none(b_s, 0.95, has),
        "tllt_p99_ms": pct_or_none(b_s, 0.99, has),
        "generation_p50_ms": pct_or_none(g_s, 0.50, gen.size > 0),
        "generation_p95_ms": pct_or_none(g_s, 0.95, gen.size > 0),
        "per_token_p50_ms": pct_or_none(p_s, 0.50, per.size > 0),
        "per_token_p95_ms": pct_or_none(p_s, 0.95, per.size > 0),
        "avg_tokens_per_response": (sum_ctok / success) if has else None,
    })
    res["window"] = {"start": t0, "end": t1, "seconds": duration}
    res["requests"] = {"total": n, "success": success}
    for key in ("network_rtt_p50_ms", "network_rtt_p95_ms", "s3_avg_MBps", "cache_hit_ratio"):
        res[key] = None
    return res


def window_series(cols: Dict[str, np.ndarray], window_s: float = 1.0) -> Dict[str, Any]:
    """py_port.window_series with np.bincount (extension A12, parity unpinned)."""
    start = cols["start_ms"]
    if start.size == 0:
        return py_port.window_series({"start_ms": [], "latency_ms": [], "status": [],
                                      "completion_tokens": []}, window_s)
    s0 = float(start.min())
    t0, t1 = window_bounds(start, cols["latency_ms"])
    bins = np.floor((start - s0) / (window_s * 1000.0)).astype(np.int64)
    nb = max(1, int(math.ceil((t1 - t0) / window_s)), int(bins.max()) + 1)
    ok = cols["status"] == SUCCESS
    req = np.bincount(bins, minlength=nb)
    suc = np.bincount(bins[ok], minlength=nb)
    # integer-exact token sums per bin (np.bincount weights would go through float64)
    tok = np.zeros(nb, dtype=np.int64)
    np.add.at(tok, bins[ok], cols["completion_tokens"][ok])
    return {"window_s": window_s, "requests": req.tolist(), "success": suc.tolist(),
            "completion_tokens": tok.tolist(),
            "rps": [c / window_s for c in suc.tolist()],
            "tokens_per_sec": [c / window_s for c in tok.tolist()]}


def _kept_sorted(ts: np.ndarray, watts: np.ndarray, t0: float, t1: float):
    """collector.py:138-141: drop missing (nan) watts, clip by membership, stable sort."""
    m = ~np.isnan(watts) & (ts >= t0) & (ts <= t1)
    t, w = ts[m], watts[m]
    order = np.argsort(t, kind="stable")
    return t[order], w[order]


def trapezoid_terms(t: np.ndarray, w: np.ndarray) -> np.ndarray:
    """Per-segment areas in the reference's expression order; 0 where dt<=0."""
    dt_h = (t[1:] - t[:-1]) / 3600.0
    area = ((w[:-1] + w[1:]) / 2.0) * dt_h
    return np.where(dt_h > 0, area, 0.0)


def trapezoid_wh(ts: np.ndarray, watts: np.ndarray, t0: float, t1: float,
                 exact: bool = False) -> float:
    """collector.py:133-149 (missing watts == nan).  Sequential fp64 accumulation via
    np.cumsum (strictly left-to-right) reproduces the reference's `wh +=` bit for bit;
    exact=True returns math.fsum of the same terms (the yardstick at 1e8 samples)."""
    if ts.size == 0 or t1 <= t0:
        return 0.0
    t, w = _kept_sorted(ts, watts, t0, t1)
    if t.size < 2:
        return 0.0
    terms = trapezoid_terms(t, w)
    terms = terms[(t[1:] - t[:-1]) / 3600.0 > 0]
    if terms.size == 0:
        return 0.0
    if exact:
        return math.fsum(terms.tolist())
    return float(np.cumsum(terms)[-1])


def interval_energy(ts: np.ndarray, watts: np.ndarray, start_ms: np.ndarray,
                    latency_ms: np.ndarray) -> np.ndarray:
    """Extension A16: E_i = trapezoid over samples with s_i<=ts<=e_i (membership clip, no
    edge interpolation).  Prefix sums in long double keep the difference of two large
    prefixes accurate to ~1e-15 of the trace total."""
    m = ~np.isnan(watts)
    t, w = ts[m], watts[m]
    order = np.argsort(t, kind="stable")
    t, w = t[order], w[order]
    s = start_ms / 1000.0
    e = (start_ms + latency_ms) / 1000.0
    out = np.zeros(s.shape, dtype=np.float64)
    if t.size < 2:
        return out
    terms = trapezoid_terms(t, w).astype(np.longdouble)
    prefix = np.concatenate([np.zeros(1, dtype=np.longdouble), np.cumsum(terms)])
    lb = np.searchsorted(t, s, side="left")
    ub = np.searchsorted(t, e, side="right")
    ok = (ub - lb >= 2) & (e > s)
    hi = np.clip(ub - 1, 0, t.size - 1)
    out[ok] = (prefix[hi[ok]] - prefix[lb[ok]]).astype(np.float64)
    return out
