"""Runs the UNMODIFIED reference (Siryoos/kserve-vllm-mini) on in-memory rows.  TEST INFRASTRUCTURE ONLY.

`install()` copies the five source files of the reference's run-analysis path from
/root/reference into oracle/_ref/ (git-ignored: they never enter the history; gpurun ships the
directory to the GPU box, where /root/reference does not exist) - the file-copy equivalent of
`pip install --target`; the reference has no setup.py / pyproject to install from.
`__graft_entry__.build()` calls it in the build container.

`analyze_rows()` calls the reference's own functions in the order of `analyze.main()`
(/root/reference/analyze.py:477-595) on rows already parsed into the list-of-dicts form
`read_requests_csv` returns (analyze.py:34-56), leaving out what is not arithmetic on the rows:
CSV parsing, kubectl, Prometheus, the io-probe passthrough and the two file writers.  That is
the same slice of work the GPU arm of bench.py is timed on (columns resident, results as a dict).

Only tests/ and bench.py's CPU legs may import this module (see oracle/__init__.py).
"""

from __future__ import annotations

import importlib.util
import math
import os
import shutil
import statistics as stats
import sys
from typing import Any, Dict, List, Optional, Sequence

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DST = os.path.join(HERE, "_ref")
REF_SRC = "/root/reference"
FILES = ("analyze.py", "cost_estimator.py", "cost_calculator.py", os.path.join("energy", "collector.py"),
         os.path.join("energy", "__init__.py"))


def install(src: str = REF_SRC, dst: str = REF_DST) -> bool:
    """Copy the reference's path files into oracle/_ref/ when the reference tree is present.
    Returns True when oracle/_ref/ is complete afterwards."""
    if os.path.isdir(src):
        for rel in FILES:
            s, d = os.path.join(src, rel), os.path.join(dst, rel)
            if not os.path.exists(s):
                continue
            os.makedirs(os.path.dirname(d), exist_ok=True)
            if not os.path.exists(d) or open(s, "rb").read() != open(d, "rb").read():
                shutil.copyfile(s, d)
    return available(dst)


def available(dst: str = REF_DST) -> bool:
    return all(os.path.exists(os.path.join(dst, rel)) for rel in FILES)


_mods: Dict[str, Any] = {}


def module(name: str = "analyze"):
    """Import oracle/_ref/<name>.py under a private module name (the repo root has drop-in
    scripts of the same names, which must not be shadowed or picked up instead)."""
    if name not in _mods:
        path = os.path.join(REF_DST, name + ".py")
        spec = importlib.util.spec_from_file_location("_kvm_reference_" + name, path)
        mod = importlib.util.module_from_spec(spec)
        old = sys.dont_write_bytecode
        sys.dont_write_bytecode = True
        try:
            spec.loader.exec_module(mod)
        finally:
            sys.dont_write_bytecode = old
        _mods[name] = mod
    return _mods[name]


COLS = ("start_ms", "latency_ms", "ttfb_ms", "tllt_ms", "status", "completion_tokens", "total_tokens")


def rows_from_columns(cols: Dict[str, Sequence], n: Optional[int] = None) -> List[dict]:
    """Columns -> the list of dicts analyze.read_requests_csv produces (every numeric cell a float,
    analyze.py:46-53).  Not part of any timed region."""
    n = len(cols["status"]) if n is None else n
    lists = {k: (cols[k][:n].tolist() if hasattr(cols[k], "tolist") else list(cols[k][:n])) for k in COLS}
    rows = []
    for i in range(n):
        rows.append({"id": str(i), "scheduled_ms": "", "start_ms": float(lists["start_ms"][i]),
                     "ttfb_ms": float(lists["ttfb_ms"][i]), "tllt_ms": float(lists["tllt_ms"][i]),
                     "latency_ms": float(lists["latency_ms"][i]), "status": float(lists["status"][i]),
                     "prompt_tokens": 0.0, "completion_tokens": float(lists["completion_tokens"][i]),
                     "total_tokens": float(lists["total_tokens"][i]), "error": "", "trace_id": ""})
    return rows


def analyze_rows(rows: List[dict], cold_start_times: Sequence[float]) -> Dict[str, Any]:
    """analyze.main() from `window_bounds` to the results dict (analyze.py:477-595), every step a
    call into the unmodified reference module."""
    A = module("analyze")
    start, end = A.window_bounds(rows)
    rows = A.classify_requests_cold_warm(rows, list(cold_start_times))
    ok = [r for r in rows if str(int(r.get("status", 0))) == "200"]
    lats = [r.get("latency_ms", float("nan")) for r in ok]
    ttfs = [r.get("ttfb_ms", float("nan")) for r in ok if not math.isnan(r.get("ttfb_ms", float("nan")))]
    success = len(ok)
    total = len(rows)
    error_rate = (total - success) / total if total else None
    duration = end - start if end > start else None
    throughput = (success / duration) if duration and duration > 0 else None
    completion_tokens = [r.get("completion_tokens", 0.0) or 0.0 for r in ok]
    tokens_per_sec = (sum(completion_tokens) / duration) if duration and duration > 0 else None
    cold_warm = A.compute_cold_warm_metrics(rows)
    timing = A.compute_token_timing_analysis(rows)
    return {
        "p50_ms": A.percentile(lats, 0.50), "p95_ms": A.percentile(lats, 0.95), "p99_ms": A.percentile(lats, 0.99),
        "throughput_rps": throughput, "tokens_per_sec": tokens_per_sec, "error_rate": error_rate,
        "cold_start_count": len(cold_start_times),
        "time_to_first_token_ms": stats.mean(ttfs) if ttfs else None,
        "ttft_p50_ms": A.percentile(ttfs, 0.50) if ttfs else None,
        "ttft_p95_ms": A.percentile(ttfs, 0.95) if ttfs else None,
        **cold_warm, **timing,
        "window": {"start": start, "end": end, "seconds": duration},
        "requests": {"total": total, "success": success},
    }


def cost_rows(rows: List[dict], total_cost: float) -> Dict[str, Any]:
    """cost_estimator.main()'s reductions over the rows (cost_estimator.py:418-422,448-454)."""
    CE = module("cost_estimator")
    ok = [r for r in rows if str(int(r.get("status", 0))) == "200"]
    success = len(ok)
    total_tokens = sum(float(r.get("total_tokens", 0.0) or 0.0) for r in ok)
    out = {"cost_per_request": (total_cost / success) if success > 0 else None,
           "cost_per_1k_tokens": (total_cost / total_tokens * 1000.0) if total_tokens > 0 else None}
    out.update(CE.calculate_cold_warm_costs(rows, total_cost, success, total_tokens))
    return out
