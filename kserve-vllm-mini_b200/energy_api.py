"""Drop-in for the reference's energy/collector.py `integrate` path (and its importable
helpers), with the sample arithmetic on the GPU, plus the per-request interval join.

Reference interface mirrored (file:line):
  PowerSample(ts, watts)                         energy/collector.py:125-130
  read_requests_csv(path) -> rows                energy/collector.py:95-111
  run_window_bounds(rows) -> (t0, t1)            energy/collector.py:114-122
  trapezoidal_wh(samples, t0, t1) -> Wh          energy/collector.py:133-149
  load_power_samples(path) -> [PowerSample]      energy/collector.py:152-177
  integrate_energy(args) -> rc                   energy/collector.py:254-381
  main(): collect | integrate sub-commands       energy/collector.py:384-435
The `collect` sub-command polls Prometheus live (network I/O, out of scope of the device
path); it is kept as plain host code so the CLI contract is complete.
"""

from __future__ import annotations

import argparse
import csv
import ctypes as C
import json
import math
import os
import signal
import sys
import time
import urllib.parse
import urllib.request
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _native as N
from . import columns, engine

PROM_QUERIES = [
    'sum(DCGM_FI_DEV_POWER_USAGE{{namespace="{ns}",pod=~"{pod_re}"}})',
    'sum(nvidia_dcgm_power_usage_watts{{namespace="{ns}",pod=~"{pod_re}"}})',
    'sum(nvidia_gpu_power_watts{{namespace="{ns}",pod=~"{pod_re}"}})',
]


@dataclass
class PowerSample:
    """One power sample at timestamp `ts` with value `watts` (or None)."""
    ts: float
    watts: Optional[float]


# --------------------------------------------------------------------------- device trace
class PowerTrace:
    """A power trace in HBM: ts / watts fp64 columns (NaN = null reading), kept both in load
    order (the baseline idle tax needs the load-order sample steps) and ascending."""

    def __init__(self, ts: np.ndarray, watts: np.ndarray, device: Optional[torch.device] = None):
        self.device = device or engine.require_cuda()
        self.lib = N.lib()
        self.m = int(ts.shape[0])
        self.ts_loaded = torch.from_numpy(np.ascontiguousarray(ts, dtype=np.float64)).to(self.device)
        self.watts_loaded = torch.from_numpy(np.ascontiguousarray(watts, dtype=np.float64)).to(self.device)
        self._ws = torch.empty(self.lib.kvm_energy_workspace_bytes(self.m), dtype=torch.uint8, device=self.device)
        self._stats_d = torch.zeros(C.sizeof(N.PowerStatsC), dtype=torch.uint8, device=self.device)
        self.ts = self.ts_loaded
        self.watts = self.watts_loaded
        self._prefix: Optional[Tuple[torch.Tensor, torch.Tensor]] = None
        base = self.scan(-math.inf, math.inf)
        self.ts_min, self.ts_max = base.ts_min, base.ts_max
        self.n_valid = int(base.n_valid)
        self.step_sum = (base.step_sum_hi, base.step_sum_lo)
        if self.m > 1 and base.n_unsorted:
            self._sort()

    @staticmethod
    def from_samples(samples: Sequence[PowerSample], device=None) -> "PowerTrace":
        ts = np.fromiter((s.ts for s in samples), dtype=np.float64, count=len(samples))
        w = np.fromiter((math.nan if s.watts is None else s.watts for s in samples), dtype=np.float64,
                        count=len(samples))
        return PowerTrace(ts, w, device)

    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    def scan(self, t0: float, t1: float, ids: Optional[torch.Tensor] = None) -> N.PowerStatsC:
        """Load-order facts of the trace relative to [t0, t1] (kvm_power_scan)."""
        N.check(self.lib.kvm_power_scan(self.ts_loaded.data_ptr(), self.watts_loaded.data_ptr(), self.m,
                                        float(t0), float(t1), ids.data_ptr() if ids is not None else None,
                                        self._stats_d.data_ptr(), self._ws.data_ptr(), self._ws.numel(),
                                        self._stream()), "kvm_power_scan")
        return N.PowerStatsC.from_buffer_copy(self._stats_d.cpu().numpy().tobytes())

    def _sort(self) -> None:
        """Stable sort by ts (collector.py:141)."""
        ws = torch.empty(self.lib.kvm_sort_pairs_workspace_bytes(self.m), dtype=torch.uint8, device=self.device)
        ts_s, w_s = torch.empty_like(self.ts_loaded), torch.empty_like(self.watts_loaded)
        N.check(self.lib.kvm_sort_pairs_f64(self.ts_loaded.data_ptr(), self.watts_loaded.data_ptr(),
                                            ts_s.data_ptr(), w_s.data_ptr(), self.m, ws.data_ptr(), ws.numel(),
                                            self._stream()), "kvm_sort_pairs_f64")
        self.ts, self.watts = ts_s, w_s

    def trapezoid_dd(self, windows: Sequence[Tuple[float, float]]) -> List[Tuple[float, float]]:
        """Wh over each window as the device's double-double (hi, lo) pair - what the multi-GPU path
        exchanges so that the ranks' values can be added exactly, in rank order (dist_energy.py)."""
        k = len(windows)
        flat = (C.c_double * (2 * k))(*[float(x) for w in windows for x in w])
        out = torch.zeros(2 * k, dtype=torch.float64, device=self.device)
        N.check(self.lib.kvm_trapezoid_wh(self.ts.data_ptr(), self.watts.data_ptr(), self.m, flat, k,
                                          out.data_ptr(), self._ws.data_ptr(), self._ws.numel(),
                                          self._stream()), "kvm_trapezoid_wh")
        o = out.cpu().tolist()
        return [(o[2 * i], o[2 * i + 1]) for i in range(k)]

    def trapezoid(self, windows: Sequence[Tuple[float, float]]) -> List[float]:
        """Wh over each (t0, t1) window in one pass (kvm_trapezoid_wh)."""
        k = len(windows)
        flat = (C.c_double * (2 * k))(*[float(x) for w in windows for x in w])
        out = torch.zeros(2 * k, dtype=torch.float64, device=self.device)
        N.check(self.lib.kvm_trapezoid_wh(self.ts.data_ptr(), self.watts.data_ptr(), self.m, flat, k,
                                          out.data_ptr(), self._ws.data_ptr(), self._ws.numel(),
                                          self._stream()), "kvm_trapezoid_wh")
        o = out.cpu().tolist()
        return [o[2 * i] + o[2 * i + 1] for i in range(k)]

    def idle_power_upper_median(self, t0: float, t1: float) -> Tuple[Optional[float], N.PowerStatsC]:
        """`sorted(outside)[len(outside)//2]` over non-null samples outside [t0,t1]
        (collector.py:326-333) via the exact select."""
        ids = torch.empty(max(self.m, 1), dtype=torch.uint8, device=self.device)
        st = self.scan(t0, t1, ids)
        if st.n_outside_valid == 0:
            return None, st
        ops = engine.ColumnOps(self.device)
        res = ops.order_stats(self.watts_loaded, [(1, N.RANK_UPPER_MEDIAN, 0.0, 0)],
                              member_mode=N.MEMBER_SET_IDS, member=ids, nsets=1)
        return float(res[0]["value"]), st

    def prefix(self) -> Tuple[torch.Tensor, torch.Tensor]:
        if self._prefix is None:
            p_hi = torch.empty(max(self.m, 1), dtype=torch.float64, device=self.device)
            p_lo = torch.empty(max(self.m, 1), dtype=torch.float64, device=self.device)
            N.check(self.lib.kvm_energy_prefix(self.ts.data_ptr(), self.watts.data_ptr(), self.m,
                                               p_hi.data_ptr(), p_lo.data_ptr(), self._ws.data_ptr(),
                                               self._ws.numel(), self._stream()), "kvm_energy_prefix")
            self._prefix = (p_hi, p_lo)
        return self._prefix

    def interval_energy(self, start_ms: torch.Tensor, latency_ms: torch.Tensor,
                        out: Optional[torch.Tensor] = None, accumulate: bool = False) -> torch.Tensor:
        """Extension A16: Wh of this trace inside every request interval (kvm_interval_energy)."""
        n = int(start_ms.shape[0])
        if out is None:
            out = torch.zeros(n, dtype=torch.float64, device=self.device)
            accumulate = False
        p_hi, p_lo = self.prefix()
        N.check(self.lib.kvm_interval_energy(self.ts.data_ptr(), self.watts.data_ptr(), p_hi.data_ptr(),
                                             p_lo.data_ptr(), self.m, start_ms.data_ptr(), latency_ms.data_ptr(),
                                             n, out.data_ptr(), 1 if accumulate else 0, self._stream()),
                "kvm_interval_energy")
        return out


def interval_energy_multi(traces: Sequence[PowerTrace], start_ms: torch.Tensor, latency_ms: torch.Tensor
                          ) -> torch.Tensor:
    """Per-request Wh summed over several GPU traces in index order (PromQL `sum`, collector.py:45)."""
    n = int(start_ms.shape[0])
    if not traces:
        return torch.zeros(n, dtype=torch.float64, device=start_ms.device)
    lib, dev = traces[0].lib, traces[0].device
    out = torch.zeros(n, dtype=torch.float64, device=dev)
    for b in range(0, len(traces), N.KVM_MAX_TRACES):      # eight traces per pass over the intervals
        chunk = traces[b:b + N.KVM_MAX_TRACES]
        table = (N.PowerTraceC * len(chunk))()
        for k, tr in enumerate(chunk):
            p_hi, p_lo = tr.prefix()
            table[k] = N.PowerTraceC(tr.ts.data_ptr(), tr.watts.data_ptr(), p_hi.data_ptr(), p_lo.data_ptr(), tr.m)
        N.check(lib.kvm_interval_energy_multi(table, len(chunk), start_ms.data_ptr(), latency_ms.data_ptr(), n,
                                              out.data_ptr(), 1 if b else 0, traces[0]._stream()),
                "kvm_interval_energy_multi")
    return out


# --------------------------------------------------------------------------- importable helpers
def read_requests_csv(req_csv: str) -> List[dict]:
    """energy/collector.py:95-111 - its own reader: start/latency default 0, tokens optional,
    status left as read."""
    rows: List[dict] = []
    with open(req_csv, newline="") as f:
        for row in csv.DictReader(f):
            try:
                row["start_ms"] = float(row.get("start_ms", 0) or 0)
                row["latency_ms"] = float(row.get("latency_ms", 0) or 0)
                for tk in ("total_tokens", "completion_tokens"):
                    if tk in row and row[tk] not in (None, "", "NaN"):
                        row[tk] = float(row[tk])
            except Exception:
                pass
            rows.append(row)
    return rows


def _rows_to_log(rows: List[dict]) -> columns.RequestLog:
    """Rows of read_requests_csv -> the columns the scan needs (ttfb/tllt absent)."""
    n = len(rows)
    start = np.empty(n)
    lat = np.empty(n)
    status = np.empty(n, dtype=np.int64)
    ttok = np.zeros(n, dtype=np.int64)
    for i, r in enumerate(rows):
        s, l = r.get("start_ms", 0.0), r.get("latency_ms", 0.0)
        if not isinstance(s, (int, float)) or not isinstance(l, (int, float)):
            raise TypeError(f"unsupported operand type(s) for +: non-numeric start/latency in row {i}")
        start[i], lat[i] = s, l
        status[i] = int(r.get("status", 0))          # collector.py:350: int() of the raw cell
        if status[i] == 200:
            t = r.get("total_tokens", 0.0) or 0.0      # collector.py:352
            if not isinstance(t, (int, float)) or t != t or float(t) != math.floor(t):
                raise TypeError(f"unsupported operand type(s) for +: total_tokens of row {i}")
            ttok[i] = int(t)
    nan = np.full(n, np.nan)
    return columns.columns_from_arrays(start_ms=start, latency_ms=lat, ttfb_ms=nan, tllt_ms=nan, status=status,
                                       completion_tokens=np.zeros(n, dtype=np.int64), total_tokens=ttok)


def _scan_rows(rows: List[dict]):
    log = _rows_to_log(rows)
    analyzer = engine.LogAnalyzer()
    return analyzer.scan(engine.DeviceLog.from_host(log, analyzer.device), ())


def run_window_bounds(rows: List[dict]) -> Tuple[float, float]:
    """Start and end (seconds) of the run (collector.py:114-122); (0,0) without rows."""
    if not rows:
        return (0.0, 0.0)
    st = _scan_rows(rows)
    return st.min_start_ms / 1000.0, st.max_end_ms / 1000.0


def trapezoidal_wh(samples: List[PowerSample], t0: float, t1: float) -> float:
    """Integrate power (W) over [t0, t1] with the trapezoidal rule -> Wh (collector.py:133-149)."""
    if not samples or t1 <= t0:
        return 0.0
    return PowerTrace.from_samples(samples).trapezoid([(t0, t1)])[0]


def _samples_to_columns(handle) -> List[Tuple[np.ndarray, np.ndarray]]:
    lib = N.lib()
    try:
        out = []
        for k in range(int(lib.kvm_samples_series(handle))):
            n = int(lib.kvm_samples_count(handle, k))
            ts, val = np.empty(n, dtype=np.float64), np.empty(n, dtype=np.float64)
            N.check(lib.kvm_samples_copy(handle, k, ts.ctypes.data, val.ctypes.data), "kvm_samples_copy")
            out.append((ts, val))
        return out
    finally:
        lib.kvm_samples_free(handle)


def load_power_columns(path: str, threads: Optional[int] = None) -> Tuple[np.ndarray, np.ndarray]:
    """power.json -> (ts, watts) fp64 columns, NaN = null watts: what `load_power_samples`
    (collector.py:152-177) yields, without one PowerSample object per sample.  The native scanner
    (csrc/kvm_samples.cu) takes the file when it is plain; anything else goes through json."""
    handle = C.c_void_p()
    rc = N.lib().kvm_power_json_parse(os.fsencode(path), threads or min(32, os.cpu_count() or 1), C.byref(handle))
    if rc == N.KVM_OK:
        cols = _samples_to_columns(handle)
        return cols[0] if cols else (np.empty(0), np.empty(0))
    samples = load_power_samples(path)      # raises what the reference raises on malformed input
    ts = np.fromiter((s.ts for s in samples), dtype=np.float64, count=len(samples))
    w = np.fromiter((math.nan if s.watts is None else s.watts for s in samples), dtype=np.float64, count=len(samples))
    return ts, w


def prom_result_columns(body: bytes, threads: Optional[int] = None) -> Optional[List[Tuple[np.ndarray, np.ndarray]]]:
    """A Prometheus query / query_range response body -> [(ts, values)] per series (values: the
    quoted numbers, NaN / +-Inf included), or None when the scanner declines (use json.loads)."""
    handle = C.c_void_p()
    rc = N.lib().kvm_prom_result_parse(body, len(body), threads or min(32, os.cpu_count() or 1), C.byref(handle))
    if rc != N.KVM_OK:
        return None
    return _samples_to_columns(handle)


def load_power_samples(path: str) -> List[PowerSample]:
    """collector.py:152-177: bare list or {"samples": [...]}; `ts_s` accepted for `ts`."""
    with open(path) as f:
        data = json.load(f)
    items = data if isinstance(data, list) else (data["samples"] if isinstance(data, dict) and "samples" in data else [])
    return [PowerSample(ts=float(it.get("ts_s", it.get("ts", 0))),
                        watts=(float(it["watts"]) if it.get("watts") is not None else None)) for it in items]


def write_json(path: str, obj: Dict[str, Any]) -> None:
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "w") as f:
        json.dump(obj, f, indent=2)


def merge_results(run_dir: str, fields: Dict[str, Any]) -> None:
    """collector.py:187-200."""
    path = os.path.join(run_dir, "results.json")
    prev: Dict[str, Any] = {}
    if os.path.exists(path):
        try:
            with open(path) as f:
                prev = json.load(f)
        except Exception:
            prev = {}
    prev.update(fields)
    with open(path, "w") as f:
        json.dump(prev, f, indent=2)


# --------------------------------------------------------------------------- integrate
def integrate_core(stats, trace: Optional[PowerTrace], include_warmup: bool, idle_tax: Optional[str]
                   ) -> Optional[Dict[str, Any]]:
    """energy.json body from the log scan + the device trace (collector.py:289-366).
    None == the reference's `return 2` (invalid window)."""
    if stats is None:
        t0, t1 = 0.0, 0.0
    else:
        t0, t1 = stats.min_start_ms / 1000.0, stats.max_end_ms / 1000.0
    has = trace is not None and trace.m > 0
    if include_warmup and has:
        t0, t1 = trace.ts_min, trace.ts_max
    if t1 <= t0:
        return None
    wh_idle: Optional[float] = None
    if has:
        windows = [(t0, t1)]
        smin, smax = trace.ts_min, trace.ts_max
        if idle_tax == "series":
            # collector.py:315-322; unused windows are given an empty range
            windows.append((smin, min(t0, smax)) if smin < t0 else (0.0, 0.0))
            windows.append((max(smin, t1), smax) if smax > t1 else (0.0, 0.0))
        wh = trace.trapezoid(windows)
        wh_active = wh[0]
        if idle_tax == "series":
            before = wh[1] if smin < t0 else 0.0
            after = wh[2] if smax > t1 else 0.0
            wh_idle = before + after
        elif idle_tax == "baseline":
            p_idle, st = trace.idle_power_upper_median(t0, t1)
            if p_idle is not None:
                if trace.m > 1:
                    avg_step = (trace.step_sum[0] + trace.step_sum[1]) / (trace.m - 1)
                else:
                    avg_step = t1 - t0
                idle_dur_h = (int(st.n_outside) * avg_step) / 3600.0
                wh_idle = p_idle * idle_dur_h
    else:
        wh_active = 0.0
        if idle_tax == "series":
            wh_idle = None
    success = int(stats.n_success) if stats is not None else 0
    total_tokens = float(int(stats.sum_total_tokens)) if success > 0 else 0
    return {
        "Wh_active": wh_active,
        "Wh_idle_tax": wh_idle,
        "Wh_per_1k_tokens_active": ((wh_active / total_tokens) * 1000.0) if total_tokens > 0 else None,
        "Wh_per_request_active": (wh_active / success) if success > 0 else None,
        "window": {"start": t0, "end": t1},
    }


def integrate_energy(args: argparse.Namespace) -> int:
    run_dir = args.run_dir
    req_csv = os.path.join(run_dir, "requests.csv")
    power_path = args.power or os.path.join(run_dir, "power.json")
    energy_out = os.path.join(run_dir, "energy.json")
    if not os.path.exists(req_csv):
        print(f"ERROR: requests.csv not found at {req_csv}", file=sys.stderr)
        return 2
    if not os.path.exists(power_path):
        print(f"WARNING: power.json not found at {power_path}; energy fields will be null", file=sys.stderr)
        write_json(energy_out, {"Wh_active": None, "Wh_idle_tax": None, "Wh_per_1k_tokens_active": None,
                                "Wh_per_request_active": None, "warnings": ["Missing power samples"]})
        if args.merge_results:
            merge_results(run_dir, {"energy_wh_active": None, "energy_wh_idle_tax": None,
                                    "energy_wh_per_1k_tokens": None, "energy_wh_per_request": None})
        return 0

    rows = read_requests_csv(req_csv)
    stats = _scan_rows(rows) if rows else None
    ts_col, w_col = load_power_columns(power_path)      # columnar: no PowerSample object per sample
    trace = PowerTrace(ts_col, w_col) if ts_col.size else None
    energy = integrate_core(stats, trace, args.include_warmup, args.idle_tax)
    if energy is None:
        print("ERROR: Invalid integration window", file=sys.stderr)
        return 2
    energy["samples"] = os.path.basename(power_path)
    write_json(energy_out, energy)
    print(json.dumps(energy, indent=2))
    if args.merge_results:
        merge_results(run_dir, {"energy_wh_active": energy["Wh_active"],
                                "energy_wh_idle_tax": energy["Wh_idle_tax"],
                                "energy_wh_per_1k_tokens": energy["Wh_per_1k_tokens_active"],
                                "energy_wh_per_request": energy["Wh_per_request_active"]})
        print(f"Merged energy fields into {os.path.join(run_dir, 'results.json')}")
    return 0


# --------------------------------------------------------------------------- collect (host I/O)
def _prom_instant(prom_url: str, query: str) -> Optional[float]:
    url = urllib.parse.urljoin(prom_url, "/api/v1/query") + "?" + urllib.parse.urlencode({"query": query})
    try:
        with urllib.request.urlopen(url, timeout=10) as resp:
            result = json.loads(resp.read()).get("data", {}).get("result", [])
        vals = []
        for series in result:
            v = series.get("value", [None, None])[1]
            if v is not None:
                try:
                    vals.append(float(v))
                except Exception:
                    pass
        return sum(vals) / len(vals) if vals else None
    except Exception:
        return None


def collect_power(args: argparse.Namespace) -> int:
    """collector.py:203-251: poll DCGM power through Prometheus, write power.json."""
    if not args.prom_url:
        print("ERROR: --prom-url is required for collect", file=sys.stderr)
        return 2
    pod_re = f"{args.service}-predictor-.*"
    queries = [q.format(ns=args.namespace, pod_re=pod_re) for q in PROM_QUERIES]
    samples: List[PowerSample] = []
    stop = {"flag": False}
    signal.signal(signal.SIGINT, lambda *_: stop.__setitem__("flag", True))
    begin = time.time()
    duration = float(args.duration) if args.duration else 0.0
    interval = max(0.5, float(args.interval or 1.0))
    while True:
        ts = time.time()
        watts = None
        for q in queries:
            watts = _prom_instant(args.prom_url, q)
            if watts is not None:
                break
        samples.append(PowerSample(ts=ts, watts=watts))
        if args.verbose:
            print(json.dumps({"ts_s": ts, "watts": watts}))
        if stop["flag"] or (duration and (ts - begin) >= duration):
            break
        time.sleep(interval)
    if not args.out:
        print("ERROR: --out required to write power samples", file=sys.stderr)
        return 2
    write_json(args.out, {"samples": [s.__dict__ for s in samples], "interval_s": interval})
    print(f"Wrote power samples: {args.out} ({len(samples)} points)")
    return 0


def main(argv: Optional[List[str]] = None) -> int:
    ap = argparse.ArgumentParser(description="Energy collector and integrator")
    sub = ap.add_subparsers(dest="cmd", required=True)
    c = sub.add_parser("collect", help="Collect DCGM power samples via Prometheus")
    c.add_argument("--namespace", required=True)
    c.add_argument("--service", required=True)
    c.add_argument("--prom-url", required=True)
    c.add_argument("--interval", type=float, default=1.0, help="Sampling interval seconds")
    c.add_argument("--duration", type=float, default=0.0, help="Duration seconds (0=until Ctrl-C)")
    c.add_argument("--out", required=True, help="Output path for power samples JSON")
    c.add_argument("--verbose", action="store_true")
    i = sub.add_parser("integrate", help="Integrate Wh aligned to benchmark window")
    i.add_argument("--run-dir", required=True, help="Run directory containing requests.csv")
    i.add_argument("--power", help="Path to power samples JSON (default: run-dir/power.json)")
    i.add_argument("--include-warmup", action="store_true", help="Integrate across full sample span")
    i.add_argument("--idle-tax", choices=["baseline", "series"], default=None)
    i.add_argument("--merge-results", action="store_true", help="Merge energy fields into results.json")
    args = ap.parse_args(argv)
    if args.cmd == "collect":
        return collect_power(args)
    if args.cmd == "integrate":
        return integrate_energy(args)
    return 1
