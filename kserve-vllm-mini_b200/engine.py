"""Device side of the run analysis: columnar log in HBM -> raw statistics.

Host code here only owns buffers (torch tensors), builds the query list and launches the
C ABI (ctypes).  All arithmetic on the records happens in libkvmini_b200.so; all final
scalar expressions (interpolation, ratios, histogram edges) are evaluated by
`report.py` in Python floats, in the reference's order.
"""

from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _native as N
from .columns import F64_COLS, I64_COLS, RequestLog


def _stream_ptr(device: torch.device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def require_cuda() -> torch.device:
    if not torch.cuda.is_available():
        raise N.NativeError("no CUDA device: the run analysis has no CPU path "
                            "(kserve-vllm-mini_b200 runs on sm_100a only)")
    N.lib()
    return torch.device("cuda", torch.cuda.current_device())


# --------------------------------------------------------------------------- device log
@dataclass
class DeviceLog:
    """Seven device-resident columns (fp64 x4, int64 x3) of one request log."""

    cols: Dict[str, torch.Tensor]
    n: int
    device: torch.device

    @staticmethod
    def from_host(log: RequestLog, device: Optional[torch.device] = None,
                  pinned: bool = False) -> "DeviceLog":
        device = device or require_cuda()
        cols = {}
        for k in F64_COLS + I64_COLS:
            t = torch.from_numpy(np.ascontiguousarray(log.cols[k]))
            if pinned:
                t = t.pin_memory()
            cols[k] = t.to(device, non_blocking=pinned)
        return DeviceLog(cols=cols, n=log.n, device=device)

    @staticmethod
    def from_tensors(cols: Dict[str, torch.Tensor]) -> "DeviceLog":
        n = int(cols["status"].shape[0])
        for k in F64_COLS:
            assert cols[k].dtype == torch.float64 and cols[k].is_contiguous() and cols[k].shape[0] == n, k
        for k in I64_COLS:
            assert cols[k].dtype == torch.int64 and cols[k].is_contiguous() and cols[k].shape[0] == n, k
        return DeviceLog(cols=dict(cols), n=n, device=cols["status"].device)

    def c_struct(self) -> N.RequestLogC:
        c = self.cols
        return N.RequestLogC(c["start_ms"].data_ptr(), c["latency_ms"].data_ptr(),
                             c["ttfb_ms"].data_ptr(), c["tllt_ms"].data_ptr(),
                             c["status"].data_ptr(), c["completion_tokens"].data_ptr(),
                             c["total_tokens"].data_ptr(), self.n)


# --------------------------------------------------------------------------- queries
@dataclass(frozen=True)
class Quantile:
    """One percentile() call site: a multiset (union of base sets) and a p."""
    name: str
    set_mask: int
    p: float


LAT_ALL = (1 << N.SET_LAT_COLD) | (1 << N.SET_LAT_WARM)
TTFB_ALL = (1 << N.SET_TTFB_COLD) | (1 << N.SET_TTFB_WARM)

# The 23 distinct (multiset, p) pairs behind analyze.py's 25 percentile() calls
# (ttft_p50/p95 and ttfb_p50/p95 rank the same multiset - SURVEY.md A2).
ANALYZE_QUANTILES: Tuple[Quantile, ...] = (
    Quantile("lat_p50", LAT_ALL, 0.50), Quantile("lat_p95", LAT_ALL, 0.95), Quantile("lat_p99", LAT_ALL, 0.99),
    Quantile("cold_lat_p50", 1 << N.SET_LAT_COLD, 0.50), Quantile("cold_lat_p95", 1 << N.SET_LAT_COLD, 0.95),
    Quantile("cold_lat_p99", 1 << N.SET_LAT_COLD, 0.99),
    Quantile("warm_lat_p50", 1 << N.SET_LAT_WARM, 0.50), Quantile("warm_lat_p95", 1 << N.SET_LAT_WARM, 0.95),
    Quantile("warm_lat_p99", 1 << N.SET_LAT_WARM, 0.99),
    Quantile("ttfb_p50", TTFB_ALL, 0.50), Quantile("ttfb_p95", TTFB_ALL, 0.95), Quantile("ttfb_p99", TTFB_ALL, 0.99),
    Quantile("cold_ttfb_p50", 1 << N.SET_TTFB_COLD, 0.50), Quantile("cold_ttfb_p95", 1 << N.SET_TTFB_COLD, 0.95),
    Quantile("warm_ttfb_p50", 1 << N.SET_TTFB_WARM, 0.50), Quantile("warm_ttfb_p95", 1 << N.SET_TTFB_WARM, 0.95),
    Quantile("tllt_p50", 1 << N.SET_TLLT, 0.50), Quantile("tllt_p95", 1 << N.SET_TLLT, 0.95),
    Quantile("tllt_p99", 1 << N.SET_TLLT, 0.99),
    Quantile("gen_p50", 1 << N.SET_GEN, 0.50), Quantile("gen_p95", 1 << N.SET_GEN, 0.95),
    Quantile("pertok_p50", 1 << N.SET_PERTOK, 0.50), Quantile("pertok_p95", 1 << N.SET_PERTOK, 0.95),
)


@dataclass
class OrderStat:
    """What the device returns for one quantile: the two neighbouring order statistics."""
    n: int
    v_lo: float
    v_hi: float
    status: int = 0

    def interpolate(self, p: float) -> float:
        """analyze.py:65-81 given (n, vals[lo], vals[hi])."""
        if self.n == 0:
            return float("nan")
        if self.n == 1:
            return self.v_lo
        pos = max(0.0, min(p, 1.0)) * (self.n - 1)
        lo = int(math.floor(pos))
        hi = int(math.ceil(pos))
        if lo == hi:
            return self.v_lo
        frac = pos - lo
        return self.v_lo + frac * (self.v_hi - self.v_lo)


def build_queries(quantiles: Sequence[Quantile]) -> "C.Array":
    arr = (N.QueryC * (2 * len(quantiles)))()
    for i, q in enumerate(quantiles):
        arr[2 * i] = N.QueryC(q.set_mask, N.RANK_LERP_LO, q.p, 0)
        arr[2 * i + 1] = N.QueryC(q.set_mask, N.RANK_LERP_HI, q.p, 0)
    return arr


def results_to_orderstats(res: np.ndarray, quantiles: Sequence[Quantile]) -> Dict[str, OrderStat]:
    out = {}
    for i, q in enumerate(quantiles):
        lo, hi = res[2 * i], res[2 * i + 1]
        out[q.name] = OrderStat(int(lo["n"]), float(lo["value"]), float(hi["value"]), int(lo["status"]))
    return out


RESULT_DTYPE = np.dtype([("value", "<f8"), ("n", "<i8"), ("rank", "<i8"), ("status", "<i4"), ("pad", "<i4")])
assert RESULT_DTYPE.itemsize == C.sizeof(N.QueryResultC)


# --------------------------------------------------------------------------- stage timing
class StageTimer:
    """CUDA-event timing of individual launches on the launching stream (bench.py uses it
    for the per-kernel roofline).  `with timer.stage("scan"): launch(...)`."""

    def __init__(self, device: torch.device):
        self.device = device
        self.events: List[Tuple[str, torch.cuda.Event, torch.cuda.Event]] = []

    class _Ctx:
        def __init__(self, owner, name):
            self.owner, self.name = owner, name

        def __enter__(self):
            self.a = torch.cuda.Event(enable_timing=True)
            self.b = torch.cuda.Event(enable_timing=True)
            self.a.record(torch.cuda.current_stream(self.owner.device))

        def __exit__(self, *exc):
            self.b.record(torch.cuda.current_stream(self.owner.device))
            self.owner.events.append((self.name, self.a, self.b))

    def stage(self, name: str):
        return StageTimer._Ctx(self, name)

    def totals_ms(self) -> Dict[str, Tuple[float, int]]:
        """{stage: (total ms, launches)}; call after a synchronize."""
        out: Dict[str, Tuple[float, int]] = {}
        for name, a, b in self.events:
            t, c = out.get(name, (0.0, 0))
            out[name] = (t + a.elapsed_time(b), c + 1)
        return out


class _NoTimer:
    class _Ctx:
        def __enter__(self):
            return None

        def __exit__(self, *exc):
            return None

    _ctx = _Ctx()

    def stage(self, name: str):
        return self._ctx


# --------------------------------------------------------------------------- raw analysis
@dataclass
class RawAnalysis:
    """Everything the device produced for one log; consumed by report.py."""
    stats: N.LogStatsC
    order: Dict[str, OrderStat]
    hist_counts: np.ndarray                 # int64 [4][num_bins]: ttfb, tllt, generation, per-token
    num_bins: int
    series: Optional[np.ndarray] = None     # int64 [3][nb]: requests, success, completion tokens
    series_window_s: float = 1.0
    flags: Optional[torch.Tensor] = None    # uint8 [n] on device (KVM_F_*)
    timings_ms: Dict[str, float] = field(default_factory=dict)

    def key_to_float(self, key: int) -> float:
        bits = key ^ (0x8000000000000000 if (key >> 63) else 0xFFFFFFFFFFFFFFFF)
        return float(np.array([bits], dtype=np.uint64).view(np.float64)[0])

    def col_range(self, col: int) -> Optional[Tuple[float, float]]:
        kmin, kmax = int(self.stats.sets.kmin[col]), int(self.stats.sets.kmax[col])
        if kmin > kmax:
            return None
        # zero bounds are widened to key(-0.0)/key(+0.0) on the device (kvm_log_device.cuh kmin());
        # the reference's min()/max() print 0.0 there: x + 0.0 turns -0.0 into 0.0, nothing else
        return self.key_to_float(kmin) + 0.0, self.key_to_float(kmax) + 0.0


class LogAnalyzer:
    """Reusable buffers + launch sequence for one log (workspaces and output staging are cached,
    so a steady-state call allocates nothing and synchronises once)."""

    # layout of the device "outbox" that comes back in ONE device->host copy
    _OFF_STATS = 0
    _OFF_RESULTS = 512
    _OFF_HIST = _OFF_RESULTS + N.KVM_MAX_QUERIES * 32
    _OFF_DROPPED = _OFF_HIST + N.KVM_MAX_HISTS * N.KVM_HIST_MAX_BINS * 8
    _OFF_STATUS = _OFF_DROPPED + 8
    _OUTBOX = _OFF_DROPPED + 64

    REFINE_MIN_ROWS = 1 << 25   # KVM_REFINE_MIN_ROWS
    REFINE_MIN_KEYS = 1 << 19   # KVM_REFINE_MIN_KEYS

    def __init__(self, device: Optional[torch.device] = None, stream: Optional[torch.cuda.Stream] = None):
        self.device = device or require_cuda()
        self.lib = N.lib()
        # own stream (optional): lets two analyzers overlap on the GPU - the single-CTA planning /
        # resolving kernels of one analysis run beside the streaming passes of the other
        self.stream = stream
        assert C.sizeof(N.LogStatsC) <= self._OFF_RESULTS and C.sizeof(N.QueryResultC) == 32
        self._scan_ws = torch.empty(max(int(self.lib.kvm_log_scan_workspace_bytes()),
                                        int(self.lib.kvm_log_analyze_fused_scan_workspace_bytes())),
                                    dtype=torch.uint8, device=self.device)
        self._out_d = torch.zeros(self._OUTBOX, dtype=torch.uint8, device=self.device)
        self._out_h = torch.zeros(self._OUTBOX, dtype=torch.uint8).pin_memory()
        base = self._out_d.data_ptr()
        self._stats_ptr = base + self._OFF_STATS
        self._results_ptr = base + self._OFF_RESULTS
        self._hist_ptr = base + self._OFF_HIST
        self._dropped_ptr = base + self._OFF_DROPPED
        # views used by subclasses / tests
        self._stats_d = self._out_d[self._OFF_STATS:self._OFF_STATS + C.sizeof(N.LogStatsC)]
        self._results_d = self._out_d[self._OFF_RESULTS:self._OFF_HIST]
        self._hist_ws = torch.empty(self.lib.kvm_log_histograms_workspace_bytes(256), dtype=torch.uint8, device=self.device)
        self._sel_ws: Optional[torch.Tensor] = None
        self._sel_cap = 0
        self._flags: Optional[torch.Tensor] = None
        self._empty_cold = torch.zeros(1, dtype=torch.float64, device=self.device)
        self._cold_key: Optional[tuple] = None
        self._cold_d = self._empty_cold
        self._series_d: Optional[torch.Tensor] = None
        self._series_h: Optional[torch.Tensor] = None
        self._series_cap = 1 << 16
        self._query_cache: Dict[int, tuple] = {}
        self._graphs: Dict[tuple, tuple] = {}
        self._done = torch.cuda.Event()
        self.launches = 0   # kernels launched by the last analyze() (for bench accounting)
        self.timer = _NoTimer()   # bench.py swaps in a StageTimer
        # one-pass form (kvm_log_fused_pass: scan + histograms in one kernel) whenever the columns
        # allow it (32-byte aligned); False: always the two-pass form (scan, then histogram pass)
        self.use_fused = True
        self.last_path = ""
        self.refine_rounds: Optional[int] = None            # None: by log size (_refine_rounds)
        self.refine_min_keys: int = self.REFINE_MIN_KEYS    # a round works only above this many candidates

    # -- buffers
    def _ensure(self, n: int, arena: Optional[int] = None) -> None:
        if self._flags is None or self._flags.shape[0] < n + 2:
            self._flags = torch.empty(n + 2, dtype=torch.uint8, device=self.device)
            self._digits = torch.empty(n + 2, dtype=torch.int64, device=self.device)   # 8 B/row scratch
        cap = arena if arena is not None else int(self.lib.kvm_select_default_arena(n))
        if self._sel_ws is None or self._sel_cap < cap:
            self._sel_ws = torch.empty(self.lib.kvm_select_workspace_bytes(cap), dtype=torch.uint8, device=self.device)
            self._sel_cap = cap

    def close(self) -> None:
        """Release the captured CUDA graphs (and what they pin).  A ShardedAnalyzer must be closed
        before torch.distributed.destroy_process_group(): tearing the communicator down under
        live graphs that contain its collectives hangs."""
        self._graphs.clear()
        torch.cuda.synchronize(self.device)

    def _refine_rounds(self, n_rows: int) -> int:
        """Refine rounds to launch (kvm_select_plan_refine): `refine_rounds` when set, else one
        round for logs of REFINE_MIN_ROWS rows and more (include/kvmini_b200.h)."""
        if self.refine_rounds is not None:
            return int(self.refine_rounds)
        return 1 if n_rows >= self.REFINE_MIN_ROWS else 0

    def _fused_ok(self, dlog: "DeviceLog", fused_hists: bool) -> bool:
        """The one-pass form needs 16-byte aligned columns and the histograms folded into the select."""
        if not (self.use_fused and fused_hists and dlog.n >= 1 and not getattr(self, "_fused_off", False)):
            return False
        if (dlog.cols["start_ms"].data_ptr(), dlog.n) in getattr(self, "_unordered_logs", ()):
            return False   # seen before: rows out of time order, the series needs the exact minimum first
        return all(dlog.cols[k].data_ptr() % 16 == 0 for k in F64_COLS + I64_COLS)

    def _cold_device(self, cold_times: Sequence[float]) -> Tuple[torch.Tensor, int]:
        key = tuple(sorted(float(c) for c in cold_times))
        if key != self._cold_key:
            self._cold_d = (torch.tensor(key, dtype=torch.float64, device=self.device) if key else self._empty_cold)
            self._cold_key = key
        return self._cold_d, len(key)

    def _queries(self, quantiles: Sequence[Quantile]):
        hit = self._query_cache.get(id(quantiles))
        if hit is None or hit[0] is not quantiles:
            hit = (quantiles, build_queries(quantiles))
            self._query_cache[id(quantiles)] = hit
        return hit[1]

    def _read_outbox(self) -> bytes:
        """One D2H copy of stats + results + histogram counts + dropped counter, then sync."""
        self._out_h.copy_(self._out_d, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return self._out_h.numpy().tobytes()

    def _launch_scan(self, dlog: DeviceLog, cold_times: Sequence[float], cold_window_s: float = 30.0,
                     cold_mask: Optional[torch.Tensor] = None) -> None:
        """K1, stream-ordered: stats land in the outbox, flags in self._flags."""
        self._ensure(dlog.n)
        cold_d, nc = self._cold_device(cold_times)
        log_c = dlog.c_struct()
        with self.timer.stage("log_scan"):
            N.check(self.lib.kvm_log_scan(C.byref(log_c), cold_d.data_ptr(), nc, float(cold_window_s),
                                          cold_mask.data_ptr() if cold_mask is not None else None,
                                          self._flags.data_ptr(), self._stats_ptr,
                                          self._scan_ws.data_ptr(), self._scan_ws.numel(),
                                          _stream_ptr(self.device)), "kvm_log_scan")
        self.launches += 1 + (1 if nc and cold_mask is None else 0)

    def scan(self, dlog: DeviceLog, cold_times: Sequence[float], cold_window_s: float = 30.0,
             cold_mask: Optional[torch.Tensor] = None) -> N.LogStatsC:
        """K1: one pass, returns the stats struct (synchronises).  `cold_mask` (uint8 per row)
        replaces the time-based classification (cost_estimator's is_cold_start column)."""
        self._launch_scan(dlog, cold_times, cold_window_s, cold_mask)
        return N.LogStatsC.from_buffer_copy(self._read_outbox()[:C.sizeof(N.LogStatsC)])

    @staticmethod
    def hist_spec(num_bins: int) -> N.HistSpecC:
        """The four histograms of analyze.py:149-152 as a kvm_hist_spec."""
        hs = N.HistSpecC()
        hs.nhist, hs.num_bins = 4, num_bins
        for i, (col, mask) in enumerate(((N.COL_TTFB, TTFB_ALL), (N.COL_TLLT, 1 << N.SET_TLLT),
                                         (N.COL_GEN, 1 << N.SET_GEN), (N.COL_PERTOK, 1 << N.SET_PERTOK))):
            hs.col[i], hs.set_mask[i] = col, mask
        return hs

    def _enqueue(self, dlog: DeviceLog, cold_times: Sequence[float], num_bins: int,
                 series_window_s: Optional[float], quantiles: Sequence[Quantile], arena: int,
                 hist_big: Optional[torch.Tensor]) -> None:
        """Everything of one analysis, stream-ordered, ending with the outbox copy (no sync).
        This is the sequence a CUDA graph captures."""
        lib, tm, dev = self.lib, self.timer, self.device
        queries = self._queries(quantiles)
        nq = len(queries)
        fused_hists = hist_big is None
        hspec = self.hist_spec(num_bins) if fused_hists else None
        log_c = dlog.c_struct()
        st = _stream_ptr(dev)
        fused = self._fused_ok(dlog, fused_hists)
        if not fused:
            self._launch_scan(dlog, cold_times)
        extras = N.LogExtrasC()
        if series_window_s is not None and dlog.n > 0:
            extras.series_window_ms = float(series_window_s) * 1000.0
            extras.series_max_bins = int(self._series_d.shape[1])
            extras.series = self._series_d.data_ptr()
            extras.series_dropped = self._dropped_ptr
        extras.digit_records = self._digits.data_ptr()
        ws, fl = self._sel_ws.data_ptr(), self._flags.data_ptr()
        rounds = self._refine_rounds(dlog.n)
        self.last_path = "fused" if fused else "two-pass"
        if fused:
            cold_d, nc = self._cold_device(cold_times)
            with tm.stage("log_fused"):
                N.check(lib.kvm_log_fused_pass(C.byref(log_c), cold_d.data_ptr(), nc, 30.0, fl, self._stats_ptr,
                                               C.byref(extras), self._scan_ws.data_ptr(), self._scan_ws.numel(),
                                               ws, self._sel_ws.numel(), arena, st), "kvm_log_fused_pass")
            self.launches += 1 + (1 if nc else 0)
            with tm.stage("select_plan"):
                N.check(lib.kvm_select_plan_preset(self._stats_ptr + N.LogStatsC.sets.offset, queries, nq,
                                                   C.byref(hspec) if hspec is not None else None, ws,
                                                   self._sel_ws.numel(), arena, rounds, int(self.refine_min_keys), st),
                        "kvm_select_plan_preset")
            self.launches -= 1   # (no separate histogram pass; the common count below assumes one)
        else:
            with tm.stage("select_plan"):
                N.check(lib.kvm_select_plan_refine(self._stats_ptr + N.LogStatsC.sets.offset, queries, nq,
                                                   C.byref(hspec) if hspec is not None else None, ws,
                                                   self._sel_ws.numel(), arena, rounds, int(self.refine_min_keys), st),
                        "kvm_select_plan")
            with tm.stage("select_hist"):
                N.check(lib.kvm_log_select_hist(C.byref(log_c), fl, self._stats_ptr, C.byref(extras), ws, st),
                        "kvm_log_select_hist")
        with tm.stage("select_resolve"):
            N.check(lib.kvm_select_resolve(ws, st), "kvm_select_resolve")
        for _ in range(rounds):   # large logs: narrow every candidate bucket 2048x before compacting
            with tm.stage("select_refine_hist"):
                N.check(lib.kvm_log_select_refine_hist(C.byref(log_c), fl, self._digits.data_ptr(), ws, st),
                        "kvm_log_select_refine_hist")
            with tm.stage("select_refine"):
                N.check(lib.kvm_select_refine(ws, st), "kvm_select_refine")
            self.launches += 2
        with tm.stage("select_compact"):
            N.check(lib.kvm_log_select_compact(C.byref(log_c), fl, self._digits.data_ptr(), ws, st),
                    "kvm_log_select_compact")
        with tm.stage("select_finish"):
            N.check(lib.kvm_select_finish(ws, None, 0, None, self._results_ptr,
                                          self._hist_ptr if fused_hists else None, st), "kvm_select_finish")
        self.launches += 5
        if not fused_hists:
            with tm.stage("histograms"):
                N.check(lib.kvm_log_histograms(C.byref(log_c), fl, self._stats_ptr, num_bins,
                                               hist_big.data_ptr(), self._hist_ws.data_ptr(),
                                               self._hist_ws.numel(), st), "kvm_log_histograms")
            self.launches += 2
        # select status word -> outbox, then the whole outbox -> pinned host memory
        self._out_d[self._OFF_STATUS:self._OFF_STATUS + 4].copy_(self._sel_ws[0:4], non_blocking=True)
        self._out_h.copy_(self._out_d, non_blocking=True)
        if series_window_s is not None and dlog.n > 0:
            p = self._series_h.shape[1]
            self._series_h.copy_(self._series_d[:, :p], non_blocking=True)

    SERIES_PREFIX = 8192   # windows read back together with the outbox (longer series: one more copy)

    def _submit(self, dlog: DeviceLog, cold_times: Sequence[float], num_bins: int,
                series_window_s: Optional[float], quantiles: Sequence[Quantile], arena: int,
                hist_big: Optional[torch.Tensor], use_graph: bool) -> None:
        """Put one analysis on the stream (graph replay when possible) followed by the read-back
        of the series prefix and the completion event."""
        dev = self.device
        want_series = series_window_s is not None
        self.launches = 0
        if use_graph and isinstance(self.timer, _NoTimer):
            graph_key = (tuple(dlog.cols[k].data_ptr() for k in F64_COLS + I64_COLS), dlog.n,
                         tuple(sorted(float(c) for c in cold_times)), num_bins, series_window_s,
                         id(quantiles), arena, self._sel_ws.data_ptr(), self._flags.data_ptr(),
                         self._series_d.data_ptr() if want_series else 0,
                         self._refine_rounds(dlog.n), self.refine_min_keys,
                         self._fused_ok(dlog, hist_big is None))
            entry = self._graphs.get(graph_key)
            if entry is None:
                # eager run first (sets function attributes, fills caches), then capture
                self._enqueue(dlog, cold_times, num_bins, series_window_s, quantiles, arena, hist_big)
                torch.cuda.current_stream(dev).synchronize()
                g = torch.cuda.CUDAGraph()
                self.launches = 0
                with torch.cuda.graph(g):
                    self._enqueue(dlog, cold_times, num_bins, series_window_s, quantiles, arena, hist_big)
                if len(self._graphs) >= 4:
                    self._graphs.pop(next(iter(self._graphs)))
                entry = (g, self.launches, self.last_path)
                self._graphs[graph_key] = entry
            entry[0].replay()
            self.launches = entry[1]
            self.last_path = entry[2]
        else:
            self._enqueue(dlog, cold_times, num_bins, series_window_s, quantiles, arena, hist_big)
        self._done.record(torch.cuda.current_stream(dev))

    def analyze_async(self, dlog: DeviceLog, cold_times: Sequence[float] = (), num_bins: int = 20,
                      series_window_s: Optional[float] = 1.0, max_series_bins: int = 1 << 24,
                      quantiles: Sequence[Quantile] = ANALYZE_QUANTILES, use_graph: bool = False):
        """Submit an analysis and return a zero-argument callable that waits for it and builds
        the RawAnalysis.  Lets a caller post-process analysis k on the host while analysis k+1
        (on another LogAnalyzer) runs on the GPU.  One submission per analyzer at a time."""
        dev = self.device
        nq = 2 * len(quantiles)
        assert nq <= N.KVM_MAX_QUERIES
        fused_hists = num_bins <= N.KVM_HIST_MAX_BINS
        hist_big = None if fused_hists else torch.zeros((4, num_bins), dtype=torch.int64, device=dev)
        want_series = series_window_s is not None and dlog.n > 0
        win = series_window_s if want_series else None
        self._ensure(dlog.n)
        state = {"arena": self._sel_cap}

        def submit():
            if want_series and (self._series_d is None or self._series_d.shape[1] < self._series_cap):
                self._series_d = torch.empty((3, self._series_cap), dtype=torch.int64, device=dev)
                self._series_h = torch.empty((3, min(self._series_cap, self.SERIES_PREFIX)),
                                             dtype=torch.int64).pin_memory()
            if self.stream is None:
                self._submit(dlog, cold_times, num_bins, win, quantiles, state["arena"], hist_big, use_graph)
            else:
                self.stream.wait_stream(torch.cuda.current_stream(dev))   # inputs were produced there
                with torch.cuda.stream(self.stream):
                    self._submit(dlog, cold_times, num_bins, win, quantiles, state["arena"], hist_big, use_graph)

        submit()

        def result() -> RawAnalysis:
            while True:
                self._done.synchronize()
                out = self._out_h.numpy().tobytes()
                rc = int(np.frombuffer(out, dtype=np.int32, count=1, offset=self._OFF_STATUS)[0])
                stats = N.LogStatsC.from_buffer_copy(out[:C.sizeof(N.LogStatsC)])
                if rc == N.KVM_ERR_ARENA:
                    # pathological concentration: retry with room for every key of every query
                    state["arena"] = max(2 * state["arena"], dlog.n * 4 + 1024)
                    self._ensure(dlog.n, state["arena"])
                    submit()
                    continue
                if rc != N.KVM_OK:
                    raise N.NativeError(f"select failed with status {rc}")
                series = None
                if (want_series and self.last_path == "fused" and not math.isnan(stats.min_start_ms)
                        and np.frombuffer(out, dtype=np.float64, count=1, offset=self._OFF_DROPPED + 16)[0]
                        != stats.min_start_ms):
                    # the one-pass kernel binned the series relative to the start of its sample, which
                    # is not the log's minimum (rows out of time order): redo with the two-pass form
                    if not hasattr(self, "_unordered_logs"):
                        self._unordered_logs = set()
                    if len(self._unordered_logs) < 64:
                        self._unordered_logs.add((dlog.cols["start_ms"].data_ptr(), dlog.n))
                    self._fused_off = True
                    try:
                        submit()
                    finally:
                        self._fused_off = False
                    continue
                if want_series and not math.isnan(stats.min_start_ms):
                    t0, t1 = stats.min_start_ms / 1000.0, stats.max_end_ms / 1000.0
                    nb = max(1, int(math.ceil((t1 - t0) / series_window_s)) + 1)
                    dropped = int(np.frombuffer(out, dtype=np.int64, count=1, offset=self._OFF_DROPPED)[0])
                    if dropped != 0:
                        if nb > max_series_bins:
                            raise N.NativeError("window series needs more than max_series_bins windows")
                        self._series_cap = nb + 1   # now known exactly: redo with room for every window
                        submit()
                        continue
                    nb = min(nb, self._series_d.shape[1])
                    if nb <= self._series_h.shape[1]:
                        series = self._series_h[:, :nb].numpy().copy()   # came back with the outbox
                    else:
                        series = self._series_d[:, :nb].cpu().numpy()
                break
            res = np.frombuffer(out, dtype=RESULT_DTYPE, count=nq, offset=self._OFF_RESULTS)
            if fused_hists:
                hist = np.frombuffer(out, dtype=np.int64, count=4 * num_bins,
                                     offset=self._OFF_HIST).reshape(4, num_bins).copy()
            else:
                hist = hist_big.cpu().numpy()
            return RawAnalysis(stats=stats, order=results_to_orderstats(res, quantiles), hist_counts=hist,
                               num_bins=num_bins, series=series, series_window_s=series_window_s or 1.0,
                               flags=self._flags[:dlog.n])

        return result

    def analyze(self, dlog: DeviceLog, cold_times: Sequence[float] = (), num_bins: int = 20,
                series_window_s: Optional[float] = 1.0, max_series_bins: int = 1 << 24,
                quantiles: Sequence[Quantile] = ANALYZE_QUANTILES, use_graph: bool = False) -> RawAnalysis:
        """scan -> plan -> hist(+series) -> resolve -> compact -> finish, all stream-ordered, ONE
        host synchronisation.  use_graph=True replays a captured CUDA graph when the same device
        buffers are analysed again (steady-state monitoring, benchmarks)."""
        return self.analyze_async(dlog, cold_times, num_bins, series_window_s, max_series_bins,
                                  quantiles, use_graph)()


# --------------------------------------------------------------------------- single columns
class ColumnOps:
    """Generic one-column entry points behind `percentile()` / `compute_histograms()`."""

    def __init__(self, device: Optional[torch.device] = None):
        self.device = device or require_cuda()
        self.lib = N.lib()

    _P_KINDS = (N.RANK_LERP_LO, N.RANK_LERP_HI, N.RANK_UPPER_MEDIAN, N.RANK_FLOOR_PN, N.RANK_FLOOR_PNM1)

    def _sampled(self, values, n, arr, queries, member_mode, member, st) -> Optional[np.ndarray]:
        """The one-pass path (kvm_column_select_sampled); None when it does not apply or declines."""
        nq = len(queries)
        if (member_mode not in (N.MEMBER_ALL, N.MEMBER_STATUS) or nq > 8 or n >= 1 << 31
                or n < int(self.lib.kvm_column_select_sampled_min_n())
                or any(q[0] != 1 or q[1] not in self._P_KINDS for q in queries)
                ):
            return None
        fn = self.lib.kvm_column_select_sampled
        need = int(self.lib.kvm_column_select_sampled_workspace_bytes(n))
        if getattr(self, "_qs_ws", None) is None or self._qs_ws.numel() < need:
            self._qs_ws = torch.empty(need, dtype=torch.uint8, device=self.device)
        ws = self._qs_ws
        rc = fn(values.data_ptr(), n, member_mode,
                                                member.data_ptr() if member is not None else None, arr, nq,
                                                self._res_d.data_ptr(), ws.data_ptr(), ws.numel(), st)
        if rc == N.KVM_ERR_UNSUPPORTED:
            return None
        N.check(rc, "kvm_column_select_sampled")
        status_off = N.KVM_MAX_QUERIES * C.sizeof(N.QueryResultC)
        self._res_d[status_off:status_off + 4].copy_(ws[0:4], non_blocking=True)
        self._res_h.copy_(self._res_d, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        out = self._res_h.numpy().tobytes()
        if int(np.frombuffer(out, dtype=np.int32, count=1, offset=status_off)[0]) != N.KVM_OK:
            return None    # the sample missed the range / duplicates: the general select decides
        self.last_path = "sampled"
        return np.frombuffer(out, dtype=RESULT_DTYPE, count=nq).copy()

    def order_stats(self, values: torch.Tensor, queries: Sequence[Tuple[int, int, float, int]],
                    member_mode: int = N.MEMBER_ALL, member: Optional[torch.Tensor] = None,
                    nsets: int = 1, sampled: bool = True) -> np.ndarray:
        """queries: (set_mask, rank_kind, p, rank).  Returns structured array of results.
        Workspace and read-back buffers are cached; one synchronisation per call.  Few quantiles of one
        multiset go through the sampled-range digit select first (csrc/kvm_qdigit.cu: exact, or it
        declines and the general select below runs)."""
        assert values.dtype == torch.float64 and values.is_contiguous()
        n = int(values.shape[0])
        nq = len(queries)
        arr = (N.QueryC * nq)(*[N.QueryC(*q) for q in queries])
        st = _stream_ptr(self.device)
        if getattr(self, "_res_d", None) is None:
            self._res_d = torch.zeros(N.KVM_MAX_QUERIES * C.sizeof(N.QueryResultC) + 64, dtype=torch.uint8,
                                      device=self.device)
            self._res_h = torch.zeros_like(self._res_d, device="cpu").pin_memory()
        self.last_path = "general"
        if sampled and nsets == 1:
            got = self._sampled(values, n, arr, queries, member_mode, member, st)
            if got is not None:
                return got
        arena = max(int(self.lib.kvm_select_default_arena(n)), getattr(self, "_arena", 0))
        if getattr(self, "_res_d", None) is None:
            self._res_d = torch.zeros(N.KVM_MAX_QUERIES * C.sizeof(N.QueryResultC) + 64, dtype=torch.uint8,
                                      device=self.device)
            self._res_h = torch.zeros_like(self._res_d, device="cpu").pin_memory()
        status_off = N.KVM_MAX_QUERIES * C.sizeof(N.QueryResultC)
        while True:
            if getattr(self, "_ws", None) is None or self._arena < arena:
                self._ws = torch.empty(self.lib.kvm_select_workspace_bytes(arena), dtype=torch.uint8, device=self.device)
                self._arena = arena
            ws = self._ws
            N.check(self.lib.kvm_column_select(values.data_ptr(), n, member_mode,
                                               member.data_ptr() if member is not None else None, nsets,
                                               arr, nq, self._res_d.data_ptr(), ws.data_ptr(), ws.numel(),
                                               self._arena, st), "kvm_column_select")
            self._res_d[status_off:status_off + 4].copy_(ws[0:4], non_blocking=True)
            self._res_h.copy_(self._res_d, non_blocking=True)
            torch.cuda.current_stream(self.device).synchronize()
            out = self._res_h.numpy().tobytes()
            rc = int(np.frombuffer(out, dtype=np.int32, count=1, offset=status_off)[0])
            if rc == N.KVM_ERR_ARENA:
                arena = max(2 * arena, n * nq + 1024)
                continue
            if rc != N.KVM_OK:
                raise N.NativeError(f"kvm_column_select failed with status {rc}")
            break
        return np.frombuffer(out, dtype=RESULT_DTYPE, count=nq).copy()

    def quantile_stat(self, values: torch.Tensor, p: float, **kw) -> OrderStat:
        r = self.order_stats(values, [(1, N.RANK_LERP_LO, p, 0), (1, N.RANK_LERP_HI, p, 0)], **kw)
        return OrderStat(int(r[0]["n"]), float(r[0]["value"]), float(r[1]["value"]), int(r[0]["status"]))

    def nearest(self, values: torch.Tensor, target: float, member_mode: int = N.MEMBER_ALL,
                member: Optional[torch.Tensor] = None) -> Tuple[int, float]:
        """(index, |value - target|) of the member element closest to `target` (lowest index on
        ties, NaN never wins; (-1, nan) when there is none).  report_generator.py:437-454."""
        assert values.dtype == torch.float64 and values.is_contiguous()
        if getattr(self, "_near_ws", None) is None:
            self._near_ws = torch.zeros(self.lib.kvm_column_nearest_workspace_bytes(), dtype=torch.uint8,
                                        device=self.device)
            self._near_d = torch.zeros(16, dtype=torch.uint8, device=self.device)
        N.check(self.lib.kvm_column_nearest(values.data_ptr(), int(values.shape[0]), member_mode,
                                            member.data_ptr() if member is not None else None, float(target),
                                            self._near_d.data_ptr(), self._near_ws.data_ptr(), self._near_ws.numel(),
                                            _stream_ptr(self.device)), "kvm_column_nearest")
        raw = self._near_d.cpu().numpy().tobytes()
        return int(np.frombuffer(raw, dtype=np.int64, count=1)[0]), float(np.frombuffer(raw, dtype=np.float64, count=1, offset=8)[0])

    def minmax_count(self, values: torch.Tensor) -> Tuple[int, float, float]:
        """count / min / max over non-NaN values via kvm_column_stats."""
        n = int(values.shape[0])
        info_d = torch.zeros(C.sizeof(N.SetInfoC), dtype=torch.uint8, device=self.device)
        ws = torch.empty(int(self.lib.kvm_column_stats_workspace_bytes()), dtype=torch.uint8, device=self.device)
        N.check(self.lib.kvm_column_stats(values.data_ptr(), n, N.MEMBER_ALL, None, 1, info_d.data_ptr(),
                                          ws.data_ptr(), ws.numel(), _stream_ptr(self.device)), "kvm_column_stats")
        info = N.SetInfoC.from_buffer_copy(info_d.cpu().numpy().tobytes())
        cnt = int(info.count[0])
        if cnt == 0:
            return 0, float("nan"), float("nan")
        def unkey(k):
            bits = k ^ (0x8000000000000000 if (k >> 63) else 0xFFFFFFFFFFFFFFFF)
            return float(np.array([bits], dtype=np.uint64).view(np.float64)[0])
        # a zero bound travels as key(-0.0) / key(+0.0) (it covers both signed zeros for the digit
        # planner); Python's min()/max() of such a list prints 0.0 unless every zero is negative -
        # which no producer of these logs writes - so report +0.0 (x + 0.0 maps -0.0 to 0.0)
        return cnt, unkey(int(info.kmin[0])) + 0.0, unkey(int(info.kmax[0])) + 0.0

    def histogram_counts(self, values: torch.Tensor, lo: float, hi: float, num_bins: int) -> List[int]:
        counts = torch.empty(num_bins, dtype=torch.int64, device=self.device)
        ws = torch.empty(self.lib.kvm_log_histograms_workspace_bytes(num_bins), dtype=torch.uint8, device=self.device)
        N.check(self.lib.kvm_column_histogram(values.data_ptr(), int(values.shape[0]), lo, hi, num_bins,
                                              counts.data_ptr(), ws.data_ptr(), ws.numel(),
                                              _stream_ptr(self.device)), "kvm_column_histogram")
        return counts.cpu().tolist()
