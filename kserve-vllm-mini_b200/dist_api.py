"""Multi-GPU run analysis: one process per GPU, the log sharded by time range.

SURVEY.md section 8(e).  Rank r holds a contiguous time slice of the log.  Nothing on the
data path waits on another rank inside a kernel; the host exchanges small tensors between
the stages with torch.distributed (NCCL on GPUs, gloo in the CPU tests):

  K1 partials   int64 sums  -> all_reduce(SUM)      key minima / maxima -> all_reduce(MIN/MAX)
                fp sums (double-double)             -> all_gather, added in RANK ORDER on the host
  quantiles     per-set digit histograms (7 x 2048 u64) -> all_reduce(SUM); every rank then
                resolves identically, compacts its own candidates, all_gathers the candidate
                arenas and finishes identically  => exact global order statistics
  histograms / series   int64 bin counts -> all_reduce(SUM)

All integer reductions are order-independent, the fp ones have a fixed order, so every rank
ends with bit-identical results.
"""

from __future__ import annotations

import ctypes as C
import math
from fractions import Fraction
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch
import torch.distributed as dist

from . import _native as N
from . import engine

_SIGN = -(1 << 63)

# int64 fields of kvm_log_stats that add up across shards
_SUM_FIELDS = ("n_total", "n_success", "n_cold", "n_cold_success", "sum_completion_tokens",
               "sum_total_tokens", "sum_total_tokens_cold", "n_gen_rows")


def _key_to_signed(k: int) -> int:
    """u64 ordered key -> int64 with the same order (torch has no uint64 reductions)."""
    v = k ^ (1 << 63)
    return v - (1 << 64) if v >= (1 << 63) else v


def _signed_to_key(v: int) -> int:
    return (v + (1 << 64) if v < 0 else v) ^ (1 << 63)


def _exact_dd_total(pairs: Sequence[Sequence[float]]):
    """Sum of double-double partials in rank order, exactly (rational), back to (hi, lo)."""
    tot = Fraction(0)
    for hi, lo in pairs:
        tot += Fraction(hi) + Fraction(lo)
    hi = float(tot)
    lo = float(tot - Fraction(hi))
    return hi, lo


def combine_stats(local: N.LogStatsC, device: torch.device, group=None) -> N.LogStatsC:
    """Global kvm_log_stats from every rank's local one.  Works on any backend (the CPU tests
    drive it over gloo with world_size 2)."""
    world = dist.get_world_size(group)
    sums = torch.tensor([getattr(local, f) for f in _SUM_FIELDS] + list(local.sets.count),
                        dtype=torch.int64, device=device)
    dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group)
    kmin = torch.tensor([_key_to_signed(int(k)) for k in local.sets.kmin], dtype=torch.int64, device=device)
    kmax = torch.tensor([_key_to_signed(int(k)) for k in local.sets.kmax], dtype=torch.int64, device=device)
    dist.all_reduce(kmin, op=dist.ReduceOp.MIN, group=group)
    dist.all_reduce(kmax, op=dist.ReduceOp.MAX, group=group)
    inf = math.inf
    lo = torch.tensor([inf if math.isnan(local.min_start_ms) else local.min_start_ms], dtype=torch.float64, device=device)
    hi = torch.tensor([-inf if math.isnan(local.max_end_ms) else local.max_end_ms], dtype=torch.float64, device=device)
    dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=group)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=group)
    dd_local = torch.tensor([local.ttfb_sum_hi, local.ttfb_sum_lo, local.lat_sum_hi, local.lat_sum_lo],
                            dtype=torch.float64, device=device)
    dd_all = torch.empty(world * 4, dtype=torch.float64, device=device)
    dist.all_gather_into_tensor(dd_all, dd_local, group=group) if device.type == "cuda" else \
        dist.all_gather(list(dd_all.view(world, 4).unbind(0)), dd_local, group=group)
    dd = dd_all.view(world, 4).cpu().tolist()

    out = N.LogStatsC.from_buffer_copy(bytes(local))
    s = sums.cpu().tolist()
    for i, f in enumerate(_SUM_FIELDS):
        setattr(out, f, s[i])
    for b in range(N.KVM_MAX_SETS):
        out.sets.count[b] = s[len(_SUM_FIELDS) + b]
    for c, (a, b) in enumerate(zip(kmin.cpu().tolist(), kmax.cpu().tolist())):
        out.sets.kmin[c] = _signed_to_key(a)
        out.sets.kmax[c] = _signed_to_key(b)
    lo_v, hi_v = float(lo.item()), float(hi.item())
    out.min_start_ms = math.nan if lo_v == inf else lo_v
    out.max_end_ms = math.nan if hi_v == -inf else hi_v
    out.ttfb_sum_hi, out.ttfb_sum_lo = _exact_dd_total([(r[0], r[1]) for r in dd])
    out.lat_sum_hi, out.lat_sum_lo = _exact_dd_total([(r[2], r[3]) for r in dd])
    return out


class ShardedAnalyzer(engine.LogAnalyzer):
    """LogAnalyzer whose analyze() returns GLOBAL results for a log sharded over the ranks of
    the default process group (one rank per GPU)."""

    def __init__(self, device: Optional[torch.device] = None, group=None):
        super().__init__(device)
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self._bins = int(self.lib.kvm_select_bins())

    def _wrap(self, ptr: int, count: int, dtype=torch.int64) -> torch.Tensor:
        """A tensor view over `count` 8-byte words inside the select workspace."""
        off = ptr - self._sel_ws.data_ptr()
        return self._sel_ws[off: off + 8 * count].view(dtype)

    def analyze(self, dlog: engine.DeviceLog, cold_times: Sequence[float] = (), num_bins: int = 20,
                series_window_s: Optional[float] = 1.0, max_series_bins: int = 1 << 24,
                quantiles: Sequence[engine.Quantile] = engine.ANALYZE_QUANTILES) -> engine.RawAnalysis:
        self.launches = 0
        lib, tm, dev, grp = self.lib, self.timer, self.device, self.group
        local = self.scan(dlog, cold_times)
        stats = combine_stats(local, dev, grp)
        # the global stats drive plan / histogram thresholds / series origin on every rank
        self._stats_d.copy_(torch.frombuffer(bytearray(bytes(stats)), dtype=torch.uint8))
        st = engine._stream_ptr(dev)
        log_c = dlog.c_struct()
        queries = engine.build_queries(quantiles)
        nq = len(queries)
        n_global = int(stats.n_total)

        assert num_bins <= N.KVM_HIST_MAX_BINS, "sharded analysis folds the histograms into the select"
        hist_d = torch.zeros(N.KVM_MAX_HISTS * N.KVM_HIST_MAX_BINS, dtype=torch.int64, device=dev)
        hspec = self.hist_spec(num_bins)
        series_d, nb = None, 0
        extras = N.LogExtrasC()
        if series_window_s is not None and n_global > 0 and not math.isnan(stats.min_start_ms):
            t0, t1 = stats.min_start_ms / 1000.0, stats.max_end_ms / 1000.0
            nb = max(1, int(math.ceil((t1 - t0) / series_window_s)) + 1)
            if nb > max_series_bins:
                raise N.NativeError("window series needs more than max_series_bins windows")
            series_d = torch.empty((3, nb), dtype=torch.int64, device=dev)
            dropped_d = torch.empty(1, dtype=torch.int64, device=dev)
            extras.series_window_ms = float(series_window_s) * 1000.0
            extras.series_max_bins = nb
            extras.series = series_d.data_ptr()
            extras.series_dropped = dropped_d.data_ptr()

        extras.digit_records = self._digits.data_ptr()
        arena = self._sel_cap
        while True:
            ws, fl = self._sel_ws.data_ptr(), self._flags.data_ptr()
            with tm.stage("select_plan"):
                N.check(lib.kvm_select_plan(self._stats_d.data_ptr() + N.LogStatsC.sets.offset, queries, nq,
                                            C.byref(hspec), ws, self._sel_ws.numel(), arena, st), "kvm_select_plan")
            with tm.stage("select_hist"):
                N.check(lib.kvm_log_select_hist(C.byref(log_c), fl, self._stats_d.data_ptr(), C.byref(extras),
                                                ws, st), "kvm_log_select_hist")
            hist = self._wrap(lib.kvm_select_hist_ptr(ws), N.KVM_MAX_SETS * self._bins)
            with tm.stage("allreduce_hist"):
                dist.all_reduce(hist, op=dist.ReduceOp.SUM, group=grp)
            with tm.stage("select_resolve"):
                N.check(lib.kvm_select_resolve(ws, st), "kvm_select_resolve")
            with tm.stage("select_compact"):
                N.check(lib.kvm_log_select_compact(C.byref(log_c), fl, self._digits.data_ptr(), ws, st),
                        "kvm_log_select_compact")
            rc = lib.kvm_select_status(ws, st)   # synchronises; identical on every rank (global counts)
            if rc == N.KVM_ERR_ARENA:
                arena = max(2 * arena, n_global * 4 + 1024)
                self._ensure(dlog.n, arena)
                continue
            N.check(rc, "select")
            break
        self.launches += 4

        offsets = self._wrap(lib.kvm_select_bucket_offsets_ptr(ws), N.KVM_MAX_QUERIES + 1)
        used = int(offsets[N.KVM_MAX_QUERIES].item())
        counts_local = self._wrap(lib.kvm_select_bucket_counts_ptr(ws), N.KVM_MAX_QUERIES)
        counts_all = torch.empty(self.world * N.KVM_MAX_QUERIES, dtype=torch.int64, device=dev)
        arena_local = self._wrap(lib.kvm_select_arena_ptr(ws), max(used, 1))
        arena_all = torch.empty(self.world * max(used, 1), dtype=torch.int64, device=dev)
        edge = self._wrap(lib.kvm_select_edge_ptr(ws), N.KVM_MAX_HISTS * N.KVM_HIST_MAX_BINS)
        with tm.stage("allgather_candidates"):
            dist.all_gather_into_tensor(counts_all, counts_local, group=grp)
            dist.all_gather_into_tensor(arena_all, arena_local, group=grp)
            dist.all_reduce(edge, op=dist.ReduceOp.SUM, group=grp)
        ptrs = (C.c_void_p * self.world)(*[arena_all.data_ptr() + 8 * r * max(used, 1) for r in range(self.world)])
        with tm.stage("select_finish"):
            # global fine histogram + global edge counters -> global value histograms on every rank
            N.check(lib.kvm_select_finish(ws, ptrs, self.world, counts_all.data_ptr(),
                                          self._results_ptr, hist_d.data_ptr(), st), "kvm_select_finish")
        self.launches += 1
        if series_d is not None:
            dist.all_reduce(series_d, op=dist.ReduceOp.SUM, group=grp)
            dist.all_reduce(dropped_d, op=dist.ReduceOp.SUM, group=grp)
        out = self._read_outbox()

        res = np.frombuffer(out, dtype=engine.RESULT_DTYPE, count=nq, offset=self._OFF_RESULTS)
        series = None
        if series_d is not None:
            series = series_d.cpu().numpy()
            if int(dropped_d.item()) != 0:
                raise N.NativeError("kvm_log_series dropped rows: raise max_series_bins")
        return engine.RawAnalysis(stats=stats, order=engine.results_to_orderstats(res, quantiles),
                                  hist_counts=hist_d[:4 * num_bins].view(4, num_bins).cpu().numpy(),
                                  num_bins=num_bins, series=series,
                                  series_window_s=series_window_s or 1.0, flags=self._flags[:dlog.n])
