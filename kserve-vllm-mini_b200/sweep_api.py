"""Sweep roll-up: analyse many runs stored back to back in one launch (BASELINE.json
configs[2]: 4096 runs x 250k requests).

The reference does this with one `analyze.py` process per run (grid-sweep.sh:103-174 ->
bench.sh:263) and then pulls 11 keys out of each results.json (grid-sweep.sh:136-161).
Here the runs are CSR segments of one columnar log; `kvm_seg_analyze` gives every run to one
persistent CTA, which reads the run's columns once.  Runs whose candidate set does not fit on
chip are redone through the flat path, so every run's result equals `analyze_api.analyze_log`
on that run alone.
Runs are independent, so multi-GPU = whole runs per GPU and no collective.
"""

from __future__ import annotations

import ctypes as C
from typing import Any, Dict, List, Optional, Sequence

import numpy as np
import torch

from . import _native as N
from . import engine, report

# the keys grid-sweep.sh copies into its CSV row (grid-sweep.sh:137-160)
SWEEP_KEYS = ("p50_ms", "p95_ms", "p99_ms", "throughput_rps", "tokens_per_sec", "error_rate",
              "ttft_p50_ms", "ttft_p95_ms", "cost_per_request", "cost_per_1k_tokens", "gpu_util_avg")


class SweepAnalyzer:
    def __init__(self, device: Optional[torch.device] = None):
        self.device = device or engine.require_cuda()
        self.lib = N.lib()
        self._flat: Optional[engine.LogAnalyzer] = None
        self.fallback_runs = 0
        self._scratch: Optional[torch.Tensor] = None   # 8 B/row digit records of the segmented kernel

    def analyze(self, dlog: engine.DeviceLog, run_offsets: Sequence[int],
                cold_times: Optional[Sequence[Sequence[float]]] = None, num_bins: int = 20,
                quantiles: Sequence[engine.Quantile] = engine.ANALYZE_QUANTILES,
                assemble: bool = True):
        """One RawAnalysis per run (series not computed in sweep mode).  assemble=False returns the
        raw per-run arrays (stats as a structured array, results [runs, nq], status) instead."""
        dev, lib = self.device, self.lib
        offs = np.ascontiguousarray(run_offsets, dtype=np.int64)
        runs = int(offs.shape[0]) - 1
        assert runs >= 0 and int(offs[-1]) == dlog.n and int(offs[0]) == 0
        offs_d = torch.from_numpy(offs).to(dev)
        cold_off_d = cold_d = None
        if cold_times is not None:
            lens = [len(c) for c in cold_times]
            co = np.zeros(runs + 1, dtype=np.int64)
            np.cumsum(lens, out=co[1:])
            flat = np.array([t for c in cold_times for t in sorted(c)], dtype=np.float64)
            cold_off_d = torch.from_numpy(co).to(dev)
            cold_d = torch.from_numpy(flat if flat.size else np.zeros(1)).to(dev)
        queries = engine.build_queries(quantiles)
        nq = len(queries)
        flags = torch.empty(dlog.n + 2, dtype=torch.uint8, device=dev)
        need = int(lib.kvm_seg_analyze_scratch_bytes(dlog.n))
        if self._scratch is None or self._scratch.numel() < need:
            self._scratch = None
            self._scratch = torch.empty(need, dtype=torch.uint8, device=dev)
        scratch = self._scratch
        stats_d = torch.zeros(max(runs, 1) * C.sizeof(N.LogStatsC), dtype=torch.uint8, device=dev)
        res_d = torch.zeros(max(runs, 1) * nq * C.sizeof(N.QueryResultC), dtype=torch.uint8, device=dev)
        hist_d = torch.zeros((max(runs, 1), 4, num_bins), dtype=torch.int64, device=dev)
        status_d = torch.zeros(max(runs, 1), dtype=torch.int32, device=dev)
        log_c = dlog.c_struct()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record(torch.cuda.current_stream(dev))
        N.check(lib.kvm_seg_analyze(C.byref(log_c), offs_d.data_ptr(), runs,
                                    cold_off_d.data_ptr() if cold_off_d is not None else None,
                                    cold_d.data_ptr() if cold_d is not None else None, 30.0,
                                    queries, nq, num_bins, flags.data_ptr(), scratch.data_ptr(), scratch.numel(),
                                    stats_d.data_ptr(),
                                    res_d.data_ptr(), hist_d.data_ptr(), status_d.data_ptr(),
                                    engine._stream_ptr(dev)), "kvm_seg_analyze")
        ev1.record(torch.cuda.current_stream(dev))
        self._last = (flags, stats_d, res_d, hist_d, status_d)
        self._events = (ev0, ev1)
        stats_b = stats_d.cpu().numpy().tobytes()
        res = np.frombuffer(res_d.cpu().numpy().tobytes(), dtype=engine.RESULT_DTYPE).reshape(max(runs, 1), nq)
        hist = hist_d.cpu().numpy()
        status = status_d.cpu().numpy()
        self.last_kernel_ms = ev0.elapsed_time(ev1)   # device time of the one segmented launch
        if not assemble:
            stats_arr = np.frombuffer(stats_b, dtype=np.dtype(N.LogStatsC), count=runs)
            return stats_arr, res[:runs], status[:runs]
        out: List[engine.RawAnalysis] = []
        ssz = C.sizeof(N.LogStatsC)
        self.fallback_runs = 0
        for r in range(runs):
            if status[r] != 0:
                out.append(self._flat_run(dlog, int(offs[r]), int(offs[r + 1]),
                                          cold_times[r] if cold_times is not None else (), num_bins, quantiles))
                self.fallback_runs += 1
                continue
            st = N.LogStatsC.from_buffer_copy(stats_b[r * ssz:(r + 1) * ssz])
            out.append(engine.RawAnalysis(stats=st, order=engine.results_to_orderstats(res[r], quantiles),
                                          hist_counts=hist[r], num_bins=num_bins, series=None,
                                          flags=flags[int(offs[r]):int(offs[r + 1])]))
        return out

    # keys of results.json that grid-sweep.sh:137-160 copies into its CSV and that come from the
    # request log (cost_* and gpu_util_avg come from cost_estimator / Prometheus)
    SWEEP_KEYS = ("p50_ms", "p95_ms", "p99_ms", "throughput_rps", "tokens_per_sec", "error_rate",
                  "ttft_p50_ms", "ttft_p95_ms")

    def sweep_table(self, dlog: engine.DeviceLog, run_offsets: Sequence[int],
                    cold_times: Optional[Sequence[Sequence[float]]] = None) -> Dict[str, np.ndarray]:
        """The sweep CSV's log-derived columns for every run as float64 arrays, without building a
        results dict per run: the same expressions as report.core_results, evaluated over all runs
        at once (elementwise IEEE operations in the same order -> the same doubles).  Where the
        reference yields None the value is NaN and `table["is_none"][key]` is True."""
        quantiles = engine.ANALYZE_QUANTILES
        raws = self.analyze(dlog, run_offsets, cold_times, assemble=False)
        stats, res, status = raws
        runs = int(stats.shape[0])
        names = [q.name for q in quantiles]

        def interp(name: str, p: float) -> np.ndarray:      # OrderStat.interpolate, vectorised
            i = names.index(name)
            n = res[:, 2 * i]["n"].astype(np.int64)
            vlo, vhi = res[:, 2 * i]["value"], res[:, 2 * i + 1]["value"]
            pos = max(0.0, min(p, 1.0)) * (n - 1).astype(np.float64)
            lo, hi = np.floor(pos), np.ceil(pos)
            with np.errstate(invalid="ignore"):
                val = np.where(lo == hi, vlo, vlo + (pos - lo) * (vhi - vlo))
            val = np.where(n == 1, vlo, val)
            return np.where(n == 0, np.nan, val)

        total, success = stats["n_total"].astype(np.int64), stats["n_success"].astype(np.int64)
        start, end = stats["min_start_ms"] / 1000.0, stats["max_end_ms"] / 1000.0
        has_dur = end > start                                # NaN compares False, like the reference's None
        with np.errstate(invalid="ignore", divide="ignore"):
            duration = np.where(has_dur, end - start, np.nan)
            cnt = stats["sets"]["count"]
            n_ttfb = cnt[:, N.SET_TTFB_COLD] + cnt[:, N.SET_TTFB_WARM]
            table = {
                "p50_ms": interp("lat_p50", 0.50), "p95_ms": interp("lat_p95", 0.95), "p99_ms": interp("lat_p99", 0.99),
                "throughput_rps": np.where(has_dur, success / duration, np.nan),
                "tokens_per_sec": np.where(has_dur, stats["sum_completion_tokens"].astype(np.float64) / duration, np.nan),
                "error_rate": np.where(total > 0, (total - success) / np.where(total > 0, total, 1), np.nan),
                "ttft_p50_ms": np.where(n_ttfb > 0, interp("ttfb_p50", 0.50), np.nan),
                "ttft_p95_ms": np.where(n_ttfb > 0, interp("ttfb_p95", 0.95), np.nan),
            }
        none = {"p50_ms": np.zeros(runs, bool), "p95_ms": np.zeros(runs, bool), "p99_ms": np.zeros(runs, bool),
                "throughput_rps": ~has_dur, "tokens_per_sec": ~has_dur, "error_rate": total == 0,
                "ttft_p50_ms": n_ttfb <= 0, "ttft_p95_ms": n_ttfb <= 0}
        # runs the segmented kernel handed back (they did not fit on chip) go through the flat path
        self.fallback_runs = 0
        offs = np.asarray(run_offsets, dtype=np.int64)
        for r in np.nonzero(status != 0)[0].tolist():
            cold = cold_times[r] if cold_times is not None else ()
            core = report.core_results(self._flat_run(dlog, int(offs[r]), int(offs[r + 1]), cold, 20, quantiles),
                                       len(cold))
            for k in self.SWEEP_KEYS:
                none[k][r] = core[k] is None
                table[k][r] = np.nan if core[k] is None else core[k]
            self.fallback_runs += 1
        table["is_none"] = none
        return table

    def _flat_run(self, dlog, r0, r1, cold, num_bins, quantiles) -> engine.RawAnalysis:
        if self._flat is None:
            self._flat = engine.LogAnalyzer(self.device)
        sub = engine.DeviceLog(cols={k: v[r0:r1] for k, v in dlog.cols.items()}, n=r1 - r0, device=self.device)
        return self._flat.analyze(sub, cold, num_bins=num_bins, series_window_s=None, quantiles=quantiles)

    def rollup(self, dlog: engine.DeviceLog, run_offsets: Sequence[int],
               cold_times: Optional[Sequence[Sequence[float]]] = None) -> List[Dict[str, Any]]:
        """results.json core per run."""
        raws = self.analyze(dlog, run_offsets, cold_times)
        return [report.core_results(raw, len(cold_times[i]) if cold_times is not None else 0)
                for i, raw in enumerate(raws)]


# --------------------------------------------------------------------------- runs across GPUs
def partition_runs(run_offsets: Sequence[int], world: int) -> List[int]:
    """Cut `runs` into `world` contiguous groups of whole runs with row counts as even as whole
    runs allow (configs[2] on 8 GPUs = 512 runs each).  Returns world + 1
    run indices; group r = runs [cut[r], cut[r+1]).  Deterministic, the same on every rank."""
    offs = np.asarray(run_offsets, dtype=np.int64)
    runs = int(offs.shape[0]) - 1
    total = int(offs[-1] - offs[0]) if runs > 0 else 0
    cuts = [0]
    for r in range(1, world):
        target = offs[0] + (total * r) // world
        # the run boundary nearest to the ideal cut, never behind the previous cut
        j = int(np.searchsorted(offs, target, side="left"))
        if j > 0 and j <= runs and abs(int(offs[j - 1]) - int(target)) <= abs(int(offs[min(j, runs)]) - int(target)):
            j -= 1
        cuts.append(max(cuts[-1], min(j, runs)))
    cuts.append(runs)
    return cuts


def gather_sweep_tables(local: Dict[str, Any], keys: Sequence[str], counts: Sequence[int],
                        device: Optional[torch.device] = None, group=None) -> Dict[str, Any]:
    """Per-rank sweep tables (run order inside a rank, rank order = run order) -> the whole
    table on every rank.  Results travel, not data: 8 doubles and 8 flags per run, as ONE
    float64 tensor per rank (padded to the largest group; `counts[r]` = runs of rank r, known to
    every rank from partition_runs) through one all-gather."""
    import torch.distributed as dist
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    nk, width = len(keys), max(1, max(counts))
    mine = np.zeros((2 * nk, width), dtype=np.float64)
    for i, k in enumerate(keys):
        mine[i, :counts[rank]] = np.asarray(local[k], dtype=np.float64)
        mine[nk + i, :counts[rank]] = np.asarray(local["is_none"][k], dtype=np.float64)
    dev = device if device is not None else torch.device("cpu")
    loc = torch.from_numpy(mine).to(dev)
    allv = torch.empty((world,) + tuple(loc.shape), dtype=torch.float64, device=dev)
    if dev.type == "cuda":
        dist.all_gather_into_tensor(allv.view(-1), loc.view(-1), group=group)
    else:
        dist.all_gather(list(allv.unbind(0)), loc, group=group)
    allv = allv.cpu().numpy()
    out: Dict[str, Any] = {k: np.concatenate([allv[r, i, :counts[r]] for r in range(world)]) for i, k in enumerate(keys)}
    out["is_none"] = {k: np.concatenate([allv[r, nk + i, :counts[r]] for r in range(world)]) != 0.0
                      for i, k in enumerate(keys)}
    return out


class ShardedSweep:
    """configs[2] on N GPUs: rank r holds the rows of its runs only (whole runs, no collective on
    the data path) and analyses them with the segmented kernel; the per-run table is gathered."""

    def __init__(self, device: Optional[torch.device] = None, group=None):
        import torch.distributed as dist
        self.group = group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.local = SweepAnalyzer(device)

    def my_runs(self, run_offsets: Sequence[int]) -> range:
        cuts = partition_runs(run_offsets, self.world)
        return range(cuts[self.rank], cuts[self.rank + 1])

    def sweep_table(self, dlog_local: engine.DeviceLog, local_offsets: Sequence[int],
                    colds_local: Optional[Sequence[Sequence[float]]] = None,
                    counts: Optional[Sequence[int]] = None) -> Dict[str, Any]:
        """dlog_local / local_offsets / colds_local describe THIS rank's runs (offsets start at 0).
        counts[r] = number of runs of rank r (from partition_runs; default: every rank holds as
        many runs as this one)."""
        t = self.local.sweep_table(dlog_local, local_offsets, colds_local)
        if counts is None:
            counts = [len(local_offsets) - 1] * self.world
        return gather_sweep_tables(t, SweepAnalyzer.SWEEP_KEYS, counts, self.local.device, self.group)
