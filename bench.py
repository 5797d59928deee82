#!/usr/bin/env python3
"""bench.py - request-records analysed per second on B200 (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU algorithm

A "step" = one full run analysis of one synthetic request log: K1 scan (window, cold/warm,
counts, token sums) -> exact select of the 23 quantiles of analyze.py -> four 20-bin
histograms -> 1 s window series -> results.json fields + the cost-per-1K-token rollup on
the host.  N=1 workload: BASELINE.json configs[1] (10M-request log).  N>1: the log is
sharded by time range over the ranks (configs[4] layout, weak scaling, 10M rows per rank
by default) with exact global quantiles through NCCL histogram all-reduce.

Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for how each field is taken.
"""

import argparse
import collections
import json
import math
import os
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
if REPO not in sys.path:
    sys.path.insert(0, REPO)

METRIC = "request_records_analyzed_per_sec"
UNIT = "records/s"
ROW_BYTES = 56  # 4 fp64 + 3 int64 columns


def measured_peaks():
    path = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            with open(path) as f:
                return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """Samples SM clock / throttle reasons during the timed region (pynvml)."""

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.004)

    def start(self):
        if self.nv is not None:
            self._thr = threading.Thread(target=self._run, daemon=True)
            self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr is not None:
            self._thr.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


# --------------------------------------------------------------------------- CPU arm
def cpu_port_lists(cols, rows):
    """Columns as Python lists - what the reference holds after parsing (not part of the timing)."""
    return {k: v[:rows].tolist() for k, v in cols.items()}


def cpu_port_step(sub, cold):
    """One pass of the pure-Python port of analyze.py (+ cost rollup); returns seconds."""
    from oracle import py_port
    t = time.perf_counter()
    res = py_port.analyze(sub, cold)
    tokens = py_port.total_tokens_success(sub)
    py_port.cost_rollup(3600.0, 14400.0, 57600.2, 1.5, 0.04, 0.005, 0.10, res["requests"]["success"], tokens)
    return time.perf_counter() - t


def cpu_port_rate(cols, cold, rows, repeats=1):
    """records/s of the port on the first `rows` rows."""
    sub = cpu_port_lists(cols, rows)
    best = min(cpu_port_step(sub, cold) for _ in range(repeats))
    return rows / best, best


def ref_step(rows_list, cold):
    """One pass of the UNMODIFIED reference (oracle/_ref/analyze.py + cost_estimator.py functions,
    driven by oracle/ref_run.py in the order of analyze.main()) on rows already parsed."""
    from oracle import ref_run
    t = time.perf_counter()
    res = ref_run.analyze_rows(rows_list, cold)
    ref_run.cost_rows(rows_list, 2.0)
    return time.perf_counter() - t, res


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on the host cores.
    oracle/_ref/ (the unmodified analyze.py / cost_estimator.py, installed by build()) when it
    travelled with the repo - kind "reference" - else the pure-Python port - kind "port".  The
    reference is single-threaded by construction: 1 core.  Each step analyses a bounded sample
    (`rows_per_step`, stated in config) of the same synthetic log."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from kvmini_b200 import synth
    from oracle import ref_run
    use_ref = ref_run.available() and not args.port
    n = args.rows
    budget_s = 100.0

    def prep(cols, rows):
        return ref_run.rows_from_columns(cols, rows) if use_ref else cpu_port_lists(cols, rows)

    def one(sub, cold):
        return ref_step(sub, cold)[0] if use_ref else cpu_port_step(sub, cold)

    # calibrate on 20k rows, then size each step so that the whole run takes ~budget_s
    cal_cols = synth.request_columns(20_000, seed=1234)
    t0 = float(cal_cols["start_ms"].min()) / 1000.0
    cal = prep(cal_cols, 20_000)
    rate = 20_000 / min(one(cal, [t0 + 5.0]) for _ in range(2))
    rows = int(min(n, max(10_000, rate * budget_s / (args.steps + args.warmup))))
    cols = synth.request_columns(rows, seed=1234, span_ms=3.6e6 * rows / n)
    t0 = float(cols["start_ms"].min()) / 1000.0
    cold = [t0 + 5.0]
    sub = prep(cols, rows)
    for _ in range(args.warmup):
        one(sub, cold)
    dt = sum(one(sub, cold) for _ in range(args.steps))
    value = rows * args.steps / dt
    cfg = workload_config(args, args.gpus)
    cfg["rows_per_step"] = rows
    cfg["note"] = (f"CPU arm: each step analyses a {rows}-row log drawn like the first {rows} rows of the "
                   f"{n}-row workload (same distribution and request rate); rows_per_gpu names the GPU arm's size")
    what = ("oracle/_ref/analyze.py (the unmodified reference: window_bounds, classify_requests_cold_warm, "
            "compute_cold_warm_metrics, compute_token_timing_analysis, percentile) + cost_estimator."
            "calculate_cold_warm_costs on rows already parsed" if use_ref else
            "oracle/py_port.analyze (pure-Python restatement of analyze.py + cost rollup)")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": cfg,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": "reference" if use_ref else "port",
                         "sample": f"{rows} rows per step (of the {n}-row workload); {what}; single thread"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def workload_config(args, n_gpus):
    return {
        "workload": ("configs[1]: synthetic 10M-request log, full analyze (p50/p95/p99 latency+TTFT, cold/warm, "
                     "4 histograms, tokens/sec, 1s series) + cost per 1K tokens" if n_gpus == 1 else
                     "configs[4] layout: log sharded by time range, exact global quantiles (digit histograms "
                     "summed across ranks, candidates exchanged) + 1s-window series"),
        "rows_per_gpu": args.rows, "rows_total": args.rows * n_gpus,
        "bytes_per_row": ROW_BYTES,
        "l2": (f"inputs ({args.rows * ROW_BYTES / 1e6:.0f} MB per GPU) exceed the 126 MB L2; no flush needed"
               if args.rows * ROW_BYTES > 126e6 else
               f"inputs ({args.rows * ROW_BYTES / 1e6:.0f} MB per GPU) FIT in the 126 MB L2: not a valid roofline run"),
        "parallelism": f"time-range shards x{n_gpus}" if n_gpus > 1 else "single GPU",
    }


# --------------------------------------------------------------------------- parity / extra blocks
def _same(a, b, rel=None):
    if isinstance(a, dict) and isinstance(b, dict):
        return list(a) == list(b) and all(_same(a[k], b[k], rel) for k in a)
    if isinstance(a, (list, tuple)) and isinstance(b, (list, tuple)):
        return len(a) == len(b) and all(_same(x, y, rel) for x, y in zip(a, b))
    if isinstance(a, float) and isinstance(b, float):
        if math.isnan(a) and math.isnan(b):
            return True
        return a == b or (rel is not None and abs(a - b) <= rel * abs(b))
    return a == b and type(a) is type(b)


def parity_block(core, series, host_cols, cold):
    """The result of the last timed step against the numpy oracle on the same columns (outside
    every timed region).  Bit-exact on every key except the mean TTFT (<= 1e-12 relative)."""
    from oracle import np_port
    t = time.perf_counter()
    exp = np_port.analyze(host_cols, cold)
    bad = [k for k in core if k not in exp or not _same(core[k], exp[k], 1e-12 if k == "time_to_first_token_ms" else None)]
    exp_series = np_port.window_series(host_cols)
    if not _same(series, exp_series):
        bad.append("timeseries")
    return {"oracle": "np_port (numpy restatement of analyze.py, pinned to the reference in tests/test_oracle.py)",
            "rows": int(len(host_cols["status"])), "keys_compared": len(core) + 1, "equal": not bad,
            "mismatches": bad[:8], "oracle_seconds": round(time.perf_counter() - t, 2)}


def percentile_path_block(device, peak, sizes=(10_000_000, 100_000_000), steps=50):
    """The 16 B/record percentile path: p50/p95/p99 of the status-200 latencies,
    kvm_column_select_sampled and - beside it - the general kvm_column_select, each timed as a
    CUDA-graph replay of its launch sequence with CUDA events, columns resident (160 MB / 1.6 GB:
    larger than L2).  frac = 16 B x n / time / measured HBM peak.  Results of the two are compared
    with each other and with a torch.sort of the same column (ranks exact)."""
    import ctypes as C
    import numpy as np
    import torch
    from kvmini_b200 import _native as N
    from kvmini_b200 import engine, synth
    lib = N.lib()
    out = {}
    q = [(1, k, p, 0) for p in (0.5, 0.95, 0.99) for k in (N.RANK_LERP_LO, N.RANK_LERP_HI)]
    arr = (N.QueryC * len(q))(*[N.QueryC(*x) for x in q])
    for n in sizes:
        c = synth.request_columns_device(n, device, seed=1234)
        lat, st = c["latency_ms"], c["status"]
        del c
        res = torch.zeros(len(q) * 32, dtype=torch.uint8, device=device)
        ws_s = torch.empty(lib.kvm_column_select_sampled_workspace_bytes(n), dtype=torch.uint8, device=device)
        cap = int(lib.kvm_select_default_arena(n))
        ws_g = torch.empty(lib.kvm_select_workspace_bytes(cap), dtype=torch.uint8, device=device)
        stream = torch.cuda.Stream(device)

        def sampled():
            N.check(lib.kvm_column_select_sampled(lat.data_ptr(), n, N.MEMBER_STATUS, st.data_ptr(), arr, len(q),
                                                  res.data_ptr(), ws_s.data_ptr(), ws_s.numel(),
                                                  torch.cuda.current_stream(device).cuda_stream))

        def general():
            N.check(lib.kvm_column_select(lat.data_ptr(), n, N.MEMBER_STATUS, st.data_ptr(), 1, arr, len(q),
                                          res.data_ptr(), ws_g.data_ptr(), ws_g.numel(), cap,
                                          torch.cuda.current_stream(device).cuda_stream))

        kept = torch.sort(lat[(st == 200) & ~torch.isnan(lat)]).values
        m = int(kept.numel())
        want = []
        for p in (0.5, 0.95, 0.99):
            pos = p * (m - 1)
            want += [float(kept[int(math.floor(pos))]), float(kept[int(math.ceil(pos))])]
        del kept
        entry = {"rows": n, "algorithmic_bytes": 16.0 * n}
        with torch.cuda.stream(stream):
            for name, call, ws in (("sampled", sampled, ws_s), ("general", general, ws_g)):
                for _ in range(3):
                    call()
                stream.synchronize()
                if int(ws[0:4].view(torch.int32).item()) != 0:
                    entry[name] = {"declined": True}
                    continue
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=stream):
                    call()
                for _ in range(3):
                    g.replay()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                for _ in range(steps):
                    g.replay()
                e1.record(stream)
                stream.synchronize()
                ms = e0.elapsed_time(e1) / steps
                got = np.frombuffer(res.cpu().numpy().tobytes(), dtype=engine.RESULT_DTYPE, count=len(q))
                entry[name] = {"ms": ms, "gbs": 16.0 * n / ms / 1e6, "frac": 16.0 * n / ms / 1e6 / peak,
                               "equal_to_sorted_ranks": [float(x) for x in got["value"]] == want
                               and all(int(x) == m for x in got["n"])}
        # the same call with several in flight (own workspace, result block and stream each): what
        # a monitor pays per column when it asks for the quantiles of several columns / logs at
        # once - the single-CTA phases of one call overlap the streaming pass of the others
        if "ms" in entry.get("sampled", {}):
            depth, rounds = int(os.environ.get("KVM_PP_DEPTH", "8")), max(8, steps // 2)
            lanes = []
            for _ in range(depth):
                w = torch.empty_like(ws_s)
                r = torch.zeros_like(res)
                sst = torch.cuda.Stream(device)
                with torch.cuda.stream(sst):
                    def call(w=w, r=r):
                        N.check(lib.kvm_column_select_sampled(lat.data_ptr(), n, N.MEMBER_STATUS, st.data_ptr(), arr, len(q),
                                                              r.data_ptr(), w.data_ptr(), w.numel(),
                                                              torch.cuda.current_stream(device).cuda_stream))
                    for _ in range(2):
                        call()
                    sst.synchronize()
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g, stream=sst):
                        call()
                    g.replay()
                    sst.synchronize()
                lanes.append((sst, g, w, r))
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(device)
            e0.record(stream)
            for sst, _, _, _ in lanes:
                sst.wait_event(e0)
            for _ in range(rounds):
                for sst, g, _, _ in lanes:
                    with torch.cuda.stream(sst):
                        g.replay()
            for sst, _, _, _ in lanes:
                stream.wait_stream(sst)
            e1.record(stream)
            stream.synchronize()
            ms = e0.elapsed_time(e1) / (rounds * depth)
            ok = True
            for _, _, w, r in lanes:
                got = np.frombuffer(r.cpu().numpy().tobytes(), dtype=engine.RESULT_DTYPE, count=len(q))
                ok = ok and int(w[0:4].view(torch.int32).item()) == 0 and [float(x) for x in got["value"]] == want
            entry["sampled_in_flight"] = {"ms_per_call": ms, "gbs": 16.0 * n / ms / 1e6, "frac": 16.0 * n / ms / 1e6 / peak,
                                            "calls_in_flight": depth, "equal_to_sorted_ranks": ok}
            del lanes
        out["rows_%g" % n] = entry
        del lat, st, ws_s, ws_g, res
        torch.cuda.empty_cache()
    return out


def cfg5_block(args, device, rank, world, dist, cold):
    """BASELINE.json configs[4] in the driver's own line: 1.25e8 rows per GPU (1e9 at 8 GPUs), the
    log sharded by time range, generated on the device (no host copy: 7 GB per GPU).  Every rank
    runs the sharded analysis; then the SAME shard is analysed by the single-GPU analyzer in the
    same run (what one GPU does alone on 1/world of the log) for the scaling ratio.  Parity at this
    size: a rank certificate - for every reported latency order statistic v with rank k, all ranks
    count their success rows below / up to v; the counts, summed, must bracket k (an exact
    order statistic of the WHOLE log, independent of any oracle)."""
    import torch
    from kvmini_b200 import dist_api, engine, report, synth
    rows = args.cfg5_rows
    cols = synth.request_columns_device(rows, device, seed=4321 + rank, t_epoch_ms=1.7e12, t_offset_ms=rank * 3.6e6)
    dlog = engine.DeviceLog.from_tensors(cols)
    steps = max(3, min(args.steps, 10))
    sharded = dist_api.ShardedAnalyzer(device, exchange=args.exchange)
    for _ in range(3):
        raw = sharded.analyze(dlog, cold, use_graph=not args.eager)
    dist.barrier()
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        raw = sharded.analyze(dlog, cold, use_graph=not args.eager)
    e1.record()
    dist.barrier()
    torch.cuda.synchronize(device)
    ms = e0.elapsed_time(e1) / steps
    core = report.core_results(raw, len(cold))
    sharded.close()
    # one GPU alone on its shard (same run, same data)
    single = engine.LogAnalyzer(device)
    for _ in range(3):
        single.analyze(dlog, cold, use_graph=True)
    torch.cuda.synchronize(device)
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s0.record()
    for _ in range(steps):
        single.analyze(dlog, cold, use_graph=True)
    s1.record()
    torch.cuda.synchronize(device)
    ms_single = s0.elapsed_time(s1) / steps
    single.close()
    # rank certificate of the global latency order statistics
    ok_lat = cols["latency_ms"][(cols["status"] == 200) & ~torch.isnan(cols["latency_ms"])]
    checks = []
    for name, p in (("lat_p50", 0.50), ("lat_p95", 0.95), ("lat_p99", 0.99)):
        st = raw.order[name]
        for v, fn in ((st.v_lo, math.floor), (st.v_hi, math.ceil)):
            checks.append((float(v), fn, p))
    cnt = torch.tensor([[int((ok_lat < v).sum()), int((ok_lat <= v).sum())] for v, _, _ in checks] + [[ok_lat.numel(), 0]],
                       dtype=torch.int64, device=device)
    dist.all_reduce(cnt)
    t = torch.tensor([ms, ms_single], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    n_glob = int(cnt[-1, 0])
    certified = all(int(cnt[i, 0]) <= int(fn(p * (n_glob - 1))) < int(cnt[i, 1]) for i, (_, fn, p) in enumerate(checks))
    del cols, dlog, ok_lat
    torch.cuda.empty_cache()
    ms, ms_single = float(t[0]), float(t[1])
    return {"workload": "configs[4]: 1e9-record log sharded by time range over 8 GPUs (rows_per_gpu x n_gpus here)",
            "rows_per_gpu": rows, "rows_total": rows * world, "steps": steps, "ms_per_step": ms,
            "records_per_s": rows * world / ms * 1e3,
            "one_gpu_on_one_shard_ms": ms_single, "one_gpu_records_per_s": rows / ms_single * 1e3,
            "speedup_vs_one_gpu_on_one_shard": (rows * world / ms) / (rows / ms_single),
            "parity": {"method": "rank certificate of lat p50/p95/p99 (lo and hi order statistics): global counts "
                                 "below / up to the reported value bracket the rank", "success_rows": n_glob,
                       "equal": bool(certified)},
            "check": {"p50_ms": core["p50_ms"], "p99_ms": core["p99_ms"]}}


def sweep_block(args, device, rank, world, dist):
    """BASELINE.json configs[2] in the driver's own line: whole runs per GPU (SURVEY.md section
    8e: 4096 runs x 250k rows = 512 runs per GPU at 8 GPUs; `--sweep-runs` per GPU here), generated
    on the device, analysed by the segmented kernel (kvm_seg_analyze), per-run tables gathered in
    run order.  Roofline on 56 B/row (every column read once).  Parity: two runs of rank 0 are
    copied to the host and compared with the oracle, all 8 sweep columns bit for bit."""
    import numpy as np
    import torch
    from kvmini_b200 import engine, sweep_api, synth
    per, rows = args.sweep_runs, 250_000
    n = per * rows
    g = {k: torch.empty(n, dtype=torch.float64, device=device) for k in ("start_ms", "latency_ms", "ttfb_ms", "tllt_ms")}
    g.update({k: torch.empty(n, dtype=torch.int64, device=device) for k in ("status", "completion_tokens", "total_tokens")})
    for r0 in range(0, per, 64):
        r1 = min(per, r0 + 64)
        gr = rank * per + r0
        c = synth.request_columns_device(rows * (r1 - r0), device, seed=1000 + gr, span_ms=rows * 25.0 * (r1 - r0),
                                         mu=5.0 + 1.5 * ((gr * 37) % 100) / 100.0, sigma=0.3 + 0.5 * ((gr * 17) % 100) / 100.0,
                                         fail_p=0.05 * ((gr * 7) % 100) / 100.0, t_epoch_ms=1.7e12 + gr * rows * 25.0)
        for k in g:
            g[k][r0 * rows:r1 * rows] = c[k]
        del c
    offs = (np.arange(per + 1, dtype=np.int64) * rows).tolist()
    starts = g["start_ms"][::rows].cpu().numpy() / 1000.0
    colds = [[float(starts[r]) + 1.0 + (r % 5)] if r % 3 else [] for r in range(per)]
    dlog = engine.DeviceLog.from_tensors(g)
    if dist is not None:
        sh = sweep_api.ShardedSweep(device)
        local, table_fn = sh.local, (lambda: sh.sweep_table(dlog, offs, colds))
    else:
        local = sweep_api.SweepAnalyzer(device)
        table_fn = lambda: local.sweep_table(dlog, offs, colds)
    table_fn()
    steps, ms_tab, ms_k = 3, 0.0, 0.0
    for _ in range(steps):
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(device)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        table = table_fn()
        e1.record()
        torch.cuda.synchronize(device)
        ms_tab += e0.elapsed_time(e1) / steps
        ms_k += local.last_kernel_ms / steps
    t = torch.tensor([ms_tab, ms_k], dtype=torch.float64, device=device)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    out = None
    if rank == 0:
        from oracle import np_port   # the checker
        equal = True
        for r in (1, per - 1):
            sub = {k: v[offs[r]:offs[r + 1]].cpu().numpy() for k, v in g.items()}
            ref = np_port.analyze(sub, colds[r])
            for k in sweep_api.SweepAnalyzer.SWEEP_KEYS:
                got = None if bool(table["is_none"][k][r]) else float(table[k][r])
                equal = equal and (got == ref[k])
        peak, _ = measured_peaks()
        tot = n * world
        out = {"workload": "configs[2]: 4096 runs x 250k rows, whole runs per GPU (runs_per_gpu x n_gpus here)",
               "runs_per_gpu": per, "runs_total": per * world, "rows_total": tot,
               "ms_kernel": float(t[1]), "ms_table_incl_host_and_gather": float(t[0]),
               "records_per_s_kernel": tot / float(t[1]) * 1e3, "records_per_s_table": tot / float(t[0]) * 1e3,
               "roofline": {"bound": "hbm", "kernel": "seg_analyze", "algorithmic_bytes_per_row": 56,
                            "achieved_per_gpu": 56.0 * n / float(t[1]) / 1e6, "peak": peak, "unit": "GB/s",
                            "frac": 56.0 * n / float(t[1]) / 1e6 / peak},
               "fallback_runs_rank0": local.fallback_runs,
               "parity": {"method": "runs 1 and runs_per_gpu-1 of rank 0 vs oracle/np_port.analyze, 8 sweep columns, bit for bit",
                          "equal": bool(equal)}}
    del g, dlog
    torch.cuda.empty_cache()
    return out


# --------------------------------------------------------------------------- our arm
def run_ours(args):
    import numpy as np
    import torch

    from kvmini_b200 import _native as N
    from kvmini_b200 import columns, engine, report, synth
    from kvmini_b200 import cost_api

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus > 1 and world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} needs torchrun with {args.gpus} ranks (WORLD_SIZE={world})")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=device)

    n = args.rows
    # synthetic shard of this rank: cfg2 distribution, shard r covers hour r of the log
    if args.device_data:
        # generated on the device (sizes whose host copy would not fit comfortably); the host
        # copy for the e2e leg is taken from it once, outside every timed region
        resident = synth.request_columns_device(n, device, seed=1234 + rank, t_epoch_ms=1.7e12,
                                                t_offset_ms=rank * 3.6e6)
        host = None
        pinned = {k: v.cpu().pin_memory() for k, v in resident.items()} if not args.no_e2e else None
    else:
        host = synth.request_columns(n, seed=1234 + rank, t_epoch_ms=1.7e12 + rank * 3.6e6)
        log = columns.columns_from_arrays(**host)
        pinned = {k: torch.from_numpy(v).pin_memory() for k, v in log.cols.items()}
        resident = {k: v.to(device) for k, v in pinned.items()}
    cold = [1.7e9 + 5.0, 1.7e9 + 1800.0]
    pricing = cost_api.UnitPricing(1.5, {}, 0.04, 0.005, 0.10, True, False)
    rsecs = {"gpu_seconds": 3600.0, "cpu_core_seconds": 14400.0, "mem_gib_seconds": 57600.2}

    if world > 1:
        from kvmini_b200 import dist_api
        analyzer = dist_api.ShardedAnalyzer(device, exchange=args.exchange)
    else:
        analyzer = engine.LogAnalyzer(device)

    graphs = not args.eager   # CUDA-graph replay of the launch sequence (N>1: kernels AND collectives)
    dlog_resident = engine.DeviceLog.from_tensors(resident)
    staging = ({k: torch.empty_like(v) for k, v in resident.items()} if pinned is not None else None)
    dlog_staging = engine.DeviceLog.from_tensors(staging) if staging is not None else None

    def finish(raw):
        core = report.core_results(raw, len(cold))
        cost = cost_api.rollup(raw, pricing, rsecs, 1.5, None)
        return core, cost, raw

    def step(dlog, use_graph):
        raw = analyzer.analyze(dlog, cold, use_graph=use_graph)
        return finish(raw)

    # two analyzers with their own buffers: the host turns analysis k into results.json fields
    # while analysis k+1 is on the GPU (single-GPU path)
    pipe = None
    if world == 1:
        # (three: while the host collects analysis k and submits k+3, k+1 and k+2 are both on the GPU -
        # with two, the GPU holds a single analysis for the ~100 us the host needs per step)
        pipe = tuple(engine.LogAnalyzer(device, stream=torch.cuda.Stream(device)) for _ in range(max(2, args.pipeline_depth)))
    elif not args.no_pipeline:
        # N>1: only with the peer-memory exchange (no collective library call on the data path, so
        # two analyses can be in flight); every rank takes the same branch
        analyzer.analyze(engine.DeviceLog.from_tensors(resident), cold)
        if analyzer.exchange != "nccl":
            pipe = tuple(dist_api.ShardedAnalyzer(device, exchange="peer", stream=torch.cuda.Stream(device))
                         for _ in range(max(2, args.pipeline_depth)))
    pipelined = pipe is not None
    if args.refine_min_keys is not None:
        for an in (analyzer,) + tuple(pipe or ()):
            an.refine_min_keys = args.refine_min_keys

    def run_pipelined(dlog, k, before_submit=None, depth=None):
        """`dlog`: one DeviceLog, or one per analyzer of the pair (the e2e leg: each analyzer has
        its own staging columns; before_submit(slot) fills them)."""
        depth = len(pipe) if depth is None else min(depth, len(pipe))
        pair = pipe[:depth]
        dlogs = dlog if isinstance(dlog, (tuple, list)) else (dlog,) * depth
        pending, out, launches = collections.deque(), None, 0
        for it in range(k):
            slot = it % depth          # its previous analysis (it - depth) has been collected below
            an = pair[slot]
            if before_submit is not None:
                before_submit(slot)
            pending.append(an.analyze_async(dlogs[slot], cold, use_graph=graphs))
            launches += an.launches
            if len(pending) == depth:
                out = finish(pending.popleft()())
        while pending:
            out = finish(pending.popleft()())
        for an in pair:   # join the side streams before the caller records its end event
            torch.cuda.current_stream(device).wait_stream(an.stream)
        return out, launches

    def e2e_step():
        for k, v in pinned.items():
            staging[k].copy_(v, non_blocking=True)   # H2D from pinned host memory, inside the timed region
        return step(dlog_staging, graphs)

    # pipelined e2e: every analyzer of the pair owns a staging copy of the columns; the H2D copy of
    # step k+1 (current stream) runs while step k computes on its analyzer's stream.  Slot s is
    # refilled only after the analysis that used it has been collected (run_pipelined's order).
    staging2 = dlog_staging2 = None
    if pinned is not None and pipelined:
        # (the e2e leg is PCIe-bound: two analyses in flight are enough, more staging copies only add traffic)
        staging2 = (staging, {k: torch.empty_like(v) for k, v in resident.items()})
        dlog_staging2 = tuple(engine.DeviceLog.from_tensors(s) for s in staging2)

    def e2e_fill(slot):
        for k, v in pinned.items():
            staging2[slot][k].copy_(v, non_blocking=True)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(device)

    # ---- warm-up (also captures the graphs)
    for _ in range(args.warmup):
        step(dlog_resident, graphs)
    if pipelined:
        run_pipelined(dlog_resident, max(2 * args.warmup, 2 * len(pipe)))   # every analyzer captures and replays its graph once
    barrier()

    # ---- timed region 1: inputs resident in HBM
    sampler = ClockSampler(local)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    launches = 0
    if pipelined:
        (core, cost, raw), launches = run_pipelined(dlog_resident, args.steps)
    else:
        for _ in range(args.steps):
            core, cost, raw = step(dlog_resident, graphs)
            launches += analyzer.launches
    ev1.record()
    barrier()
    ms_total = ev0.elapsed_time(ev1)

    # ---- same K steps launched eagerly with a CUDA-event pair around every launch: per-kernel
    # durations for the roofline (events cannot be read inside a graph replay)
    timer = engine.StageTimer(device)
    analyzer.timer = timer
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k0.record()
    for _ in range(args.steps):
        step(dlog_resident, False)
    k1.record()
    barrier()
    clocks = sampler.stop()
    ms_eager = k0.elapsed_time(k1)
    stage_ms = timer.totals_ms()
    analyzer.timer = engine._NoTimer()

    # ---- timed region 2: end to end from pinned host buffers
    ms_e2e = float("nan")
    e2e_steps = max(3, min(args.steps, 20))
    if pinned is not None:
        if dlog_staging2 is not None:
            run_pipelined(dlog_staging2, 4, e2e_fill, depth=2)
        else:
            e2e_step()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        if dlog_staging2 is not None:
            (core, cost, raw), _ = run_pipelined(dlog_staging2, e2e_steps, e2e_fill, depth=2)
        else:
            for _ in range(e2e_steps):
                core, cost, raw = e2e_step()
        e1.record()
        barrier()
        ms_e2e = e0.elapsed_time(e1)

    if dist is not None:
        t = torch.tensor([ms_total, ms_e2e, ms_eager], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total, ms_e2e, ms_eager = float(t[0]), float(t[1]), float(t[2])

    cfg5 = None
    if world > 1 and not args.no_cfg5:
        cfg5 = cfg5_block(args, device, rank, world, dist, cold)
    sweep = None
    if not args.no_sweep:
        sweep = sweep_block(args, device, rank, world, dist)
    host_all = None
    if rank == 0 and host is not None and not args.no_parity:
        # the whole (sharded) log on the host of rank 0: every shard is regenerated from its seed
        import numpy as np
        shards = [host] + [synth.request_columns(n, seed=1234 + r, t_epoch_ms=1.7e12 + r * 3.6e6)
                           for r in range(1, world)]
        host_all = ({k: np.concatenate([sh[k] for sh in shards]) for k in host} if world > 1 else host)
        del shards
    if rank == 0:
        peak, peak_src = measured_peaks()
        # algorithmic bytes per launch of each streaming kernel (DESIGN.md "Kernels")
        # (56 B/row of input read once + the 1 B/row flag output; the 8-byte
        # digit records the one-pass kernel writes are workspace and earn no credit)
        alg = {"log_fused": 57.0 * n, "log_scan": 57.0 * n, "select_hist": 49.0 * n, "select_compact": 9.0 * n,
               "select_refine_hist": 9.0 * n}
        kernels = {}
        for name, (ms, cnt) in stage_ms.items():
            avg = ms / cnt
            k = {"avg_ms": avg, "launches": cnt}
            if name in alg:
                k["algorithmic_bytes"] = alg[name]
                k["achieved_gbs"] = alg[name] / (avg * 1e-3) / 1e9
                k["frac"] = k["achieved_gbs"] / peak
            kernels[name] = k
        dom = max((k for k in kernels if k in alg), key=lambda k: kernels[k]["avg_ms"])
        traffic = None
        try:   # dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu capture
            with open(os.path.join(REPO, "profiles", "ncu_traffic.json")) as f:
                tj = json.load(f)
            if tj.get("rows") == n:
                traffic = tj["kernels"].get(dom)
        except Exception:
            traffic = None
        d2h = (raw.hist_counts.nbytes + (raw.series.nbytes if raw.series is not None else 0)
               + 48 * 32 + 312)
        value = n * world * args.steps / (ms_total * 1e-3)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, world),
            "e2e": ({"value": n * world * e2e_steps / (ms_e2e * 1e-3), "unit": UNIT, "steps": e2e_steps,
                     "h2d_bytes_per_step": ROW_BYTES * n, "d2h_bytes_per_step": int(d2h)}
                    if pinned is not None else None),
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": kernels[dom]["achieved_gbs"],
                         "peak": peak, "unit": "GB/s", "frac": kernels[dom]["frac"], "traffic": traffic,
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": kernels[dom]["algorithmic_bytes"]},
            "kernels": kernels,
            "launch_mode": (f"timed region: CUDA-graph replay of the kernel sequence, {len(pipe) if pipe else 1} analyses in flight on "
                            "their own streams (host post-processing of step k and the few-CTA kernels of one "
                            "analysis overlap the streaming passes of the others: the streaming kernels leave "
                            "kvm_tail_slots CTAs out of their grids); per-kernel durations from "
                            "the same K steps launched eagerly on one stream with CUDA events"
                            if pipelined and world == 1 else
                            ("timed region: CUDA-graph replay (kernels + exchanges), " +
                             (f"{len(pipe)} analyses in flight on their own streams, " if pipelined else "one analysis at a time, ") +
                             "one host sync per step; per-kernel durations from the same K steps launched eagerly "
                             "with CUDA events")),
            "exchange": (None if world == 1 else
                         ("peer memory over NVLink (kvm_peer_* kernels)" if analyzer.exchange != "nccl"
                          else "NCCL collectives")),
            "ms_per_step_eager_instrumented": ms_eager / args.steps,
            "effective_gbs_whole_step": ROW_BYTES * n / (ms_total / args.steps * 1e-3) / 1e9,
            "check": {"p50_ms": core["p50_ms"], "p99_ms": core["p99_ms"],
                      "cost_per_1k_tokens": cost["cost_per_1k_tokens"]},
        }
        if not args.no_parity and host_all is not None:
            line["parity"] = parity_block(core, report.trim_series(raw), host_all, cold)
        if cfg5 is not None:
            line["cfg5"] = cfg5
        if sweep is not None:
            line["sweep"] = sweep
        if world == 1 and not args.no_percentile_path:
            line["percentile_path"] = percentile_path_block(device, peak)
        if world == 1 and not args.no_cpu_baseline and host is not None:
            rows = min(n, 3_000_000)
            rate, secs = cpu_port_rate({k: v for k, v in host.items()}, cold, rows)
            line["cpu_baseline"] = {
                "value": rate, "unit": UNIT, "cores": 1, "kind": "port",
                "sample": f"first {rows} of {n} rows, one pass of oracle/py_port.analyze + cost rollup "
                          f"({secs:.1f} s, single thread: the reference is single-threaded)"}
        print(json.dumps(line), flush=True)
    if dist is not None:
        # graphs that hold captured collectives must be gone before the communicator is torn down
        # (destroying the group with them alive hung this stack)
        analyzer.close()
        for an in pipe or ():
            an.close()
        barrier()
        dist.destroy_process_group()


def default_depth(world):
    return 6 if world <= 2 else 3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=10_000_000, help="rows per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle comparison of the timed log")
    ap.add_argument("--no-cfg5", action="store_true", help="N>1: skip the 1.25e8-rows-per-GPU block")
    ap.add_argument("--cfg5-rows", type=int, default=125_000_000, help="rows per GPU of the cfg5 block")
    ap.add_argument("--pipeline-depth", type=int, default=0,
                    help="analyses kept in flight on their own streams (>= 2); 0 = 6 on one or two GPUs, 3 beyond "
                         "(measured: with more ranks the spinning barrier kernels of many analyses in flight "
                         "get in each other's way)")
    ap.add_argument("--no-sweep", action="store_true", help="skip the configs[2] (sweep roll-up) block")
    ap.add_argument("--sweep-runs", type=int, default=512, help="runs per GPU of the sweep block (250k rows each)")
    ap.add_argument("--no-percentile-path", action="store_true",
                    help="N=1: skip the 16 B/record percentile-path block")
    ap.add_argument("--port", action="store_true",
                    help="--impl reference: time the pure-Python port even when oracle/_ref/ is present")
    ap.add_argument("--device-data", action="store_true", help="generate the shard on the GPU (large --rows)")
    ap.add_argument("--no-e2e", action="store_true", help="with --device-data: skip the host-buffer leg")
    ap.add_argument("--eager", action="store_true", help="launch every step eagerly (no CUDA-graph replay)")
    ap.add_argument("--no-pipeline", action="store_true", help="N>1: one analysis in flight at a time")
    ap.add_argument("--refine-min-keys", type=int, default=None,
                    help="candidate total above which a refine round runs (default: library policy)")
    ap.add_argument("--exchange", default="auto", choices=["auto", "peer", "nccl"],
                    help="N>1: how the ranks exchange intermediate results")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.pipeline_depth <= 0:
        args.pipeline_depth = default_depth(int(os.environ.get("WORLD_SIZE", "1")))
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
