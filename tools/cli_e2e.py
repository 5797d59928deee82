#!/usr/bin/env python3
"""File-to-file timing of the drop-in analysis on a generated requests.csv (default 1M rows):
native CSV ingest -> H2D -> GPU analysis -> results + classified CSV, next to the pure-Python
restatement of the reference (oracle/py_port) on the same file, same box, one core.

  python tools/cli_e2e.py [--rows 1000000] [--dir /tmp/kvm_cli]
"""
import argparse
import csv
import json
import os
import sys
import time

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)

import numpy as np


def write_csv(path, n):
    from kvmini_b200 import synth
    c = synth.request_columns(n, seed=3)
    with open(path, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["id", "scheduled_ms", "start_ms", "ttfb_ms", "tllt_ms", "latency_ms", "status", "prompt_tokens",
                    "completion_tokens", "total_tokens", "error", "trace_id"])
        st = c["status"]
        for i in range(n):
            ok = st[i] == 200
            w.writerow([f"req-{i}", "%.3f" % (i * 0.36), "%.3f" % c["start_ms"][i],
                        "%.3f" % c["ttfb_ms"][i] if ok else "", "%.3f" % c["tllt_ms"][i] if ok else "",
                        "%.3f" % c["latency_ms"][i], int(st[i]), 12, int(c["completion_tokens"][i]) if ok else "",
                        int(c["total_tokens"][i]) if ok else "", "" if ok else "HTTP 500", f"{i:016x}"])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=1_000_000)
    ap.add_argument("--dir", default="/tmp/kvm_cli")
    ap.add_argument("--no-cpu", action="store_true")
    a = ap.parse_args()
    import torch
    from kvmini_b200 import _native as N
    from kvmini_b200 import analyze_api, columns, engine, report
    os.makedirs(a.dir, exist_ok=True)
    src = os.path.join(a.dir, "requests.csv")
    write_csv(src, a.rows)
    size_mb = os.path.getsize(src) / 1e6
    analyzer = engine.LogAnalyzer()
    analyzer.analyze(engine.DeviceLog.from_host(columns.load_requests_csv(src), analyzer.device), [])   # warm-up
    t = [time.perf_counter()]
    log = columns.load_requests_csv(src, keep_rows=True)
    columns.check_analyzable(log)
    t.append(time.perf_counter())
    dlog = engine.DeviceLog.from_host(log, analyzer.device)
    cold = [float(log.cols["start_ms"].min()) / 1000.0 + 5.0]
    raw = analyzer.analyze(dlog, cold)
    core = report.core_results(raw, len(cold))
    series = report.trim_series(raw)
    is_cold = (raw.flags.cpu().numpy() & N.F_COLD) != 0
    torch.cuda.synchronize()
    t.append(time.perf_counter())
    analyze_api._write_classified_csv(os.path.join(a.dir, "requests_classified.csv"), log, is_cold)
    with open(os.path.join(a.dir, "results.json"), "w") as f:
        json.dump(dict(core, timeseries=series), f, indent=2)
    t.append(time.perf_counter())
    out = {"mode": "cli_e2e", "rows": a.rows, "csv_mb": size_mb, "native_ingest": log.native is not None,
           "ingest_s": t[1] - t[0], "h2d_analyze_results_s": t[2] - t[1], "write_outputs_s": t[3] - t[2],
           "total_s": t[3] - t[0], "records_per_s_file_to_file": a.rows / (t[3] - t[0]), "p95_ms": core["p95_ms"]}
    if not a.no_cpu:
        from oracle import py_port
        c0 = time.perf_counter()
        rows = columns.read_rows(src)                 # csv.DictReader + float(): analyze.py:34-56 restated
        c1 = time.perf_counter()
        plog = columns.columns_from_rows(rows)        # row dicts -> columns (not timed on the reference side:
        cols = {k: v.tolist() for k, v in plog.cols.items()}   # it works on the dicts directly)
        c1b = time.perf_counter()
        ref = py_port.analyze(cols, cold)
        c2 = time.perf_counter()
        c2 -= c1b - c1
        assert ref["p95_ms"] == core["p95_ms"] and ref["p50_ms"] == core["p50_ms"]
        out["cpu_port"] = {"parse_s": c1 - c0, "analyze_s": c2 - c1, "total_s": c2 - c0,
                           "records_per_s": a.rows / (c2 - c0), "cores": 1,
                           "note": "classified-CSV writing not included on this side"}
        out["speedup_file_to_file"] = (c2 - c0) / (t[3] - t[0])
    print(json.dumps(out))


if __name__ == "__main__":
    main()
