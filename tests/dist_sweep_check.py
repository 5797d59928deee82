#!/usr/bin/env python3
"""Run under torchrun (one rank per GPU): the sweep roll-up with whole runs per GPU == the same
runs analysed on one GPU (and, for a few runs, == the oracle).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
      --master-port 29513 tests/dist_sweep_check.py [runs] [rows_per_run] [--bench]
--bench: configs[2] share per GPU (512 runs x 250k rows), device-generated, timing + counts only."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    bench = "--bench" in sys.argv
    runs = int(args[0]) if args else 37
    rows = int(args[1]) if len(args) > 1 else 6000
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from kvmini_b200 import engine, sweep_api, synth
    keys = sweep_api.SweepAnalyzer.SWEEP_KEYS
    sh = sweep_api.ShardedSweep(dev)
    failures = []
    if not bench:
        cols, offsets, colds = synth.sweep_columns(runs=runs, rows_per_run=rows)
        cols["ttfb_ms"][::53] = np.nan
        colds = [c.tolist() for c in colds]
        mine = sh.my_runs(offsets)
        lo, hi = int(offsets[mine.start]), int(offsets[mine.stop])
        local_cols = {k: torch.from_numpy(np.ascontiguousarray(v[lo:hi])).to(dev) for k, v in cols.items()}
        local_offs = (np.asarray(offsets[mine.start:mine.stop + 1]) - lo).tolist()
        cuts = sweep_api.partition_runs(offsets, world)
        got = sh.sweep_table(engine.DeviceLog.from_tensors(local_cols), local_offs, colds[mine.start:mine.stop],
                             counts=[cuts[r + 1] - cuts[r] for r in range(world)])
        if rank == 0:
            whole = {k: torch.from_numpy(v).to(dev) for k, v in cols.items()}
            exp = sweep_api.SweepAnalyzer(dev).sweep_table(engine.DeviceLog.from_tensors(whole), offsets.tolist(), colds)
            for k in keys:
                same = np.array_equal(got[k], exp[k], equal_nan=True) and np.array_equal(got["is_none"][k], exp["is_none"][k])
                if not same:
                    failures.append(k)
            from oracle import np_port
            for r in (0, runs // 2, runs - 1):
                sub = {k: v[offsets[r]:offsets[r + 1]] for k, v in cols.items()}
                ref = np_port.analyze(sub, colds[r])
                for k in ("p50_ms", "p95_ms", "p99_ms", "ttft_p50_ms", "ttft_p95_ms", "error_rate"):
                    if float(got[k][r]) != ref[k]:
                        failures.append((r, k, float(got[k][r]), ref[k]))
    else:
        per = runs // world if runs >= world else 1
        n = per * rows
        g = {k: torch.empty(n, dtype=torch.float64, device=dev) for k in ("start_ms", "latency_ms", "ttfb_ms", "tllt_ms")}
        g.update({k: torch.empty(n, dtype=torch.int64, device=dev) for k in ("status", "completion_tokens", "total_tokens")})
        chunk = 64
        for r0 in range(0, per, chunk):
            r1 = min(per, r0 + chunk)
            c = synth.request_columns_device(rows * (r1 - r0), dev, seed=1000 + rank * per + r0, span_ms=rows * 25.0 * (r1 - r0),
                                             t_epoch_ms=1.7e12 + (rank * per + r0) * rows * 25.0)
            for k in g:
                g[k][r0 * rows:r1 * rows] = c[k]
            del c
        offs = (np.arange(per + 1, dtype=np.int64) * rows).tolist()
        dlog = engine.DeviceLog.from_tensors(g)
        sh.sweep_table(dlog, offs)                      # warm-up
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        table = sh.sweep_table(dlog, offs)
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1), sh.local.last_kernel_ms], device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        if len(table["p50_ms"]) != per * world or not np.all(table["p50_ms"] <= table["p99_ms"]):
            failures.append("table shape / order")
        if rank == 0:
            tot = per * world * rows
            print(json.dumps({"mode": "sweep_sharded", "n_gpus": world, "runs": per * world, "rows_per_run": rows, "rows": tot,
                              "ms_table_max_over_ranks": float(ms[0]), "ms_kernel_max_over_ranks": float(ms[1]),
                              "records_per_s_table": tot / float(ms[0]) * 1e3, "records_per_s_kernel": tot / float(ms[1]) * 1e3,
                              "fallback_runs_rank0": sh.local.fallback_runs}))
    bad = torch.tensor([len(failures)], device=dev)
    dist.all_reduce(bad)
    if rank == 0:
        print("DIST_SWEEP_CHECK OK" if int(bad) == 0 else f"DIST_SWEEP_CHECK FAILED {failures[:6]}")
    dist.destroy_process_group()
    sys.exit(0 if int(bad) == 0 else 1)


if __name__ == "__main__":
    main()
