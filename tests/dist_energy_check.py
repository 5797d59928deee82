#!/usr/bin/env python3
"""Run under torchrun (one rank per GPU): the multi-GPU energy path == the oracle on all traces.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
      --master-port 29512 tests/dist_energy_check.py [samples] [intervals] [--bench]
trace-sharded: rank r holds trace r; window Wh and per-interval energies must equal the rank-ordered
sums of the numpy oracle (<= 1e-12 relative; empty intervals exactly 0).  time-sharded: one trace
cut into `world` shards with a one-sample halo == the whole trace.  --bench: configs[3] sizes
(86.4M samples per rank, 5e7 intervals), device-generated, timing only + size-independent checks."""
import json
import math
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
    m = int(args[0]) if args else 400_000
    n = int(args[1]) if len(args) > 1 else 50_000
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from kvmini_b200 import dist_energy, energy_api, synth
    from oracle import np_port
    failures = []
    if not bench:
        hz = 100.0
        traces = [synth.power_trace(m, gpu=g, hz=hz, null_p=0.002) for g in range(world)]
        span_ms = m / hz * 1000.0
        cols = synth.request_columns(n, seed=21, span_ms=span_ms * 1.04, t_epoch_ms=1.7e12 - 0.02 * span_ms)
        start, lat = cols["start_ms"], cols["latency_ms"] * 3.0
        tr = energy_api.PowerTrace(*traces[rank], device=dev)
        sh = dist_energy.TraceShardedEnergy(tr, dev)
        t0 = float(traces[0][0][m // 10])
        wins = [(t0, t0 + 0.5 * m / hz), (t0 + 1.0, t0 + 2.0)]
        got = sh.window_wh(wins)
        exp = [math.fsum(np_port.trapezoid_wh(ts, w, a, b, exact=True) for ts, w in traces) for a, b in wins]
        for g, e in zip(got, exp):
            if not abs(g - e) <= 1e-12 * abs(e):
                failures.append(("window", g, e))
        e_all = sh.interval_energy(torch.from_numpy(start).to(dev), torch.from_numpy(lat).to(dev)).cpu().numpy()
        ref = np.zeros(n)
        for ts, w in traces:
            ref = ref + np_port.interval_energy(ts, w, start, lat)
        rel = np.abs(e_all - ref) / np.maximum(np.abs(ref), 1e-300)
        if not (np.all((rel <= 1e-12) | (e_all == ref)) and np.array_equal(e_all == 0.0, ref == 0.0)):
            failures.append(("intervals", float(rel.max())))
        # every rank holds the same bits
        chk = torch.from_numpy(e_all).to(dev)
        first = chk.clone()
        dist.broadcast(first, 0)
        if not torch.equal(first, chk):
            failures.append("ranks disagree")
        # time-sharded
        ts, w = synth.power_trace(m, gpu=9, hz=hz, null_p=0.01)
        cut = m // world
        lo, hi = rank * cut, ((rank + 1) * cut if rank < world - 1 else m)
        ts_r, w_r = ts[lo:hi], w[lo:hi]
        nn = np.flatnonzero(~np.isnan(w_r))
        last = (float(ts_r[nn[-1]]), float(w_r[nn[-1]])) if len(nn) else (0.0, math.nan)
        tsh = dist_energy.TimeShardedEnergy(ts_r, w_r, dev, lambda a, b: energy_api.PowerTrace(a, b, device=dev), last)
        wins = [(float(ts[0]), float(ts[-1])), (float(ts[max(cut - 70, 0)]), float(ts[min(cut + 70, m - 1)]))]
        got = tsh.window_wh(wins)
        exp = [np_port.trapezoid_wh(ts, w, a, b, exact=True) for a, b in wins]
        for g, e in zip(got, exp):
            if not abs(g - e) <= 1e-12 * abs(e):
                failures.append(("time-sharded", g, e))
    else:
        m, n = 86_400_000, 50_000_000
        g = torch.Generator(device=dev)
        g.manual_seed(77 + rank)
        k = torch.arange(m, dtype=torch.float64, device=dev)
        t0 = 1.7e9
        ts = t0 + k / 1000.0 + torch.rand(m, generator=g, dtype=torch.float64, device=dev) * 5e-5
        util = 0.5 + 0.5 * torch.sin(2.0 * math.pi * (k / 1000.0) / 3600.0 + rank)
        w = (400.0 + 300.0 * util + torch.randn(m, generator=g, dtype=torch.float64, device=dev) * 15.0).clamp_(60.0, 1000.0)
        del k, util
        g.manual_seed(5)
        start = torch.sort(t0 * 1000.0 + torch.rand(n, generator=g, dtype=torch.float64, device=dev) * (m * 1.0)).values
        lat = torch.exp(torch.randn(n, generator=g, dtype=torch.float64, device=dev) * 0.6 + 5.5)

        class DevTrace(energy_api.PowerTrace):
            def __init__(self, ts, w):
                self.device, self.lib, self.m = dev, energy_api.N.lib(), int(ts.shape[0])
                self.ts_loaded = self.ts = ts
                self.watts_loaded = self.watts = w
                self._ws = torch.empty(self.lib.kvm_energy_workspace_bytes(self.m), dtype=torch.uint8, device=dev)
                self._stats_d = torch.zeros(128, dtype=torch.uint8, device=dev)
                self._prefix = None

        tr = DevTrace(ts, w)
        sh = dist_energy.TraceShardedEnergy(tr, dev)
        win = [(t0 + 0.1 * m / 1000.0, t0 + 0.9 * m / 1000.0)]
        tr.prefix()

        def step():
            wh = sh.window_wh(win)
            e = sh.interval_energy(start, lat)
            return wh, e

        step()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        steps = 3
        e0.record()
        for _ in range(steps):
            wh, e = step()
        e1.record()
        dist.barrier()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / steps], dtype=torch.float64, device=dev)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        # property: the interval energies of intervals that tile a span add up to (about) the window
        # energy of that span - checked on rank 0's own trace alone elsewhere; here: finite, >= 0
        if not (bool(torch.isfinite(e).all()) and bool((e >= 0).all()) and wh[0] > 0):
            failures.append("bench sanity")
        if rank == 0:
            print(json.dumps({"mode": "energy_dist", "config": "configs[3]: one 86.4M-sample trace per rank, 5e7 intervals",
                              "n_gpus": world, "ms_per_step": float(ms[0]),
                              "samples_per_s": world * m / float(ms[0]) * 1e3, "intervals_per_s": n / float(ms[0]) * 1e3,
                              "Wh_window_all_gpus": wh[0],
                              "step": "window trapezoid of every trace (rank-ordered exact sum) + per-interval join of "
                                      "every trace + all-to-all of the slices + rank-ordered add + all-gather"}))
    flag = torch.tensor([len(failures)], device=dev)
    dist.all_reduce(flag)
    if rank == 0:
        print("DIST_ENERGY_CHECK", "OK" if int(flag.item()) == 0 else "FAIL", f"world={world} samples={m} intervals={n}")
    if failures:
        print(f"rank {rank}:", failures[:6])
    dist.destroy_process_group()
    sys.exit(1 if failures else 0)


if __name__ == "__main__":
    main()
