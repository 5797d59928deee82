"""N>1 host path on CPU: two gloo ranks combine their shard statistics exactly like the NCCL
ranks do on the GPU box (dist_api.combine_stats), and the sharding helpers keep order."""
import math
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import REPO


def np_key(v: np.ndarray) -> np.ndarray:
    b = v.view(np.uint64)
    return np.where(b >> np.uint64(63), ~b, b | np.uint64(1 << 63))


def local_stats(cols, cold):
    """kvm_log_stats of one shard, computed with numpy (test-side stand-in for K1)."""
    from kvmini_b200 import _native as N
    from oracle import np_port
    s = N.LogStatsC()
    ok = cols["status"] == 200
    is_cold = np_port.classify_cold(cols["start_ms"], cold)
    s.n_total, s.n_success = len(ok), int(ok.sum())
    s.n_cold, s.n_cold_success = int(is_cold.sum()), int((is_cold & ok).sum())
    s.sum_completion_tokens = int(cols["completion_tokens"][ok].sum())
    s.sum_total_tokens = int(cols["total_tokens"][ok].sum())
    s.sum_total_tokens_cold = int(cols["total_tokens"][ok & is_cold].sum())
    lat, tf, tl, ct = cols["latency_ms"], cols["ttfb_ms"], cols["tllt_ms"], cols["completion_tokens"]
    gmask = ok & ~np.isnan(tf) & ~np.isnan(tl) & (ct > 0)
    s.n_gen_rows = int(gmask.sum())
    gen = np.where(gmask, tl - tf, np.nan)
    per = gen / np.where(ct > 0, ct, 1)
    sets = [(lat, ok & is_cold, 0), (lat, ok & ~is_cold, 0), (tf, ok & is_cold, 1), (tf, ok & ~is_cold, 1),
            (tl, ok, 2), (gen, gmask, 3), (per, gmask, 4)]
    kmin = [2**64 - 1] * 8
    kmax = [0] * 8
    for b, (v, m, c) in enumerate(sets):
        vv = v[m & ~np.isnan(v)]
        s.sets.count[b] = len(vv)
        s.sets.set_col[b] = c
        if len(vv):
            k = np_key(np.ascontiguousarray(vv))
            kmin[c], kmax[c] = min(kmin[c], int(k.min())), max(kmax[c], int(k.max()))
    for c in range(8):
        s.sets.kmin[c], s.sets.kmax[c] = kmin[c], kmax[c]
    s.sets.nsets, s.sets.ncols = 7, 5
    s.min_start_ms = float(np.nanmin(cols["start_ms"]))
    s.max_end_ms = float(np.nanmax(cols["start_ms"] + cols["latency_ms"]))
    tt = tf[ok & ~np.isnan(tf)]
    s.ttfb_sum_hi = math.fsum(tt.tolist())
    s.ttfb_sum_lo = 0.0
    s.lat_sum_hi = math.fsum(lat[ok & ~np.isnan(lat)].tolist())
    s.lat_sum_lo = 0.0
    return s


def _worker(rank, world, port, ret):
    sys.path.insert(0, REPO)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from kvmini_b200 import dist_api, synth
    full = synth.request_columns(20_000, seed=3, span_ms=400_000.0)
    full["ttfb_ms"][::97] = np.nan
    cold = [float(full["start_ms"].min()) / 1000.0 + 7.0, float(full["start_ms"].min()) / 1000.0 + 300.0]
    lo, hi = rank * 20_000 // world, (rank + 1) * 20_000 // world
    shard = {k: v[lo:hi] for k, v in full.items()}
    got = dist_api.combine_stats(local_stats(shard, cold), torch.device("cpu"))
    exp = local_stats(full, cold)
    bad = []
    for f in dist_api._SUM_FIELDS + ("min_start_ms", "max_end_ms"):
        if getattr(got, f) != getattr(exp, f):
            bad.append(f)
    for b in range(8):
        if got.sets.count[b] != exp.sets.count[b] or got.sets.kmin[b] != exp.sets.kmin[b] \
                or got.sets.kmax[b] != exp.sets.kmax[b]:
            bad.append(f"set{b}")
    if abs((got.ttfb_sum_hi + got.ttfb_sum_lo) - exp.ttfb_sum_hi) > 1e-12 * abs(exp.ttfb_sum_hi):
        bad.append("ttfb_sum")
    ret[rank] = bad
    dist.destroy_process_group()


def test_combine_stats_world2_gloo():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(2, port, ret), nprocs=2, join=True)
    assert dict(ret) == {0: [], 1: []}


def test_key_sign_mapping_keeps_order():
    from kvmini_b200 import dist_api
    keys = [0, 1, (1 << 63) - 1, 1 << 63, (1 << 64) - 1, 0x4060000000000000 ^ (1 << 63)]
    signed = [dist_api._key_to_signed(k) for k in keys]
    assert signed == sorted(signed) or sorted(keys) == [k for _, k in sorted(zip(signed, keys))]
    assert [dist_api._signed_to_key(v) for v in signed] == keys
    assert all(-(1 << 63) <= v < (1 << 63) for v in signed)


def test_rank_ordered_dd_total_is_exact():
    from kvmini_b200 import dist_api
    parts = [(1e16, 1.0), (-1e16, 0.25), (3.0, 1e-20)]
    hi, lo = dist_api._exact_dd_total(parts)
    assert hi == 4.25 and abs(lo - 1e-20) < 1e-35


# ---- multi-GPU energy host logic (dist_energy.py) over gloo, numpy stand-ins for the kernels ----
class _NpTrace:
    """Test-side stand-in for energy_api.PowerTrace: same two methods, oracle arithmetic."""

    def __init__(self, ts, watts):
        self.ts, self.watts = np.asarray(ts, dtype=np.float64), np.asarray(watts, dtype=np.float64)

    def trapezoid_dd(self, windows):
        from oracle import np_port
        return [(np_port.trapezoid_wh(self.ts, self.watts, a, b, exact=True), 0.0) for a, b in windows]

    def interval_energy(self, start_ms, latency_ms):
        from oracle import np_port
        return torch.from_numpy(np_port.interval_energy(self.ts, self.watts, start_ms.numpy(), latency_ms.numpy()))


def _energy_worker(rank, world, port, ret):
    sys.path.insert(0, REPO)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from kvmini_b200 import dist_energy, synth
    from oracle import np_port
    dev = torch.device("cpu")
    bad = []
    # (1) trace-sharded: rank r owns trace r; windows and per-interval energies are rank-ordered sums
    traces = [synth.power_trace(20_000, gpu=g, hz=50.0, null_p=0.01) for g in range(world)]
    cols = synth.request_columns(3001, seed=4, span_ms=20_000 / 50.0 * 1000.0)
    start, lat = torch.from_numpy(cols["start_ms"]), torch.from_numpy(cols["latency_ms"] * 5.0)
    sh = dist_energy.TraceShardedEnergy(_NpTrace(*traces[rank]), dev)
    t0 = float(traces[0][0][100])
    wins = [(t0, t0 + 200.0), (t0 + 10.0, t0 + 11.0)]
    got = sh.window_wh(wins)
    exp = [math.fsum(np_port.trapezoid_wh(ts, w, a, b, exact=True) for ts, w in traces) for a, b in wins]
    if not all(abs(g - e) <= 1e-15 * abs(e) for g, e in zip(got, exp)):
        bad.append(("window_wh", got, exp))
    e = sh.interval_energy(start, lat).numpy()
    ref = np.zeros(len(cols["start_ms"]))
    for ts, w in traces:
        ref = ref + np_port.interval_energy(ts, w, cols["start_ms"], cols["latency_ms"] * 5.0)
    if not np.array_equal(e, ref):     # same additions in the same (rank) order: bit-equal
        bad.append(("interval_energy", float(np.abs(e - ref).max())))
    # (2) time-sharded: one trace cut in `world` pieces, the cut inside a run of nulls
    ts, w = synth.power_trace(30_000, gpu=7, hz=20.0, null_p=0.0)
    cut = 30_000 // world
    w[cut - 3:cut + 2] = np.nan                    # the segment across the cut spans 6 samples
    lo, hi = rank * cut, (rank + 1) * cut if rank < world - 1 else 30_000
    ts_r, w_r = ts[lo:hi], w[lo:hi]
    nn = np.flatnonzero(~np.isnan(w_r))
    last = (float(ts_r[nn[-1]]), float(w_r[nn[-1]])) if len(nn) else (0.0, math.nan)
    tsh = dist_energy.TimeShardedEnergy(ts_r, w_r, dev, _NpTrace, last)
    if (rank == 0) != (tsh.halo is None):
        bad.append(("halo", tsh.halo))
    wins = [(float(ts[0]), float(ts[-1])), (float(ts[cut - 50]), float(ts[cut + 50])), (float(ts[5]), float(ts[50]))]
    got = tsh.window_wh(wins)
    exp = [np_port.trapezoid_wh(ts, w, a, b, exact=True) for a, b in wins]
    if not all(abs(g - e) <= 1e-14 * abs(e) for g, e in zip(got, exp)):
        bad.append(("time_sharded", got, exp))
    ret[rank] = bad
    dist.destroy_process_group()


def test_energy_sharding_world2_gloo():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_energy_worker, args=(2, port, ret), nprocs=2, join=True)
    assert dict(ret) == {0: [], 1: []}


def test_halo_skips_shards_without_samples():
    from kvmini_b200 import dist_energy
    nan = math.nan
    last = [(10.0, 5.0), (0.0, nan), (30.0, 7.0), (0.0, nan)]
    assert dist_energy.halo_for_rank(last, 0) is None
    assert dist_energy.halo_for_rank(last, 1) == (10.0, 5.0)
    assert dist_energy.halo_for_rank(last, 2) == (10.0, 5.0)      # shard 1 had no non-null sample
    assert dist_energy.halo_for_rank(last, 3) == (30.0, 7.0)
    assert dist_energy.rank_ordered_total([(1e16, 1.0), (-1e16, 0.5)]) == 1.5


# ---- sweep roll-up across ranks (sweep_api.partition_runs / gather_sweep_tables) ----------------
def test_partition_runs_whole_runs_balanced_and_ordered():
    from kvmini_b200 import sweep_api
    offs = np.arange(4097, dtype=np.int64) * 250_000              # configs[2]
    assert sweep_api.partition_runs(offs, 8) == [512 * r for r in range(9)]
    rng = np.random.default_rng(2)
    sizes = rng.integers(0, 5000, 301)
    sizes[17] = 0
    sizes[100] = 400_000                                           # one run larger than an even share
    offs = np.concatenate([[0], np.cumsum(sizes)])
    for world in (1, 2, 3, 8):
        cuts = sweep_api.partition_runs(offs, world)
        assert len(cuts) == world + 1 and cuts[0] == 0 and cuts[-1] == 301
        assert all(a <= b for a, b in zip(cuts, cuts[1:]))         # contiguous groups of whole runs, in order
        rows = [int(offs[b] - offs[a]) for a, b in zip(cuts, cuts[1:])]
        assert sum(rows) == int(offs[-1])
        assert max(rows) <= int(offs[-1]) // world + 400_000       # never worse than one run off the even share
    assert sweep_api.partition_runs([0], 4) == [0, 0, 0, 0, 0]
    assert sweep_api.partition_runs([0, 10], 4)[-1] == 1


def _sweep_worker(rank, world, port, ret):
    sys.path.insert(0, REPO)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from kvmini_b200 import sweep_api
    keys = sweep_api.SweepAnalyzer.SWEEP_KEYS
    offs = np.concatenate([[0], np.cumsum([5, 0, 7, 1000, 3, 900, 2])])
    cuts = sweep_api.partition_runs(offs, world)
    mine = range(cuts[rank], cuts[rank + 1])
    # stand-in for the per-run analysis: a value that names (key, run)
    local = {k: np.array([1000.0 * i + r for r in mine]) for i, k in enumerate(keys)}
    local["is_none"] = {k: np.array([(r + i) % 3 == 0 for r in mine], dtype=bool) for i, k in enumerate(keys)}
    got = sweep_api.gather_sweep_tables(local, keys, [cuts[r + 1] - cuts[r] for r in range(world)])
    runs = len(offs) - 1
    bad = [k for i, k in enumerate(keys)
           if got[k].tolist() != [1000.0 * i + r for r in range(runs)]
           or got["is_none"][k].tolist() != [(r + i) % 3 == 0 for r in range(runs)]]
    ret[rank] = bad
    dist.destroy_process_group()


def test_sweep_tables_gather_in_run_order_world2_gloo():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_sweep_worker, args=(2, port, ret), nprocs=2, join=True)
    assert dict(ret) == {0: [], 1: []}
