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
