"""Segmented (sweep) analysis: every run equals the oracle on that run alone."""
import numpy as np
import pytest

from conftest import diff_keys

pytestmark = pytest.mark.gpu
MEAN_TOL = {"time_to_first_token_ms": 1e-12}


def _check(cols, offsets, colds, expect_fallback=None):
    import torch
    from kvmini_b200 import engine, sweep_api
    from oracle import np_port
    dev = {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in cols.items()}
    sw = sweep_api.SweepAnalyzer()
    got = sw.rollup(engine.DeviceLog.from_tensors(dev), offsets, colds)
    assert len(got) == len(offsets) - 1
    for r in range(len(offsets) - 1):
        sub = {k: v[offsets[r]:offsets[r + 1]] for k, v in cols.items()}
        exp = np_port.analyze(sub, colds[r] if colds is not None else [])
        d = diff_keys(got[r], {k: exp[k] for k in got[r]}, rel=MEAN_TOL)
        assert d == [], (r, d[:5])
    if expect_fallback is not None:
        assert (sw.fallback_runs > 0) == expect_fallback
    return sw


def test_sweep_config3_shape_small():
    from kvmini_b200 import synth
    cols, offsets, colds = synth.sweep_columns(runs=40, rows_per_run=20_000)
    _check(cols, offsets.tolist(), [c.tolist() for c in colds], expect_fallback=False)


def test_sweep_ragged_runs_with_nans_and_empty_runs():
    from kvmini_b200 import synth
    rng = np.random.default_rng(4)
    sizes = [1, 0, 2, 3, 999, 0, 12_345, 250_001, 7]
    cols = synth.request_columns(sum(sizes), seed=31, span_ms=5.0e6)
    n = sum(sizes)
    cols["ttfb_ms"][rng.integers(0, n, n // 50)] = np.nan
    cols["latency_ms"][rng.integers(0, n, n // 70)] = np.nan
    cols["completion_tokens"][rng.integers(0, n, n // 60)] = 0
    cols["status"][-7:] = 500                                  # last run: all failures
    offsets = np.concatenate([[0], np.cumsum(sizes)]).tolist()
    colds = []
    for r, sz in enumerate(sizes):
        if sz == 0:
            colds.append([])
            continue
        t0 = float(cols["start_ms"][offsets[r]]) / 1000.0
        colds.append([t0 + 1.0, t0 + 100.0] if r % 2 == 0 else [])
    # empty runs: the reference exits 1 on an empty CSV, so only non-empty runs are compared
    import torch
    from kvmini_b200 import engine, sweep_api
    from oracle import np_port
    dev = {k: torch.from_numpy(v).cuda() for k, v in cols.items()}
    sw = sweep_api.SweepAnalyzer()
    got = sw.rollup(engine.DeviceLog.from_tensors(dev), offsets, colds)
    for r, sz in enumerate(sizes):
        if sz == 0:
            assert got[r]["requests"] == {"total": 0, "success": 0}
            continue
        sub = {k: v[offsets[r]:offsets[r + 1]] for k, v in cols.items()}
        exp = np_port.analyze(sub, colds[r])
        assert diff_keys(got[r], {k: exp[k] for k in got[r]}, rel=MEAN_TOL) == [], r


def test_sweep_concentrated_run_falls_back_to_flat_path():
    from kvmini_b200 import synth
    cols, offsets, colds = synth.sweep_columns(runs=3, rows_per_run=60_000)
    cols["latency_ms"][60_000:120_000:2] = 30000.0   # run 1: half the latencies identical -> bucket > on-chip arena
    _check(cols, offsets.tolist(), [c.tolist() for c in colds], expect_fallback=True)


def test_sweep_table_equals_per_run_rollup():
    """The vectorised sweep columns (grid-sweep.sh:137-160) are the same doubles as the per-run dicts,
    including empty runs, all-failure runs, runs without ttfb values and a run on the flat path."""
    import math
    import torch
    from kvmini_b200 import engine, sweep_api, synth
    rng = np.random.default_rng(8)
    sizes = [5000, 0, 1, 2, 70_000, 3000, 60_000, 999]
    cols = synth.request_columns(sum(sizes), seed=41, span_ms=4.0e6)
    n = sum(sizes)
    offsets = np.concatenate([[0], np.cumsum(sizes)]).tolist()
    cols["ttfb_ms"][rng.integers(0, n, n // 40)] = np.nan
    cols["status"][offsets[5]:offsets[6]] = 500                        # all failures
    cols["ttfb_ms"][offsets[7]:offsets[8]] = np.nan                    # success rows without ttfb
    cols["latency_ms"][offsets[6]:offsets[7]:2] = 30000.0              # concentrated -> flat path
    cols["start_ms"][offsets[2]:offsets[3]] = 1.7e12                   # one row: end > start still holds
    colds = [[float(cols["start_ms"][offsets[r]]) / 1000.0 + 1.0] if sizes[r] else [] for r in range(len(sizes))]
    dev = {k: torch.from_numpy(v).cuda() for k, v in cols.items()}
    dlog = engine.DeviceLog.from_tensors(dev)
    sw = sweep_api.SweepAnalyzer()
    per_run = sw.rollup(dlog, offsets, colds)
    table = sw.sweep_table(dlog, offsets, colds)
    assert sw.fallback_runs == 1
    for r in range(len(sizes)):
        for k in sweep_api.SweepAnalyzer.SWEEP_KEYS:
            exp, got, none = per_run[r][k], float(table[k][r]), bool(table["is_none"][k][r])
            if exp is None:
                assert none and math.isnan(got), (r, k)
            elif isinstance(exp, float) and math.isnan(exp):
                assert not none and math.isnan(got), (r, k)
            else:
                assert not none and got == exp, (r, k, got, exp)
