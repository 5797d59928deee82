"""Parity AT THE SIZES BASELINE.json names (the small-size tests in test_gpu_*.py cannot see a
bug that only shows when a candidate bucket outgrows a CTA, a refine round runs with its
default thresholds, or a histogram flush overflows a 32-bit counter):

  configs[1]   10,000,000-row log, full analyze            == numpy oracle, every key
  refine path  2^25 + 1 rows: the refine round switches on by itself (default thresholds)
  configs[2]   256 runs x 250,000 rows through kvm_seg_analyze, every run == oracle on that run
  configs[3]   1.2e7 power samples x 1e6 request intervals x 2 traces == searchsorted/cumsum oracle
  percentile   16 B/record path at 1e7 rows == numpy sort ranks

Bit-exact for counts / ranks / order statistics / histogram counts / flags; <= 1e-12 relative
for the mean TTFT and energies (BASELINE.json north_star).  The oracle at these sizes is
oracle/np_port (numpy), itself pinned to the reference in tests/test_oracle.py.
"""
import math

import numpy as np
import pytest

from conftest import diff_keys, same

pytestmark = pytest.mark.gpu
MEAN_TOL = {"time_to_first_token_ms": 1e-12}


def _check_log(cols, cold, analyzer=None, expect_refine=None):
    from kvmini_b200 import _native as N
    from kvmini_b200 import analyze_api, columns, engine, report
    from oracle import np_port
    an = analyzer or engine.LogAnalyzer()
    if expect_refine is not None:
        assert (an._refine_rounds(len(cols["status"])) == 1) == expect_refine
    log = columns.columns_from_arrays(**cols)
    got, raw = analyze_api.analyze_log(log, cold, an)
    exp = np_port.analyze(cols, cold)
    assert diff_keys(got, {k: exp[k] for k in got}, rel=MEAN_TOL) == []
    flags = raw.flags.cpu().numpy()
    assert np.array_equal((flags & N.F_COLD) != 0, np_port.classify_cold(cols["start_ms"], cold))
    assert np.array_equal((flags & N.F_SUCCESS) != 0, cols["status"] == 200)
    assert same(report.trim_series(raw), np_port.window_series(cols))
    return got


def test_config2_ten_million_rows_equals_oracle():
    """The log bench.py times (same generator, seed and cold starts)."""
    from kvmini_b200 import synth
    cols = synth.request_columns(10_000_000, seed=1234, t_epoch_ms=1.7e12)
    got = _check_log(cols, [1.7e9 + 5.0, 1.7e9 + 1800.0], expect_refine=False)
    assert got["requests"]["total"] == 10_000_000 and got["cold_count"] > 0


def test_refine_path_with_default_thresholds_equals_oracle():
    """2^25 + 1 rows: LogAnalyzer switches the refine round on by itself (odd row count: the tail
    row takes the scalar path).  NaNs and zero-token rows included."""
    from kvmini_b200 import synth
    n = (1 << 25) + 1
    cols = synth.request_columns(n, seed=4321, t_epoch_ms=1.7e12, span_ms=7.2e6)
    rng = np.random.default_rng(1)
    cols["ttfb_ms"][rng.integers(0, n, n // 200)] = np.nan
    cols["completion_tokens"][rng.integers(0, n, n // 300)] = 0
    _check_log(cols, [1.7e9 + 30.0, 1.7e9 + 3000.0, 1.7e9 + 7000.0], expect_refine=True)


def test_config3_sweep_256_runs_of_250k_rows_equals_oracle_per_run():
    import torch
    from kvmini_b200 import engine, sweep_api, synth
    from oracle import np_port
    cols, offsets, colds = synth.sweep_columns(runs=256, rows_per_run=250_000)
    offsets = offsets.tolist()
    colds = [c.tolist() for c in colds]
    dev = {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in cols.items()}
    sw = sweep_api.SweepAnalyzer()
    got = sw.rollup(engine.DeviceLog.from_tensors(dev), offsets, colds)
    assert len(got) == 256
    for r in range(256):
        sub = {k: v[offsets[r]:offsets[r + 1]] for k, v in cols.items()}
        exp = np_port.analyze(sub, colds[r])
        d = diff_keys(got[r], {k: exp[k] for k in got[r]}, rel=MEAN_TOL)
        assert d == [], (r, d[:5])


def test_config4_energy_join_twelve_million_samples_one_million_intervals():
    import torch
    from kvmini_b200 import energy_api, synth
    from oracle import np_port
    m, n, hz = 12_000_000, 1_000_000, 1000.0
    traces = [synth.power_trace(m, gpu=g, hz=hz, null_p=1e-4) for g in range(2)]
    span_ms = m / hz * 1000.0
    cols = synth.request_columns(n, seed=77, span_ms=span_ms * 1.02, t_epoch_ms=1.7e12 - 0.01 * span_ms)
    dev = torch.device("cuda", 0)
    s_d, l_d = torch.from_numpy(cols["start_ms"]).to(dev), torch.from_numpy(cols["latency_ms"]).to(dev)
    trs = [energy_api.PowerTrace(ts, w) for ts, w in traces]
    got = energy_api.interval_energy_multi(trs, s_d, l_d).cpu().numpy()
    exp = sum(np_port.interval_energy(ts, w, cols["start_ms"], cols["latency_ms"]) for ts, w in traces)
    rel = np.abs(got - exp) / np.maximum(np.abs(exp), 1e-300)
    assert np.all((rel <= 1e-12) | (got == exp)), float(rel.max())
    assert np.array_equal(got == 0.0, exp == 0.0)
    # whole-window energy (the reference's one-interval case) at this size, against fsum too
    ts, w = traces[0]
    t0, t1 = float(ts[m // 10]), float(ts[9 * m // 10])
    wh = trs[0].trapezoid([(t0, t1)])[0]
    ref = np_port.trapezoid_wh(ts, w, t0, t1)
    exact = np_port.trapezoid_wh(ts, w, t0, t1, exact=True)
    assert abs(wh - ref) <= 1e-12 * abs(ref)
    assert abs(wh - exact) <= 1e-14 * abs(exact)


@pytest.mark.parametrize("n", [10_000_000, 10_000_001])
def test_percentile_path_ten_million_rows_equals_sorted_ranks(n):
    """The 16 B/record path (kvm_column_select_sampled through ColumnOps.order_stats)."""
    import torch
    from kvmini_b200 import _native as N
    from kvmini_b200 import engine, synth
    cols = synth.request_columns(n, seed=1234, t_epoch_ms=1.7e12)
    lat, st = cols["latency_ms"], cols["status"]
    dev = torch.device("cuda", 0)
    ops = engine.ColumnOps(dev)
    q = [(1, k, p, 0) for p in (0.5, 0.95, 0.99) for k in (N.RANK_LERP_LO, N.RANK_LERP_HI)]
    r = ops.order_stats(torch.from_numpy(lat).to(dev), q, member_mode=N.MEMBER_STATUS,
                        member=torch.from_numpy(st).to(dev))
    assert ops.last_path == "sampled"
    kept = np.sort(lat[(st == 200) & ~np.isnan(lat)])
    assert (r["n"] == kept.size).all() and (r["status"] == 0).all()
    for i, (p, fn) in enumerate((p, fn) for p in (0.5, 0.95, 0.99) for fn in (math.floor, math.ceil)):
        rank = int(fn(p * (kept.size - 1)))
        assert int(r["rank"][i]) == rank
        assert float(r["value"][i]) == float(kept[rank])
