"""Energy path: sort, trapezoid, idle tax, interval join, and the collector.py drop-in.
fp tolerance: <= 1e-12 relative (north_star); indices / counts / order statistics exact."""
import json
import math
import os
import shutil
import subprocess
import sys

import numpy as np
import pytest

from conftest import GOLDEN, REPO, diff_keys, load_json, same

pytestmark = pytest.mark.gpu
TOL = 1e-12


def close(a, b, tol=TOL):
    if a is None or b is None:
        return a is b
    return abs(a - b) <= tol * max(abs(b), 1e-300) or a == b


def test_reference_unit_tests_through_the_drop_in():
    # the reference's tests/test_energy.py, verbatim assertions, against our module
    from energy.collector import PowerSample, run_window_bounds, trapezoidal_wh
    s = [PowerSample(ts=0.0, watts=100.0), PowerSample(ts=5.0, watts=100.0), PowerSample(ts=10.0, watts=100.0)]
    assert abs(trapezoidal_wh(s, 0.0, 10.0) - (100.0 * (10.0 / 3600.0))) < 1e-6
    s = [PowerSample(ts=0.0, watts=None), PowerSample(ts=5.0, watts=100.0),
         PowerSample(ts=10.0, watts=100.0), PowerSample(ts=12.0, watts=None)]
    assert abs(trapezoidal_wh(s, 0.0, 12.0) - (100.0 * (5.0 / 3600.0))) < 1e-6
    t0, t1 = run_window_bounds([{"start_ms": 1000.0, "latency_ms": 500.0}, {"start_ms": 4000.0, "latency_ms": 1000.0}])
    assert t0 == 1.0 and t1 == 5.0
    assert run_window_bounds([]) == (0.0, 0.0)


def test_trapezoid_golden(known_answers):
    from energy.collector import PowerSample, trapezoidal_wh
    for case in known_answers["trapezoid"]:
        samples = [PowerSample(ts=t, watts=w) for t, w in zip(case["ts"], case["watts"])]
        got = trapezoidal_wh(samples, case["t0"], case["t1"])
        assert close(got, case["expected"]), (got, case["expected"], len(samples))


@pytest.mark.parametrize("m,seed,shuffle", [(1, 1, False), (2, 2, False), (1000, 3, True), (300_000, 4, True),
                                            (2_000_003, 5, False)])
def test_sort_pairs_is_stable_and_exact(m, seed, shuffle):
    import torch
    from kvmini_b200 import _native as N
    rng = np.random.default_rng(seed)
    keys = 1.7e9 + np.round(rng.uniform(0, 5000.0, m), 2 if shuffle else 6)   # many duplicate keys when rounded
    if not shuffle:
        keys = np.sort(keys)
        if m > 3:
            keys[m // 2:] = keys[m // 2:][::-1].copy()
    if m > 10:
        keys[rng.integers(0, m, 5)] = -3.5        # negative keys too
    vals = np.arange(m, dtype=np.float64)          # payload = original index -> stability check
    lib = N.lib()
    dev = torch.device("cuda", 0)
    k_in, v_in = torch.from_numpy(keys).to(dev), torch.from_numpy(vals).to(dev)
    k_out, v_out = torch.empty_like(k_in), torch.empty_like(v_in)
    ws = torch.empty(lib.kvm_sort_pairs_workspace_bytes(m), dtype=torch.uint8, device=dev)
    N.check(lib.kvm_sort_pairs_f64(k_in.data_ptr(), v_in.data_ptr(), k_out.data_ptr(), v_out.data_ptr(), m,
                                   ws.data_ptr(), ws.numel(), torch.cuda.current_stream(dev).cuda_stream))
    order = np.argsort(keys, kind="stable")
    assert np.array_equal(k_out.cpu().numpy(), keys[order])
    assert np.array_equal(v_out.cpu().numpy(), vals[order])


@pytest.mark.parametrize("m,null_p,shuffled", [(50_000, 0.0, False), (400_000, 0.01, False), (400_000, 0.3, True)])
def test_trapezoid_and_idle_tax_vs_oracle(m, null_p, shuffled):
    from kvmini_b200 import energy_api, synth
    from oracle import np_port
    ts, w = synth.power_trace(m, hz=100.0, null_p=null_p)
    if shuffled:
        perm = np.random.default_rng(1).permutation(m)
        ts, w = ts[perm], w[perm]
    tr = energy_api.PowerTrace(ts, w)
    span = m / 100.0
    t0, t1 = 1.7e9 + 0.2 * span, 1.7e9 + 0.7 * span
    windows = [(t0, t1), (float(ts.min()), t0), (t1, float(ts.max())), (5.0, 1.0)]
    got = tr.trapezoid(windows)
    for g, (a, b) in zip(got, windows):
        exact = np_port.trapezoid_wh(ts, w, a, b, exact=True)
        ref = np_port.trapezoid_wh(ts, w, a, b)
        assert close(g, ref), (g, ref)
        assert close(g, exact, 1e-14)          # ours is a compensated sum: at least as close to fsum as the reference
    # baseline idle tax pieces (collector.py:324-347)
    p_idle, st = tr.idle_power_upper_median(t0, t1)
    outside = w[~np.isnan(w) & ((ts < t0) | (ts > t1))]
    assert p_idle == sorted(outside.tolist())[len(outside) // 2]
    assert int(st.n_outside) == int(((ts < t0) | (ts > t1)).sum())
    steps = np.diff(ts)
    assert close(tr.step_sum[0] + tr.step_sum[1], math.fsum(steps.tolist()), 1e-15)


@pytest.mark.parametrize("mode,argv", [
    ("plain", []), ("series", ["--idle-tax", "series", "--merge-results"]),
    ("baseline", ["--idle-tax", "baseline"]), ("warmup", ["--include-warmup"])])
def test_collector_cli_drop_in(mode, argv, tmp_path):
    d = os.path.join(GOLDEN, "cfg1")
    run = tmp_path / "run"
    run.mkdir()
    shutil.copy(os.path.join(d, "requests.csv"), run / "requests.csv")
    shutil.copy(os.path.join(d, "power.json"), run / "power.json")
    shutil.copy(os.path.join(d, "expected_results_cost.json"), run / "results.json")
    r = subprocess.run([sys.executable, os.path.join(REPO, "energy", "collector.py"), "integrate",
                        "--run-dir", str(run)] + argv, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    got = load_json(run / "energy.json")
    exp = load_json(os.path.join(d, f"expected_energy_{mode}.json"))
    fp = {"Wh_active": TOL, "Wh_idle_tax": TOL, "Wh_per_1k_tokens_active": TOL, "Wh_per_request_active": TOL}
    assert diff_keys(got, exp, rel=fp) == []
    if mode == "series":
        res = load_json(run / "results.json")
        exp_res = load_json(os.path.join(d, "expected_results_final.json"))
        fp2 = {"energy_wh_active": TOL, "energy_wh_idle_tax": TOL, "energy_wh_per_1k_tokens": TOL,
               "energy_wh_per_request": TOL}
        assert diff_keys(res, exp_res, rel=fp2) == []
        assert "Merged energy fields into" in r.stdout
        # consumer acceptance: the reference's gate passed on its own outputs; same fields here
        gate = load_json(os.path.join(d, "expected_gate.json"))
        assert gate["returncode"] == 0 and res["p95_ms"] <= 1200


def test_collector_cli_error_paths(tmp_path):
    script = os.path.join(REPO, "energy", "collector.py")
    r = subprocess.run([sys.executable, script, "integrate", "--run-dir", str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 2 and "requests.csv not found" in r.stderr
    shutil.copy(os.path.join(GOLDEN, "ci5", "requests.csv"), tmp_path / "requests.csv")
    r = subprocess.run([sys.executable, script, "integrate", "--run-dir", str(tmp_path), "--merge-results"],
                       capture_output=True, text=True)
    assert r.returncode == 0 and "power.json not found" in r.stderr
    e = load_json(tmp_path / "energy.json")
    assert e["Wh_active"] is None and e["warnings"] == ["Missing power samples"]
    assert load_json(tmp_path / "results.json")["energy_wh_active"] is None


@pytest.mark.parametrize("m,n,null_p", [(20_000, 300, 0.01), (1_000_000, 200_000, 1e-4), (200_000, 50_000, 0.2)])
def test_interval_energy_join_vs_oracle(m, n, null_p):
    """Extension A16 (parity unpinned by the reference; the one-interval case is trapezoidal_wh)."""
    import torch
    from kvmini_b200 import energy_api, synth
    from oracle import np_port
    hz = 100.0
    traces = [synth.power_trace(m, gpu=g, hz=hz, null_p=null_p) for g in range(2)]
    span_ms = m / hz * 1000.0
    cols = synth.request_columns(n, seed=9, span_ms=span_ms * 1.05, t_epoch_ms=1.7e12 - 0.02 * span_ms)
    dev = torch.device("cuda", 0)
    s_d, l_d = torch.from_numpy(cols["start_ms"]).to(dev), torch.from_numpy(cols["latency_ms"]).to(dev)
    trs = [energy_api.PowerTrace(ts, w) for ts, w in traces]
    got = energy_api.interval_energy_multi(trs, s_d, l_d).cpu().numpy()
    exp = sum(np_port.interval_energy(ts, w, cols["start_ms"], cols["latency_ms"]) for ts, w in traces)
    assert (exp > 0).sum() > 0.8 * n
    rel = np.abs(got - exp) / np.maximum(np.abs(exp), 1e-300)
    assert np.all((rel <= TOL) | (got == exp)), float(rel.max())
    assert np.array_equal(got == 0.0, exp == 0.0)   # empty intervals agree exactly
    # one interval == the reference function itself
    from energy.collector import PowerSample, trapezoidal_wh
    ts0, w0 = traces[0]
    i = int(np.argmax(exp))
    s, e = cols["start_ms"][i] / 1000.0, (cols["start_ms"][i] + cols["latency_ms"][i]) / 1000.0
    lo, hi = np.searchsorted(ts0, s - 1), np.searchsorted(ts0, e + 1)
    samples = [PowerSample(ts=float(t), watts=(None if math.isnan(x) else float(x)))
               for t, x in zip(ts0[lo:hi], w0[lo:hi])]
    one = trs[0].interval_energy(s_d[i:i + 1].clone(), l_d[i:i + 1].clone()).item()
    assert close(one, trapezoidal_wh(samples, s, e))


def test_division_free_quotient_is_the_ieee_quotient():
    """The energy kernels evaluate (ts1 - ts0) / 3600.0 (collector.py:146) without the hardware
    division (reciprocal multiply + exact fma residual + a proof check, else the real division):
    bit-identical to x / 3600.0 on 2e7 doubles of every magnitude, sign and special value."""
    import torch
    from kvmini_b200 import _native as N
    rng = np.random.default_rng(3600)
    parts = [rng.uniform(0.0, 2e-3, 4_000_000),                                # sample steps at 1 kHz
             rng.uniform(0.0, 10.0, 4_000_000),
             np.ldexp(rng.uniform(1.0, 2.0, 4_000_000), rng.integers(-1070, 1020, 4_000_000)),
             -np.ldexp(rng.uniform(1.0, 2.0, 2_000_000), rng.integers(-300, 300, 2_000_000)),
             3600.0 * np.ldexp(rng.integers(1, 1 << 53, 2_000_000).astype(np.float64), rng.integers(-80, 20, 2_000_000)),
             (3600.0 * rng.integers(1, 1 << 40, 2_000_000)).astype(np.float64),  # exact quotients
             np.nextafter(3600.0 * rng.integers(1, 1 << 40, 1_000_000).astype(np.float64), np.inf),
             np.array([0.0, -0.0, np.inf, -np.inf, np.nan, 5e-324, 1e-308, 1.7976931348623157e308, 3600.0, 1.0])]
    x = np.concatenate(parts)
    d = torch.from_numpy(x).cuda()
    out = torch.empty_like(d)
    N.check(N.lib().kvm_selfcheck_div3600(d.data_ptr(), out.data_ptr(), x.size, torch.cuda.current_stream().cuda_stream))
    got = out.cpu().numpy()
    with np.errstate(all="ignore"):
        exp = x / 3600.0
    assert np.array_equal(got.view(np.uint64)[~np.isnan(exp)], exp.view(np.uint64)[~np.isnan(exp)])
    assert np.isnan(got[np.isnan(exp)]).all()


def test_interval_join_unsorted_intervals_irregular_trace_and_many_traces():
    """The join searches from an interpolated guess: intervals in random order, a trace with a long
    gap and bursts (guess far off), 9 traces (two passes of <= 8), intervals outside the trace."""
    import torch
    from kvmini_b200 import energy_api, synth
    from oracle import np_port
    rng = np.random.default_rng(8)
    m, n = 300_000, 60_000
    traces = []
    for g in range(9):
        ts, w = synth.power_trace(m, gpu=g, hz=50.0, null_p=0.02)
        if g == 1:
            ts = np.sort(np.concatenate([ts[: m // 3], ts[m // 3:] + 5000.0]))      # a 5000 s hole
        if g == 2:
            ts = np.sort(ts[0] + (ts - ts[0]) ** 1.3 / (ts[-1] - ts[0]) ** 0.3)     # accelerating clock
        traces.append((ts, w))
    span_ms = m / 50.0 * 1000.0
    cols = synth.request_columns(n, seed=3, span_ms=span_ms * 1.2, t_epoch_ms=1.7e12 - 0.1 * span_ms)
    perm = rng.permutation(n)
    start, lat = cols["start_ms"][perm].copy(), cols["latency_ms"][perm].copy() * 20.0
    dev = torch.device("cuda", 0)
    trs = [energy_api.PowerTrace(ts, w) for ts, w in traces]
    got = energy_api.interval_energy_multi(trs, torch.from_numpy(start).to(dev), torch.from_numpy(lat).to(dev)).cpu().numpy()
    exp = np.zeros(n)
    for ts, w in traces:            # trace order, like the device
        exp = exp + np_port.interval_energy(ts, w, start, lat)
    rel = np.abs(got - exp) / np.maximum(np.abs(exp), 1e-300)
    assert np.all((rel <= TOL) | (got == exp)), float(rel.max())
    assert (exp > 0).sum() > n // 2
