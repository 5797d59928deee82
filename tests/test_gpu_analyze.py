"""Parity of the CUDA analyze path (through the C ABI / the drop-in functions) against the
oracle and the reference's golden fixtures.  Counts, ranks, order statistics, histogram
counts and everything derived from them: bit-exact.  Mean TTFT: <= 1e-12 relative
(BASELINE.json north_star)."""
import csv
import datetime as dt
import json
import math
import os
import shutil
import subprocess
import sys

import numpy as np
import pytest

from conftest import FAKE_BIN, GOLDEN, REPO, diff_keys, load_json, same, unj

pytestmark = pytest.mark.gpu

MEAN_TOL = {"time_to_first_token_ms": 1e-12}


@pytest.fixture(scope="module")
def api():
    import analyze
    return analyze


@pytest.fixture(scope="module")
def analyzer():
    from kvmini_b200 import engine
    return engine.LogAnalyzer()


def test_native_library_is_loaded():
    from kvmini_b200 import _native
    assert _native.lib().kvm_sm_count() >= 100  # B200: 148
    loaded = open("/proc/self/maps").read()
    assert "libkvmini_b200.so" in loaded


def test_percentile_golden(api, known_answers):
    for case in known_answers["percentile"]:
        vals, exp = unj(case["values"]), unj(case["expected"])
        assert same(api.percentile(vals, case["p"]), exp), (len(vals), case["p"])


def test_histogram_golden(api, known_answers):
    for case in known_answers["histogram"]:
        vals, exp = unj(case["values"]), unj(case["expected"])
        assert same(api.compute_histograms(vals, case["num_bins"]), exp), case["num_bins"]


def _cold_from_pods(path):
    out = []
    for p in load_json(path)["items"]:
        st = p["status"]["containerStatuses"][0]["state"]["running"]["startedAt"]
        out.append(dt.datetime.fromisoformat(st.replace("Z", "+00:00")).timestamp())
    return out


@pytest.mark.parametrize("name,expected,kube", [
    ("ci5", "expected_results.json", None),
    ("cfg1", "expected_results_analyze.json", "kube"),
    ("repro100", "expected_results_analyze.json", None),
    ("edge5k", "expected_results_analyze.json", "kube"),
])
def test_analyze_log_matches_reference(name, expected, kube, analyzer):
    from kvmini_b200 import analyze_api, columns
    d = os.path.join(GOLDEN, name)
    exp = load_json(os.path.join(d, expected))
    log = columns.load_requests_csv(os.path.join(d, "requests.csv"))
    cold = _cold_from_pods(os.path.join(d, kube, "pods.json")) if kube else []
    cold = [c for c in cold if exp["window"]["start"] <= c <= exp["window"]["end"]]   # analyze.py:383-388
    got, _raw = analyze_api.analyze_log(log, cold, analyzer)
    exp_core = {k: v for k, v in exp.items() if k in got}
    assert diff_keys(got, exp_core, rel=MEAN_TOL) == []


@pytest.mark.parametrize("name,expected,kube", [
    ("ci5", "expected_results.json", None),
    ("cfg1", "expected_results_analyze.json", "kube"),
    ("repro100", "expected_results_analyze.json", None),
    ("edge5k", "expected_results_analyze.json", "kube"),
])
def test_analyze_cli_drop_in(name, expected, kube, tmp_path):
    """`python analyze.py --run-dir ...` writes the reference's results.json and
    requests_classified.csv (byte-identical CSV; JSON equal key by key, same order, plus the
    additive `timeseries` key)."""
    d = os.path.join(GOLDEN, name)
    run = tmp_path / "run"
    run.mkdir()
    shutil.copy(os.path.join(d, "requests.csv"), run / "requests.csv")
    env = dict(os.environ)
    env["PATH"] = FAKE_BIN + os.pathsep + env["PATH"]
    if kube:
        env["KVM_FAKE_KUBECTL_DIR"] = os.path.join(d, kube)
    r = subprocess.run([sys.executable, os.path.join(REPO, "analyze.py"), "--run-dir", str(run),
                        "--namespace", "ml-prod", "--service", "demo-llm"], env=env,
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    got = load_json(run / "results.json")
    exp = load_json(os.path.join(d, expected))
    assert "timeseries" in got
    ts = got.pop("timeseries")
    assert sum(ts["requests"]) == exp["requests"]["total"]
    assert sum(ts["success"]) == exp["requests"]["success"]
    assert diff_keys(got, exp, rel=MEAN_TOL) == []
    assert json.loads(r.stdout)["p95_ms"] == exp["p95_ms"] or math.isnan(exp["p95_ms"])
    assert (run / "requests_classified.csv").read_bytes() == \
        open(os.path.join(d, "expected_classified.csv"), "rb").read()


def test_cli_empty_csv_exits_1(tmp_path):
    (tmp_path / "requests.csv").write_text("id,start_ms,latency_ms,status\n")
    r = subprocess.run([sys.executable, os.path.join(REPO, "analyze.py"), "--run-dir", str(tmp_path),
                        "--namespace", "n", "--service", "s"], capture_output=True, text=True)
    assert r.returncode == 1 and "No request rows found" in r.stderr


def test_cli_success_row_without_ttfb_raises_like_reference(tmp_path):
    # math.isnan("") -> TypeError
    (tmp_path / "requests.csv").write_text(
        "id,start_ms,ttfb_ms,tllt_ms,latency_ms,status,completion_tokens,total_tokens\n"
        "1,1000.0,,,150.0,200,5,10\n")
    r = subprocess.run([sys.executable, os.path.join(REPO, "analyze.py"), "--run-dir", str(tmp_path),
                        "--namespace", "n", "--service", "s"], capture_output=True, text=True)
    assert r.returncode != 0 and "TypeError: must be real number, not str" in r.stderr


def _random_log(n, seed, nan_rows=0, zero_tok_rows=0, span_ms=900_000.0, **kw):
    from kvmini_b200 import synth
    cols = synth.request_columns(n, seed=seed, span_ms=span_ms, **kw)
    rng = np.random.default_rng(seed + 1000)
    if nan_rows:
        idx = rng.integers(0, n, nan_rows)
        cols["ttfb_ms"][idx] = np.nan
        idx = rng.integers(0, n, nan_rows)
        cols["tllt_ms"][idx] = np.nan
        idx = rng.integers(0, n, nan_rows)
        cols["latency_ms"][idx] = np.nan
    if zero_tok_rows:
        cols["completion_tokens"][rng.integers(0, n, zero_tok_rows)] = 0
    return cols


def _check_against_oracle(cols, cold, analyzer):
    from kvmini_b200 import analyze_api, columns, report
    from oracle import np_port
    log = columns.columns_from_arrays(**cols)
    got, raw = analyze_api.analyze_log(log, cold, analyzer)
    exp = np_port.analyze(cols, cold)
    assert diff_keys(got, {k: exp[k] for k in got}, rel=MEAN_TOL) == []
    # is_cold flags, row by row
    from kvmini_b200 import _native as N
    flags = raw.flags.cpu().numpy()
    assert np.array_equal((flags & N.F_COLD) != 0, np_port.classify_cold(cols["start_ms"], cold))
    assert np.array_equal((flags & N.F_SUCCESS) != 0, cols["status"] == 200)
    # extension A12
    series = report.trim_series(raw)
    assert same(series, np_port.window_series(cols))
    return got


@pytest.mark.parametrize("n,seed", [(1, 11), (2, 12), (3, 13), (31, 14), (1000, 15), (4097, 16),
                                    (100_000, 17), (1_000_001, 18)])
def test_random_logs_vs_oracle(n, seed, analyzer):
    cols = _random_log(n, seed, nan_rows=n // 50, zero_tok_rows=n // 40)
    t0 = float(cols["start_ms"].min()) / 1000.0
    cold = [t0 + 1.0, t0 + 120.0, t0 + 121.0, t0 + 700.0] if n > 3 else []
    _check_against_oracle(cols, cold, analyzer)


def test_all_failures(analyzer):
    cols = _random_log(5000, 21)
    cols["status"][:] = 500
    got = _check_against_oracle(cols, [], analyzer)
    assert math.isnan(got["p50_ms"]) and got["throughput_rps"] == 0.0 and got["ttft_p50_ms"] is None


def test_constant_latency_and_heavy_duplicates(analyzer):
    cols = _random_log(200_000, 22)
    cols["latency_ms"][:] = 123.456                      # one distinct key: width 0
    cols["ttfb_ms"][cols["status"] == 200] = 7.0
    cols["tllt_ms"][::2] = 30000.0                       # half of the rows share one value
    _check_against_oracle(cols, [float(cols["start_ms"].min()) / 1000.0 + 3.0], analyzer)


def test_negative_and_mixed_sign_values(analyzer):
    cols = _random_log(50_000, 23)
    rng = np.random.default_rng(5)
    cols["tllt_ms"] = np.where(cols["status"] == 200, rng.normal(0.0, 100.0, 50_000), np.nan)  # gen < 0 too
    _check_against_oracle(cols, [], analyzer)


def test_unaligned_columns_take_the_scalar_path(analyzer):
    """A shard that starts at an odd row has 8-byte- but not 16-byte-aligned columns."""
    import torch
    from kvmini_b200 import engine, report
    from oracle import np_port
    cols = _random_log(100_001, 24, nan_rows=50)
    dev = {k: torch.from_numpy(v).cuda() for k, v in cols.items()}
    sliced = {k: v[1:] for k, v in dev.items()}
    assert sliced["latency_ms"].data_ptr() % 16 == 8
    dlog = engine.DeviceLog(cols=sliced, n=100_000, device=sliced["status"].device)
    raw = analyzer.analyze(dlog, [])
    got = report.core_results(raw, 0)
    exp = np_port.analyze({k: v[1:] for k, v in cols.items()}, [])
    assert diff_keys(got, {k: exp[k] for k in got}, rel=MEAN_TOL) == []


def test_many_cold_starts_use_global_table(analyzer):
    cols = _random_log(20_000, 25, span_ms=3.0e6)
    t0 = float(cols["start_ms"].min()) / 1000.0
    cold = [t0 + 2.0 * i for i in range(1500)]           # > shared-memory staging limit
    _check_against_oracle(cols, cold, analyzer)


def test_rank_dialects_match_python(api):
    """F4: the other percentile dialects of the repo on one column."""
    import torch
    from kvmini_b200 import _native as N
    from kvmini_b200 import engine
    rng = np.random.default_rng(3)
    v = np.round(rng.lognormal(5.0, 1.0, 9999), 3)
    s = sorted(v.tolist())
    ops = engine.ColumnOps()
    res = ops.order_stats(torch.from_numpy(v).cuda(), [
        (1, N.RANK_FLOOR_PN, 0.95, 0), (1, N.RANK_FLOOR_PNM1, 0.95, 0), (1, N.RANK_UPPER_MEDIAN, 0.0, 0),
        (1, N.RANK_ABSOLUTE, 0.0, 17), (1, N.RANK_FLOOR_PN, 1.0, 0)])
    n = len(s)
    assert res[0]["value"] == s[int(0.95 * n)]          # runners/backends/tgi/invoke.sh:208-210
    assert res[1]["value"] == s[int(0.95 * (n - 1))]    # scripts/fairness_dual_tenant.py:59-65
    assert res[2]["value"] == s[n // 2]                 # energy/collector.py:333
    assert res[3]["value"] == s[17]
    assert res[4]["status"] == N.Q_OUT_OF_RANGE         # sorted(x)[int(1.0*n)] -> IndexError


def test_status_member_mode_is_the_16_byte_path(api):
    import torch
    from kvmini_b200 import _native as N
    from kvmini_b200 import engine
    from oracle import np_port
    cols = _random_log(300_000, 26)
    ops = engine.ColumnOps()
    lat, st = torch.from_numpy(cols["latency_ms"]).cuda(), torch.from_numpy(cols["status"]).cuda()
    for p in (0.5, 0.95, 0.99):
        stat = ops.quantile_stat(lat, p, member_mode=N.MEMBER_STATUS, member=st)
        exp = np_port.percentile(cols["latency_ms"][cols["status"] == 200], p)
        assert stat.interpolate(p) == exp


# ---- refine rounds (large logs narrow every candidate bucket before the compact pass) --------
def _refining_analyzer(rounds):
    from kvmini_b200 import engine
    an = engine.LogAnalyzer()
    an.refine_rounds = rounds
    an.refine_min_keys = 0          # refine whenever a bucket exists, however small the log
    return an


@pytest.mark.parametrize("rounds", [1, 2, 3])
@pytest.mark.parametrize("n,seed", [(3, 30), (1000, 31), (100_000, 32), (1_000_001, 33)])
def test_refine_rounds_vs_oracle(n, seed, rounds):
    cols = _random_log(n, seed, nan_rows=n // 50, zero_tok_rows=n // 40)
    t0 = float(cols["start_ms"].min()) / 1000.0
    cold = [t0 + 1.0, t0 + 120.0, t0 + 700.0] if n > 3 else []
    _check_against_oracle(cols, cold, _refining_analyzer(rounds))


@pytest.mark.parametrize("rounds", [1, 4])
def test_refine_with_duplicates_and_mixed_signs(rounds):
    an = _refining_analyzer(rounds)
    cols = _random_log(200_000, 34)
    cols["latency_ms"][:] = 123.456
    cols["ttfb_ms"][cols["status"] == 200] = 7.0
    cols["tllt_ms"][::2] = 30000.0                       # a refined bucket that cannot shrink
    _check_against_oracle(cols, [float(cols["start_ms"].min()) / 1000.0 + 3.0], an)
    cols = _random_log(50_000, 35)
    rng = np.random.default_rng(6)
    cols["tllt_ms"] = np.where(cols["status"] == 200, rng.normal(0.0, 100.0, 50_000), np.nan)
    _check_against_oracle(cols, [], an)
    cols = _random_log(60_000, 36)
    cols["latency_ms"] = np.round(cols["latency_ms"])   # few distinct values, narrow key range
    _check_against_oracle(cols, [], an)


def test_refined_and_plain_select_agree_on_a_large_log(analyzer):
    """40M rows on the device (above KVM_REFINE_MIN_ROWS, so refinement is on by default):
    order statistics, counts, histograms are identical with and without the refine round."""
    import torch
    from kvmini_b200 import engine, report
    n = 40_000_000
    g = torch.Generator(device="cuda").manual_seed(7)
    dev = torch.device("cuda")
    start = torch.rand(n, generator=g, device=dev, dtype=torch.float64) * 3.6e6 + 1.7e12
    ttfb = torch.rand(n, generator=g, device=dev, dtype=torch.float64) * 180.0 + 20.0
    gen = torch.empty(n, device=dev, dtype=torch.float64).exponential_(1.0 / 900.0, generator=g)
    tok = torch.randint(1, 512, (n,), generator=g, device=dev, dtype=torch.int64)
    status = torch.where(torch.rand(n, generator=g, device=dev) < 0.97, 200, 500).to(torch.int64)
    cols = {"start_ms": start, "ttfb_ms": ttfb, "tllt_ms": ttfb + gen, "latency_ms": ttfb + gen + 1.5,
            "status": status, "completion_tokens": tok, "total_tokens": tok + 100}
    dlog = engine.DeviceLog.from_tensors(cols)
    cold = [1.7e9 + 5.0, 1.7e9 + 1800.0]
    plain = engine.LogAnalyzer()
    plain.refine_rounds = 0
    auto = engine.LogAnalyzer()
    assert auto._refine_rounds(n) == 1
    a = report.core_results(plain.analyze(dlog, cold), len(cold))
    b = report.core_results(auto.analyze(dlog, cold), len(cold))
    assert diff_keys(a, b) == []
    # the median really is an order statistic of the data: count below / at it on the device
    lat_ok = cols["latency_ms"][status == 200]
    p50 = b["p50_ms"]
    below, upto = int((lat_ok < p50).sum()), int((lat_ok <= p50).sum())
    k = 0.5 * (lat_ok.numel() - 1)
    assert below <= math.ceil(k) and upto > math.floor(k)


# ---- F2: exemplar search (report_generator.py:437-454) ---------------------------------------
def _nearest_py(lat, status, target):
    """The reference's expression: first element of a stable sort by |latency - target|."""
    rows = [(abs(float(v) - target), i) for i, (v, s) in enumerate(zip(lat, status)) if s == 200 and not math.isnan(v)]
    return sorted(rows, key=lambda x: x[0])[0][1] if rows else -1


@pytest.mark.parametrize("n", [1, 2, 7, 1000, 100_001, 1_000_000])
def test_nearest_matches_stable_sort(n):
    import torch
    from kvmini_b200 import _native as N
    from kvmini_b200 import engine
    ops = engine.ColumnOps()
    rng = np.random.default_rng(n)
    lat = np.round(rng.lognormal(5.5, 0.6, n), 1)            # coarse values: many exact ties
    status = np.where(rng.random(n) < 0.9, 200, 500).astype(np.int64)
    lat[rng.integers(0, n, max(1, n // 50))] = np.nan
    target = float(np.nanpercentile(lat, 95)) if n > 2 else 100.0
    d_lat, d_st = torch.from_numpy(lat).cuda(), torch.from_numpy(status).cuda()
    idx, dist = ops.nearest(d_lat, target, member_mode=N.MEMBER_STATUS, member=d_st)
    exp = _nearest_py(lat, status, target)
    assert idx == exp
    if exp >= 0:
        assert dist == abs(lat[exp] - target)
    # unaligned views take the scalar path
    if n > 8:
        idx2, _ = ops.nearest(d_lat[1:], target, member_mode=N.MEMBER_STATUS, member=d_st[1:])
        assert idx2 == _nearest_py(lat[1:], status[1:], target)
    # all elements are members
    idx3, _ = ops.nearest(d_lat, target)
    assert idx3 == _nearest_py(lat, np.full(n, 200), target)


def test_nearest_without_members_and_trace_lookup(tmp_path):
    import torch
    from kvmini_b200 import _native as N
    from kvmini_b200 import analyze_api, columns, engine
    ops = engine.ColumnOps()
    lat = torch.tensor([1.0, 2.0, float("nan")], dtype=torch.float64).cuda()
    st = torch.tensor([500, 500, 200], dtype=torch.int64).cuda()
    assert ops.nearest(lat, 1.5, member_mode=N.MEMBER_STATUS, member=st)[0] == -1
    src = os.path.join(REPO, "tests", "golden", "cfg1", "requests.csv")
    log = columns.load_requests_csv(src, keep_rows=True)
    with open(os.path.join(REPO, "tests", "golden", "cfg1", "expected_results_analyze.json")) as f:
        p95 = json.load(f)["p95_ms"]
    got = analyze_api.exemplar_trace_id(log, p95, ops)
    rows = []
    with open(src, newline="") as f:                     # report_generator.py:440-451 verbatim in spirit
        for row in csv.DictReader(f):
            if row.get("status") == "200" and row.get("latency_ms"):
                rows.append((abs(float(row["latency_ms"]) - p95), row.get("trace_id")))
    assert got == sorted(rows, key=lambda x: x[0])[0][1]


# ---- sampled-range digit select (csrc/kvm_qdigit.cu) -------------------------------------------
def _ref_stat(vals, kind, p):
    """Order statistic the reference's expressions index (analyze.py:65-81 and the dialects)."""
    from kvmini_b200 import _native as N
    n = len(vals)
    pc = min(max(p, 0.0), 1.0)
    rank = {N.RANK_LERP_LO: lambda: int(math.floor(pc * (n - 1))), N.RANK_LERP_HI: lambda: int(math.ceil(pc * (n - 1))),
            N.RANK_UPPER_MEDIAN: lambda: n // 2, N.RANK_FLOOR_PN: lambda: int(p * n),
            N.RANK_FLOOR_PNM1: lambda: int(p * (n - 1))}[kind]()
    return vals[rank] if 0 <= rank < n else None


@pytest.mark.parametrize("shape", ["lognormal", "sorted", "reversed", "periodic", "rounded", "two_values", "few_members"])
@pytest.mark.parametrize("n", [1 << 18, 1_000_003, 5_000_000])
def test_sampled_select_is_exact_or_declines(shape, n):
    import torch
    from kvmini_b200 import _native as N
    from kvmini_b200 import engine
    ops = engine.ColumnOps()
    rng = np.random.default_rng(n % 1000 + len(shape))
    lat = rng.lognormal(5.5, 0.6, n)
    if shape == "sorted":
        lat.sort()
    elif shape == "reversed":
        lat = np.sort(lat)[::-1].copy()
    elif shape == "periodic":
        lat = 100.0 + (np.arange(n) % 4096) * 0.25 + rng.random(n) * 1e-3     # period = the sampling stratum count
    elif shape == "rounded":
        lat = np.round(lat, 0)                                                # ~1500 distinct values
    elif shape == "two_values":
        lat = np.where(rng.random(n) < 0.5, 10.0, 20.0)
    status = np.where(rng.random(n) < (0.001 if shape == "few_members" else 0.97), 200, 500).astype(np.int64)
    lat[rng.integers(0, n, n // 100)] = np.nan
    d_lat, d_st = torch.from_numpy(lat).cuda(), torch.from_numpy(status).cuda()
    ok = np.sort(lat[(status == 200) & ~np.isnan(lat)])
    q = [(1, k, p, 0) for p in (0.5, 0.95, 0.99) for k in (N.RANK_LERP_LO, N.RANK_LERP_HI)]
    q += [(1, N.RANK_UPPER_MEDIAN, 0.0, 0), (1, N.RANK_FLOOR_PN, 0.95, 0)]
    got = ops.order_stats(d_lat, q, member_mode=N.MEMBER_STATUS, member=d_st, sampled=True)
    path = ops.last_path
    for (mask, kind, p, _), g in zip(q, got):
        assert int(g["n"]) == len(ok)
        assert float(g["value"]) == _ref_stat(ok, kind, p), (shape, kind, p, path)
    if shape in ("lognormal", "sorted", "reversed", "periodic"):
        assert path == "sampled"          # continuous data in any order takes the one-pass path
    # and the general select agrees bit for bit
    ref = ops.order_stats(d_lat, q, member_mode=N.MEMBER_STATUS, member=d_st, sampled=False)
    assert ops.last_path == "general" and np.array_equal(got["value"], ref["value"]) and np.array_equal(got["rank"], ref["rank"])


def test_sampled_select_edges():
    import torch
    from kvmini_b200 import _native as N
    from kvmini_b200 import engine
    ops = engine.ColumnOps()
    n = 1 << 20
    rng = np.random.default_rng(77)
    lat = rng.normal(0.0, 50.0, n)                     # mixed signs
    d = torch.from_numpy(lat).cuda()
    srt = np.sort(lat)
    q = [(1, k, p, 0) for p in (0.0, 1.0, 0.001, 0.9999) for k in (N.RANK_LERP_LO, N.RANK_LERP_HI)]
    got = ops.order_stats(d, q, sampled=True)          # extremes: catch-all bins -> may decline, still exact
    for (m, kind, p, _), g in zip(q, got):
        assert float(g["value"]) == _ref_stat(srt, kind, p)
    # unaligned view, member mode ALL
    got = ops.order_stats(d[1:], [(1, N.RANK_LERP_LO, 0.5, 0), (1, N.RANK_LERP_HI, 0.5, 0)], sampled=True)
    s1 = np.sort(lat[1:])
    assert float(got[0]["value"]) == _ref_stat(s1, N.RANK_LERP_LO, 0.5)
    assert float(got[1]["value"]) == _ref_stat(s1, N.RANK_LERP_HI, 0.5)
    # no member at all: declines (too few sample members), the general path reports the empty set
    st = torch.full((n,), 500, dtype=torch.int64).cuda()
    got = ops.order_stats(d, [(1, N.RANK_LERP_LO, 0.5, 0)], member_mode=N.MEMBER_STATUS, member=st, sampled=True)
    assert int(got[0]["n"]) == 0 and int(got[0]["status"]) == N.Q_EMPTY
    # host-side refusals
    lib = N.lib()
    ws = torch.empty(lib.kvm_column_select_sampled_workspace_bytes(n), dtype=torch.uint8, device="cuda")
    res = torch.zeros(8 * 32, dtype=torch.uint8, device="cuda")
    absolute = (N.QueryC * 1)(N.QueryC(1, N.RANK_ABSOLUTE, 0.0, 5))
    assert lib.kvm_column_select_sampled(d.data_ptr(), n, N.MEMBER_ALL, None, absolute, 1, res.data_ptr(),
                                         ws.data_ptr(), ws.numel(), None) == N.KVM_ERR_UNSUPPORTED



def test_sampled_select_dirty_workspace_and_graph_replay():
    """The sampled select clears its own accumulators inside its first kernel (no memset) behind a
    per-call token: a workspace full of garbage, the same workspace reused for other data, and a
    CUDA-graph replay of one captured call (same token every time) must all give exact answers."""
    import ctypes as C
    import torch
    from kvmini_b200 import _native as N
    from kvmini_b200.engine import RESULT_DTYPE
    lib = N.lib()
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(5)
    sizes = (1 << 18, 300_007, 1_000_000, 2_000_019)
    q = (N.QueryC * 6)(*[N.QueryC(1, k, p, 0) for p in (0.5, 0.95, 0.99) for k in (N.RANK_LERP_LO, N.RANK_LERP_HI)])
    ws = torch.empty(lib.kvm_column_select_sampled_workspace_bytes(max(sizes)), dtype=torch.uint8, device=dev)
    res = torch.zeros(6 * 32, dtype=torch.uint8, device=dev)
    st = torch.cuda.current_stream(dev).cuda_stream

    def expect(v, s):
        ok = np.sort(v[(s == 200) & ~np.isnan(v)])
        out = []
        for p in (0.5, 0.95, 0.99):
            pos = p * (len(ok) - 1)
            out += [ok[int(math.floor(pos))], ok[int(math.ceil(pos))]]
        return out, len(ok)

    def read():
        torch.cuda.synchronize()
        assert int(ws[0:4].view(torch.int32).item()) == 0, "declined"
        return np.frombuffer(res.cpu().numpy().tobytes(), dtype=RESULT_DTYPE, count=6)

    for rep, n in enumerate(sizes + sizes[::-1]):
        v = np.round(rng.lognormal(5.0 + 0.2 * rep, 0.5, n), 3)
        s = np.where(rng.random(n) < 0.9, 200, 500).astype(np.int64)
        d_v, d_s = torch.from_numpy(v).to(dev), torch.from_numpy(s).to(dev)
        if rep % 2 == 0:
            ws.random_(0, 256)            # garbage everywhere, the handshake word included
        N.check(lib.kvm_column_select_sampled(d_v.data_ptr(), n, N.MEMBER_STATUS, d_s.data_ptr(), q, 6,
                                              res.data_ptr(), ws.data_ptr(), ws.numel(), st))
        want, m = expect(v, s)
        got = read()
        assert [float(x) for x in got["value"]] == want and all(int(x) == m for x in got["n"]), (rep, n)
    # one captured call replayed on changing data (in place): same token on every replay
    n = 1_000_000
    d_v = torch.empty(n, dtype=torch.float64, device=dev)
    d_s = torch.empty(n, dtype=torch.int64, device=dev)
    side = torch.cuda.Stream(dev)
    with torch.cuda.stream(side):
        def call():
            N.check(lib.kvm_column_select_sampled(d_v.data_ptr(), n, N.MEMBER_STATUS, d_s.data_ptr(), q, 6,
                                                  res.data_ptr(), ws.data_ptr(), ws.numel(),
                                                  torch.cuda.current_stream(dev).cuda_stream))
        d_v.fill_(1.0)
        d_s.fill_(200)
        call()
        side.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            call()
        for rep in range(5):
            v = np.round(rng.lognormal(4.0 + rep, 0.7, n), 3)
            s = np.where(rng.random(n) < 0.8, 200, 503).astype(np.int64)
            d_v.copy_(torch.from_numpy(v))
            d_s.copy_(torch.from_numpy(s))
            g.replay()
            g.replay()       # back to back: the second replay must wait for its own clearing
            side.synchronize()
            want, m = expect(v, s)
            got = read()
            assert [float(x) for x in got["value"]] == want and all(int(x) == m for x in got["n"]), rep


# ---- one-pass form (csrc/kvm_fused.cu): sampled digit parameters, catch-all end bins -----------
def _extreme_quantiles():
    from kvmini_b200 import _native as N
    from kvmini_b200 import engine
    qs = []
    for name, mask in (("lat", engine.LAT_ALL), ("ttfb", engine.TTFB_ALL), ("tllt", 1 << N.SET_TLLT),
                       ("gen", 1 << N.SET_GEN), ("pertok", 1 << N.SET_PERTOK)):
        for p in (0.0, 0.00001, 0.5, 0.9999, 1.0):
            qs.append(engine.Quantile(f"{name}_{p}", mask, p))
    return tuple(qs[:24])      # 48 queries at most


@pytest.mark.parametrize("shape", ["outliers", "unsorted", "sample_sees_failures_only", "two_clusters", "tiny"])
def test_fused_pass_equals_two_pass_and_oracle(shape):
    """The one-pass kernel takes its digit parameters from a 512-row sample; whatever lies outside
    the sampled range lands in the two end bins.  Extreme quantiles (p = 0, 1e-5, 0.9999, 1) force
    ranks into those bins; the answers must equal the two-pass form and a numpy sort."""
    import torch
    from kvmini_b200 import _native as N
    from kvmini_b200 import analyze_api, columns, engine, report
    from oracle import np_port
    n = 300_000 if shape != "tiny" else 37
    cols = _random_log(n, 91, nan_rows=n // 60, zero_tok_rows=n // 50)
    rng = np.random.default_rng(17)
    if shape == "outliers":
        idx = rng.integers(0, n, 40)
        cols["latency_ms"][idx] = 30000.0 + rng.random(40)            # time-outs far above the sampled range
        cols["tllt_ms"][idx] = 29999.0
        cols["ttfb_ms"][rng.integers(0, n, 40)] = 1e-6                # and far below it
    elif shape == "unsorted":
        perm = rng.permutation(n)
        cols = {k: v[perm].copy() for k, v in cols.items()}
    elif shape == "sample_sees_failures_only":
        cols["status"][:] = 500
        ok = rng.integers(0, n, 2000)                                 # a sparse set of successes
        cols["status"][ok] = 200
    elif shape == "two_clusters":
        half = cols["status"] == 200
        cols["latency_ms"] = np.where(rng.random(n) < 0.5, 10.0, 1.0e9) + rng.random(n) * 1e-3
    t0 = float(np.nanmin(cols["start_ms"])) / 1000.0
    cold = [t0 + 2.0, t0 + 300.0]
    quantiles = _extreme_quantiles()
    dev = {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in cols.items()}
    dlog = engine.DeviceLog.from_tensors(dev)
    one = engine.LogAnalyzer()
    two = engine.LogAnalyzer()
    two.use_fused = False
    ra = one.analyze(dlog, cold, quantiles=quantiles)
    rb = two.analyze(dlog, cold, quantiles=quantiles)
    assert two.last_path == "two-pass"
    assert one.last_path == ("two-pass" if shape == "unsorted" else "fused")     # unsorted: the series needs the true minimum
    for f in ("n_total", "n_success", "n_cold", "n_cold_success", "sum_completion_tokens", "sum_total_tokens",
              "sum_total_tokens_cold", "n_gen_rows", "min_start_ms", "max_end_ms"):
        assert same(getattr(ra.stats, f), getattr(rb.stats, f)), f
    for hi, lo in (("ttfb_sum_hi", "ttfb_sum_lo"), ("lat_sum_hi", "lat_sum_lo")):     # double-double sums: same to ~1e-30
        va = getattr(ra.stats, hi) + getattr(ra.stats, lo)
        vb = getattr(rb.stats, hi) + getattr(rb.stats, lo)
        assert va == vb or abs(va - vb) <= 1e-15 * abs(vb), hi
    assert list(ra.stats.sets.count) == list(rb.stats.sets.count)
    assert list(ra.stats.sets.kmin)[:5] == list(rb.stats.sets.kmin)[:5] and list(ra.stats.sets.kmax)[:5] == list(rb.stats.sets.kmax)[:5]
    for q in quantiles:
        a, b = ra.order[q.name], rb.order[q.name]
        assert (a.n, a.status) == (b.n, b.status), q.name
        assert same(a.v_lo, b.v_lo) and same(a.v_hi, b.v_hi), (q.name, a, b)
    assert np.array_equal(ra.hist_counts, rb.hist_counts)
    assert np.array_equal(ra.flags.cpu().numpy(), rb.flags.cpu().numpy())
    assert same(report.trim_series(ra), report.trim_series(rb))
    # and against numpy for the latency column
    lat_ok = np.sort(cols["latency_ms"][(cols["status"] == 200) & ~np.isnan(cols["latency_ms"])])
    for p in (0.0, 0.00001, 0.5, 0.9999, 1.0):
        st = ra.order[f"lat_{p}"]
        if len(lat_ok):
            pos = p * (len(lat_ok) - 1)
            assert st.v_lo == lat_ok[int(math.floor(pos))] and st.v_hi == lat_ok[int(math.ceil(pos))], p
    # the default analysis through the public API: every results.json key against the oracle
    log = columns.columns_from_arrays(**cols)
    got, _ = analyze_api.analyze_log(log, cold, one)
    exp = np_port.analyze(cols, cold)
    assert diff_keys(got, {k: exp[k] for k in got}, rel=MEAN_TOL) == []


def test_fused_pass_graph_replay_and_repeat():
    import torch
    from kvmini_b200 import engine, report
    cols = _random_log(500_003, 5)
    dev = {k: torch.from_numpy(v).cuda() for k, v in cols.items()}
    dlog = engine.DeviceLog.from_tensors(dev)
    t0 = float(cols["start_ms"].min()) / 1000.0
    an = engine.LogAnalyzer()
    first = report.core_results(an.analyze(dlog, [t0 + 5.0]), 1)
    assert an.last_path == "fused"
    for _ in range(4):
        again = report.core_results(an.analyze(dlog, [t0 + 5.0], use_graph=True), 1)
        assert an.last_path == "fused" and diff_keys(again, first) == []


def test_tail_slots_change_grids_not_results():
    """kvm_set_tail_slots: the streaming kernels leave that many CTAs out of their machine-filling
    grids (room for the few-CTA kernels of another analysis in flight).  A log large enough for a
    full grid gives bit-identical results with 0, 16 and 40 slots - and with three analyses in
    flight on their own streams."""
    import torch
    from kvmini_b200 import engine, report
    from kvmini_b200 import _native as N
    lib = N.lib()
    cols = _random_log(3_000_001, 9)
    dev = {k: torch.from_numpy(v).cuda() for k, v in cols.items()}
    dlog = engine.DeviceLog.from_tensors(dev)
    t0 = float(cols["start_ms"].min()) / 1000.0
    try:
        outs = []
        for slots in (0, 16, 40):
            N.check(lib.kvm_set_tail_slots(slots), "kvm_set_tail_slots")
            an = engine.LogAnalyzer()
            outs.append(report.core_results(an.analyze(dlog, [t0 + 5.0]), 1))
            assert an.last_path == "fused"
        assert diff_keys(outs[1], outs[0]) == [] and diff_keys(outs[2], outs[0]) == []
        N.check(lib.kvm_set_tail_slots(16), "kvm_set_tail_slots")
        ring = [engine.LogAnalyzer(stream=torch.cuda.Stream()) for _ in range(3)]
        pending = [an.analyze_async(dlog, [t0 + 5.0], use_graph=False) for an in ring]
        for p in pending:
            assert diff_keys(report.core_results(p(), 1), outs[0]) == []
    finally:
        lib.kvm_set_tail_slots(16)
