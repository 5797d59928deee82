"""Native CSV ingest / classified-CSV writer (csrc/kvm_csv.cu) against the
csv-module path of columns.py (which restates analyze.py:34-56,597-604).  CPU only."""
import csv
import io
import math
import os
import random

import numpy as np
import pytest

import kvmini_b200  # noqa: F401  (root shim puts the package on the path)
from kvmini_b200 import analyze_api, columns

HEADER = ["id", "scheduled_ms", "start_ms", "ttfb_ms", "tllt_ms", "latency_ms", "status", "prompt_tokens",
          "completion_tokens", "total_tokens", "error", "trace_id"]


def _same_log(a, b):
    assert a.n == b.n and a.present == b.present and a.header == b.header
    for k in a.cols:
        assert np.array_equal(a.cols[k], b.cols[k], equal_nan=True), k
        # bit-exact, not just equal: -0.0 vs 0.0, nan payloads aside
        if a.cols[k].dtype == np.float64:
            fa, fb = a.cols[k], b.cols[k]
            m = ~np.isnan(fa)
            assert np.array_equal(fa[m].view(np.int64), fb[m].view(np.int64)), k
    assert {k: v.tolist() for k, v in a.unparsed.items()} == {k: v.tolist() for k, v in b.unparsed.items()}
    if a.is_cold_csv is None:
        assert b.is_cold_csv is None
    else:
        assert np.array_equal(a.is_cold_csv, b.is_cold_csv)


def _write(path, rows, header=HEADER, lineterminator="\r\n", trailing=True):
    buf = io.StringIO()
    w = csv.writer(buf, lineterminator=lineterminator)
    w.writerow(header)
    w.writerows(rows)
    text = buf.getvalue()
    if not trailing:
        text = text[: -len(lineterminator)]
    with open(path, "w", newline="") as f:
        f.write(text)


def _rows(n, seed, odd_cells=()):
    rnd = random.Random(seed)
    out = []
    for i in range(n):
        ok = rnd.random() > 0.05
        lat = round(rnd.lognormvariate(5.5, 0.6), 3)
        row = [f"req-{i}", "%.3f" % (i * 25.0), "%.3f" % (1.7e12 + i * 25.0 + rnd.random()),
               "%.3f" % (0.2 * lat) if ok else "", "%.3f" % (0.98 * lat) if ok else "", "%.3f" % lat,
               "200" if ok else "500", "12", str(rnd.randint(16, 128)) if ok else "", str(140) if ok else "",
               "" if ok else 'HTTP 500, "boom"\nsecond line', f"t{i:08x}"]
        out.append(row)
    for (r, c, text) in odd_cells:
        out[r][HEADER.index(c)] = text
    return out


@pytest.mark.parametrize("term,trailing", [("\r\n", True), ("\n", True), ("\n", False)])
def test_plain_file_matches_csv_module_path(tmp_path, term, trailing):
    p = str(tmp_path / "requests.csv")
    _write(p, _rows(3000, 1), lineterminator=term, trailing=trailing)
    nat = columns._load_native(p, keep_handle=False)
    assert nat is not None
    _same_log(nat, columns.load_requests_csv(p, native=False))


def test_irregular_cells_go_through_pythons_float(tmp_path):
    odd = [(3, "latency_ms", " 12.5 "), (4, "latency_ms", "1e3"), (5, "latency_ms", "+5"), (6, "latency_ms", "1_0"),
           (7, "latency_ms", "inf"), (8, "latency_ms", "-Infinity"), (9, "latency_ms", "nan"),
           (10, "latency_ms", "NaN"), (11, "latency_ms", "abc"), (12, "latency_ms", "1e999"),
           (13, "latency_ms", "1e-999"), (14, "latency_ms", "0x10"), (15, "latency_ms", ".5"),
           (16, "latency_ms", "5."), (17, "latency_ms", "-0.0"), (18, "latency_ms", "123456789012345678901234567890"),
           (19, "latency_ms", "0.1e-2"), (20, "latency_ms", "1E5"), (21, "latency_ms", "١٢"),
           (22, "ttfb_ms", "NaN"), (23, "ttfb_ms", "nan"), (24, "tllt_ms", " "), (25, "start_ms", "1.7e12"),
           (26, "status", "200.0"), (27, "status", " 200 "), (28, "status", "2e2"), (29, "status", "404.9"),
           (30, "completion_tokens", "17.0"), (31, "completion_tokens", "17.5"), (32, "completion_tokens", "NaN"),
           (33, "completion_tokens", " 9 "), (34, "total_tokens", "1e2"), (35, "total_tokens", "0"),
           (36, "total_tokens", "-3"), (37, "total_tokens", "abc"), (38, "completion_tokens", "inf")]
    rows = _rows(60, 2, odd)
    for r in range(26, 39):
        rows[r][HEADER.index("status")] = rows[r][HEADER.index("status")] if r < 30 else "200"
    p = str(tmp_path / "requests.csv")
    _write(p, rows)
    nat = columns._load_native(p, keep_handle=False)
    assert nat is not None
    _same_log(nat, columns.load_requests_csv(p, native=False))


@pytest.mark.parametrize("cell,exc", [("", ValueError), ("NaN", ValueError), ("abc", ValueError), ("nan", ValueError),
                                      ("inf", OverflowError)])
def test_status_cells_raise_like_the_reference(tmp_path, cell, exc):
    rows = _rows(20, 3, [(7, "status", cell)])
    p = str(tmp_path / "requests.csv")
    _write(p, rows)
    with pytest.raises(exc) as e_py:
        columns.load_requests_csv(p, native=False)
    with pytest.raises(exc) as e_nat:
        columns.load_requests_csv(p, native=True)
    assert str(e_py.value) == str(e_nat.value)


def test_files_outside_the_model_fall_back(tmp_path):
    base = _rows(10, 4)
    # ragged row
    p = str(tmp_path / "ragged.csv")
    _write(p, base)
    with open(p, "a", newline="") as f:
        f.write("x,0,1,2,3,4,200,1,2,3,,t,extra\r\n")
    assert columns._load_native(p, keep_handle=False) is None
    _same_log(columns.load_requests_csv(p), columns.load_requests_csv(p, native=False))
    with open(p, "a", newline="") as f:
        f.write("x,1,2\r\n")          # short row: status is None -> the reference's int(None)
    for native in (True, False):
        with pytest.raises(TypeError):
            columns.load_requests_csv(p, native=native)
    # bare quote inside an unquoted field
    p = str(tmp_path / "bare.csv")
    _write(p, base)
    with open(p, "a", newline="") as f:
        f.write('r,0,1,2,3,4,200,1,2,3,ab"c,t\r\n')
    assert columns._load_native(p, keep_handle=False) is None
    # lone carriage return as row terminator
    p = str(tmp_path / "cr.csv")
    _write(p, base, lineterminator="\r")
    assert columns._load_native(p, keep_handle=False) is None
    # duplicate header names
    p = str(tmp_path / "dup.csv")
    _write(p, [r + ["1"] for r in base], header=HEADER + ["status"])
    assert columns._load_native(p, keep_handle=False) is None
    _same_log(columns.load_requests_csv(p), columns.load_requests_csv(p, native=False))


def test_blank_lines_missing_columns_and_cold_column(tmp_path):
    rows = _rows(50, 5)
    hdr = [h for h in HEADER if h not in ("tllt_ms", "total_tokens")] + ["is_cold_start"]
    keep = [HEADER.index(h) for h in hdr[:-1]]
    rnd = random.Random(9)
    data = [[r[i] for i in keep] + [rnd.choice(["True", "true", "1", "False", "0", "", "TRUE"])] for r in rows]
    p = str(tmp_path / "requests.csv")
    _write(p, data, header=hdr)
    text = open(p, newline="").read().replace("\r\n", "\r\n\r\n", 7)   # blank lines are skipped by csv
    with open(p, "w", newline="") as f:
        f.write(text)
    nat = columns._load_native(p, keep_handle=False)
    assert nat is not None and nat.n == 50
    _same_log(nat, columns.load_requests_csv(p, native=False))
    with open(p, "w", newline="") as f:
        f.write("\r\n" + text)      # DictReader: a blank first line means "no field names"
    assert columns._load_native(p, keep_handle=False) is None


def test_empty_and_header_only_files(tmp_path):
    p = str(tmp_path / "requests.csv")
    open(p, "w").close()
    _same_log(columns.load_requests_csv(p), columns.load_requests_csv(p, native=False))
    with open(p, "w", newline="") as f:
        f.write(",".join(HEADER) + "\r\n")
    a = columns.load_requests_csv(p)
    assert a.n == 0
    _same_log(a, columns.load_requests_csv(p, native=False))


def _classified_by_csv_module(path, out, is_cold):
    log = columns.load_requests_csv(path, keep_rows=True, native=False)
    analyze_api._write_classified_csv(out, log, is_cold)


@pytest.mark.parametrize("with_cold_col", [False, True])
def test_classified_csv_is_byte_identical(tmp_path, with_cold_col):
    rnd = random.Random(11)
    specials = [1e16, 1e15, 9999999999999998.0, 1e-5, 1e-4, 0.0001234, 123456789012345678.0, 5e-324, 1.7976931348623157e308,
                0.1, 1 / 3, 2.5e-7, 100.0, 1e22, 1e21, 123456.789e3, -0.0, 0.0, 7.0, 1700000000123.456]
    rows = _rows(400, 6)
    for i, v in enumerate(specials):
        rows[i][HEADER.index("latency_ms")] = repr(v)
        rows[i + 30][HEADER.index("start_ms")] = "%.6e" % v
        rows[i + 60][HEADER.index("prompt_tokens")] = "%.17g" % (v * rnd.random())
    for i in range(100, 400):
        rows[i][HEADER.index("ttfb_ms")] = repr(rnd.uniform(-1e6, 1e6) * 10.0 ** rnd.randint(-12, 12))
    rows[5][HEADER.index("tllt_ms")] = "NaN"
    rows[6][HEADER.index("error")] = 'plain, with comma'
    rows[7][HEADER.index("error")] = 'has "quotes"'
    rows[8][HEADER.index("trace_id")] = ' leading space'
    hdr = HEADER + (["is_cold_start"] if with_cold_col else [])
    if with_cold_col:
        rows = [r + ["old"] for r in rows]
    p = str(tmp_path / "requests.csv")
    _write(p, rows, header=hdr)
    is_cold = np.array([rnd.random() < 0.3 for _ in rows])
    out_nat, out_py = str(tmp_path / "nat.csv"), str(tmp_path / "py.csv")
    log = columns.load_requests_csv(p, keep_rows=True)
    assert log.native is not None
    analyze_api._write_classified_csv(out_nat, log, is_cold)
    _classified_by_csv_module(p, out_py, is_cold)
    assert open(out_nat, "rb").read() == open(out_py, "rb").read()


def test_classified_csv_with_irregular_cell_falls_back_to_csv_module(tmp_path):
    rows = _rows(30, 7, [(3, "latency_ms", " 12.5 "), (4, "latency_ms", "abc")])
    p = str(tmp_path / "requests.csv")
    _write(p, rows)
    is_cold = np.zeros(30, dtype=bool)
    log = columns.load_requests_csv(p, keep_rows=True)
    assert log.native is not None and not log.native.write_classified(str(tmp_path / "x.csv"), is_cold)
    out_nat, out_py = str(tmp_path / "nat.csv"), str(tmp_path / "py.csv")
    analyze_api._write_classified_csv(out_nat, log, is_cold)
    _classified_by_csv_module(p, out_py, is_cold)
    assert open(out_nat, "rb").read() == open(out_py, "rb").read()


def test_float_repr_matches_python_on_random_doubles(tmp_path):
    rng = np.random.default_rng(5)
    bits = rng.integers(0, 1 << 63, 20000, dtype=np.int64)
    vals = bits.view(np.float64)
    vals = vals[np.isfinite(vals)]
    vals = np.concatenate([vals, -vals[:2000], 10.0 ** np.arange(-30, 30), np.arange(0, 50, dtype=np.float64)])
    rows = [[f"r{i}", "0", repr(float(v)), "1", "2", "3", "200", "1", "2", "3", "", "t"] for i, v in enumerate(vals)]
    p = str(tmp_path / "requests.csv")
    _write(p, rows)
    log = columns.load_requests_csv(p, keep_rows=True)
    assert np.array_equal(log.cols["start_ms"].view(np.int64), vals.view(np.int64))   # parse: bit-exact
    out = str(tmp_path / "c.csv")
    assert log.native.write_classified(out, np.zeros(len(rows), dtype=bool))
    with open(out, newline="") as f:
        got = [r["start_ms"] for r in csv.DictReader(f)]
    assert got == [repr(float(v)) for v in vals]
