"""requests.csv -> columnar request log (host side of the ingest).

The reference keeps every row as a dict and coerces eight named cells with `float()`
when the cell is not None / "" / "NaN" (analyze.py:34-56).  Here the same coercion
produces seven typed columns (the only ones the analysis reads):

  start_ms, latency_ms, ttfb_ms, tllt_ms   float64   (nan where the cell holds `nan`)
  status, completion_tokens, total_tokens  int64

Cells the reference would have left as strings are remembered per column, so that the
analysis raises where the reference raises instead of silently
treating them as missing.
"""

from __future__ import annotations

import csv
import ctypes as C
import math
import os
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Sequence

import numpy as np

# analyze.py:40-49 - the cells read_requests_csv coerces
COERCED = ("start_ms", "ttfb_ms", "tllt_ms", "latency_ms", "status",
           "prompt_tokens", "completion_tokens", "total_tokens")
_UNCOERCED_CELLS = (None, "", "NaN")

F64_COLS = ("start_ms", "latency_ms", "ttfb_ms", "tllt_ms")
I64_COLS = ("status", "completion_tokens", "total_tokens")
SUCCESS = 200


def coerce_cell(cell: Any) -> Any:
    """analyze.py:50-54: float(cell) unless the cell is None/""/"NaN" or not parseable."""
    if cell in _UNCOERCED_CELLS:
        return cell
    try:
        return float(cell)
    except Exception:
        return cell


def read_rows(path: str) -> List[dict]:
    """Drop-in for `analyze.read_requests_csv` (analyze.py:34-56): list of row dicts with
    the eight known numeric cells coerced to float."""
    rows: List[dict] = []
    with open(path, newline="") as f:
        for row in csv.DictReader(f):
            for k in COERCED:
                if k in row:
                    row[k] = coerce_cell(row[k])
            rows.append(row)
    return rows


@dataclass
class RequestLog:
    """Columnar request log on the host.  `cols` holds the seven typed numpy columns;
    `present` says which CSV columns existed; `unparsed[k]` lists row indices whose cell
    in column k stayed a non-float (what the reference would trip over)."""

    n: int
    cols: Dict[str, np.ndarray]
    present: Dict[str, bool]
    unparsed: Dict[str, np.ndarray] = field(default_factory=dict)
    header: Optional[List[str]] = None
    raw_rows: Optional[List[dict]] = None      # kept only when the classified CSV is wanted
    is_cold_csv: Optional[np.ndarray] = None   # `is_cold_start` column, when present (A18)
    native: Optional[Any] = None               # open NativeCsv handle (fast path with keep_rows)
    source_path: Optional[str] = None

    def success_mask(self) -> np.ndarray:
        return self.cols["status"] == SUCCESS


def _status_to_int(v: Any) -> int:
    """`int(r.get("status", 0))` (analyze.py:127 and 8 more call sites), same exceptions."""
    return int(v)  # ValueError on ""/"abc"/nan, TypeError on None, OverflowError on inf


def columns_from_rows(rows: Sequence[dict], keep_rows: bool = False,
                      header: Optional[List[str]] = None) -> RequestLog:
    """Row dicts (already coerced) -> RequestLog, raising where the reference's
    analysis would raise on the same rows."""
    n = len(rows)
    present = {k: (n > 0 and k in rows[0]) for k in F64_COLS + I64_COLS}
    cols: Dict[str, np.ndarray] = {k: np.full(n, np.nan) for k in F64_COLS}
    cols.update({k: np.zeros(n, dtype=np.int64) for k in I64_COLS})
    bad: Dict[str, List[int]] = {k: [] for k in F64_COLS + I64_COLS}

    status = cols["status"]
    for i, r in enumerate(rows):
        status[i] = _status_to_int(r.get("status", 0))
    ok = status == SUCCESS

    for k in F64_COLS:
        default = 0.0 if k in ("start_ms", "latency_ms") else math.nan
        col = cols[k]
        for i, r in enumerate(rows):
            v = r.get(k, default)
            if isinstance(v, float):
                col[i] = v
            elif isinstance(v, int) and not isinstance(v, bool):
                col[i] = float(v)
            else:
                bad[k].append(i)

    for k in ("completion_tokens", "total_tokens"):
        col = cols[k]
        for i, r in enumerate(rows):
            if not ok[i]:
                continue  # token cells of failed rows are never read (analyze.py:502-511)
            v = r.get(k, 0.0) or 0.0   # analyze.py:503,508: `or 0.0`
            if isinstance(v, (int, float)) and not isinstance(v, bool) and v == v \
                    and not math.isinf(v) and float(v) == math.floor(v):
                col[i] = int(v)
            else:
                bad[k].append(i)

    cold_csv = None
    if n > 0 and "is_cold_start" in rows[0]:
        # cost_estimator.py:108-111
        cold_csv = np.fromiter((r.get("is_cold_start") in ("True", "true", "1", True) for r in rows),
                               dtype=bool, count=n)

    return RequestLog(n=n, cols=cols, present=present,
                      unparsed={k: np.asarray(v, dtype=np.int64) for k, v in bad.items() if v},
                      header=header, raw_rows=list(rows) if keep_rows else None,
                      is_cold_csv=cold_csv)


class NativeCsv:
    """An open kvm_csv handle (csrc/kvm_csv.cu): mmap of the file + row index."""

    def __init__(self, path: str):
        from . import _native as N
        self._N, self.lib = N, N.lib()
        h = C.c_void_p()
        N.check(self.lib.kvm_csv_open(os.fsencode(path), C.byref(h)), "kvm_csv_open")
        self.h = h
        self.path = path
        self.rows = int(self.lib.kvm_csv_rows(h))
        self.flags = int(self.lib.kvm_csv_flags(h))
        self.header = [self.lib.kvm_csv_col_name(h, i).decode("utf-8")
                       for i in range(int(self.lib.kvm_csv_ncols(h)))]

    def close(self) -> None:
        if self.h is not None:
            self.lib.kvm_csv_close(self.h)
            self.h = None

    __del__ = close

    def column(self, name: str):
        """(values float64[n], kinds uint8[n]) of the named column, or None when absent."""
        c = int(self.lib.kvm_csv_find_col(self.h, name.encode()))
        if c < 0:
            return None
        val = np.empty(self.rows, dtype=np.float64)
        kind = np.empty(self.rows, dtype=np.uint8)
        self._N.check(self.lib.kvm_csv_parse_f64(self.h, c, val.ctypes.data, kind.ctypes.data, 0), "kvm_csv_parse_f64")
        return val, kind

    def columns(self, names: Sequence[str]) -> Dict[str, tuple]:
        """{name: (values, kinds)} for the names that exist, parsed in one pass over the file."""
        idx = {k: int(self.lib.kvm_csv_find_col(self.h, k.encode())) for k in names}
        have = [k for k in names if idx[k] >= 0]
        out = {k: (np.empty(self.rows, dtype=np.float64), np.empty(self.rows, dtype=np.uint8)) for k in have}
        if have:
            cols = np.asarray([idx[k] for k in have], dtype=np.int32)
            vals = (C.c_void_p * len(have))(*[out[k][0].ctypes.data for k in have])
            kinds = (C.c_void_p * len(have))(*[out[k][1].ctypes.data for k in have])
            self._N.check(self.lib.kvm_csv_parse_columns(self.h, cols.ctypes.data, len(have), vals, kinds, 0),
                          "kvm_csv_parse_columns")
        return out

    def cell(self, row: int, name: str) -> str:
        c = int(self.lib.kvm_csv_find_col(self.h, name.encode()))
        n = int(self.lib.kvm_csv_cell(self.h, row, c, None, 0))
        buf = C.create_string_buffer(max(n, 1))
        self.lib.kvm_csv_cell(self.h, row, c, buf, n)
        return buf.raw[:n].decode("utf-8")

    def matches(self, name: str, options: Sequence[str]) -> np.ndarray:
        c = int(self.lib.kvm_csv_find_col(self.h, name.encode()))
        out = np.zeros(self.rows, dtype=np.uint8)
        arr = (C.c_char_p * len(options))(*[o.encode() for o in options])
        self._N.check(self.lib.kvm_csv_match_any(self.h, c, arr, len(options), out.ctypes.data, 0), "kvm_csv_match_any")
        return out.astype(bool)

    def write_classified(self, out_path: str, is_cold: np.ndarray) -> bool:
        """analyze.py:597-604 natively; False when a coerced cell needs Python's float() (caller
        then rewrites the file through the csv module)."""
        cold = np.ascontiguousarray(is_cold, dtype=np.uint8)
        cols = [int(self.lib.kvm_csv_find_col(self.h, k.encode())) for k in COERCED]
        cols = np.asarray([c for c in cols if c >= 0], dtype=np.int32)
        rc = self.lib.kvm_csv_write_classified(self.h, os.fsencode(out_path), cold.ctypes.data, cols.ctypes.data,
                                               int(cols.size), 0)
        if rc == self._N.KVM_ERR_UNSUPPORTED:
            return False
        self._N.check(rc, "kvm_csv_write_classified")
        return True


def _cell_value(kind: int, native: NativeCsv, row: int, name: str) -> Any:
    """What the reference's row dict holds for a cell the fast path did not parse."""
    from . import _native as N
    if kind == N.CSV_CELL_EMPTY:
        return ""
    if kind == N.CSV_CELL_NAN_TEXT:
        return "NaN"
    return coerce_cell(native.cell(row, name))


def _load_native(path: str, keep_handle: bool) -> Optional[RequestLog]:
    """Columnar ingest through csrc/kvm_csv.cu; None when the file needs the csv-module path."""
    from . import _native as N
    nat = NativeCsv(path)
    if (nat.flags & N.CSV_NEEDS_CSV_MODULE) or len(set(nat.header)) != len(nat.header):
        nat.close()
        return None
    n = nat.rows
    present = {k: (n > 0 and k in nat.header) for k in F64_COLS + I64_COLS}
    cols: Dict[str, np.ndarray] = {k: np.full(n, np.nan) for k in F64_COLS}
    cols.update({k: np.zeros(n, dtype=np.int64) for k in I64_COLS})
    bad: Dict[str, List[int]] = {k: [] for k in F64_COLS + I64_COLS}

    parsed = nat.columns([k for k in F64_COLS + I64_COLS if present[k]])
    # status first: `int(r.get("status", 0))`, exceptions raised for the first offending row
    if present["status"]:
        val, kind = parsed["status"]
        plain = (kind == N.CSV_CELL_NUMBER) & (np.abs(val) < 2.0 ** 62)
        cols["status"][plain] = np.trunc(val[plain]).astype(np.int64)
        for i in np.nonzero(~plain)[0].tolist():
            v = val[i] if kind[i] == N.CSV_CELL_NUMBER else _cell_value(int(kind[i]), nat, i, "status")
            cols["status"][i] = _status_to_int(v if not isinstance(v, np.floating) else float(v))
    ok = cols["status"] == SUCCESS

    for k in F64_COLS:
        if not present[k]:
            if k in ("start_ms", "latency_ms"):
                cols[k][:] = 0.0
            continue
        val, kind = parsed[k]
        col = cols[k]
        plain = kind == N.CSV_CELL_NUMBER
        col[plain] = val[plain]
        for i in np.nonzero(kind == N.CSV_CELL_IRREGULAR)[0].tolist():
            v = coerce_cell(nat.cell(i, k))
            if isinstance(v, float):
                col[i] = v
            else:
                bad[k].append(i)
        bad[k] = sorted(bad[k] + np.nonzero((kind == N.CSV_CELL_EMPTY) | (kind == N.CSV_CELL_NAN_TEXT))[0].tolist())

    for k in ("completion_tokens", "total_tokens"):
        if not present[k]:
            continue
        val, kind = parsed[k]
        col = cols[k]
        integral = (kind == N.CSV_CELL_NUMBER) & (val == np.floor(val)) & (np.abs(val) < 2.0 ** 62)
        take = ok & integral
        col[take] = val[take].astype(np.int64)
        # "" -> `or 0.0` -> 0 (already 0); everything else goes through the reference's expression
        rest = ok & ~integral & (kind != N.CSV_CELL_EMPTY)
        for i in np.nonzero(rest)[0].tolist():
            v = float(val[i]) if kind[i] == N.CSV_CELL_NUMBER else _cell_value(int(kind[i]), nat, i, k)
            v = v or 0.0
            if isinstance(v, (int, float)) and not isinstance(v, bool) and v == v \
                    and not math.isinf(v) and float(v) == math.floor(v):
                col[i] = int(v)
            else:
                bad[k].append(i)

    cold_csv = nat.matches("is_cold_start", ("True", "true", "1")) if (n > 0 and "is_cold_start" in nat.header) else None
    log = RequestLog(n=n, cols=cols, present=present,
                     unparsed={k: np.asarray(v, dtype=np.int64) for k, v in bad.items() if v},
                     header=list(nat.header) if nat.header else None, raw_rows=None, is_cold_csv=cold_csv)
    if keep_handle:
        log.native = nat
    else:
        nat.close()
    return log


def load_requests_csv(path: str, keep_rows: bool = False, native: bool = True) -> RequestLog:
    """requests.csv -> RequestLog.  `native`: columnar fast path (csrc/kvm_csv.cu) with the
    csv-module path below as the fallback for files it does not model; both give the same log.
    keep_rows: keep what the classified-CSV writer needs (the open native handle, or the row dicts)."""
    if native:
        log = _load_native(path, keep_handle=keep_rows)
        if log is not None:
            log.source_path = path
            return log
    with open(path, newline="") as f:
        reader = csv.DictReader(f)
        header = reader.fieldnames
        rows = []
        for row in reader:
            for k in COERCED:
                if k in row:
                    row[k] = coerce_cell(row[k])
            rows.append(row)
    return columns_from_rows(rows, keep_rows=keep_rows, header=list(header) if header else None)


def columns_from_arrays(**arrays: np.ndarray) -> RequestLog:
    """Wrap already-columnar data (synthetic generators, .npy sidecars)."""
    n = int(arrays["status"].shape[0])
    cols = {}
    for k in F64_COLS:
        cols[k] = np.ascontiguousarray(arrays[k], dtype=np.float64)
    for k in I64_COLS:
        cols[k] = np.ascontiguousarray(arrays[k], dtype=np.int64)
    return RequestLog(n=n, cols=cols, present={k: True for k in F64_COLS + I64_COLS})


def check_analyzable(log: RequestLog) -> None:
    """Raise what analyze.py raises on rows it cannot digest."""
    ok = log.success_mask()
    for k in ("start_ms", "latency_ms"):
        if k in log.unparsed:
            raise TypeError(f"unsupported operand type(s) for +: non-numeric {k} cell "
                            f"in row {int(log.unparsed[k][0])}")
    for k in ("ttfb_ms", "tllt_ms"):
        if k in log.unparsed:
            hit = log.unparsed[k][ok[log.unparsed[k]]]
            if hit.size:
                # analyze.py:493 -> math.isnan("") on a status-200 row
                raw = None
                if log.raw_rows is not None:
                    raw = log.raw_rows[int(hit[0])].get(k)
                kind = "NoneType" if raw is None and log.raw_rows is not None else "str"
                raise TypeError(f"must be real number, not {kind}")
    for k in ("completion_tokens", "total_tokens"):
        if k in log.unparsed:
            raise ValueError(f"{k}: non-integer token count in row {int(log.unparsed[k][0])} "
                             "is not representable in the int64 column layout")
