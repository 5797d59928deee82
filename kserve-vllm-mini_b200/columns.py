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
import math
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


def load_requests_csv(path: str, keep_rows: bool = False) -> RequestLog:
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
