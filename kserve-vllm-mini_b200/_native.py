"""ctypes binding of libkvmini_b200.so (the C ABI declared in include/kvmini_b200.h).

There is no CPU fallback: if the shared library is missing or a call fails, the caller
gets an exception.  torch tensors appear here only as device buffers (`.data_ptr()`).
"""

from __future__ import annotations

import ctypes as C
import os
from typing import Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libkvmini_b200.so")

KVM_OK = 0
KVM_ERR_ARENA = -4
KVM_ERR_UNSUPPORTED = -5
CSV_CELL_NUMBER, CSV_CELL_EMPTY, CSV_CELL_NAN_TEXT, CSV_CELL_IRREGULAR = range(4)
CSV_NEEDS_CSV_MODULE = 7
KVM_MAX_SETS = 8
KVM_MAX_COLS = 8
KVM_MAX_QUERIES = 48
KVM_MAX_HISTS = 4
KVM_HIST_MAX_BINS = 64
KVM_PEER_MAX_WORLD = 16

# row flags (kvm_log_scan)
F_SUCCESS, F_COLD, F_TOKENS, F_GEN = 0x01, 0x02, 0x04, 0x08
# base sets / columns of the request log
SET_LAT_COLD, SET_LAT_WARM, SET_TTFB_COLD, SET_TTFB_WARM, SET_TLLT, SET_GEN, SET_PERTOK = range(7)
COL_LAT, COL_TTFB, COL_TLLT, COL_GEN, COL_PERTOK = range(5)
# rank kinds
RANK_LERP_LO, RANK_LERP_HI, RANK_ABSOLUTE, RANK_UPPER_MEDIAN, RANK_FLOOR_PN, RANK_FLOOR_PNM1 = range(6)
# member modes
MEMBER_ALL, MEMBER_SET_IDS, MEMBER_STATUS = range(3)
# query status
Q_OK, Q_EMPTY, Q_OUT_OF_RANGE = range(3)


class NativeError(RuntimeError):
    """A libkvmini_b200 entry point returned an error code."""


class RequestLogC(C.Structure):
    _fields_ = [("start_ms", C.c_void_p), ("latency_ms", C.c_void_p), ("ttfb_ms", C.c_void_p),
                ("tllt_ms", C.c_void_p), ("status", C.c_void_p), ("completion_tokens", C.c_void_p),
                ("total_tokens", C.c_void_p), ("n", C.c_int64)]


class SetInfoC(C.Structure):
    _fields_ = [("count", C.c_int64 * KVM_MAX_SETS), ("kmin", C.c_uint64 * KVM_MAX_COLS),
                ("kmax", C.c_uint64 * KVM_MAX_COLS), ("set_col", C.c_int32 * KVM_MAX_SETS),
                ("nsets", C.c_int32), ("ncols", C.c_int32)]


class LogStatsC(C.Structure):
    _fields_ = [("n_total", C.c_int64), ("n_success", C.c_int64), ("n_cold", C.c_int64),
                ("n_cold_success", C.c_int64), ("sum_completion_tokens", C.c_int64),
                ("sum_total_tokens", C.c_int64), ("sum_total_tokens_cold", C.c_int64),
                ("n_gen_rows", C.c_int64), ("min_start_ms", C.c_double), ("max_end_ms", C.c_double),
                ("ttfb_sum_hi", C.c_double), ("ttfb_sum_lo", C.c_double),
                ("lat_sum_hi", C.c_double), ("lat_sum_lo", C.c_double), ("sets", SetInfoC)]


class PowerStatsC(C.Structure):
    _fields_ = [("m", C.c_int64), ("n_valid", C.c_int64), ("n_outside", C.c_int64),
                ("n_outside_valid", C.c_int64), ("n_unsorted", C.c_int64), ("ts_min", C.c_double),
                ("ts_max", C.c_double), ("step_sum_hi", C.c_double), ("step_sum_lo", C.c_double)]


class PowerTraceC(C.Structure):
    _fields_ = [("ts", C.c_void_p), ("watts", C.c_void_p), ("p_hi", C.c_void_p), ("p_lo", C.c_void_p),
                ("m", C.c_int64)]


KVM_MAX_TRACES = 8


class HistSpecC(C.Structure):
    _fields_ = [("nhist", C.c_int32), ("num_bins", C.c_int32), ("col", C.c_int32 * 4),
                ("set_mask", C.c_uint32 * 4)]


class LogExtrasC(C.Structure):
    _fields_ = [("num_bins", C.c_int32), ("pad", C.c_int32), ("hist_counts", C.c_void_p),
                ("series_window_ms", C.c_double), ("series_max_bins", C.c_int64),
                ("series", C.c_void_p), ("series_dropped", C.c_void_p), ("digit_records", C.c_void_p)]


class QueryC(C.Structure):
    _fields_ = [("set_mask", C.c_uint32), ("rank_kind", C.c_int32), ("p", C.c_double),
                ("rank", C.c_int64)]


class QueryResultC(C.Structure):
    _fields_ = [("value", C.c_double), ("n", C.c_int64), ("rank", C.c_int64),
                ("status", C.c_int32), ("pad", C.c_int32)]


_lib: Optional[C.CDLL] = None


def _declare(lib: C.CDLL) -> None:
    vp, i32, i64, f64, sz = C.c_void_p, C.c_int32, C.c_int64, C.c_double, C.c_size_t
    P = C.POINTER
    sigs = {
        "kvm_last_error": (C.c_char_p, []),
        "kvm_abi_version": (C.c_int, []),
        "kvm_sm_count": (C.c_int, []),
        "kvm_log_scan_workspace_bytes": (sz, []),
        "kvm_log_scan": (C.c_int, [P(RequestLogC), vp, i32, f64, vp, vp, vp, vp, sz, vp]),
        "kvm_log_stats_pack_sums": (i32, []),
        "kvm_log_stats_pack_mins": (i32, []),
        "kvm_log_stats_pack": (C.c_int, [vp, vp, vp, vp]),
        "kvm_log_stats_unpack": (C.c_int, [vp, vp, vp, vp]),
        "kvm_log_stats_pack_words": (i32, []),
        "kvm_log_stats_pack_words_into": (C.c_int, [vp, vp, vp]),
        "kvm_log_stats_combine": (C.c_int, [vp, vp, i32, vp]),
        "kvm_select_workspace_bytes": (sz, [i64]),
        "kvm_select_default_arena": (i64, [i64]),
        "kvm_log_select": (C.c_int, [P(RequestLogC), vp, vp, P(QueryC), i32, vp, P(LogExtrasC), vp, sz, i64, vp]),
        "kvm_log_analyze_fused_scan_workspace_bytes": (sz, []),
        "kvm_log_fused_pass": (C.c_int, [P(RequestLogC), vp, i32, f64, vp, vp, P(LogExtrasC), vp, sz, vp, sz, i64, vp]),
        "kvm_log_fused_preset_words": (i32, []),
        "kvm_log_fused_sample": (C.c_int, [P(RequestLogC), vp, vp]),
        "kvm_log_fused_preset_combine": (C.c_int, [vp, i32, i64, vp, vp]),
        "kvm_log_fused_pass_preset": (C.c_int, [P(RequestLogC), vp, i32, f64, vp, vp, P(LogExtrasC), vp, sz, vp, sz, i64, vp, vp]),
        "kvm_select_plan_preset": (C.c_int, [vp, P(QueryC), i32, P(HistSpecC), vp, sz, i64, i32, i64, vp]),
        "kvm_log_analyze_fused": (C.c_int, [P(RequestLogC), vp, i32, f64, vp, vp, P(QueryC), i32, vp, P(LogExtrasC),
                                            vp, sz, vp, sz, i64, i32, i64, vp]),
        "kvm_column_select": (C.c_int, [vp, i64, i32, vp, i32, P(QueryC), i32, vp, vp, sz, i64, vp]),
        "kvm_column_select_sampled": (C.c_int, [vp, i64, i32, vp, P(QueryC), i32, vp, vp, sz, vp]),
        "kvm_column_select_sampled_workspace_bytes": (sz, [i64]),
        "kvm_column_select_sampled_min_n": (i64, []),
        "kvm_column_select_static": (C.c_int, [vp, i64, i32, vp, P(QueryC), i32, vp, vp]),
        "kvm_column_select_static_max_n": (i64, []),
        "kvm_select_status": (C.c_int, [vp, vp]),
        "kvm_column_stats": (C.c_int, [vp, i64, i32, vp, i32, vp, vp, sz, vp]),
        "kvm_select_hist_ptr": (vp, [vp]),
        "kvm_select_edge_ptr": (vp, [vp]),
        "kvm_select_bins": (i32, []),
        "kvm_select_arena_ptr": (vp, [vp]),
        "kvm_select_bucket_counts_ptr": (vp, [vp]),
        "kvm_select_bucket_offsets_ptr": (vp, [vp]),
        "kvm_select_info_ptr": (vp, [vp]),
        "kvm_select_require_capacity": (C.c_int, [vp, i64, vp]),
        "kvm_select_post_words": (i32, []),
        "kvm_select_post_pack": (C.c_int, [vp, vp, vp, vp]),
        "kvm_select_post_unpack": (C.c_int, [vp, vp, i32, vp, vp, vp]),
        "kvm_select_plan": (C.c_int, [vp, P(QueryC), i32, P(HistSpecC), vp, sz, i64, vp]),
        "kvm_select_plan_refine": (C.c_int, [vp, P(QueryC), i32, P(HistSpecC), vp, sz, i64, i32, i64, vp]),
        "kvm_log_select_refine_hist": (C.c_int, [P(RequestLogC), vp, vp, vp, vp]),
        "kvm_select_refine": (C.c_int, [vp, vp]),
        "kvm_select_hist2_ptr": (vp, [vp]),
        "kvm_select_hist2_words": (i64, []),
        "kvm_column_select_hist": (C.c_int, [vp, i64, i32, vp, vp, vp]),
        "kvm_log_select_hist": (C.c_int, [P(RequestLogC), vp, vp, P(LogExtrasC), vp, vp]),
        "kvm_select_resolve": (C.c_int, [vp, vp]),
        "kvm_column_select_compact": (C.c_int, [vp, i64, i32, vp, vp, vp]),
        "kvm_log_select_compact": (C.c_int, [P(RequestLogC), vp, vp, vp, vp]),
        "kvm_select_finish": (C.c_int, [vp, P(vp), i32, vp, vp, vp, vp]),
        "kvm_peer_flag_bytes": (i64, []),
        "kvm_peer_put": (C.c_int, [vp, i64, i64, P(vp), i32, i32, vp]),
        "kvm_peer_add_u64": (C.c_int, [vp, i64, i64, P(vp), i32, i32, vp]),
        "kvm_peer_barrier": (C.c_int, [i64, P(vp), i32, i32, vp]),
        "kvm_peer_put_candidates": (C.c_int, [vp, i64, i64, P(vp), i32, i32, vp]),
        "kvm_peer_set_timeout_s": (f64, [f64]),
        "kvm_column_stats_workspace_bytes": (sz, []),
        "kvm_column_nearest_workspace_bytes": (sz, []),
        "kvm_column_nearest": (C.c_int, [vp, i64, i32, vp, f64, vp, vp, sz, vp]),
        "kvm_csv_open": (C.c_int, [C.c_char_p, P(vp)]),
        "kvm_csv_close": (C.c_int, [vp]),
        "kvm_csv_rows": (i64, [vp]),
        "kvm_csv_ncols": (i32, [vp]),
        "kvm_csv_flags": (i32, [vp]),
        "kvm_csv_col_name": (C.c_char_p, [vp, i32]),
        "kvm_csv_find_col": (i32, [vp, C.c_char_p]),
        "kvm_csv_parse_f64": (C.c_int, [vp, i32, vp, vp, i32]),
        "kvm_csv_parse_columns": (C.c_int, [vp, vp, i32, P(vp), P(vp), i32]),
        "kvm_csv_cell": (i64, [vp, i64, i32, vp, i64]),
        "kvm_csv_match_any": (C.c_int, [vp, i32, P(C.c_char_p), i32, vp, i32]),
        "kvm_csv_write_classified": (C.c_int, [vp, C.c_char_p, vp, vp, i32, i32]),
        "kvm_power_json_parse": (C.c_int, [C.c_char_p, i32, P(vp)]),
        "kvm_prom_result_parse": (C.c_int, [C.c_char_p, i64, i32, P(vp)]),
        "kvm_samples_series": (i32, [vp]),
        "kvm_samples_count": (i64, [vp, i32]),
        "kvm_samples_copy": (C.c_int, [vp, i32, vp, vp]),
        "kvm_samples_free": (None, [vp]),
        "kvm_log_histograms_workspace_bytes": (sz, [i32]),
        "kvm_log_histograms": (C.c_int, [P(RequestLogC), vp, vp, i32, vp, vp, sz, vp]),
        "kvm_column_histogram": (C.c_int, [vp, i64, f64, f64, i32, vp, vp, sz, vp]),
        "kvm_log_series": (C.c_int, [P(RequestLogC), vp, vp, f64, i64, vp, vp, vp]),
        "kvm_log_fused_phase_us": (C.c_int, [vp, vp]),
        "kvm_set_tail_slots": (C.c_int, [i32]),
        "kvm_seg_analyze_scratch_bytes": (sz, [i64]),
        "kvm_seg_phase_times": (C.c_int, [vp, i32]),
        "kvm_seg_analyze": (C.c_int, [P(RequestLogC), vp, i64, vp, vp, f64, P(QueryC), i32, i32, vp, vp, sz, vp, vp, vp, vp, vp]),
        "kvm_sort_pairs_workspace_bytes": (sz, [i64]),
        "kvm_sort_pairs_f64": (C.c_int, [vp, vp, vp, vp, i64, vp, sz, vp]),
        "kvm_energy_workspace_bytes": (sz, [i64]),
        "kvm_power_scan": (C.c_int, [vp, vp, i64, f64, f64, vp, vp, vp, sz, vp]),
        "kvm_trapezoid_wh": (C.c_int, [vp, vp, i64, P(f64), i32, vp, vp, sz, vp]),
        "kvm_energy_prefix": (C.c_int, [vp, vp, i64, vp, vp, vp, sz, vp]),
        "kvm_interval_energy": (C.c_int, [vp, vp, vp, vp, i64, vp, vp, i64, vp, i32, vp]),
        "kvm_ordered_sum_f64": (C.c_int, [vp, i32, i64, i64, vp, vp]),
        "kvm_selfcheck_div3600": (C.c_int, [vp, vp, i64, vp]),
        "kvm_interval_energy_multi": (C.c_int, [P(PowerTraceC), i32, vp, vp, i64, vp, i32, vp]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    lib._kvm_declared = tuple(sigs)


def lib() -> C.CDLL:
    """Load the shared library (once).  Fails loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise NativeError(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; "
                "g.build()'` (or `make -C kserve-vllm-mini_b200/csrc`). There is no CPU fallback.")
        loaded = C.CDLL(LIB_PATH)
        _declare(loaded)
        _lib = loaded
    return _lib


def check(rc: int, what: str = "") -> int:
    if rc != KVM_OK:
        msg = lib().kvm_last_error().decode("utf-8", "replace")
        raise NativeError(f"{what or 'libkvmini_b200'} failed (rc={rc}): {msg}")
    return rc
