// kvm_log_device.cuh - device functions shared by the flat and the segmented (per-run)
// kernels: per-row semantics of the request log, key digits, rank formulas, histogram bins.
// Everything lives in an anonymous namespace (one copy per translation unit).
#pragma once

#include "kvm_common.cuh"

namespace {

constexpr int BINS_LOG2 = 11;          // select: bins per base set
constexpr int BINS = 1 << BINS_LOG2;

// 64-bit slots of one partial.  [0,NSUM) summed, [NSUM,NSUM+NMIN) min, then max, then dd pairs.
enum {
  S_SUCCESS = 0, S_COLD, S_COLD_SUCCESS, S_CTOK, S_TTOK, S_TTOK_COLD, S_GEN_ROWS,
  S_CNT0,                      // 7 set counts
  NSUM = S_CNT0 + KVM_LOG_NSETS,
  M_KMIN0 = NSUM,              // 5 column minima + start
  M_START = M_KMIN0 + KVM_LOG_NCOLS,
  NMIN_END = M_START + 1,
  X_KMAX0 = NMIN_END,          // 5 column maxima + end
  X_END = X_KMAX0 + KVM_LOG_NCOLS,
  NMAX_END = X_END + 1,
  D_TTFB = NMAX_END,           // dd pairs (hi, lo)
  D_LAT = D_TTFB + 2,
  NSLOTS = D_LAT + 2
};

// Per-thread accumulator of the scan.  Row counts are 32-bit (a thread sees far fewer than
// 2^31 rows), token sums 64-bit, ranges tracked as doubles (fmin/fmax skip NaN) and turned
// into ordered keys only in the reduction.
struct LogAcc {
  int32_t cnt[NSUM];        // counters; the three token slots are unused here
  int64_t tok[3];           // S_CTOK, S_TTOK, S_TTOK_COLD
  double vmin[KVM_LOG_NCOLS + 1];
  double vmax[KVM_LOG_NCOLS + 1];
  kvm_dd ttfb, lat;

  __device__ __forceinline__ void init() {
    const double inf = __longlong_as_double(0x7ff0000000000000ll);
#pragma unroll
    for (int i = 0; i < NSUM; ++i) cnt[i] = 0;
    tok[0] = tok[1] = tok[2] = 0;
#pragma unroll
    for (int i = 0; i <= KVM_LOG_NCOLS; ++i) {
      vmin[i] = inf;
      vmax[i] = -inf;
    }
    ttfb = {0.0, 0.0};
    lat = {0.0, 0.0};
  }
  __device__ __forceinline__ void range(int c, double v) {
    vmin[c] = fmin(vmin[c], v);
    vmax[c] = fmax(vmax[c], v);
  }
  __device__ __forceinline__ int64_t sum(int s) const {
    return s == S_CTOK ? tok[0] : s == S_TTOK ? tok[1] : s == S_TTOK_COLD ? tok[2] : (int64_t)cnt[s];
  }
  // ordered keys of the range; an untouched range gives kmin > kmax.  A zero bound is widened
  // to cover both signed zeros (fmin/fmax do not order -0.0 against +0.0).
  __device__ __forceinline__ uint64_t kmin(int c) const {
    return vmin[c] == 0.0 ? kvm_key(-0.0) : kvm_key(vmin[c]);
  }
  __device__ __forceinline__ uint64_t kmax(int c) const {
    return vmax[c] == 0.0 ? kvm_key(0.0) : kvm_key(vmax[c]);
  }
};

// Cold-start classification (analyze.py:402-419): row is cold iff some cold start c has
// c <= start_ms/1000.0 <= c + window.  x/1000.0 is monotone in x, so each cold start is turned
// once into the closed interval [lo, hi] of start_ms values that satisfy the reference
// expression (bisection over the ordered doubles, evaluating that expression), and rows are
// classified with comparisons only.  With sorted cold starts lo[] and hi[] are both ascending,
// so the last interval with lo <= start decides.
struct ColdTable {
  const double* lo;
  const double* hi;
  int n;
};

// smallest key whose value v satisfies pred(v) (pred monotone: false ... false true ... true),
// searched by galloping from `guess`; the key range is all finite doubles.
template <class Pred>
__device__ __forceinline__ uint64_t first_key_where(double guess, Pred pred) {
  const uint64_t kmin = kvm_key(-1.7976931348623157e308), kmax = kvm_key(1.7976931348623157e308);
  uint64_t g = kvm_key(guess);
  g = g < kmin ? kmin : (g > kmax ? kmax : g);
  uint64_t a, z;  // pred(a) false, pred(z) true
  if (pred(kvm_unkey(g))) {
    z = g;
    uint64_t step = 1;
    for (;;) {
      if (z == kmin) return kmin;
      a = (z - kmin > step) ? z - step : kmin;
      if (!pred(kvm_unkey(a))) break;
      z = a;
      step <<= 1;
    }
  } else {
    a = g;
    uint64_t step = 1;
    for (;;) {
      if (a == kmax) return ~0ull;  // never true
      z = (kmax - a > step) ? a + step : kmax;
      if (pred(kvm_unkey(z))) break;
      a = z;
      step <<= 1;
    }
  }
  while (z - a > 1) {
    const uint64_t mid = a + ((z - a) >> 1);
    if (pred(kvm_unkey(mid))) z = mid; else a = mid;
  }
  return z;
}

__device__ __forceinline__ void cold_interval(double c, double window, double* lo, double* hi) {
  const double c_end = c + window;                       // analyze.py:415
  // smallest s with s/1000.0 >= c
  const uint64_t k_lo = first_key_where(c * 1000.0, [&](double s) { return s / 1000.0 >= c; });
  *lo = k_lo == ~0ull ? __longlong_as_double(0x7ff0000000000000ll) : kvm_unkey(k_lo);
  // largest s with s/1000.0 <= c_end == predecessor of the smallest s with s/1000.0 > c_end
  const uint64_t k_hi = first_key_where(c_end * 1000.0, [&](double s) { return s / 1000.0 > c_end; });
  *hi = k_hi == ~0ull ? 1.7976931348623157e308 : kvm_unkey(k_hi - 1);
}

__device__ __forceinline__ bool is_cold(double start_ms, const ColdTable& t) {
  int lo = 0, hi = t.n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (t.lo[mid] <= start_ms) lo = mid + 1; else hi = mid;
  }
  return lo > 0 && start_ms <= t.hi[lo - 1];
}

__device__ __forceinline__ uint8_t process_row(LogAcc& a, double start, double lat, double ttfb,
                                               double tllt, int64_t status, int64_t ctok,
                                               int64_t ttok, const ColdTable& cold, int cold_override) {
  const bool ok = status == 200;
  const bool cold_row = cold_override >= 0 ? cold_override != 0 : (cold.n > 0 && is_cold(start, cold));
  a.vmin[KVM_LOG_NCOLS] = fmin(a.vmin[KVM_LOG_NCOLS], start);        // NaN skipped by fmin/fmax
  a.vmax[KVM_LOG_NCOLS] = fmax(a.vmax[KVM_LOG_NCOLS], start + lat);  // analyze.py:187
  a.cnt[S_COLD] += cold_row;
  uint8_t f = (ok ? KVM_F_SUCCESS : 0) | (cold_row ? KVM_F_COLD : 0) | (ctok > 0 ? KVM_F_TOKENS : 0);
  if (ok) {
    a.cnt[S_SUCCESS] += 1;
    a.cnt[S_COLD_SUCCESS] += cold_row;
    a.tok[0] += ctok;
    a.tok[1] += ttok;
    a.tok[2] += cold_row ? ttok : 0;
    if (!kvm_isnan(lat)) {
      a.cnt[S_CNT0 + KVM_SET_LAT_COLD] += cold_row;
      a.cnt[S_CNT0 + KVM_SET_LAT_WARM] += !cold_row;
      a.range(KVM_COL_LAT, lat);
      a.lat = kvm_dd_add(a.lat, lat);
    }
    const bool has_ttfb = !kvm_isnan(ttfb), has_tllt = !kvm_isnan(tllt);
    if (has_ttfb) {
      a.cnt[S_CNT0 + KVM_SET_TTFB_COLD] += cold_row;
      a.cnt[S_CNT0 + KVM_SET_TTFB_WARM] += !cold_row;
      a.range(KVM_COL_TTFB, ttfb);
      a.ttfb = kvm_dd_add(a.ttfb, ttfb);
    }
    if (has_tllt) {
      a.cnt[S_CNT0 + KVM_SET_TLLT] += 1;
      a.range(KVM_COL_TLLT, tllt);
    }
    if (has_ttfb && has_tllt && ctok > 0) {
      f |= KVM_F_GEN;
      a.cnt[S_GEN_ROWS] += 1;
      const double g = tllt - ttfb;  // analyze.py:144
      if (!kvm_isnan(g)) {
        a.cnt[S_CNT0 + KVM_SET_GEN] += 1;
        a.range(KVM_COL_GEN, g);
        const double pt = g / (double)ctok;  // analyze.py:146
        if (!kvm_isnan(pt)) {
          a.cnt[S_CNT0 + KVM_SET_PERTOK] += 1;
          a.range(KVM_COL_PERTOK, pt);
        }
      }
    }
  }
  return f;
}

// Warp-level reduction of an accumulator into one row of 64-bit slots (lane 0 writes).
__device__ __forceinline__ void acc_warp_reduce(const LogAcc& a, uint64_t* slots, int lane) {
#pragma unroll
  for (int s = 0; s < NSUM; ++s) {
    const int64_t v = kvm_warp_sum(a.sum(s));
    if (lane == 0) slots[s] = (uint64_t)v;
  }
#pragma unroll
  for (int c = 0; c <= KVM_LOG_NCOLS; ++c) {
    const uint64_t lo = kvm_warp_min(a.kmin(c)), hi = kvm_warp_max(a.kmax(c));
    if (lane == 0) {
      slots[M_KMIN0 + c] = lo;
      slots[X_KMAX0 + c] = hi;
    }
  }
  const kvm_dd t = kvm_warp_sum(a.ttfb), l = kvm_warp_sum(a.lat);
  if (lane == 0) {
    slots[D_TTFB] = (uint64_t)__double_as_longlong(t.hi);
    slots[D_TTFB + 1] = (uint64_t)__double_as_longlong(t.lo);
    slots[D_LAT] = (uint64_t)__double_as_longlong(l.hi);
    slots[D_LAT + 1] = (uint64_t)__double_as_longlong(l.lo);
  }
}

// Final slots -> kvm_log_stats (thread 0 of whoever holds the totals).
__device__ __forceinline__ void stats_from_slots(const uint64_t* fin, int64_t n, kvm_log_stats& r) {
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  r.n_total = n;
  r.n_success = (int64_t)fin[S_SUCCESS];
  r.n_cold = (int64_t)fin[S_COLD];
  r.n_cold_success = (int64_t)fin[S_COLD_SUCCESS];
  r.sum_completion_tokens = (int64_t)fin[S_CTOK];
  r.sum_total_tokens = (int64_t)fin[S_TTOK];
  r.sum_total_tokens_cold = (int64_t)fin[S_TTOK_COLD];
  r.n_gen_rows = (int64_t)fin[S_GEN_ROWS];
  // start / end: an untouched (or all-NaN) range has min > max
  r.min_start_ms = fin[M_START] > kvm_key(1.7976931348623157e308) ? nan : kvm_unkey(fin[M_START]);
  r.max_end_ms = fin[X_END] < kvm_key(-1.7976931348623157e308) ? nan : kvm_unkey(fin[X_END]);
  r.ttfb_sum_hi = __longlong_as_double((long long)fin[D_TTFB]);
  r.ttfb_sum_lo = __longlong_as_double((long long)fin[D_TTFB + 1]);
  r.lat_sum_hi = __longlong_as_double((long long)fin[D_LAT]);
  r.lat_sum_lo = __longlong_as_double((long long)fin[D_LAT + 1]);
  for (int s = 0; s < KVM_MAX_SETS; ++s) {
    r.sets.count[s] = s < KVM_LOG_NSETS ? (int64_t)fin[S_CNT0 + s] : 0;
    r.sets.set_col[s] = 0;
  }
  for (int c = 0; c < KVM_MAX_COLS; ++c) {
    const bool used = c < KVM_LOG_NCOLS && fin[M_KMIN0 + c] <= fin[X_KMAX0 + c];
    r.sets.kmin[c] = used ? fin[M_KMIN0 + c] : ~0ull;
    r.sets.kmax[c] = used ? fin[X_KMAX0 + c] : 0ull;
  }
  r.sets.set_col[KVM_SET_LAT_COLD] = KVM_COL_LAT;
  r.sets.set_col[KVM_SET_LAT_WARM] = KVM_COL_LAT;
  r.sets.set_col[KVM_SET_TTFB_COLD] = KVM_COL_TTFB;
  r.sets.set_col[KVM_SET_TTFB_WARM] = KVM_COL_TTFB;
  r.sets.set_col[KVM_SET_TLLT] = KVM_COL_TLLT;
  r.sets.set_col[KVM_SET_GEN] = KVM_COL_GEN;
  r.sets.set_col[KVM_SET_PERTOK] = KVM_COL_PERTOK;
  r.sets.nsets = KVM_LOG_NSETS;
  r.sets.ncols = KVM_LOG_NCOLS;
}

__device__ __forceinline__ int64_t resolve_rank(int kind, double p, int64_t n, int64_t given) {
  switch (kind) {
    case KVM_RANK_LERP_LO: {  // analyze.py:75-76
      double pos = fmax(0.0, fmin(p, 1.0)) * (double)(n - 1);
      return (int64_t)floor(pos);
    }
    case KVM_RANK_LERP_HI: {  // analyze.py:77
      double pos = fmax(0.0, fmin(p, 1.0)) * (double)(n - 1);
      return (int64_t)ceil(pos);
    }
    case KVM_RANK_ABSOLUTE: return given;
    case KVM_RANK_UPPER_MEDIAN: return n / 2;               // collector.py:333
    case KVM_RANK_FLOOR_PN: return (int64_t)(p * (double)n);  // tgi/invoke.sh:208-210
    case KVM_RANK_FLOOR_PNM1: return (int64_t)(p * (double)(n - 1));
  }
  return -1;
}

// Bin 0 also takes every key below kbase and bin BINS-1 every key beyond the last bin: with
// digit parameters taken from the exact column range neither happens; with parameters estimated
// from a sample (csrc/kvm_fused.cu) the two end bins are catch-alls.
__device__ __forceinline__ uint32_t digit_of(uint64_t key, uint64_t kbase, int shift) {
  if (key < kbase) return 0u;
  uint64_t d = (key - kbase) >> shift;
  return d < (uint64_t)BINS ? (uint32_t)d : (uint32_t)(BINS - 1);
}
// x >> s for s in [0, 64]
__device__ __forceinline__ uint64_t shr64(uint64_t x, int s) { return s >= 64 ? 0ull : x >> s; }
__device__ __forceinline__ int bit_width64(uint64_t x) { return x ? 64 - __clzll((long long)x) : 0; }

// Calls f(set, col, key) for every base set the row belongs to (KVM_SET_* semantics).
template <class F>
__device__ __forceinline__ void log_row_keys(uint8_t fl, double lat, double ttfb, double tllt,
                                             int64_t ctok, F&& f) {
  if (!(fl & KVM_F_SUCCESS)) return;
  const bool cold = fl & KVM_F_COLD;
  if (!kvm_isnan(lat)) f(cold ? KVM_SET_LAT_COLD : KVM_SET_LAT_WARM, KVM_COL_LAT, kvm_key(lat));
  if (!kvm_isnan(ttfb)) f(cold ? KVM_SET_TTFB_COLD : KVM_SET_TTFB_WARM, KVM_COL_TTFB, kvm_key(ttfb));
  if (!kvm_isnan(tllt)) f(KVM_SET_TLLT, KVM_COL_TLLT, kvm_key(tllt));
  if (fl & KVM_F_GEN) {
    const double g = tllt - ttfb;  // analyze.py:144
    if (!kvm_isnan(g)) {
      f(KVM_SET_GEN, KVM_COL_GEN, kvm_key(g));
      const double pt = g / (double)ctok;  // analyze.py:146
      if (!kvm_isnan(pt)) f(KVM_SET_PERTOK, KVM_COL_PERTOK, kvm_key(pt));
    }
  }
}

__device__ __forceinline__ int ref_bin(double v, double lo, double w, int nb) {
  // analyze.py:110 - same operations, same order
  const double x = (v - lo) / w;
  long long i = (long long)x;  // int() truncates toward zero
  return (int)(i < (long long)(nb - 1) ? i : (long long)(nb - 1));
}

__device__ __forceinline__ int bin_of(uint64_t key, const uint64_t* thr, int nb) {
  // number of thresholds <= key  (thr non-decreasing, nb-1 entries)
  int lo = 0, hi = nb - 1;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (thr[mid] <= key) lo = mid + 1; else hi = mid;
  }
  return lo;
}


// smallest key in [klo, khi] whose reference bin index reaches b (klo < khi); ~0 if none does.
// The edge sits within a few ulps of lo + b*w, so the search gallops outward from that guess
// and bisects the bracket it finds (a handful of evaluations instead of 64).
__device__ __forceinline__ uint64_t bin_threshold(uint64_t klo, uint64_t khi, int nb, int b) {
  const double lo = kvm_unkey(klo), hi = kvm_unkey(khi);
  const double w = (hi - lo) / (double)nb;  // analyze.py:104
  if (ref_bin(hi, lo, w, nb) < b) return ~0ull;
  uint64_t g = kvm_key(lo + (double)b * w);
  g = g < klo ? klo : (g > khi ? khi : g);
  uint64_t a, z;  // f(a) < b <= f(z)
  if (ref_bin(kvm_unkey(g), lo, w, nb) >= b) {
    z = g;
    uint64_t step = 1;
    for (;;) {
      a = (z - klo > step) ? z - step : klo;
      if (a == klo || ref_bin(kvm_unkey(a), lo, w, nb) < b) break;
      z = a;
      step <<= 1;
    }
  } else {
    a = g;
    uint64_t step = 1;
    for (;;) {
      z = (khi - a > step) ? a + step : khi;
      if (z == khi || ref_bin(kvm_unkey(z), lo, w, nb) >= b) break;
      a = z;
      step <<= 1;
    }
  }
  while (z - a > 1) {
    const uint64_t mid = a + ((z - a) >> 1);
    if (ref_bin(kvm_unkey(mid), lo, w, nb) >= b) z = mid; else a = mid;
  }
  return z;
}

}  // namespace
