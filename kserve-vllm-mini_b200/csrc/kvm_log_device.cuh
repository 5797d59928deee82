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

struct LogAcc {
  int64_t sum[NSUM];
  uint64_t kmin[KVM_LOG_NCOLS + 1];
  uint64_t kmax[KVM_LOG_NCOLS + 1];
  kvm_dd ttfb, lat;

  __device__ __forceinline__ void init() {
#pragma unroll
    for (int i = 0; i < NSUM; ++i) sum[i] = 0;
#pragma unroll
    for (int i = 0; i <= KVM_LOG_NCOLS; ++i) {
      kmin[i] = ~0ull;
      kmax[i] = 0ull;
    }
    ttfb = {0.0, 0.0};
    lat = {0.0, 0.0};
  }
  __device__ __forceinline__ void range(int c, double v) {
    uint64_t k = kvm_key(v);
    kmin[c] = k < kmin[c] ? k : kmin[c];
    kmax[c] = k > kmax[c] ? k : kmax[c];
  }
};

__device__ __forceinline__ bool is_cold(double t_s, const double* cold, int nc, double window) {
  // largest c <= t decides (c + window is monotone in c)
  int lo = 0, hi = nc;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (cold[mid] <= t_s) lo = mid + 1; else hi = mid;
  }
  return lo > 0 && t_s <= cold[lo - 1] + window;
}

__device__ __forceinline__ uint8_t process_row(LogAcc& a, double start, double lat, double ttfb,
                                               double tllt, int64_t status, int64_t ctok,
                                               int64_t ttok, const double* cold, int nc,
                                               double window, int cold_override) {
  const bool ok = status == 200;
  const bool cold_row = cold_override >= 0 ? cold_override != 0
                                           : (nc > 0 && is_cold(start / 1000.0, cold, nc, window));
  if (!kvm_isnan(start)) {
    uint64_t ks = kvm_key(start);
    a.kmin[KVM_LOG_NCOLS] = ks < a.kmin[KVM_LOG_NCOLS] ? ks : a.kmin[KVM_LOG_NCOLS];
  }
  const double end = start + lat;
  if (!kvm_isnan(end)) {
    uint64_t ke = kvm_key(end);
    a.kmax[KVM_LOG_NCOLS] = ke > a.kmax[KVM_LOG_NCOLS] ? ke : a.kmax[KVM_LOG_NCOLS];
  }
  a.sum[S_COLD] += cold_row;
  uint8_t f = (ok ? KVM_F_SUCCESS : 0) | (cold_row ? KVM_F_COLD : 0) | (ctok > 0 ? KVM_F_TOKENS : 0);
  if (ok) {
    a.sum[S_SUCCESS] += 1;
    a.sum[S_COLD_SUCCESS] += cold_row;
    a.sum[S_CTOK] += ctok;
    a.sum[S_TTOK] += ttok;
    a.sum[S_TTOK_COLD] += cold_row ? ttok : 0;
    if (!kvm_isnan(lat)) {
      a.sum[S_CNT0 + (cold_row ? KVM_SET_LAT_COLD : KVM_SET_LAT_WARM)] += 1;
      a.range(KVM_COL_LAT, lat);
      a.lat = kvm_dd_add(a.lat, lat);
    }
    const bool has_ttfb = !kvm_isnan(ttfb), has_tllt = !kvm_isnan(tllt);
    if (has_ttfb) {
      a.sum[S_CNT0 + (cold_row ? KVM_SET_TTFB_COLD : KVM_SET_TTFB_WARM)] += 1;
      a.range(KVM_COL_TTFB, ttfb);
      a.ttfb = kvm_dd_add(a.ttfb, ttfb);
    }
    if (has_tllt) {
      a.sum[S_CNT0 + KVM_SET_TLLT] += 1;
      a.range(KVM_COL_TLLT, tllt);
    }
    if (has_ttfb && has_tllt && ctok > 0) {
      f |= KVM_F_GEN;
      a.sum[S_GEN_ROWS] += 1;
      const double g = tllt - ttfb;  // analyze.py:144
      if (!kvm_isnan(g)) {
        a.sum[S_CNT0 + KVM_SET_GEN] += 1;
        a.range(KVM_COL_GEN, g);
        const double pt = g / (double)ctok;  // analyze.py:146
        if (!kvm_isnan(pt)) {
          a.sum[S_CNT0 + KVM_SET_PERTOK] += 1;
          a.range(KVM_COL_PERTOK, pt);
        }
      }
    }
  }
  return f;
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

__device__ __forceinline__ uint32_t digit_of(uint64_t key, uint64_t kbase, int shift) {
  uint64_t d = (key - kbase) >> shift;
  return d < (uint64_t)BINS ? (uint32_t)d : (uint32_t)(BINS - 1);
}

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


// smallest key in [klo, khi] whose reference bin index reaches b (klo < khi); ~0 if none does
__device__ __forceinline__ uint64_t bin_threshold(uint64_t klo, uint64_t khi, int nb, int b) {
  const double lo = kvm_unkey(klo), hi = kvm_unkey(khi);
  const double w = (hi - lo) / (double)nb;  // analyze.py:104
  if (ref_bin(hi, lo, w, nb) < b) return ~0ull;
  uint64_t a = klo, z = khi;  // f(a) < b <= f(z)
  while (z - a > 1) {
    const uint64_t mid = a + ((z - a) >> 1);
    if (ref_bin(kvm_unkey(mid), lo, w, nb) >= b) z = mid; else a = mid;
  }
  return z;
}

}  // namespace
