// kvm_fused_row.cuh - the one-pass row of the request log, shared by the flat one-pass kernel
// (kvm_fused.cu) and the per-run kernel of the sweep roll-up (kvm_seg.cu): the 512-row sample,
// the digit parameters taken from the sample's hull, the branch-free row (flags, counters, sums,
// digit record, shared-memory histogram counts) and its rare exact route.
// Reference semantics per row: analyze.py:183-189, :402-419, :496-514, :68,:89,:143-146.
#pragma once

#include "kvm_common.cuh"
#include "kvm_log_device.cuh"

namespace {

constexpr int NC = KVM_LOG_NCOLS;

// Per value column: exact key range so far, the sampled hull, digit parameters.
struct FusedCols {
  unsigned long long kmin[NC], kmax[NC];     // exact key range of every value column (CTA level)
  uint64_t kbase[NC];
  int32_t shift[NC];
  int32_t fast_ok;
  uint4 par[NC];   // fast-path parameters per column: kbase high word, inside floor, inside width, shift - 32
};

__device__ __forceinline__ void fused_cols_publish(FusedCols& S, int c);

// u64 min / max on shared memory, called on the rare path only
__device__ __forceinline__ void range_update(unsigned long long* lo, unsigned long long* hi, uint64_t key) {
  if (key < *reinterpret_cast<volatile unsigned long long*>(lo)) atomicMin(lo, (unsigned long long)key);
  if (key > *reinterpret_cast<volatile unsigned long long*>(hi)) atomicMax(hi, (unsigned long long)key);
}

// Sites of one row whose key lies outside the sampled hull (bit c of `rare` = value column c),
// redone exactly: digit with both clamps, exact range, histogram count, record field.
__device__ __noinline__ uint64_t fused_fix_sites(FusedCols& S, uint32_t* sh, uint32_t rare, uint32_t f, double la,
                                                 double tf, double tl, long long ct, uint64_t r) {
  const bool cold_row = f & KVM_F_COLD;
  const double g = tl - tf;
  for (int c = 0; c < NC; ++c) {
    if (!((rare >> c) & 1u)) continue;
    double v;
    int set;
    if (c == KVM_COL_LAT) { v = la; set = cold_row ? KVM_SET_LAT_COLD : KVM_SET_LAT_WARM; }
    else if (c == KVM_COL_TTFB) { v = tf; set = cold_row ? KVM_SET_TTFB_COLD : KVM_SET_TTFB_WARM; }
    else if (c == KVM_COL_TLLT) { v = tl; set = KVM_SET_TLLT; }
    else if (c == KVM_COL_GEN) { v = g; set = KVM_SET_GEN; }
    else { v = g / (double)ct; set = KVM_SET_PERTOK; }
    const uint64_t key = kvm_key(v);
    const uint32_t d = digit_of(key, S.kbase[c], S.shift[c]);
    range_update(&S.kmin[c], &S.kmax[c], key);
    fused_cols_publish(S, c);   // keys below this one are inside from now on
    atomicAdd(&sh[set * BINS + d], 1u);
    r = (r & ~(0xFFFull << (12 * c))) | ((uint64_t)d << (12 * c));
  }
  return r;
}

struct FAcc {
  int32_t n_success, n_cold, n_cold_success, n_gen;
  int64_t ctok, ttok, ttok_cold;
  double st_min, end_max;
  double ttfb_s, ttfb_e, lat_s, lat_e;   // sum and accumulated rounding error (two_sum)

  __device__ __forceinline__ void init() {
    n_success = n_cold = n_cold_success = n_gen = 0;
    ctok = ttok = ttok_cold = 0;
    st_min = __longlong_as_double(0x7ff0000000000000ll);
    end_max = -st_min;
    ttfb_s = ttfb_e = lat_s = lat_e = 0.0;
  }
};

__device__ __forceinline__ void dd_acc(double& s, double& e, double x) {
  const double t = s + x;
  const double bb = t - s;
  e += (s - (t - bb)) + (x - bb);
  s = t;
}

// The NSAMPLE-row sample (a multiple of 32, at least 64): NSAMPLE / 32 regions spread evenly over
// the log (the first row and the last row included), 32 pseudo-random rows inside a 4096-row
// window of each.  Per value column the
// complemented minimum and the maximum key of the sampled MEMBERS (0 / 0: none), and the minimum
// sampled start_ms.  Thread t of `nthreads` takes rows t, t + nthreads, ...
constexpr int FU_PRESET_WORDS = 2 * NC + 1;   // [inv_min[NC] | max[NC] | bits of the start reference]
template <int NSAMPLE>
__device__ __forceinline__ void fused_sample_rows(const kvm_request_log& log, int t, int nthreads,
                                                  uint64_t (&s_inv)[NC], uint64_t (&s_mx)[NC], double& s_st) {
  const int64_t n = log.n;
  if (n < 1) return;
  const int64_t win = n < 4096 ? n : 4096;
  for (int sidx = t; sidx < NSAMPLE; sidx += nthreads) {
    const int64_t region = sidx >> 5;
    const int64_t off = n > win ? region * ((n - win) / (NSAMPLE / 32 - 1)) : 0;
    int64_t i = off + (int64_t)(((uint32_t)sidx * 2654435761u >> 7) % (uint32_t)win);
    if (sidx == 0) i = 0;
    if (sidx == NSAMPLE - 1) i = n - 1;
    const double la = __ldg(log.latency_ms + i), tf = __ldg(log.ttfb_ms + i), tl = __ldg(log.tllt_ms + i);
    const double st = __ldg(log.start_ms + i);
    const bool ok = __ldg(log.status + i) == 200;
    const long long ct = __ldg(log.completion_tokens + i);
    s_st = fmin(s_st, st);
    auto take = [&](int c, bool mem, double v) {
      if (mem) {
        const uint64_t key = kvm_key(v);
        s_inv[c] = ~key > s_inv[c] ? ~key : s_inv[c];
        s_mx[c] = key > s_mx[c] ? key : s_mx[c];
      }
    };
    take(KVM_COL_LAT, ok && la == la, la);
    take(KVM_COL_TTFB, ok && tf == tf, tf);
    take(KVM_COL_TLLT, ok && tl == tl, tl);
    const bool gen = ok && tf == tf && tl == tl && ct > 0;
    const double g = tl - tf;
    take(KVM_COL_GEN, gen && g == g, g);
    const double pt = g / (double)ct;
    take(KVM_COL_PERTOK, gen && g == g && pt == pt, pt);
  }
}

// Digit parameters of column c from the sample's hull (inv = complemented sampled minimum, 0:
// the sample holds no member of the column).  For a column wide enough (shift >= 32) the base is
// rounded down to a multiple of 2^32, which makes a digit a function of the key's HIGH word alone.
// Returns the significant bits of the span.
__device__ __forceinline__ int fused_cols_setup(FusedCols& S, int c, uint64_t inv, uint64_t mx) {
  int width;
  if (inv == 0ull) {   // no member of this column in the sample: coarse bins over all keys
    S.kbase[c] = 0ull;
    S.shift[c] = 64 - BINS_LOG2;
    width = 64;
    S.kmin[c] = ~0ull;
    S.kmax[c] = 0ull;
  } else {
    // the bins span the smallest power of two that holds the hull; the hull sits in the middle of
    // that span, so what the sample missed on either side still lands in a bin of its own
    const uint64_t mn = ~inv, range = mx - mn;
    width = range ? bit_width64(range) : 1;
    uint64_t lo = mn;
    if (width >= BINS_LOG2 + 32) {
      if (width < 64) {
        const uint64_t span = 1ull << width, slack = span - 1ull - range;
        const uint64_t back = slack >> 1;
        lo = (mn > back ? mn - back : 0ull) & ~0xFFFFFFFFull;
        if (mx - lo >= span) ++width;   // (no slack to speak of: the rounding of the base cost a bit)
      } else {
        lo = mn & ~0xFFFFFFFFull;
      }
    }
    S.kbase[c] = lo;
    S.shift[c] = width > BINS_LOG2 ? width - BINS_LOG2 : 0;
    S.kmin[c] = mn;   // sampled values are members: the exact range starts from their hull
    S.kmax[c] = mx;
  }
  return width;
}

// Fast-path parameters of column c (shared memory: one 16-byte load per site).  With every shift
// >= 32 the digit is a 32-bit shift of x = (high word of the key) - (high word of kbase), and a key
// is "inside" when x lies STRICTLY between the high words of the column's smallest and largest key
// so far (one subtraction, one unsigned comparison).  Inside keys cannot move the exact range;
// everything else takes the exact route, which republishes these parameters when it widens the
// range - so after the sample only record-setting keys (and the two boundary words) are rare.
// The inside interval never leaves the span of the bins, hence an inside digit needs no clamp.
__device__ __forceinline__ void fused_cols_publish(FusedCols& S, int c) {
  uint4 P;
  P.x = (uint32_t)(S.kbase[c] >> 32);
  P.y = 0u;
  P.z = 0u;   // nothing is inside: every site takes the exact route
  P.w = S.fast_ok ? (uint32_t)(S.shift[c] - 32) : 0u;
  if (S.fast_ok) {
    const uint64_t kb = S.kbase[c];
    const int sh = S.shift[c];
    const uint64_t cover = sh + BINS_LOG2 >= 64 ? ~0ull : (((uint64_t)BINS << sh) - 1ull);
    const uint64_t top_key = kb + cover < kb ? ~0ull : kb + cover;
    const uint64_t kmin = *reinterpret_cast<volatile unsigned long long*>(&S.kmin[c]);
    const uint64_t kmax = *reinterpret_cast<volatile unsigned long long*>(&S.kmax[c]);
    const uint64_t lo = kmin > kb ? kmin : kb, hi = kmax < top_key ? kmax : top_key;
    if (lo <= hi) {
      const uint64_t a = (lo - kb) >> 32, t = (hi - kb) >> 32;
      if (t > a + 1) {
        P.y = (uint32_t)(a + 1);
        P.z = (uint32_t)(t - a - 1);
      }
    }
  }
  S.par[c] = P;
}
__device__ __forceinline__ void fused_cols_fast_params(FusedCols& S) {   // one thread, after the set-up
  bool fast_ok = true;
  for (int c = 0; c < NC; ++c) fast_ok = fast_ok && S.shift[c] >= 32;
  S.fast_ok = fast_ok ? 1 : 0;
  for (int c = 0; c < NC; ++c) fused_cols_publish(S, c);
}

// The fast row: branch-free; sites whose key is not inside (fused_cols_publish) are only noted in
// `rare` and redone exactly by fused_fix_sites.  Returns
// the row's digit record (five 12-bit fields, 0xFFF: not a member of that column).
__device__ __forceinline__ uint64_t fused_row(FAcc& a, FusedCols& S, uint32_t* sh, const ColdTable& cold,
                                              double cold_first, double cold_last, double st, double la, double tf,
                                              double tl, long long status, long long ct, long long tt,
                                              uint32_t& flag_out) {
  const bool ok = status == 200;
  bool cold_row = false;
  if (st >= cold_first && st <= cold_last) cold_row = is_cold(st, cold);
  a.st_min = fmin(a.st_min, st);
  a.end_max = fmax(a.end_max, st + la);
  const bool okc = ok && cold_row;
  a.n_cold += cold_row;
  a.n_success += ok;
  a.n_cold_success += okc;
  a.ctok += ok ? ct : 0;
  a.ttok += ok ? tt : 0;
  a.ttok_cold += okc ? tt : 0;
  const bool has_tf = tf == tf, has_tl = tl == tl;
  const bool m_la = ok && la == la, m_tf = ok && has_tf, m_tl = ok && has_tl;
  const bool gen = m_tf && has_tl && ct > 0;
  const double g = tl - tf;                 // analyze.py:144
  const bool m_g = gen && g == g;
  // analyze.py:146.  The quotient is only used where m_g holds (then ct > 0).  A zero numerator
  // or divisor would send the division through its slow subroutine on every non-streaming row
  // (tllt == ttfb): x/ct is x itself for x = +-0, so those rows divide 1.0 by 1.0 instead.
  const bool dv = m_g && g != 0.0;
  const double quo = (dv ? g : 1.0) / (dv ? (double)ct : 1.0);
  const double pt = dv ? quo : g;
  const bool m_pt = m_g && pt == pt;
  a.n_gen += gen;
  dd_acc(a.lat_s, a.lat_e, m_la ? la : 0.0);
  dd_acc(a.ttfb_s, a.ttfb_e, m_tf ? tf : 0.0);
  uint32_t rare = 0;
  auto site = [&](bool m, uint32_t* hist, int col, double v) -> uint32_t {
    const uint32_t hi = (uint32_t)__double2hiint(v);
    const uint32_t khi = hi ^ ((uint32_t)((int32_t)hi >> 31) | 0x80000000u);   // high word of the ordered key
    const uint4 P = S.par[col];
    const uint32_t x = khi - P.x;
    const bool out = (x - P.y) >= P.z;
    const uint32_t d = x >> P.w;
    if (m && !out) atomicAdd(hist + d, 1u);
    rare |= (m && out) ? (1u << col) : 0u;
    return m ? (out ? 0u : d) : 0xFFFu;
  };
  uint32_t* hcw = sh + (cold_row ? 0 : BINS);   // the cold / warm pair of a column: two consecutive sets
  const uint32_t d0 = site(m_la, hcw + KVM_SET_LAT_COLD * BINS, KVM_COL_LAT, la);
  const uint32_t d1 = site(m_tf, hcw + KVM_SET_TTFB_COLD * BINS, KVM_COL_TTFB, tf);
  const uint32_t d2 = site(m_tl, sh + KVM_SET_TLLT * BINS, KVM_COL_TLLT, tl);
  const uint32_t d3 = site(m_g, sh + KVM_SET_GEN * BINS, KVM_COL_GEN, g);
  const uint32_t d4 = site(m_pt, sh + KVM_SET_PERTOK * BINS, KVM_COL_PERTOK, pt);
  const uint32_t f = (ok ? KVM_F_SUCCESS : 0) | (cold_row ? KVM_F_COLD : 0) | (ct > 0 ? KVM_F_TOKENS : 0) |
                     (gen ? KVM_F_GEN : 0);
  flag_out = f;
  const uint32_t rlo = d0 | (d1 << 12) | (d2 << 24), rhi = (d2 >> 8) | (d3 << 4) | (d4 << 16);
  uint64_t r = ((uint64_t)rhi << 32) | rlo;
  if (rare) r = fused_fix_sites(S, sh, rare, f, la, tf, tl, ct, r);
  return r | ((uint64_t)cold_row << 60);   // bit 60 (REC_COLD_BIT): the compact pass needs no flag byte
}

// Per-thread accumulators -> one warp's row of 64-bit slots (slot layout of kvm_log_device.cuh;
// the set counts and the column ranges are filled in by the caller).  Lane 0 writes.
__device__ __forceinline__ void facc_warp_reduce(const FAcc& a, uint64_t* r, int lane) {
  auto wsum = [&](int64_t v) { return (uint64_t)kvm_warp_sum(v); };
  const uint64_t v_succ = wsum(a.n_success), v_cold = wsum(a.n_cold), v_cs = wsum(a.n_cold_success);
  const uint64_t v_gen = wsum(a.n_gen), v_ct = wsum(a.ctok), v_tt = wsum(a.ttok), v_tc = wsum(a.ttok_cold);
  // start / end as ordered keys (untouched: +inf / -inf, which stats_from_slots maps to NaN)
  const uint64_t v_st = kvm_warp_min(kvm_key(a.st_min)), v_en = kvm_warp_max(kvm_key(a.end_max));
  auto dd_of = [](double s, double e) {
    const double hi = s + e;
    return kvm_dd{hi, e - (hi - s)};
  };
  const kvm_dd tfs = kvm_warp_sum(dd_of(a.ttfb_s, a.ttfb_e)), las = kvm_warp_sum(dd_of(a.lat_s, a.lat_e));
  if (lane == 0) {
    r[S_SUCCESS] = v_succ; r[S_COLD] = v_cold; r[S_COLD_SUCCESS] = v_cs; r[S_CTOK] = v_ct;
    r[S_TTOK] = v_tt; r[S_TTOK_COLD] = v_tc; r[S_GEN_ROWS] = v_gen;
    r[M_START] = v_st; r[X_END] = v_en;
    r[D_TTFB] = (uint64_t)__double_as_longlong(tfs.hi); r[D_TTFB + 1] = (uint64_t)__double_as_longlong(tfs.lo);
    r[D_LAT] = (uint64_t)__double_as_longlong(las.hi); r[D_LAT + 1] = (uint64_t)__double_as_longlong(las.lo);
  }
}

__device__ __forceinline__ uint64_t slot_combine(int slot, uint64_t acc, uint64_t v) {
  if (slot < NSUM) return (uint64_t)((int64_t)acc + (int64_t)v);
  if (slot < NMIN_END) return v < acc ? v : acc;
  return v > acc ? v : acc;
}
__device__ __forceinline__ kvm_dd dd_from_bits(uint64_t hi, uint64_t lo) {
  return kvm_dd{__longlong_as_double((long long)hi), __longlong_as_double((long long)lo)};
}

}  // namespace
