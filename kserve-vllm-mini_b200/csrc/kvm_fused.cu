// kvm_fused.cu - ONE pass over the request log that does the work of kvm_log_scan (K1+K7) AND of
// the select's histogram pass (K2/K6/K8): window bounds, cold/warm flags, counts, token sums,
// double-double sums, exact key ranges, the seven per-set digit histograms, the 8-byte digit
// records and the 1-s window series - 56 B/row read, 9 B/row written (57 + 49 B/row in two passes
// before).
//
// The histogram pass used to need the exact key range of every value column (digit =
// (key - kmin) >> shift), i.e. a pass of its own.  Here every CTA draws the SAME 512 rows (16
// regions x 32 rows inside a 4096-row window: few pages, the sample's cost is TLB misses) and
// takes the digit parameters from the sample's range: kbase = sampled minimum, shift so that the
// sampled range spans <= 2048 bins.  Keys outside the sampled range fall into the two end bins
// (digit_of clamps), which the select treats as catch-alls (kvm_select.cu: plan marks them for
// exact histogram binning, resolve spans their buckets with the true range) - so the result is
// exact for any input, the sample only decides how evenly the bins are filled.  A sampled value
// is a real member of its column, hence values inside the sampled hull can never be the column's
// extreme: the exact range costs one extra comparison per value and a rare shared-memory update.
//
// Replaces (reference file:line) analyze.py:183-189, :402-419, :496-514, :68,:89,:143 filters,
// :581, the histogram side of :59-81 / :84-122 (via kvm_select.cu) and extension A12.

#include <algorithm>
#include <cstddef>

#include "kvm_common.cuh"
#include "kvm_log_device.cuh"
#include "kvm_fused_row.cuh"
#include "kvm_select_ws.cuh"

namespace {

constexpr int FU_THREADS = 384;   // x 2 CTAs per SM at <= 85 registers: 24 warps per SM hide the row arithmetic's latencies
constexpr int FU_WARPS = FU_THREADS / 32;
constexpr int FU_COLD_SMEM = 512;
constexpr int FU_MAX_GRID = 148 * 8;

struct FusedWs {   // lives in the scan workspace (kvm_log_scan_workspace_bytes)
  unsigned int ticket;
  unsigned int pad[3];
  // phase time stamps of the last pass (globaltimer ns; cleared with the ticket): [0] ~first CTA entry
  // (as ~t), [1] last "sample + set-up done", [2] last "stream done", [3] last "flushed", [4] end
  unsigned long long dbg[8];
  // totals of the integer / range slots, accumulated by every CTA with atomics on a cleared block
  // (minima stored complemented, so that every slot is a sum or a maximum over zero)
  unsigned long long fin[NMAX_END];
  unsigned int pad2[2 * (64 - 10 - NMAX_END) + 2];
  alignas(16) uint64_t partial[1];  // [grid][NSLOTS]
};

__device__ __forceinline__ unsigned long long fu_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

struct FusedSeries {
  double window_ms;
  int64_t max_bins;
  unsigned long long* series;   // [3][max_bins]
  unsigned long long* dropped;  // [0] rows outside [0, max_bins); [2] = bits of the reference start the bins are relative to
};

struct FusedSmem {
  uint64_t red[FU_WARPS][NSLOTS];
  uint64_t fin[NSLOTS];
  FusedCols fc;
  unsigned long long s_inv[FU_WARPS][NC], s_max[FU_WARPS][NC];
  double s_start[FU_WARPS];
  double series_ref;
  double cold_lo[FU_COLD_SMEM], cold_hi[FU_COLD_SMEM];
  int is_last;
};

// A shard's sample record, for the ranks to agree on one hull before the pass (sharded logs).
__global__ void __launch_bounds__(512) fused_sample_kernel(kvm_request_log log, unsigned long long* __restrict__ record) {
  __shared__ unsigned long long r_inv[16][NC], r_mx[16][NC];
  __shared__ double r_st[16];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  uint64_t s_inv[NC], s_mx[NC];
  double s_st = __longlong_as_double(0x7ff0000000000000ll);
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    s_inv[c] = 0ull;
    s_mx[c] = 0ull;
  }
  fused_sample_rows<512>(log, t, 512, s_inv, s_mx, s_st);
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const uint64_t a = kvm_warp_max(s_inv[c]), b = kvm_warp_max(s_mx[c]);
    if (lane == 0) {
      r_inv[warp][c] = a;
      r_mx[warp][c] = b;
    }
  }
#pragma unroll
  for (int m = 16; m; m >>= 1) s_st = fmin(s_st, __shfl_xor_sync(KVM_FULL, s_st, m));
  if (lane == 0) r_st[warp] = s_st;
  __syncthreads();
  if (t < NC) {
    unsigned long long inv = 0, mx = 0;
    for (int w = 0; w < 16; ++w) {
      inv = r_inv[w][t] > inv ? r_inv[w][t] : inv;
      mx = r_mx[w][t] > mx ? r_mx[w][t] : mx;
    }
    record[t] = inv;
    record[NC + t] = mx;
  }
  if (t == 32) {
    double s0 = r_st[0];
    for (int w = 1; w < 16; ++w) s0 = fmin(s0, r_st[w]);
    record[2 * NC] = (unsigned long long)__double_as_longlong(s0);
  }
}

// records of all ranks (stride words apart) -> the hull every rank's pass uses
__global__ void fused_preset_combine_kernel(const unsigned long long* __restrict__ records, int world, int64_t stride,
                                            unsigned long long* __restrict__ preset) {
  const int t = threadIdx.x;
  if (t < 2 * NC) {
    unsigned long long a = 0;
    for (int r = 0; r < world; ++r) {
      const unsigned long long v = records[(size_t)r * stride + t];
      a = v > a ? v : a;
    }
    preset[t] = a;
  } else if (t == 2 * NC) {
    double s0 = __longlong_as_double(0x7ff0000000000000ll);
    for (int r = 0; r < world; ++r) s0 = fmin(s0, __longlong_as_double((long long)records[(size_t)r * stride + t]));
    preset[t] = (unsigned long long)__double_as_longlong(s0);
  }
}

template <bool COLD_SMEM_TABLE, bool SERIES>
__global__ void __launch_bounds__(FU_THREADS, 2)
log_fused_kernel(kvm_request_log log, const double* __restrict__ cold_lo_g, const double* __restrict__ cold_hi_g,
                 int nc, uint8_t* __restrict__ flags, uint64_t* __restrict__ recs, kvm_log_stats* __restrict__ out,
                 FusedWs* fws, SelWs* ws, FusedSeries sa, const unsigned long long* __restrict__ preset, int vec) {
  extern __shared__ __align__(16) unsigned char fraw[];
  uint32_t* sh = reinterpret_cast<uint32_t*>(fraw);   // [KVM_LOG_NSETS][BINS]
  FusedSmem& S = *reinterpret_cast<FusedSmem*>(fraw + sizeof(uint32_t) * KVM_LOG_NSETS * BINS);
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int64_t n = log.n;

  // ---- sample: 512 rows, the same in every CTA (or, for a sharded log, the hull that all ranks
  // have agreed on before the pass: `preset`)
  uint64_t s_inv[NC], s_mx[NC];
  double s_st = __longlong_as_double(0x7ff0000000000000ll);
  const bool use_preset = preset != nullptr;
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    s_inv[c] = 0ull;
    s_mx[c] = 0ull;
  }
  if (!use_preset) fused_sample_rows<FU_THREADS>(log, t, FU_THREADS, s_inv, s_mx, s_st);   // one row per thread
  // (shared memory set-up overlaps the sample's loads)
  for (int i = t; i < KVM_LOG_NSETS * BINS; i += FU_THREADS) sh[i] = 0;
  ColdTable cold{COLD_SMEM_TABLE ? S.cold_lo : cold_lo_g, COLD_SMEM_TABLE ? S.cold_hi : cold_hi_g, nc};
  kvm_pdl_wait();   // the cold-interval table is written (and the accumulators cleared) by the kernel in front
  if (t == 0) atomicMax(&fws->dbg[0], ~fu_now());
  if (COLD_SMEM_TABLE)
    for (int i = t; i < nc; i += FU_THREADS) {
      S.cold_lo[i] = cold_lo_g[i];
      S.cold_hi[i] = cold_hi_g[i];
    }
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const uint64_t a = kvm_warp_max(s_inv[c]), b = kvm_warp_max(s_mx[c]);
    if (lane == 0) {
      S.s_inv[warp][c] = a;
      S.s_max[warp][c] = b;
    }
  }
#pragma unroll
  for (int m = 16; m; m >>= 1) s_st = fmin(s_st, __shfl_xor_sync(KVM_FULL, s_st, m));
  if (lane == 0) S.s_start[warp] = s_st;
  __syncthreads();
  if (t < NC) {
    uint64_t inv = 0, mx = 0;
    if (use_preset) {
      inv = preset[t];
      mx = preset[NC + t];
    } else {
      for (int w = 0; w < FU_WARPS; ++w) {
        inv = S.s_inv[w][t] > inv ? S.s_inv[w][t] : inv;
        mx = S.s_max[w][t] > mx ? S.s_max[w][t] : mx;
      }
    }
    ColParam cp;
    cp.width = fused_cols_setup(S.fc, t, inv, mx);
    cp.kbase = S.fc.kbase[t];
    cp.shift = S.fc.shift[t];
    if (blockIdx.x == 0) ws->col[t] = cp;
  }
  if (t == 32) {
    double s0 = S.s_start[0];
    for (int w = 1; w < FU_WARPS; ++w) s0 = fmin(s0, S.s_start[w]);
    if (use_preset) s0 = __longlong_as_double((long long)preset[2 * NC]);
    S.series_ref = s0;
    if (SERIES && blockIdx.x == 0) sa.dropped[2] = (unsigned long long)__double_as_longlong(s0);
  }
  __syncthreads();
  if (t == 0) fused_cols_fast_params(S.fc);
  __syncthreads();
  if (t == 0) atomicMax(&fws->dbg[1], fu_now());
  const double s0 = S.series_ref;
  const double cold_first = nc > 0 ? cold.lo[0] : 0.0, cold_last = nc > 0 ? cold.hi[nc - 1] : -1.0;

  FAcc a;
  a.init();
  auto row_fast = [&](double st, double la, double tf, double tl, long long status, long long ct, long long tt,
                      uint32_t& flag_out) -> uint64_t {
    return fused_row(a, S.fc, sh, cold, cold_first, cold_last, st, la, tf, tl, status, ct, tt, flag_out);
  };
  auto series_row = [&](long long bin, bool ok, long long tok) {
    if (bin >= 0) {
      atomicAdd(&sa.series[bin], 1ull);
      if (ok) {
        atomicAdd(&sa.series[sa.max_bins + bin], 1ull);
        atomicAdd(&sa.series[2 * sa.max_bins + bin], (unsigned long long)tok);
      }
    } else if (bin == -2) {
      atomicAdd(sa.dropped, 1ull);
    }
  };
  auto sbin = [&](double st) -> long long {
    const double b = floor((st - s0) / sa.window_ms);
    if (b >= 0.0 && b < (double)sa.max_bins) return (long long)b;
    return b == b ? -2 : -1;
  };

  // ---- the stream: two consecutive rows per thread and iteration, one 128-bit load per column
  // (a shard cut at an odd row has columns that are only 8-byte aligned: one row per thread then)
  if (!vec) {
    for (int64_t i = (int64_t)blockIdx.x * FU_THREADS + t; i < n; i += (int64_t)gridDim.x * FU_THREADS) {
      uint32_t f;
      const double st = kvm_ld1(log.start_ms + i);
      const long long ct = kvm_ld1(log.completion_tokens + i);
      recs[i] = row_fast(st, kvm_ld1(log.latency_ms + i), kvm_ld1(log.ttfb_ms + i), kvm_ld1(log.tllt_ms + i),
                         kvm_ld1(log.status + i), ct, kvm_ld1(log.total_tokens + i), f);
      flags[i] = (uint8_t)f;
      if (SERIES) series_row(sbin(st), f & KVM_F_SUCCESS, ct);
    }
  }
  const int64_t npairs = vec ? (n >> 1) : 0;
  for (int64_t base = (int64_t)blockIdx.x * FU_THREADS + (t & ~31); base < npairs;
       base += (int64_t)gridDim.x * FU_THREADS) {
    const int64_t u = base + lane;
    const bool live = u < npairs;
    const int64_t i = u << 1;
    uint32_t f0 = 0, f1 = 0;
    double sx = 0.0, sy = 0.0;
    long long t0 = 0, t1 = 0;
    if (live) {
      const double2 qs = kvm_ld2(log.start_ms + i), ql = kvm_ld2(log.latency_ms + i);
      const double2 qf = kvm_ld2(log.ttfb_ms + i), qt = kvm_ld2(log.tllt_ms + i);
      const longlong2 qx = kvm_ld2(log.status + i), qc = kvm_ld2(log.completion_tokens + i);
      const longlong2 qo = kvm_ld2(log.total_tokens + i);
      const uint64_t r0 = row_fast(qs.x, ql.x, qf.x, qt.x, qx.x, qc.x, qo.x, f0);
      const uint64_t r1 = row_fast(qs.y, ql.y, qf.y, qt.y, qx.y, qc.y, qo.y, f1);
      *reinterpret_cast<ulonglong2*>(recs + i) = make_ulonglong2(r0, r1);
      *reinterpret_cast<uint16_t*>(flags + i) = (uint16_t)(f0 | (f1 << 8));
      if (SERIES) {
        sx = qs.x;
        sy = qs.y;
        t0 = qc.x;
        t1 = qc.y;
      }
    }
    if (SERIES) {
      const bool k0 = f0 & KVM_F_SUCCESS, k1 = f1 & KVM_F_SUCCESS;
      // One division per warp iteration instead of two per lane: lane 0 bins the first row of
      // the warp's 64, lane 31 the last one (every other lane its first row).  x -> floor((x - s0) / w)
      // is monotone, so when both ends share a window and every start of the warp lies between
      // them, all 64 rows are in that window (the usual case: logs are time-ordered and a window
      // holds thousands of rows).
      const double first = __shfl_sync(KVM_FULL, sx, 0), last = __shfl_sync(KVM_FULL, sy, 31);
      const long long bx = sbin(lane == 31 ? sy : sx);
      const long long lead = __shfl_sync(KVM_FULL, bx, 0), tail = __shfl_sync(KVM_FULL, bx, 31);
      const bool inside = sx >= first && sx <= last && sy >= first && sy <= last;
      if (__all_sync(KVM_FULL, live && inside) && lead == tail && lead >= 0) {
        const int nok = __reduce_add_sync(KVM_FULL, (int)k0 + (int)k1);
        const int64_t tsum = kvm_warp_sum((int64_t)((k0 ? t0 : 0) + (k1 ? t1 : 0)));
        if (lane == 0) {
          atomicAdd(&sa.series[lead], 64ull);
          if (nok) {
            atomicAdd(&sa.series[sa.max_bins + lead], (unsigned long long)nok);
            atomicAdd(&sa.series[2 * sa.max_bins + lead], (unsigned long long)tsum);
          }
        }
      } else if (live) {   // a window boundary inside the warp's rows, or rows out of order
        series_row(lane == 31 ? sbin(sx) : bx, k0, t0);
        series_row(lane == 31 ? bx : sbin(sy), k1, t1);
      }
    }
  }
  if (vec && blockIdx.x == 0 && t < (int)(n & 1)) {   // tail row
    const int64_t i = (n & ~1ll) + t;
    uint32_t f;
    const uint64_t r = row_fast(log.start_ms[i], log.latency_ms[i], log.ttfb_ms[i], log.tllt_ms[i], log.status[i],
                           log.completion_tokens[i], log.total_tokens[i], f);
    recs[i] = r;
    flags[i] = (uint8_t)f;
    if (SERIES) series_row(sbin(log.start_ms[i]), f & KVM_F_SUCCESS, log.completion_tokens[i]);
  }

  // ---- per-thread accumulators -> warp -> CTA partial (slot layout of kvm_log_device.cuh)
  facc_warp_reduce(a, S.red[warp], lane);
  __syncthreads();
  if (t == 0) atomicMax(&fws->dbg[2], fu_now());
  // flush the histograms (64-bit global atomics on the non-zero bins) and count the sets on the way
  {
    uint32_t setsum[KVM_LOG_NSETS];
#pragma unroll
    for (int s = 0; s < KVM_LOG_NSETS; ++s) setsum[s] = 0;
#pragma unroll
    for (int s = 0; s < KVM_LOG_NSETS; ++s)
      for (int b = t; b < BINS; b += FU_THREADS) {
        const uint32_t c = sh[s * BINS + b];
        if (c) {
          atomicAdd(reinterpret_cast<unsigned long long*>(&ws->hist[s][b]), (unsigned long long)c);
          setsum[s] += c;
        }
      }
#pragma unroll
    for (int s = 0; s < KVM_LOG_NSETS; ++s) {
      const uint32_t v = __reduce_add_sync(KVM_FULL, setsum[s]);
      if (lane == 0) S.red[warp][S_CNT0 + s] = v;
    }
  }
  __syncthreads();
  auto dd_bits = [](uint64_t hi, uint64_t lo) { return dd_from_bits(hi, lo); };
  auto combine = [](int slot, uint64_t acc, uint64_t v) -> uint64_t { return slot_combine(slot, acc, v); };
  uint64_t* mine = fws->partial + (size_t)blockIdx.x * NSLOTS;
  if (t < NMAX_END) {
    const int s = t;
    uint64_t acc;
    if (s >= M_KMIN0 && s < M_KMIN0 + NC) acc = S.fc.kmin[s - M_KMIN0];
    else if (s >= X_KMAX0 && s < X_KMAX0 + NC) acc = S.fc.kmax[s - X_KMAX0];
    else {
      acc = S.red[0][s];
      for (int w = 1; w < FU_WARPS; ++w) acc = combine(s, acc, S.red[w][s]);
    }
    if (s < NSUM) {
      if (acc) atomicAdd(&fws->fin[s], (unsigned long long)acc);
    } else {
      atomicMax(&fws->fin[s], (unsigned long long)(s < NMIN_END ? ~acc : acc));
    }
  } else if (t == NMAX_END || t == NMAX_END + 1) {
    const int s = t == NMAX_END ? D_TTFB : D_LAT;
    kvm_dd acc = dd_bits(S.red[0][s], S.red[0][s + 1]);
    for (int w = 1; w < FU_WARPS; ++w) acc = kvm_dd_add(acc, dd_bits(S.red[w][s], S.red[w][s + 1]));
    mine[s] = (uint64_t)__double_as_longlong(acc.hi);
    mine[s + 1] = (uint64_t)__double_as_longlong(acc.lo);
  }
  __threadfence();
  __syncthreads();
  if (t == 0) atomicMax(&fws->dbg[3], fu_now());
  if (t == 0) S.is_last = atomicAdd(&fws->ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!S.is_last) return;
  __threadfence();
  // ---- last CTA: the integer / range totals are complete; the two double-double sums are added
  // in a fixed shape over the CTA partials (the same bits whatever order the CTAs finished in)
  const uint64_t* part = fws->partial;
  if (t < NMAX_END) {
    const unsigned long long v = __ldcg(&fws->fin[t]);
    S.fin[t] = (t >= NSUM && t < NMIN_END) ? ~v : v;
  }
  if (warp < 2) {
    const int s = warp == 0 ? D_TTFB : D_LAT;
    kvm_dd acc{0.0, 0.0};
    for (unsigned g = lane; g < gridDim.x; g += 32) {
      const size_t o = (size_t)g * NSLOTS + s;
      acc = kvm_dd_add(acc, dd_bits(__ldcg(part + o), __ldcg(part + o + 1)));
    }
    acc = kvm_warp_sum(acc);
    if (lane == 0) {
      S.fin[s] = (uint64_t)__double_as_longlong(acc.hi);
      S.fin[s + 1] = (uint64_t)__double_as_longlong(acc.lo);
    }
  }
  __syncthreads();
  if (t == 0) {
    kvm_log_stats r;
    stats_from_slots(S.fin, n, r);
    *out = r;
    fws->ticket = 0;
    fws->dbg[4] = fu_now();
  }
}

// The launch in front of the pass: one thread per cold start turns it into the closed start_ms
// interval that the reference classifies cold, and the whole grid clears what the pass accumulates
// into - the header of the scan workspace, the select's histograms, the window series - in place
// of four memset nodes.  The pass is a programmatic dependent launch behind it: its sample and its
// shared-memory set-up run while this kernel drains.
__global__ void __launch_bounds__(256)
fused_prep_kernel(const double* __restrict__ cold_times, int nc, double window, double* __restrict__ lo,
                  double* __restrict__ hi, uint4* __restrict__ header, int header_words16, uint4* __restrict__ hist,
                  int hist_words16, unsigned long long* __restrict__ series, int64_t series_words,
                  unsigned long long* __restrict__ dropped) {
  kvm_pdl_launch_dependents();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (int64_t)gridDim.x * blockDim.x;
  const uint4 z = make_uint4(0u, 0u, 0u, 0u);
  if (i < nc) cold_interval(cold_times[i], window, lo + i, hi + i);
  if (i < header_words16) header[i] = z;
  for (int64_t k = i; k < hist_words16; k += stride) hist[k] = z;
  for (int64_t k = i; k < series_words; k += stride) series[k] = 0ull;
  if (dropped && i < 3) dropped[i] = 0ull;
}

constexpr int FU_MAX_COLD = 16384;
size_t fused_partials_bytes() { return kvm_align_up(offsetof(FusedWs, partial) + sizeof(uint64_t) * (size_t)NSLOTS * FU_MAX_GRID, 256); }
bool aligned_to(const void* p, uintptr_t a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) == 0; }

}  // namespace

// debug aid: microseconds since the first CTA entered the last one-pass kernel that used this scan
// workspace - [0] last CTA set up, [1] last CTA streamed, [2] last CTA flushed, [3] statistics out
extern "C" int kvm_log_fused_phase_us(const void* scan_workspace, double* out4) {
  KVM_REQUIRE(scan_workspace && out4, "kvm_log_fused_phase_us: null argument");
  unsigned long long h[8];
  KVM_CUDA_CHECK(cudaMemcpy(h, reinterpret_cast<const unsigned char*>(scan_workspace) + offsetof(FusedWs, dbg), sizeof(h),
                            cudaMemcpyDeviceToHost));
  const unsigned long long t0 = ~h[0];
  for (int i = 0; i < 4; ++i) out4[i] = (double)(long long)(h[i + 1] - t0) / 1000.0;
  return KVM_OK;
}

extern "C" size_t kvm_log_analyze_fused_scan_workspace_bytes(void) {
  return fused_partials_bytes() + 2 * sizeof(double) * (size_t)FU_MAX_COLD;
}

static int fused_pass_impl(const kvm_request_log* log, const double* cold_times, int32_t n_cold,
                           double cold_window_s, uint8_t* flags, kvm_log_stats* stats,
                           const kvm_log_extras* extras, void* scan_workspace, size_t scan_workspace_bytes,
                           void* select_workspace, size_t select_workspace_bytes, int64_t arena_capacity,
                           const unsigned long long* preset, void* stream);

extern "C" int kvm_log_fused_pass(const kvm_request_log* log, const double* cold_times, int32_t n_cold,
                                  double cold_window_s, uint8_t* flags, kvm_log_stats* stats,
                                  const kvm_log_extras* extras, void* scan_workspace, size_t scan_workspace_bytes,
                                  void* select_workspace, size_t select_workspace_bytes, int64_t arena_capacity,
                                  void* stream) {
  return fused_pass_impl(log, cold_times, n_cold, cold_window_s, flags, stats, extras, scan_workspace,
                         scan_workspace_bytes, select_workspace, select_workspace_bytes, arena_capacity, nullptr, stream);
}

extern "C" int32_t kvm_log_fused_preset_words(void) { return FU_PRESET_WORDS; }

extern "C" int kvm_log_fused_sample(const kvm_request_log* log, uint64_t* record, void* stream) {
  KVM_REQUIRE(log && record && log->n >= 0, "kvm_log_fused_sample: bad argument");
  fused_sample_kernel<<<1, 512, 0, (cudaStream_t)stream>>>(*log, reinterpret_cast<unsigned long long*>(record));
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_log_fused_preset_combine(const uint64_t* records, int32_t world, int64_t stride_words,
                                            uint64_t* preset, void* stream) {
  KVM_REQUIRE(records && preset && world >= 1 && stride_words >= FU_PRESET_WORDS, "kvm_log_fused_preset_combine: bad argument");
  fused_preset_combine_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const unsigned long long*>(records), world, stride_words, reinterpret_cast<unsigned long long*>(preset));
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_log_fused_pass_preset(const kvm_request_log* log, const double* cold_times, int32_t n_cold,
                                         double cold_window_s, uint8_t* flags, kvm_log_stats* stats,
                                         const kvm_log_extras* extras, void* scan_workspace, size_t scan_workspace_bytes,
                                         void* select_workspace, size_t select_workspace_bytes, int64_t arena_capacity,
                                         const uint64_t* preset, void* stream) {
  KVM_REQUIRE(preset, "kvm_log_fused_pass_preset: preset is null");
  return fused_pass_impl(log, cold_times, n_cold, cold_window_s, flags, stats, extras, scan_workspace,
                         scan_workspace_bytes, select_workspace, select_workspace_bytes, arena_capacity,
                         reinterpret_cast<const unsigned long long*>(preset), stream);
}

static int fused_pass_impl(const kvm_request_log* log, const double* cold_times, int32_t n_cold,
                           double cold_window_s, uint8_t* flags, kvm_log_stats* stats,
                           const kvm_log_extras* extras, void* scan_workspace, size_t scan_workspace_bytes,
                           void* select_workspace, size_t select_workspace_bytes, int64_t arena_capacity,
                           const unsigned long long* preset, void* stream) {
  KVM_REQUIRE(log && flags && stats && extras && scan_workspace && select_workspace,
              "kvm_log_analyze_fused: null argument");
  KVM_REQUIRE(log->n >= 0 && n_cold >= 0 && (n_cold == 0 || cold_times), "kvm_log_analyze_fused: bad sizes");
  KVM_REQUIRE(n_cold <= FU_MAX_COLD, "kvm_log_analyze_fused: %d cold starts exceed the supported %d", n_cold, FU_MAX_COLD);
  KVM_REQUIRE(extras->digit_records, "kvm_log_analyze_fused: digit_records scratch is required");
  const bool series = extras->series_window_ms > 0.0;
  if (series)
    KVM_REQUIRE(extras->series && extras->series_dropped && extras->series_max_bins >= 1,
                "kvm_log_analyze_fused: series requested without buffers");
  // the one-pass kernel streams with 128-bit loads: every column 16-byte aligned (whole logs are;
  // a shard cut at an odd row is not - the caller then uses the two-pass entry points)
  const int vec = aligned_to(log->start_ms, 16) && aligned_to(log->latency_ms, 16) && aligned_to(log->ttfb_ms, 16) &&
                  aligned_to(log->tllt_ms, 16) && aligned_to(log->status, 16) && aligned_to(log->completion_tokens, 16) &&
                  aligned_to(log->total_tokens, 16) && aligned_to(flags, 2) && aligned_to(extras->digit_records, 16);
  if ((!vec && !preset) || (log->n < 1 && !preset)) {
    kvm_set_error("kvm_log_analyze_fused: needs a non-empty log with 16-byte aligned columns; use kvm_log_scan + kvm_log_select");
    return KVM_ERR_UNSUPPORTED;
  }
  if (scan_workspace_bytes < kvm_log_analyze_fused_scan_workspace_bytes()) {
    kvm_set_error("kvm_log_analyze_fused: scan workspace %zu < %zu bytes", scan_workspace_bytes,
                  kvm_log_analyze_fused_scan_workspace_bytes());
    return KVM_ERR_WORKSPACE;
  }
  if (select_workspace_bytes < kvm_select_workspace_bytes(arena_capacity)) {
    kvm_set_error("kvm_log_analyze_fused: select workspace %zu < %zu bytes", select_workspace_bytes,
                  kvm_select_workspace_bytes(arena_capacity));
    return KVM_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  FusedWs* fws = reinterpret_cast<FusedWs*>(scan_workspace);
  SelWs* ws = reinterpret_cast<SelWs*>(select_workspace);
  double* cold_lo = reinterpret_cast<double*>(reinterpret_cast<unsigned char*>(scan_workspace) + fused_partials_bytes());
  double* cold_hi = cold_lo + FU_MAX_COLD;
  FusedSeries sa{};
  int64_t series_words = 0;
  if (series) {
    sa.window_ms = extras->series_window_ms;
    sa.max_bins = extras->series_max_bins;
    sa.series = reinterpret_cast<unsigned long long*>(extras->series);
    sa.dropped = reinterpret_cast<unsigned long long*>(extras->series_dropped);
    series_words = 3 * (int64_t)extras->series_max_bins;
  }
  {
    static_assert(offsetof(FusedWs, partial) % 16 == 0 && sizeof(ws->hist) % 16 == 0, "cleared with 16-byte stores");
    const int64_t items = std::max<int64_t>(std::max<int64_t>(n_cold, (int64_t)(sizeof(ws->hist) / 16)), series_words / 4);
    const int grid = (int)std::min<int64_t>((items + 255) / 256, 2 * (int64_t)kvm_num_sms());
    fused_prep_kernel<<<grid < 1 ? 1 : grid, 256, 0, st>>>(
        cold_times, (int)n_cold, cold_window_s, cold_lo, cold_hi, reinterpret_cast<uint4*>(fws),
        (int)(offsetof(FusedWs, partial) / 16), reinterpret_cast<uint4*>(&ws->hist[0][0]), (int)(sizeof(ws->hist) / 16),
        sa.series, series_words, sa.dropped);
  }
  const size_t smem = sizeof(uint32_t) * KVM_LOG_NSETS * BINS + sizeof(FusedSmem);
  int grid = kvm_stream_grid(log->n, FU_THREADS * 2 * 4);
  if (grid > FU_MAX_GRID) grid = FU_MAX_GRID;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(FU_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;   // behind the prep kernel: the sample and the set-up overlap it
  uint64_t* recs = reinterpret_cast<uint64_t*>(extras->digit_records);
  const double* lo_c = cold_lo;
  const double* hi_c = cold_hi;
  const bool smem_table = n_cold <= FU_COLD_SMEM;
#define KVM_LAUNCH_FUSED(T, S)                                                                                 \
  do {                                                                                                         \
    KVM_CUDA_CHECK(cudaFuncSetAttribute((const void*)log_fused_kernel<T, S>,                                   \
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));              \
    KVM_CUDA_CHECK(cudaLaunchKernelEx(&cfg, log_fused_kernel<T, S>, *log, lo_c, hi_c, (int)n_cold, flags, recs, \
                                      stats, fws, ws, sa, preset, vec));                                                    \
  } while (0)
  if (smem_table && series) KVM_LAUNCH_FUSED(true, true);
  else if (smem_table) KVM_LAUNCH_FUSED(true, false);
  else if (series) KVM_LAUNCH_FUSED(false, true);
  else KVM_LAUNCH_FUSED(false, false);
#undef KVM_LAUNCH_FUSED
  return KVM_OK;
}

static kvm_hist_spec fused_hist_spec(int32_t num_bins) {
  kvm_hist_spec hs{};
  hs.nhist = 4;
  hs.num_bins = num_bins;
  hs.col[0] = KVM_COL_TTFB; hs.col[1] = KVM_COL_TLLT; hs.col[2] = KVM_COL_GEN; hs.col[3] = KVM_COL_PERTOK;
  hs.set_mask[0] = (1u << KVM_SET_TTFB_COLD) | (1u << KVM_SET_TTFB_WARM);
  hs.set_mask[1] = 1u << KVM_SET_TLLT;
  hs.set_mask[2] = 1u << KVM_SET_GEN;
  hs.set_mask[3] = 1u << KVM_SET_PERTOK;
  return hs;
}

extern "C" int kvm_select_plan_preset(const kvm_set_info* info, const kvm_query* queries, int32_t nq,
                                      const kvm_hist_spec* hists, void* workspace, size_t workspace_bytes,
                                      int64_t arena_capacity, int32_t refine_rounds, int64_t refine_min_keys,
                                      void* stream) {
  return kvm_select_plan_internal(info, queries, nq, hists, workspace, workspace_bytes, arena_capacity,
                                  refine_rounds, refine_min_keys, 1, stream);
}

extern "C" int kvm_log_analyze_fused(const kvm_request_log* log, const double* cold_times, int32_t n_cold,
                                     double cold_window_s, uint8_t* flags, kvm_log_stats* stats,
                                     const kvm_query* queries, int32_t nq, kvm_query_result* results,
                                     const kvm_log_extras* extras, void* scan_workspace,
                                     size_t scan_workspace_bytes, void* select_workspace,
                                     size_t select_workspace_bytes, int64_t arena_capacity,
                                     int32_t refine_rounds, int64_t refine_min_keys, void* stream) {
  KVM_REQUIRE(queries && results, "kvm_log_analyze_fused: null argument");
  int rc = kvm_log_fused_pass(log, cold_times, n_cold, cold_window_s, flags, stats, extras, scan_workspace,
                              scan_workspace_bytes, select_workspace, select_workspace_bytes, arena_capacity, stream);
  if (rc) return rc;
  const uint64_t* recs = reinterpret_cast<const uint64_t*>(extras->digit_records);
  const bool hists = extras->num_bins > 0;
  KVM_REQUIRE(!hists || extras->hist_counts, "kvm_log_analyze_fused: hist_counts is null");
  const kvm_hist_spec hs = fused_hist_spec(extras->num_bins);
  rc = kvm_select_plan_internal(&stats->sets, queries, nq, hists ? &hs : nullptr, select_workspace,
                                select_workspace_bytes, arena_capacity, refine_rounds, refine_min_keys, 1, stream);
  if (rc) return rc;
  if ((rc = kvm_select_resolve(select_workspace, stream))) return rc;
  for (int r = 0; r < refine_rounds; ++r) {
    if ((rc = kvm_log_select_refine_hist(log, flags, recs, select_workspace, stream))) return rc;
    if ((rc = kvm_select_refine(select_workspace, stream))) return rc;
  }
  if ((rc = kvm_log_select_compact(log, flags, recs, select_workspace, stream))) return rc;
  return kvm_select_finish(select_workspace, nullptr, 0, nullptr, results, hists ? extras->hist_counts : nullptr, stream);
}
