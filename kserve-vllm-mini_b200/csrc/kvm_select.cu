// kvm_select.cu - K2 (+K5, K6, K8 folded in): exact multi-rank radix select over fp64 keys,
// with the 20-bin value histograms and the windowed series produced by the same two passes.
//
// Replaces analyze.py:59-81 `percentile()` (25 calls / run over 9 multisets, each a fresh NaN
// filter + list.sort), analyze.py:84-122 `compute_histograms` (4 calls / run) and the
// upper-median of energy/collector.py:333.  The kernels return exact ORDER STATISTICS and
// exact COUNTS; the host evaluates the reference's interpolation / edge expressions in Python
// floats, so the results are bit-identical.
//
// Algorithm (range-adaptive radix select):
//   plan     per value column: kbase = min key, width = bits(max-min), shift = max(0,width-11):
//            digit(key) = (key-kbase) >> shift  (< 2048).  Per query: n and rank.  Per value
//            histogram: the nb-1 bin edges as keys (bisection with the reference expression) and
//            the fine bins that contain an edge.
//   hist     ONE pass over the data: per base set a 2048-bin histogram of digit(key), privatised
//            in shared memory, flushed with 64-bit global atomics.  The log variant also bins
//            start_ms into the 1-window series in the same pass (K8).
//   resolve  per query: bin holding the rank, residual rank, candidate bucket (column, set
//            mask, bin); buckets de-duplicated, sorted by (column, bin), laid out in the arena
//   compact  ONE pass over the data: keys whose digit hits a bucket are appended to it; keys
//            whose fine bin contains a histogram edge are binned exactly (edge counters)
//   finish   one CTA per query: radix select inside its bucket; one more CTA assembles the
//            value histograms = edge counters + fine bins without an edge added wholesale
// With adaptive digits a bucket holds ~n/2048 of the populated range, so everything after
// `compact` is tiny; the two streaming passes are HBM-bound.
//
// Multi-GPU: `hist` and the edge counters are all-reduced and the arenas all-gathered by the
// host between the stages; every rank resolves / finishes identically.

#include <mutex>

#include "kvm_common.cuh"
#include "kvm_log_device.cuh"
#include "kvm_peer.cuh"
#include "kvm_select_ws.cuh"

namespace {

// ------------------------------------------------------------------ plan
constexpr int PLAN_THREADS = 512;

__global__ void __launch_bounds__(PLAN_THREADS, 2)
select_plan_kernel(SelWs* ws, const kvm_set_info* info_in, QueryPack qp, kvm_hist_spec hs,
                   int64_t arena_capacity, int refine_rounds, int64_t refine_min, int preset) {
  __shared__ kvm_set_info info;
  __shared__ ColParam cps[KVM_MAX_COLS];
  __shared__ uint64_t thr_s[NH][NB_MAX];
  __shared__ uint8_t edge_s[KVM_MAX_COLS][BINS];
  const int t = threadIdx.x;
  // the info block arrives as 8-byte words loaded in parallel (it may alias ws->info)
  static_assert(sizeof(kvm_set_info) % 8 == 0, "kvm_set_info is copied as 8-byte words");
  if (t < (int)(sizeof(kvm_set_info) / 8))
    reinterpret_cast<uint64_t*>(&info)[t] = reinterpret_cast<const uint64_t*>(info_in)[t];
  // zero what the passes accumulate into (preset: the fused pass has already filled the histograms)
  if (!preset)
    for (int i = t; i < KVM_MAX_SETS * BINS; i += PLAN_THREADS) (&ws->hist[0][0])[i] = 0;
  for (int i = t; i < NH * NB_MAX; i += PLAN_THREADS) {
    (&ws->edge20[0][0])[i] = 0;
    (&thr_s[0][0])[i] = ~0ull;
  }
  for (int i = t; i < NH * (BINS / 32); i += PLAN_THREADS) (&ws->flag20[0][0])[i] = 0;
  for (int i = t; i < KVM_MAX_COLS * BINS / 4; i += PLAN_THREADS)
    reinterpret_cast<uint32_t*>(&edge_s[0][0])[i] = 0;
  __syncthreads();
  if (t == 0) {
    ws->status = KVM_OK;
    ws->nq = qp.nq;
    ws->nbuckets = 0;
    ws->arena_capacity = arena_capacity;
    ws->arena_used = 0;
    ws->refine = 0;
    ws->refine_budget = refine_rounds;
    ws->refine_min = refine_min;
    ws->level = 0;
    ws->nparents = 0;
    ws->preset_cols = preset;
    ws->hspec = hs;
  }
  if (t < (int)(sizeof(kvm_set_info) / 8))
    reinterpret_cast<uint64_t*>(&ws->info)[t] = reinterpret_cast<const uint64_t*>(&info)[t];
  if (t < KVM_MAX_COLS) {
    ColParam cp{0, 0, 0};
    if (preset) {
      cp = ws->col[t];   // written by the fused pass from its sample
    } else if (t < info.ncols && info.kmin[t] <= info.kmax[t]) {
      const uint64_t range = info.kmax[t] - info.kmin[t];
      cp.kbase = info.kmin[t];
      cp.width = range ? 64 - __clzll((long long)range) : 0;
      cp.shift = cp.width > BINS_LOG2 ? cp.width - BINS_LOG2 : 0;
    }
    ws->col[t] = cp;
    cps[t] = cp;
    int h = -1;
    for (int k = 0; k < hs.nhist; ++k)
      if (hs.col[k] == t) h = k;
    ws->col_hist[t] = h;
  }
  if (t >= 32 && t < 32 + NH) {
    const int h = t - 32;
    int act = 0;
    if (h < hs.nhist) {
      const int c = hs.col[h];
      act = info.kmin[c] < info.kmax[c];
    }
    ws->active20[h] = act;
  }
  if (t >= 64 && t < 64 + MAXQ) {
    const int qi = t - 64;
    QState s{0, 0, 0, 0, KVM_Q_EMPTY, -1, 0, 1};
    ws->bucket_counts[qi] = 0;
    if (qi < qp.nq) {
      const kvm_query q = qp.q[qi];
      ws->queries[qi] = q;
      int64_t n = 0;
      int col = -1;
      for (int b = 0; b < info.nsets; ++b)
        if ((q.set_mask >> b) & 1u) {
          n += info.count[b];
          if (col < 0) col = info.set_col[b];
        }
      s.n = n;
      s.col = col < 0 ? 0 : col;
      if (n > 0) {
        s.rank = resolve_rank(q.rank_kind, q.p, n, q.rank);
        s.status = (s.rank < 0 || s.rank >= n) ? KVM_Q_OUT_OF_RANGE : KVM_Q_OK;
        s.resolved = s.status != KVM_Q_OK;
      }
    }
    ws->q[qi] = s;
  }
  __syncthreads();
  // histogram edges as keys + the fine bins that contain one (threads 128..)
  if (t >= 128 && t < 128 + NH * (NB_MAX - 1)) {
    const int e = t - 128, h = e / (NB_MAX - 1), b = 1 + e % (NB_MAX - 1);
    if (h < hs.nhist && b < hs.num_bins) {
      const int c = hs.col[h];
      const uint64_t klo = info.kmin[c], khi = info.kmax[c];
      if (klo < khi) {
        const uint64_t T = bin_threshold(klo, khi, hs.num_bins, b);
        thr_s[h][b - 1] = T;
        if (T != ~0ull) {
          const uint32_t d = digit_of(T, cps[c].kbase, cps[c].shift);
          atomicOr(&ws->flag20[h][d >> 5], 1u << (d & 31));
        }
      }
    }
  }
  __syncthreads();
  // per histogram: which edge (if exactly one) lives in each flagged fine bin
  if (t < hs.nhist) {
    const int h = t, c = hs.col[h];
    if (info.kmin[c] < info.kmax[c]) {
      uint32_t prev = 0xffffffffu;
      for (int b = 1; b < hs.num_bins; ++b) {
        const uint64_t T = thr_s[h][b - 1];
        if (T == ~0ull) break;
        const uint32_t d = digit_of(T, cps[c].kbase, cps[c].shift);
        edge_s[c][d] = d == prev ? 255 : (uint8_t)b;
        prev = d;
      }
      if (preset) {
        // sampled digit parameters: an end bin that holds keys from outside its nominal range
        // cannot be added wholesale - its keys are ranked against all thresholds one by one
        const ColParam cp = cps[c];
        const uint64_t top = cp.kbase + ((uint64_t)(BINS - 1) << cp.shift);
        if (info.kmin[c] < cp.kbase) {
          edge_s[c][0] = 255;
          atomicOr(&ws->flag20[h][0], 1u);
        }
        if (info.kmax[c] >= top && shr64(info.kmax[c] - top, cp.shift) != 0) {
          edge_s[c][BINS - 1] = 255;
          atomicOr(&ws->flag20[h][(BINS - 1) >> 5], 1u << ((BINS - 1) & 31));
        }
      }
    }
  }
  __syncthreads();
  // publish thresholds, edge codes and the default lookup table (no bucket yet: low byte 255)
  for (int i = t; i < NH * NB_MAX; i += PLAN_THREADS) (&ws->thr20[0][0])[i] = (&thr_s[0][0])[i];
  for (int i = t; i < KVM_MAX_COLS * BINS / 4; i += PLAN_THREADS) {
    const uint32_t e4 = reinterpret_cast<const uint32_t*>(&edge_s[0][0])[i];
    reinterpret_cast<uint32_t*>(&ws->edge_code[0][0])[i] = e4;
    uint2 t4;
    t4.x = (0x00ffu | ((e4 & 0xffu) << 8)) | ((0x00ffu | (((e4 >> 8) & 0xffu) << 8)) << 16);
    t4.y = (0x00ffu | (((e4 >> 16) & 0xffu) << 8)) | ((0x00ffu | (((e4 >> 24) & 0xffu) << 8)) << 16);
    reinterpret_cast<uint2*>(&ws->tab[0][0])[i] = t4;
  }
}

// ------------------------------------------------------------------ set/key enumeration
struct ColParams {
  uint64_t kbase[KVM_MAX_COLS];
  int32_t shift[KVM_MAX_COLS];
};

// Streams the log two rows per thread (128-bit loads) and hands every (set,col,key) to f.
template <class F>
__device__ __forceinline__ void for_each_log_key(const kvm_request_log& log,
                                                 const uint8_t* __restrict__ flags, bool vec,
                                                 F&& f) {
  const int64_t n = log.n;
  if (vec) {
    const int64_t npairs = n >> 1;
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < npairs;
         p += (int64_t)gridDim.x * blockDim.x) {
      const int64_t i = p << 1;
      const uint16_t ff = kvm_ld_u8x2(flags + i);
      if (!((ff | (ff >> 8)) & KVM_F_SUCCESS)) continue;
      const double2 la = kvm_ld2(log.latency_ms + i);
      const double2 tf = kvm_ld2(log.ttfb_ms + i);
      const double2 tl = kvm_ld2(log.tllt_ms + i);
      const longlong2 ct = kvm_ld2(log.completion_tokens + i);
      log_row_keys((uint8_t)(ff & 0xff), la.x, tf.x, tl.x, ct.x, f);
      log_row_keys((uint8_t)(ff >> 8), la.y, tf.y, tl.y, ct.y, f);
    }
    if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
      const int64_t i = n - 1;
      log_row_keys(flags[i], log.latency_ms[i], log.ttfb_ms[i], log.tllt_ms[i],
                   log.completion_tokens[i], f);
    }
  } else {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
      const uint8_t fl = kvm_ld_u8(flags + i);
      if (!(fl & KVM_F_SUCCESS)) continue;
      log_row_keys(fl, kvm_ld1(log.latency_ms + i), kvm_ld1(log.ttfb_ms + i),
                   kvm_ld1(log.tllt_ms + i), kvm_ld1(log.completion_tokens + i), f);
    }
  }
}

__device__ __forceinline__ int column_member(int mode, const void* member, int64_t i) {
  if (mode == KVM_MEMBER_ALL) return 0;
  if (mode == KVM_MEMBER_SET_IDS) return (int)kvm_ld_u8(reinterpret_cast<const uint8_t*>(member) + i);
  return kvm_ld1(reinterpret_cast<const int64_t*>(member) + i) == 200 ? 0 : 255;
}

template <class F>
__device__ __forceinline__ void for_each_column_key(const double* __restrict__ values, int64_t n,
                                                    int mode, const void* __restrict__ member,
                                                    bool vec, F&& f) {
  if (vec && mode != KVM_MEMBER_SET_IDS) {
    const int64_t npairs = n >> 1;
    const int64_t* st = reinterpret_cast<const int64_t*>(member);
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < npairs;
         p += (int64_t)gridDim.x * blockDim.x) {
      const int64_t i = p << 1;
      const double2 v = kvm_ld2(values + i);
      bool m0 = true, m1 = true;
      if (mode == KVM_MEMBER_STATUS) {
        const longlong2 s = kvm_ld2(st + i);
        m0 = s.x == 200;
        m1 = s.y == 200;
      }
      if (m0 && !kvm_isnan(v.x)) f(0, 0, kvm_key(v.x));
      if (m1 && !kvm_isnan(v.y)) f(0, 0, kvm_key(v.y));
    }
    if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
      const int64_t i = n - 1;
      const double v = values[i];
      if (column_member(mode, member, i) == 0 && !kvm_isnan(v)) f(0, 0, kvm_key(v));
    }
  } else {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
      const double v = kvm_ld1(values + i);
      const int s = column_member(mode, member, i);
      if (s < KVM_MAX_SETS && !kvm_isnan(v)) f(s, 0, kvm_key(v));
    }
  }
}

// ------------------------------------------------------------------ hist
__device__ __forceinline__ void hist_prologue(SelWs* ws, uint32_t* sh, int nsets, ColParams& cp) {
  for (int i = threadIdx.x; i < nsets * BINS; i += blockDim.x) sh[i] = 0;
  if (threadIdx.x < KVM_MAX_COLS) {
    cp.kbase[threadIdx.x] = ws->col[threadIdx.x].kbase;
    cp.shift[threadIdx.x] = ws->col[threadIdx.x].shift;
  }
  __syncthreads();
}
__device__ __forceinline__ void hist_epilogue(SelWs* ws, const uint32_t* sh, int nsets) {
  __syncthreads();
  for (int i = threadIdx.x; i < nsets * BINS; i += blockDim.x) {
    const uint32_t c = sh[i];
    if (c) atomicAdd(reinterpret_cast<unsigned long long*>(&ws->hist[0][0]) + i, (unsigned long long)c);
  }
}

// series bin of one row: >= 0 bin, -1 nothing to add (NaN start), -2 outside [0, max_bins)
__device__ __forceinline__ long long series_bin(double st, double s0, double window_ms, int64_t max_bins) {
  const double b = floor((st - s0) / window_ms);
  if (b >= 0.0 && b < (double)max_bins) return (long long)b;
  return b == b ? -2 : -1;
}
__device__ __forceinline__ void series_add_row(const SeriesArgs& sa, long long bin, bool ok, long long tok) {
  if (bin >= 0) {
    atomicAdd(&sa.series[bin], 1ull);
    if (ok) {
      atomicAdd(&sa.series[sa.max_bins + bin], 1ull);
      atomicAdd(&sa.series[2 * sa.max_bins + bin], (unsigned long long)tok);
    }
  } else if (bin == -2) {
    atomicAdd(sa.dropped, 1ull);
  }
}

// Log histogram pass; ROWS = 2 (128-bit loads) or 1 (unaligned shard).  With SERIES the same
// pass bins start_ms into the window series: rows are time-ordered, so the 32/64 rows of a
// warp iteration almost always share one window and are merged into three atomics.
template <int ROWS, bool SERIES>
__global__ void __launch_bounds__(SEL_THREADS, 2)
log_hist_kernel(kvm_request_log log, const uint8_t* __restrict__ flags, SelWs* ws, SeriesArgs sa,
                uint64_t* __restrict__ digit_records) {
  extern __shared__ uint32_t sh[];  // [KVM_LOG_NSETS][BINS]
  __shared__ ColParams cp;
  hist_prologue(ws, sh, KVM_LOG_NSETS, cp);
  // column parameters in registers: `col` is a literal at every call site of log_row_keys
  uint64_t kb[KVM_LOG_NCOLS];
  int shf[KVM_LOG_NCOLS];
#pragma unroll
  for (int c = 0; c < KVM_LOG_NCOLS; ++c) {
    kb[c] = cp.kbase[c];
    shf[c] = cp.shift[c];
  }
  // digit record of the current row: 12 bits per value column, 0xFFF = row not in any set of
  // that column.  The compact pass reads these 8 bytes instead of recomputing five keys.
  uint64_t rec = DIGITS_NONE;
  auto add = [&](int set, int col, uint64_t key) {
    const uint32_t d = digit_of(key, kb[col], shf[col]);
    atomicAdd(&sh[set * BINS + d], 1u);
    rec ^= (uint64_t)(0xFFFu ^ d) << (12 * col);
  };
  const int lane = threadIdx.x & 31;
  const int64_t n = log.n;
  const int64_t nunits = ROWS == 2 ? (n >> 1) : n;
  double s0 = 0.0;
  if (SERIES) s0 = sa.stats->min_start_ms;
  for (int64_t base = (int64_t)blockIdx.x * SEL_THREADS + (threadIdx.x & ~31); base < nunits;
       base += (int64_t)gridDim.x * SEL_THREADS) {
    const int64_t u = base + lane;
    const bool live = u < nunits;
    const int64_t i = ROWS == 2 ? (u << 1) : u;
    uint8_t f0 = 0, f1 = 0;
    long long tok0 = 0, tok1 = 0;
    if (live) {
      if (ROWS == 2) {
        const uint16_t ff = kvm_ld_u8x2(flags + i);
        f0 = (uint8_t)(ff & 0xff);
        f1 = (uint8_t)(ff >> 8);
      } else {
        f0 = kvm_ld_u8(flags + i);
      }
      if ((f0 | f1) & KVM_F_SUCCESS) {
        if (ROWS == 2) {
          const double2 la = kvm_ld2(log.latency_ms + i), tf = kvm_ld2(log.ttfb_ms + i);
          const double2 tl = kvm_ld2(log.tllt_ms + i);
          const longlong2 ct = kvm_ld2(log.completion_tokens + i);
          rec = rec_blank(f0);
          log_row_keys(f0, la.x, tf.x, tl.x, ct.x, add);
          const uint64_t rec0 = rec;
          rec = rec_blank(f1);
          log_row_keys(f1, la.y, tf.y, tl.y, ct.y, add);
          if (digit_records)
            *reinterpret_cast<ulonglong2*>(digit_records + i) = make_ulonglong2(rec0, rec);
          tok0 = ct.x;
          tok1 = ct.y;
        } else {
          const long long ct = kvm_ld1(log.completion_tokens + i);
          rec = rec_blank(f0);
          log_row_keys(f0, kvm_ld1(log.latency_ms + i), kvm_ld1(log.ttfb_ms + i), kvm_ld1(log.tllt_ms + i), ct, add);
          if (digit_records) digit_records[i] = rec;
          tok0 = ct;
        }
      } else if (digit_records) {
        if (ROWS == 2) *reinterpret_cast<ulonglong2*>(digit_records + i) = make_ulonglong2(DIGITS_NONE, DIGITS_NONE);
        else digit_records[i] = DIGITS_NONE;
      }
    }
    if (SERIES) {
      long long b0 = -1, b1 = -1;
      if (live) {
        if (ROWS == 2) {
          const double2 st = kvm_ld2(sa.start_ms + i);
          b0 = series_bin(st.x, s0, sa.window_ms, sa.max_bins);
          b1 = series_bin(st.y, s0, sa.window_ms, sa.max_bins);
        } else {
          b0 = series_bin(kvm_ld1(sa.start_ms + i), s0, sa.window_ms, sa.max_bins);
          b1 = b0;
        }
      }
      const bool ok0 = f0 & KVM_F_SUCCESS, ok1 = (ROWS == 2) && (f1 & KVM_F_SUCCESS);
      const long long lead = __shfl_sync(KVM_FULL, b0, 0);
      if (__all_sync(KVM_FULL, live && b0 == lead && b1 == lead && lead >= 0)) {
        const int nok = __popc(__ballot_sync(KVM_FULL, ok0)) + __popc(__ballot_sync(KVM_FULL, ok1));
        const int64_t tsum = kvm_warp_sum((int64_t)((ok0 ? tok0 : 0) + (ok1 ? tok1 : 0)));
        if (lane == 0) {
          atomicAdd(&sa.series[lead], (unsigned long long)(32 * ROWS));
          if (nok) {
            atomicAdd(&sa.series[sa.max_bins + lead], (unsigned long long)nok);
            atomicAdd(&sa.series[2 * sa.max_bins + lead], (unsigned long long)tsum);
          }
        }
      } else if (live) {
        series_add_row(sa, b0, ok0, tok0);
        if (ROWS == 2) series_add_row(sa, b1, ok1, tok1);
      }
    }
  }
  if (ROWS == 2 && (n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {  // odd tail row
    const int64_t i = n - 1;
    const uint8_t fl = flags[i];
    rec = rec_blank(fl);
    log_row_keys(fl, log.latency_ms[i], log.ttfb_ms[i], log.tllt_ms[i], log.completion_tokens[i], add);
    if (digit_records) digit_records[i] = rec;
    if (SERIES)
      series_add_row(sa, series_bin(sa.start_ms[i], s0, sa.window_ms, sa.max_bins), fl & KVM_F_SUCCESS,
                     log.completion_tokens[i]);
  }
  hist_epilogue(ws, sh, KVM_LOG_NSETS);
}

__global__ void __launch_bounds__(SEL_THREADS, 2)
column_hist_kernel(const double* __restrict__ values, int64_t n, int mode,
                   const void* __restrict__ member, bool vec, int nsets, SelWs* ws) {
  extern __shared__ uint32_t sh[];
  __shared__ ColParams cp;
  hist_prologue(ws, sh, nsets, cp);
  for_each_column_key(values, n, mode, member, vec, [&](int set, int col, uint64_t key) {
    atomicAdd(&sh[set * BINS + digit_of(key, cp.kbase[col], cp.shift[col])], 1u);
  });
  hist_epilogue(ws, sh, nsets);
}

// ------------------------------------------------------------------ resolve
constexpr int RES_THREADS = 512;   // (512 threads x 64 registers: fits beside ONE CTA of a streaming kernel, kvm_stream_grid)

// The global histograms are staged in shared memory; one warp per query then walks the
// (summed) histogram of its sets to the bin holding the rank.
__global__ void __launch_bounds__(RES_THREADS, 1) select_resolve_kernel(SelWs* ws) {
  extern __shared__ uint64_t hs[];  // [nsets][BINS]
  __shared__ uint32_t q_bin[MAXQ];
  __shared__ int64_t q_cnt[MAXQ];
  __shared__ QState qs[MAXQ];
  __shared__ uint32_t qmask[MAXQ];
  __shared__ ColParam cps[KVM_MAX_COLS];
  __shared__ int64_t gsum[KVM_MAX_SETS][BINS / 32];
  __shared__ int q_catch[MAXQ];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = RES_THREADS / 32;
  const int nq = ws->nq;
  const int nsets = ws->info.nsets;

  for (int i = threadIdx.x; i < nsets * BINS; i += RES_THREADS) hs[i] = (&ws->hist[0][0])[i];
  if (threadIdx.x < MAXQ) {
    qs[threadIdx.x] = ws->q[threadIdx.x];
    qmask[threadIdx.x] = ws->queries[threadIdx.x].set_mask;
    q_catch[threadIdx.x] = 0;
  }
  if (threadIdx.x < KVM_MAX_COLS) cps[threadIdx.x] = ws->col[threadIdx.x];
  __syncthreads();
  // totals of every 32-bin group of every set (one warp reduction each)
  for (int g = warp; g < nsets * (BINS / 32); g += nwarps) {
    int64_t v = (int64_t)hs[g * 32 + lane];
#pragma unroll
    for (int m = 16; m; m >>= 1) v += __shfl_xor_sync(KVM_FULL, v, m);
    if (lane == 0) gsum[g / (BINS / 32)][g % (BINS / 32)] = v;
  }
  __syncthreads();

  for (int q = warp; q < nq; q += nwarps) {
    QState s = qs[q];
    if (s.status != KVM_Q_OK) continue;
    const ColParam cp = cps[s.col];
    const uint32_t mask = qmask[q];
    if (cp.width == 0 && !ws->preset_cols) {  // one distinct key in the column
      if (lane == 0) {
        s.key = cp.kbase;
        s.resolved = 1;
        s.bucket = -1;
        qs[q] = s;
      }
      continue;
    }
    // two-level walk: 64 group totals (2 per lane) locate the 32-bin group, one scan finishes
    int64_t g0 = 0, g1 = 0;
    #pragma unroll 1
    for (int b = 0; b < nsets; ++b)
      if ((mask >> b) & 1u) {
        g0 += gsum[b][lane];
        g1 += gsum[b][32 + lane];
      }
    int64_t i0 = g0, i1 = g1;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int64_t o0 = __shfl_up_sync(KVM_FULL, i0, d), o1 = __shfl_up_sync(KVM_FULL, i1, d);
      if (lane >= d) {
        i0 += o0;
        i1 += o1;
      }
    }
    const int64_t half = __shfl_sync(KVM_FULL, i0, 31);
    i1 += half;
    const unsigned h0 = __ballot_sync(KVM_FULL, i0 > s.rank), h1 = __ballot_sync(KVM_FULL, i1 > s.rank);
    bool found = (h0 | h1) != 0;
    if (found) {
      const int gl = h0 ? __ffs(h0) - 1 : __ffs(h1) - 1;
      const int grp = h0 ? gl : 32 + gl;
      const int64_t before = h0 ? __shfl_sync(KVM_FULL, i0 - g0, gl) : __shfl_sync(KVM_FULL, i1 - g1, gl);
      const int base = grp * 32;
      int64_t c = 0;
      #pragma unroll 1
      for (int b = 0; b < nsets; ++b)
        if ((mask >> b) & 1u) c += (int64_t)hs[b * BINS + base + lane];
      int64_t incl = c;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        int64_t o = __shfl_up_sync(KVM_FULL, incl, d);
        if (lane >= d) incl += o;
      }
      const unsigned hit = __ballot_sync(KVM_FULL, before + incl > s.rank);
      found = hit != 0;
      if (hit) {
        const int l = __ffs(hit) - 1;
        const int64_t excl = __shfl_sync(KVM_FULL, incl - c, l);
        const int64_t cnt = __shfl_sync(KVM_FULL, c, l);
        if (lane == 0) {
          q_bin[q] = (uint32_t)(base + l);
          q_cnt[q] = cnt;
          s.resid = s.rank - (before + excl);
          // (sampled digit parameters: an end bin may hold keys from outside its nominal range)
          const uint64_t kmn = ws->info.kmin[s.col], kmx = ws->info.kmax[s.col];
          const uint64_t top = cp.kbase + ((uint64_t)(BINS - 1) << cp.shift);
          const bool impure = ws->preset_cols &&
                              ((base + l == 0 && kmn < cp.kbase) ||
                               (base + l == BINS - 1 && kmx >= top && shr64(kmx - top, cp.shift) != 0));
          q_catch[q] = impure ? 1 : 0;
          if (cp.shift == 0 && !impure) {  // the digit is the whole key
            s.key = cp.kbase + (uint64_t)(base + l);
            s.resolved = 1;
            s.bucket = -1;
          } else {
            s.resolved = 0;
            s.bucket = -2;  // assigned below
          }
          qs[q] = s;
        }
      }
    }
    if (!found && lane == 0) {  // histogram inconsistent with counts (caller error)
      s.status = KVM_Q_OUT_OF_RANGE;
      s.resolved = 1;
      qs[q] = s;
    }
  }
  __syncthreads();
  // ---- buckets = distinct (col, bin, mask) among the unresolved queries, in (col, bin, mask)
  // order so that buckets sharing a fine bin are contiguous.  All O(nq) loops run one thread
  // per query in parallel.
  __shared__ uint64_t bkey[MAXQ];   // sort key; ~0 = query needs no bucket
  __shared__ int is_first[MAXQ];
  __shared__ int pos_of[MAXQ];
  __shared__ int64_t off_of[MAXQ];
  __shared__ int nb_s;
  __shared__ int64_t used_s;
  const int t = threadIdx.x;
  if (t < MAXQ) {
    const QState s = qs[t];
    const bool need = t < nq && s.status == KVM_Q_OK && !s.resolved;
    bkey[t] = need ? (((uint64_t)s.col << 44) | ((uint64_t)q_bin[t] << 32) | qmask[t]) : ~0ull;
  }
  __syncthreads();
  if (t < MAXQ) {
    const uint64_t k = bkey[t];
    int first = k != ~0ull;
    #pragma unroll 1
    for (int o = 0; o < t && first; ++o) first = bkey[o] != k;
    is_first[t] = first;
  }
  __syncthreads();
  if (t < MAXQ) {
    const uint64_t k = bkey[t];
    int pos = 0;
    int64_t off = 0;
    if (k != ~0ull)
      #pragma unroll 1
      for (int o = 0; o < MAXQ; ++o)
        if (is_first[o] && bkey[o] < k) {
          ++pos;
          off += q_cnt[o];
        }
    pos_of[t] = pos;   // same for every query of the bucket
    off_of[t] = off;
  }
  if (t == 0) {
    int nb = 0;
    int64_t used = 0;
    #pragma unroll 1
    for (int o = 0; o < MAXQ; ++o)
      if (is_first[o]) {
        ++nb;
        used += q_cnt[o];
      }
    nb_s = nb;
    used_s = used;
  }
  __syncthreads();
  if (t < MAXQ && is_first[t]) {
    const QState s = qs[t];
    const ColParam cp = cps[s.col];
    const uint64_t k = bkey[t];
    // (col, bin) group: first bucket of the group carries the group's length
    int smaller_same = 0, group = 0;
    #pragma unroll 1
    for (int o = 0; o < MAXQ; ++o)
      if (is_first[o] && (bkey[o] >> 32) == (k >> 32)) {
        ++group;
        smaller_same += bkey[o] < k;
      }
    Bucket b;
    b.base = cp.kbase + ((uint64_t)q_bin[t] << cp.shift);
    b.shift = cp.shift;
    if (q_catch[t]) {
      // catch-all end bin: the bucket spans from the column's true minimum (bin 0) / up to its
      // true maximum (last bin); shift = bits needed for (key - base) inside it
      const uint64_t kmn = ws->info.kmin[s.col], kmx = ws->info.kmax[s.col];
      if (q_bin[t] == 0) {
        const uint64_t hi = cp.kbase + ((1ull << cp.shift) - 1ull);   // last key of the nominal bin
        b.base = kmn;
        b.shift = bit_width64((kmx < hi ? kmx : hi) - kmn);
      } else {
        b.shift = bit_width64(kmx - b.base);
      }
      if (b.shift == 0) b.shift = 1;   // (a bucket is always narrowed by at least one pass)
    }
    b.col = s.col;
    b.set_mask = qmask[t];
    b.bin = q_bin[t];
    b.count = q_cnt[t];
    b.run = smaller_same == 0 ? group : 0;
    b.pad = 0;
    ws->bucket[pos_of[t]] = b;
    ws->parent[pos_of[t]] = b;
    ws->child_first[pos_of[t]] = (uint8_t)pos_of[t];
    ws->child_n[pos_of[t]] = 1;
    ws->tparent[pos_of[t]] = pos_of[t];
    ws->bucket_offsets[pos_of[t]] = off_of[t];
    if (smaller_same == 0)
      ws->tab[s.col][b.bin] = (uint16_t)((uint32_t)pos_of[t] | ((uint32_t)ws->edge_code[s.col][b.bin] << 8));
  }
  if (t <= MAXQ && t >= nb_s) ws->bucket_offsets[t] = used_s;
  if (t == 0) {
    ws->nbuckets = nb_s;
    ws->nparents = nb_s;
    ws->arena_used = used_s;
    ws->level = 1;
    // every level-1 bucket has shift > 0, so a refine round can always narrow it
    const int again = ws->refine_budget > 0 && used_s > ws->refine_min;
    ws->refine = again;
    if (!again && used_s > ws->arena_capacity) ws->status = KVM_ERR_ARENA;
  }
  if (t < MAXQ) {
    QState s = qs[t];
    if (bkey[t] != ~0ull) s.bucket = pos_of[t];
    ws->q[t] = s;
  }
}

// ------------------------------------------------------------------ refine
// One refine round = a pass over the digit records that histograms, per candidate bucket, the
// next 11 bits of its keys (log_rec_kernel<ROWS, 1> -> ws->hist2), then this kernel: every
// unresolved query walks the histogram of its bucket to the sub-bin holding its residual rank
// and the bucket is replaced by that sub-bin (2048x fewer keys).  The lookup table keeps
// pointing at the level-1 buckets ("parents"); each parent owns the contiguous run of targets
// that subdivide it, and a key belongs to a target iff (key - base) >> shift == 0.
// Runs only while resolve / the previous round left ws->refine set, i.e. when the buckets
// together hold more than refine_min keys (large logs).
constexpr int REF_THREADS = 512;

__global__ void __launch_bounds__(REF_THREADS, 1) select_refine_kernel(SelWs* ws) {
  if (ws->status != KVM_OK || !ws->refine) return;
  __shared__ QState qs[MAXQ];
  __shared__ Bucket tg[MAXQ];
  __shared__ int32_t tpar[MAXQ];
  __shared__ uint32_t q_sub[MAXQ];
  __shared__ int64_t q_cnt[MAXQ];
  __shared__ uint64_t bkey[MAXQ];
  __shared__ int is_first[MAXQ];
  __shared__ int pos_of[MAXQ];
  __shared__ int64_t off_of[MAXQ];
  __shared__ int newpar[MAXQ];
  __shared__ int nb_s;
  __shared__ int64_t used_s;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5, nwarps = REF_THREADS / 32;
  const int nq = ws->nq;
  if (t < MAXQ) {
    qs[t] = ws->q[t];
    tg[t] = ws->bucket[t];
    tpar[t] = ws->tparent[t];
    bkey[t] = ~0ull;
    q_cnt[t] = 0;
  }
  __syncthreads();
  for (int q = warp; q < nq; q += nwarps) {
    QState s = qs[q];
    if (s.status != KVM_Q_OK || s.resolved || s.bucket < 0) continue;
    const Bucket T = tg[s.bucket];
    const uint64_t* __restrict__ h = ws->hist2[s.bucket];
    const int take = T.shift < BINS_LOG2 ? T.shift : BINS_LOG2;
    const int sh2 = T.shift - take;
    int64_t running = 0;
    bool found = false;
    // 16 groups of 32 bins at a time: coalesced loads issued together, group totals by 16
    // independent warp reductions (pipelined), then one scan inside the group holding the rank
    constexpr int CHUNK = 16;
    for (int cbase = 0; cbase < (1 << take) && !found; cbase += 32 * CHUNK) {
      int64_t hv[CHUNK], tot[CHUNK];
#pragma unroll
      for (int u = 0; u < CHUNK; ++u) {
        hv[u] = (int64_t)__ldcg(h + cbase + 32 * u + lane);  // rows are zero past 2^take
        tot[u] = hv[u];
      }
#pragma unroll
      for (int m = 16; m; m >>= 1) {
#pragma unroll
        for (int u = 0; u < CHUNK; ++u) tot[u] += __shfl_xor_sync(KVM_FULL, tot[u], m);
      }
      int gu = -1;
      int64_t before = running, c = 0;
#pragma unroll
      for (int u = 0; u < CHUNK; ++u) {
        if (gu < 0) {
          if (before + tot[u] > s.resid) {
            gu = u;
            c = hv[u];
          } else {
            before += tot[u];
          }
        }
      }
      if (gu < 0) {
        running = before;
        continue;
      }
      int64_t incl = c;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        int64_t o = __shfl_up_sync(KVM_FULL, incl, d);
        if (lane >= d) incl += o;
      }
      const unsigned hit = __ballot_sync(KVM_FULL, before + incl > s.resid);
      const int l = __ffs(hit) - 1;   // hit != 0: the group total exceeds the remaining rank
      const int64_t excl = __shfl_sync(KVM_FULL, incl - c, l);
      const int64_t cnt = __shfl_sync(KVM_FULL, c, l);
      if (lane == 0) {
        const uint32_t sub = (uint32_t)(cbase + 32 * gu + l);
        s.resid = s.resid - (before + excl);
        if (sh2 == 0) {  // the sub-digit completes the key
          s.key = T.base + (uint64_t)sub;
          s.resolved = 1;
          s.bucket = -1;
        } else {
          q_sub[q] = sub;
          q_cnt[q] = cnt;
          bkey[q] = ((uint64_t)(uint32_t)s.bucket << 32) | sub;
        }
        qs[q] = s;
      }
      found = true;
    }
    if (!found && lane == 0) {  // histogram inconsistent with the bucket count
      s.status = KVM_Q_OUT_OF_RANGE;
      s.resolved = 1;
      qs[q] = s;
    }
  }
  __syncthreads();
  // new targets = distinct (old target, sub-bin), in that order: children of a parent stay contiguous
  if (t < MAXQ) {
    const uint64_t k = bkey[t];
    int first = k != ~0ull;
    #pragma unroll 1
    for (int o = 0; o < t && first; ++o) first = bkey[o] != k;
    is_first[t] = first;
  }
  __syncthreads();
  if (t < MAXQ) {
    const uint64_t k = bkey[t];
    int pos = 0;
    int64_t off = 0;
    if (k != ~0ull)
      #pragma unroll 1
      for (int o = 0; o < MAXQ; ++o)
        if (is_first[o] && bkey[o] < k) {
          ++pos;
          off += q_cnt[o];
        }
    pos_of[t] = pos;
    off_of[t] = off;
  }
  if (t == 0) {
    int nb = 0;
    int64_t used = 0;
    #pragma unroll 1
    for (int o = 0; o < MAXQ; ++o)
      if (is_first[o]) {
        ++nb;
        used += q_cnt[o];
      }
    nb_s = nb;
    used_s = used;
  }
  __syncthreads();
  if (t < MAXQ && is_first[t]) {
    const int old = (int)(bkey[t] >> 32);
    const Bucket T = tg[old];
    const int sh2 = T.shift - BINS_LOG2;  // > 0 here
    Bucket b = T;
    b.base = T.base + ((uint64_t)q_sub[t] << sh2);
    b.shift = sh2;
    b.count = q_cnt[t];
    b.run = 0;
    const int p = pos_of[t];
    ws->bucket[p] = b;
    ws->bucket_offsets[p] = off_of[t];
    ws->tparent[p] = tpar[old];
    newpar[p] = tpar[old];
  }
  __syncthreads();
  if (t < MAXQ) {  // t as a level-1 bucket: the run of targets below it
    int first = 0, cnt = 0;
    #pragma unroll 1
    for (int p = nb_s - 1; p >= 0; --p)
      if (newpar[p] == t) {
        first = p;
        ++cnt;
      }
    ws->child_first[t] = (uint8_t)first;
    ws->child_n[t] = (uint8_t)cnt;
  }
  if (t <= MAXQ && t >= nb_s) ws->bucket_offsets[t] = used_s;
  if (t == 0) {
    ws->nbuckets = nb_s;
    ws->arena_used = used_s;
    ws->level = ws->level + 1;
    const int budget = ws->refine_budget - 1;
    ws->refine_budget = budget;
    const int again = budget > 0 && used_s > ws->refine_min;
    ws->refine = again;
    if (!again && used_s > ws->arena_capacity) ws->status = KVM_ERR_ARENA;
  }
  if (t < MAXQ) {
    QState s = qs[t];
    if (bkey[t] != ~0ull) s.bucket = pos_of[t];
    ws->q[t] = s;
  }
}

// ------------------------------------------------------------------ compact
constexpr int STAGE_SLOTS = 112;  // candidates a CTA stages per bucket before going to global memory

struct CompactSmem {
  ColParams cp;
  Bucket bk[MAXQ];
  int64_t bk_off[MAXQ];
  uint64_t thr[NH][NB_MAX];
  uint32_t edge[NH][NB_MAX];
  int32_t col_hist[KVM_MAX_COLS];
  int32_t num_bins;   // value histogram bins
  uint32_t stage_n[MAXQ];                 // candidates staged per bucket
  unsigned long long stage_pos[MAXQ];
  uint64_t stage[MAXQ][STAGE_SLOTS];
  uint16_t tab[KVM_LOG_NCOLS][BINS];
  // record-driven passes: bk[] holds the level-1 buckets (table targets), these the append targets
  uint64_t tbase[MAXQ];
  int64_t tcount[MAXQ];
  int32_t tshift[MAXQ];
  uint8_t cf[MAXQ];
  uint8_t cn[MAXQ];
  int32_t max_cn;
};

template <int NCOLS, class Stream>
__device__ __forceinline__ void compact_body(SelWs* ws, uint64_t* __restrict__ arena, Stream&& stream) {
  extern __shared__ __align__(16) unsigned char craw[];
  CompactSmem& S = *reinterpret_cast<CompactSmem*>(craw);
  if (ws->status != KVM_OK) return;
  if (ws->level > 1) return;  // refined buckets need the record-driven pass (host enforces it)
  if (threadIdx.x == 0) S.num_bins = ws->hspec.num_bins;
  if (threadIdx.x < KVM_MAX_COLS) {
    S.cp.kbase[threadIdx.x] = ws->col[threadIdx.x].kbase;
    S.cp.shift[threadIdx.x] = ws->col[threadIdx.x].shift;
    S.col_hist[threadIdx.x] = ws->col_hist[threadIdx.x];
  }
  for (int i = threadIdx.x; i < NCOLS * BINS / 2; i += blockDim.x)
    reinterpret_cast<uint32_t*>(&S.tab[0][0])[i] = reinterpret_cast<const uint32_t*>(&ws->tab[0][0])[i];
  for (int i = threadIdx.x; i < NH * NB_MAX; i += blockDim.x) {
    (&S.thr[0][0])[i] = (&ws->thr20[0][0])[i];
    (&S.edge[0][0])[i] = 0;
  }
  for (int i = threadIdx.x; i < MAXQ; i += blockDim.x) {
    S.bk[i] = ws->bucket[i];
    S.bk_off[i] = ws->bucket_offsets[i];
    S.stage_n[i] = 0;
  }
  __syncthreads();
  uint64_t kb[NCOLS];
  int shf[NCOLS];
#pragma unroll
  for (int c = 0; c < NCOLS; ++c) {
    kb[c] = S.cp.kbase[c];
    shf[c] = S.cp.shift[c];
  }
  stream([&](int set, int col, uint64_t key) {
    const uint32_t d = digit_of(key, kb[col], shf[col]);
    const uint32_t e = S.tab[col][d];   // one 16-bit lookup decides both rare paths
    if (e == 0x00ffu) return;
    const uint32_t j0 = e & 0xffu;
    if (j0 != 255u) {
      const int run = S.bk[j0].run;
      for (int j = j0; j < (int)j0 + run; ++j)
        if ((S.bk[j].set_mask >> set) & 1u) {
          // staged in shared memory (no round trip to L2 in the streaming loop); a bucket
          // that outgrows its stage falls back to a direct append
          const uint32_t sp = atomicAdd(&S.stage_n[j], 1u);
          if (sp < (uint32_t)STAGE_SLOTS) {
            S.stage[j][sp] = key;
          } else {
            const unsigned long long pos =
                atomicAdd(reinterpret_cast<unsigned long long*>(&ws->bucket_counts[j]), 1ull);
            if ((int64_t)pos < S.bk[j].count) arena[S.bk_off[j] + (int64_t)pos] = key;
          }
        }
    }
    const uint32_t code = e >> 8;
    if (code) {
      const int h = S.col_hist[col];
      const int bin = code == 255u ? bin_of(key, S.thr[h], S.num_bins)
                                   : (int)code - (key < S.thr[h][code - 1] ? 1 : 0);
      atomicAdd(&S.edge[h][bin], 1u);
    }
  });
  __syncthreads();
  // flush the stages: one global cursor bump per non-empty bucket, then a cooperative copy
  if (threadIdx.x < MAXQ) {
    const uint32_t c = S.stage_n[threadIdx.x] < (uint32_t)STAGE_SLOTS ? S.stage_n[threadIdx.x] : (uint32_t)STAGE_SLOTS;
    S.stage_n[threadIdx.x] = c;
    if (c) S.stage_pos[threadIdx.x] =
        atomicAdd(reinterpret_cast<unsigned long long*>(&ws->bucket_counts[threadIdx.x]), (unsigned long long)c);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < MAXQ * STAGE_SLOTS; i += blockDim.x) {
    const int j = i / STAGE_SLOTS, k = i % STAGE_SLOTS;
    if ((uint32_t)k < S.stage_n[j]) {
      const int64_t pos = (int64_t)S.stage_pos[j] + k;
      if (pos < S.bk[j].count) arena[S.bk_off[j] + pos] = S.stage[j][k];
    }
  }
  for (int i = threadIdx.x; i < NH * NB_MAX; i += blockDim.x) {
    const uint32_t c = (&S.edge[0][0])[i];
    if (c) atomicAdd(&ws->edge20[0][0] + i, (unsigned long long)c);
  }
}

// Slow path of the compact pass for one key whose fine bin is "interesting" (table entry e).
__device__ __forceinline__ void compact_hit(SelWs* ws, CompactSmem& S, uint64_t* __restrict__ arena,
                                            int set, int col, uint32_t e, uint64_t key) {
  const uint32_t j0 = e & 0xffu;
  if (j0 != 255u) {
    const int run = S.bk[j0].run;
    for (int j = j0; j < (int)j0 + run; ++j)
      if ((S.bk[j].set_mask >> set) & 1u) {
        const uint32_t sp = atomicAdd(&S.stage_n[j], 1u);
        if (sp < (uint32_t)STAGE_SLOTS) {
          S.stage[j][sp] = key;
        } else {
          const unsigned long long pos =
              atomicAdd(reinterpret_cast<unsigned long long*>(&ws->bucket_counts[j]), 1ull);
          if ((int64_t)pos < S.bk[j].count) arena[S.bk_off[j] + (int64_t)pos] = key;
        }
      }
  }
  const uint32_t code = e >> 8;
  if (code) {
    const int h = S.col_hist[col];
    const int bin = code == 255u ? bin_of(key, S.thr[h], S.num_bins)
                                 : (int)code - (key < S.thr[h][code - 1] ? 1 : 0);
    atomicAdd(&S.edge[h][bin], 1u);
  }
}

// Warp-cooperative form of compact_hit for the drain phase: all 32 lanes call it (lanes
// without an entry pass valid=false).  Lanes that append to the same bucket are merged with
// __match_any_sync so that one shared-memory atomic reserves the slots of the whole group.
__device__ __noinline__ void compact_hits_warp(SelWs* ws, CompactSmem& S, uint64_t* __restrict__ arena,
                                               bool valid, int set, int col, uint32_t e, uint64_t key) {
  const int lane = threadIdx.x & 31;
  const unsigned lt = (1u << lane) - 1u;
  const uint32_t j0 = valid ? (e & 0xffu) : 255u;
  int run = 0;
  if (j0 != 255u) run = S.bk[j0].run;
  int maxrun = run;
#pragma unroll
  for (int m = 16; m; m >>= 1) {
    const int o = __shfl_xor_sync(KVM_FULL, maxrun, m);
    maxrun = o > maxrun ? o : maxrun;
  }
  for (int r = 0; r < maxrun; ++r) {
    int j = -1;
    if (r < run && ((S.bk[j0 + r].set_mask >> set) & 1u)) j = (int)j0 + r;
    const unsigned peers = __match_any_sync(KVM_FULL, j);
    if (j >= 0) {
      const int leader = __ffs(peers) - 1;
      uint32_t base = 0;
      if (lane == leader) base = atomicAdd(&S.stage_n[j], (uint32_t)__popc(peers));
      base = __shfl_sync(peers, base, leader);
      const uint32_t sp = base + (uint32_t)__popc(peers & lt);
      if (sp < (uint32_t)STAGE_SLOTS) {
        S.stage[j][sp] = key;
      } else {
        const unsigned long long pos =
            atomicAdd(reinterpret_cast<unsigned long long*>(&ws->bucket_counts[j]), 1ull);
        if ((int64_t)pos < S.bk[j].count) arena[S.bk_off[j] + (int64_t)pos] = key;
      }
    }
  }
  const uint32_t code = valid ? (e >> 8) : 0u;
  if (code) {
    const int h = S.col_hist[col];
    const int bin = code == 255u ? bin_of(key, S.thr[h], S.num_bins)
                                 : (int)code - (key < S.thr[h][code - 1] ? 1 : 0);
    atomicAdd(&S.edge[h][bin], 1u);
  }
}

// Warp-cooperative slow path of the record-driven passes: all 32 lanes call it (lanes without
// an entry pass valid=false).  The table entry names the level-1 buckets of the key's fine bin;
// each owns a run of targets, and the key belongs to the target whose key range contains it.
// PHASE 0 (compact): append the key to the target - lanes appending to the same target are
// merged with __match_any_sync so that one shared-memory atomic reserves the slots of the group;
// keys in a fine bin with a histogram edge are also binned exactly.
// PHASE 1 (refine): count the key's next 11 bits into the target's row of ws->hist2.
// (kept OUT of line and the callers' column loops rolled: this code runs a few times per warp, and
// five inlined copies of it made the kernel 19,000 instructions long - the profile showed as many
// stall samples on fetching those once-executed instructions as in the streaming loop itself)
template <int PHASE>
__device__ __noinline__ void target_hits_warp(SelWs* ws, CompactSmem& S, uint64_t* __restrict__ arena,
                                              bool valid, int set, int col, uint32_t e, uint64_t key) {
  const int lane = threadIdx.x & 31;
  const unsigned lt = (1u << lane) - 1u;
  const uint32_t j0 = valid ? (e & 0xffu) : 255u;
  int run = 0;
  if (j0 != 255u) run = S.bk[j0].run;
  int maxrun = run;
#pragma unroll
  for (int m = 16; m; m >>= 1) {
    const int o = __shfl_xor_sync(KVM_FULL, maxrun, m);
    maxrun = o > maxrun ? o : maxrun;
  }
  const int maxc = S.max_cn;
  for (int r = 0; r < maxrun; ++r) {
    int j = -1;
    if (r < run && ((S.bk[j0 + r].set_mask >> set) & 1u)) j = (int)j0 + r;
    const int cf = j >= 0 ? (int)S.cf[j] : 0, cn = j >= 0 ? (int)S.cn[j] : 0;
    for (int cc = 0; cc < maxc; ++cc) {
      int c = -1;
      if (cc < cn && shr64(key - S.tbase[cf + cc], S.tshift[cf + cc]) == 0) c = cf + cc;
      if (PHASE == 1) {
        if (c >= 0) {
          const int sh2 = S.tshift[c] > BINS_LOG2 ? S.tshift[c] - BINS_LOG2 : 0;
          atomicAdd(reinterpret_cast<unsigned long long*>(&ws->hist2[c][(uint32_t)((key - S.tbase[c]) >> sh2)]),
                    1ull);
        }
      } else {
        const unsigned peers = __match_any_sync(KVM_FULL, c);
        if (c >= 0) {
          const int leader = __ffs(peers) - 1;
          uint32_t base = 0;
          if (lane == leader) base = atomicAdd(&S.stage_n[c], (uint32_t)__popc(peers));
          base = __shfl_sync(peers, base, leader);
          const uint32_t sp = base + (uint32_t)__popc(peers & lt);
          if (sp < (uint32_t)STAGE_SLOTS) {
            S.stage[c][sp] = key;
          } else {
            const unsigned long long pos =
                atomicAdd(reinterpret_cast<unsigned long long*>(&ws->bucket_counts[c]), 1ull);
            if ((int64_t)pos < S.tcount[c]) arena[S.bk_off[c] + (int64_t)pos] = key;
          }
        }
      }
    }
  }
  if (PHASE == 0) {
    const uint32_t code = valid ? (e >> 8) : 0u;
    if (code) {
      const int h = S.col_hist[col];
      const int bin = code == 255u ? bin_of(key, S.thr[h], S.num_bins)
                                   : (int)code - (key < S.thr[h][code - 1] ? 1 : 0);
      atomicAdd(&S.edge[h][bin], 1u);
    }
  }
}

constexpr int CQ_CAP = 128;  // per-warp queue of interesting keys

struct CompactQueue {
  uint64_t key[SEL_THREADS / 32][CQ_CAP];
  uint32_t meta[SEL_THREADS / 32][CQ_CAP];  // set | col << 8 | table entry << 16
};

// Log compact pass.  Only ~3% of the keys fall into a fine bin that holds a requested rank or
// a histogram edge, but with 10 key sites per lane-iteration almost every warp iteration has a
// few: handling them in place would serialise the warp on every site.  Instead each site
// ballots its hits into a per-warp shared-memory queue (no atomics, no divergence in the
// streaming loop) and the queue is drained 32 entries at a time with all lanes busy.
template <int ROWS>
__global__ void __launch_bounds__(SEL_THREADS, 2)
log_compact_kernel(kvm_request_log log, const uint8_t* __restrict__ flags, SelWs* ws, uint64_t* arena) {
  extern __shared__ __align__(16) unsigned char craw[];
  CompactSmem& S = *reinterpret_cast<CompactSmem*>(craw);
  CompactQueue& Q = *reinterpret_cast<CompactQueue*>(craw + ((sizeof(CompactSmem) + 15) / 16) * 16);
  if (ws->status != KVM_OK) return;
  if (threadIdx.x == 0) S.num_bins = ws->hspec.num_bins;
  if (threadIdx.x < KVM_MAX_COLS) {
    S.cp.kbase[threadIdx.x] = ws->col[threadIdx.x].kbase;
    S.cp.shift[threadIdx.x] = ws->col[threadIdx.x].shift;
    S.col_hist[threadIdx.x] = ws->col_hist[threadIdx.x];
  }
  for (int i = threadIdx.x; i < KVM_LOG_NCOLS * BINS / 2; i += SEL_THREADS)
    reinterpret_cast<uint32_t*>(&S.tab[0][0])[i] = reinterpret_cast<const uint32_t*>(&ws->tab[0][0])[i];
  for (int i = threadIdx.x; i < NH * NB_MAX; i += SEL_THREADS) {
    (&S.thr[0][0])[i] = (&ws->thr20[0][0])[i];
    (&S.edge[0][0])[i] = 0;
  }
  for (int i = threadIdx.x; i < MAXQ; i += SEL_THREADS) {
    S.bk[i] = ws->bucket[i];
    S.bk_off[i] = ws->bucket_offsets[i];
    S.stage_n[i] = 0;
  }
  __syncthreads();
  uint64_t kb[KVM_LOG_NCOLS];
  int shf[KVM_LOG_NCOLS];
#pragma unroll
  for (int c = 0; c < KVM_LOG_NCOLS; ++c) {
    kb[c] = S.cp.kbase[c];
    shf[c] = S.cp.shift[c];
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned lt = (1u << lane) - 1u;
  int qn = 0;  // warp-uniform queue length
  auto drain = [&]() {
    __syncwarp();
    for (int i0 = 0; i0 < qn; i0 += 32) {
      const int i = i0 + lane;
      const bool valid = i < qn;
      const uint32_t m = valid ? Q.meta[warp][i] : 0u;
      compact_hits_warp(ws, S, arena, valid, m & 0xff, (m >> 8) & 0xff, m >> 16, valid ? Q.key[warp][i] : 0ull);
    }
    __syncwarp();
    qn = 0;
  };
  // one key site, executed by all 32 lanes: `valid` lanes look their digit up, hits are queued
  auto site = [&](bool valid, int set, int col, uint64_t key) {
    uint32_t e = 0x00ffu;
    if (valid) e = S.tab[col][digit_of(key, kb[col], shf[col])];
    const unsigned hits = __ballot_sync(KVM_FULL, e != 0x00ffu);
    if (hits) {
      const int cnt = __popc(hits);
      if (qn + cnt > CQ_CAP) drain();
      if (e != 0x00ffu) {
        const int pos = qn + __popc(hits & lt);
        Q.key[warp][pos] = key;
        Q.meta[warp][pos] = (uint32_t)set | ((uint32_t)col << 8) | (e << 16);
      }
      qn += cnt;
    }
  };
  auto row = [&](uint8_t fl, double la, double tf, double tl, long long ct) {
    const bool ok = fl & KVM_F_SUCCESS;
    const bool cold = fl & KVM_F_COLD;
    site(ok && !kvm_isnan(la), cold ? KVM_SET_LAT_COLD : KVM_SET_LAT_WARM, KVM_COL_LAT, kvm_key(la));
    site(ok && !kvm_isnan(tf), cold ? KVM_SET_TTFB_COLD : KVM_SET_TTFB_WARM, KVM_COL_TTFB, kvm_key(tf));
    site(ok && !kvm_isnan(tl), KVM_SET_TLLT, KVM_COL_TLLT, kvm_key(tl));
    const bool gen = ok && (fl & KVM_F_GEN);
    const double g = tl - tf;                       // analyze.py:144
    const bool gv = gen && !kvm_isnan(g);
    site(gv, KVM_SET_GEN, KVM_COL_GEN, kvm_key(g));
    double pt = 0.0;
    if (gv) pt = g / (double)ct;                    // analyze.py:146
    site(gv && !kvm_isnan(pt), KVM_SET_PERTOK, KVM_COL_PERTOK, kvm_key(pt));
  };

  const int64_t n = log.n;
  const int64_t nunits = ROWS == 2 ? (n >> 1) : n;
  for (int64_t base = (int64_t)blockIdx.x * SEL_THREADS + (threadIdx.x & ~31); base < nunits;
       base += (int64_t)gridDim.x * SEL_THREADS) {
    const int64_t u = base + lane;
    const bool live = u < nunits;
    const int64_t i = ROWS == 2 ? (u << 1) : u;
    uint8_t f0 = 0, f1 = 0;
    if (live) {
      if (ROWS == 2) {
        const uint16_t ff = kvm_ld_u8x2(flags + i);
        f0 = (uint8_t)(ff & 0xff);
        f1 = (uint8_t)(ff >> 8);
      } else {
        f0 = kvm_ld_u8(flags + i);
      }
    }
    // skip the loads of a warp whose rows all failed
    if (!__any_sync(KVM_FULL, (f0 | f1) & KVM_F_SUCCESS)) continue;
    if (ROWS == 2) {
      double2 la{0.0, 0.0}, tf{0.0, 0.0}, tl{0.0, 0.0};
      longlong2 ct{0, 0};
      if ((f0 | f1) & KVM_F_SUCCESS) {
        la = kvm_ld2(log.latency_ms + i);
        tf = kvm_ld2(log.ttfb_ms + i);
        tl = kvm_ld2(log.tllt_ms + i);
        ct = kvm_ld2(log.completion_tokens + i);
      }
      row(f0, la.x, tf.x, tl.x, ct.x);
      row(f1, la.y, tf.y, tl.y, ct.y);
    } else {
      double la = 0.0, tf = 0.0, tl = 0.0;
      long long ct = 0;
      if (f0 & KVM_F_SUCCESS) {
        la = kvm_ld1(log.latency_ms + i);
        tf = kvm_ld1(log.ttfb_ms + i);
        tl = kvm_ld1(log.tllt_ms + i);
        ct = kvm_ld1(log.completion_tokens + i);
      }
      row(f0, la, tf, tl, ct);
    }
    if (qn >= 32) drain();
  }
  drain();
  if (ROWS == 2 && (n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {  // odd tail row
    const int64_t i = n - 1;
    log_row_keys(flags[i], log.latency_ms[i], log.ttfb_ms[i], log.tllt_ms[i], log.completion_tokens[i],
                 [&](int set, int col, uint64_t key) {
                   const uint32_t e = S.tab[col][digit_of(key, kb[col], shf[col])];
                   if (e != 0x00ffu) compact_hit(ws, S, arena, set, col, e, key);
                 });
  }
  __syncthreads();
  // flush the stages: one global cursor bump per non-empty bucket, then a cooperative copy
  if (threadIdx.x < MAXQ) {
    const uint32_t c = S.stage_n[threadIdx.x] < (uint32_t)STAGE_SLOTS ? S.stage_n[threadIdx.x] : (uint32_t)STAGE_SLOTS;
    S.stage_n[threadIdx.x] = c;
    if (c) S.stage_pos[threadIdx.x] =
        atomicAdd(reinterpret_cast<unsigned long long*>(&ws->bucket_counts[threadIdx.x]), (unsigned long long)c);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < MAXQ * STAGE_SLOTS; i += SEL_THREADS) {
    const int j = i / STAGE_SLOTS, k = i % STAGE_SLOTS;
    if ((uint32_t)k < S.stage_n[j]) {
      const int64_t pos = (int64_t)S.stage_pos[j] + k;
      if (pos < S.bk[j].count) arena[S.bk_off[j] + pos] = S.stage[j][k];
    }
  }
  for (int i = threadIdx.x; i < NH * NB_MAX; i += SEL_THREADS) {
    const uint32_t c = (&S.edge[0][0])[i];
    if (c) atomicAdd(&ws->edge20[0][0] + i, (unsigned long long)c);
  }
}

// ---- compact pass driven by the digit records of the hist pass --------------------------------
// Streams 9 B/row (8 B record + 1 B flags).  A row's value is fetched from its column only when
// one of its fine bins is interesting (~3% of the sites), in the dense drain phase.
struct RecQueue {
  int64_t row[SEL_THREADS / 32][CQ_CAP];
  uint32_t hitmask[SEL_THREADS / 32][CQ_CAP];  // bit c: the row's value in column c is interesting
};

// PHASE 0: the compact pass.  PHASE 1: the histogram pass of a refine round (same streaming side,
// the drain counts sub-digits instead of appending; runs only while ws->refine is set).
template <int ROWS, int PHASE>
__global__ void __launch_bounds__(SEL_THREADS, 2)
log_rec_kernel(kvm_request_log log, const uint8_t* __restrict__ flags,
               const uint64_t* __restrict__ recs, SelWs* ws, uint64_t* arena) {
  extern __shared__ __align__(16) unsigned char craw[];
  CompactSmem& S = *reinterpret_cast<CompactSmem*>(craw);
  RecQueue& Q = *reinterpret_cast<RecQueue*>(craw + ((sizeof(CompactSmem) + 15) / 16) * 16);
  if (ws->status != KVM_OK) return;
  if (PHASE == 1 && !ws->refine) return;
  if (threadIdx.x == 0) S.num_bins = ws->hspec.num_bins;
  if (threadIdx.x < KVM_MAX_COLS) {
    S.col_hist[threadIdx.x] = ws->col_hist[threadIdx.x];
    S.cp.kbase[threadIdx.x] = ws->col[threadIdx.x].kbase;
    S.cp.shift[threadIdx.x] = ws->col[threadIdx.x].shift;
  }
  for (int i = threadIdx.x; i < KVM_LOG_NCOLS * BINS / 2; i += SEL_THREADS)
    reinterpret_cast<uint32_t*>(&S.tab[0][0])[i] = reinterpret_cast<const uint32_t*>(&ws->tab[0][0])[i];
  for (int i = threadIdx.x; i < NH * NB_MAX; i += SEL_THREADS) {
    (&S.thr[0][0])[i] = (&ws->thr20[0][0])[i];
    (&S.edge[0][0])[i] = 0;
  }
  if (threadIdx.x == 0) S.max_cn = 0;
  __syncthreads();
  {
    const int npar = ws->nparents;
    for (int i = threadIdx.x; i < MAXQ; i += SEL_THREADS) {
      const Bucket tg = ws->bucket[i];
      S.bk[i] = ws->parent[i];
      S.bk_off[i] = ws->bucket_offsets[i];
      S.tbase[i] = tg.base;
      S.tshift[i] = tg.shift;
      S.tcount[i] = tg.count;
      const int cn = i < npar ? ws->child_n[i] : 0;
      S.cf[i] = ws->child_first[i];
      S.cn[i] = (uint8_t)cn;
      S.stage_n[i] = 0;
      if (cn) atomicMax(&S.max_cn, cn);
    }
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned lt = (1u << lane) - 1u;
  int qn = 0;  // warp-uniform queue length
  // Drain: every lane takes one queued row; the five value columns are visited in a uniform loop
  // (the column is a compile-time constant inside), lanes whose row is not interesting in that
  // column sit the step out.
  auto drain = [&]() {
    __syncwarp();
    for (int k0 = 0; k0 < qn; k0 += 32) {
      const int k = k0 + lane;
      const bool have = k < qn;
      const int64_t qr = have ? Q.row[warp][k] : 0;
      const int64_t r = qr & ((1ll << 62) - 1);
      const bool cold = (qr >> 62) & 1;               // the record carries the row's cold flag
      const uint32_t hm = have ? Q.hitmask[warp][k] : 0u;
      // every value the row needs is requested before the first is used (one memory latency a batch)
      double la = 0.0, tf = 0.0, tl = 0.0;
      long long ct = 1;
      if (hm & 0x01u) la = log.latency_ms[r];
      if (hm & 0x1Eu) {            // ttfb / tllt feed four of the five columns
        tf = log.ttfb_ms[r];
        tl = log.tllt_ms[r];
      }
      if (hm & 0x10u) ct = log.completion_tokens[r];
#pragma unroll 1
      for (int c = 0; c < KVM_LOG_NCOLS; ++c) {
        const bool valid = (hm >> c) & 1u;
        if (!__any_sync(KVM_FULL, valid)) continue;
        double v = 0.0;
        int set = 0;
        if (valid) {
          if (c == KVM_COL_LAT) {
            v = la;
            set = cold ? KVM_SET_LAT_COLD : KVM_SET_LAT_WARM;
          } else if (c == KVM_COL_TTFB) {
            v = tf;
            set = cold ? KVM_SET_TTFB_COLD : KVM_SET_TTFB_WARM;
          } else if (c == KVM_COL_TLLT) {
            v = tl;
            set = KVM_SET_TLLT;
          } else if (c == KVM_COL_GEN) {
            v = tl - tf;                                        // analyze.py:144
            set = KVM_SET_GEN;
          } else {
            v = (tl - tf) / (double)ct;                         // analyze.py:146
            set = KVM_SET_PERTOK;
          }
        }
        const uint64_t key = kvm_key(v);
        uint32_t e = 0x00ffu;
        if (valid) e = S.tab[c][digit_of(key, S.cp.kbase[c], S.cp.shift[c])];
        target_hits_warp<PHASE>(ws, S, arena, valid, set, c, e, key);
      }
    }
    __syncwarp();
    qn = 0;
  };
  // Streaming side: five table lookups per row OR-ed into a hit mask, ONE ballot per row.
  auto row = [&](int64_t r, uint64_t rec) {
    // branch-free: the lookup is always performed (index masked into the table), a digit of
    // 0xFFF ("row not in a set of this column") is rejected arithmetically
    uint32_t hm = 0;
#pragma unroll
    for (int c = 0; c < KVM_LOG_NCOLS; ++c) {
      const uint32_t d = (uint32_t)(rec >> (12 * c)) & 0xFFFu;
      const uint32_t e = S.tab[c][d & (BINS - 1)];
      const uint32_t none = PHASE == 1 ? ((e & 0xffu) == 0xffu) : (e == 0x00ffu);
      hm |= (uint32_t)((d != 0xFFFu) & !none) << c;
    }
    const unsigned hits = __ballot_sync(KVM_FULL, hm != 0);
    if (hits) {
      const int cnt = __popc(hits);
      if (qn + cnt > CQ_CAP) drain();
      if (hm) {
        const int pos = qn + __popc(hits & lt);
        Q.row[warp][pos] = r | (int64_t)((rec >> REC_COLD_BIT) & 1ull) << 62;
        Q.hitmask[warp][pos] = hm;
      }
      qn += cnt;
    }
  };
  const int64_t n = log.n;
  const int64_t nunits = ROWS == 2 ? (n >> 1) : n;
  // REC_K units per lane and iteration, all loads issued before the first is used (a warp's load
  // covers 512 consecutive bytes); a warp takes 32 * REC_K consecutive units
  constexpr int REC_K = 4;
  for (int64_t base = ((int64_t)blockIdx.x * (SEL_THREADS / 32) + warp) * (32 * REC_K); base < nunits;
       base += (int64_t)gridDim.x * SEL_THREADS * REC_K) {
    uint64_t r0[REC_K], r1[REC_K];
#pragma unroll
    for (int k = 0; k < REC_K; ++k) {
      const int64_t u = base + k * 32 + lane;
      r0[k] = DIGITS_NONE;
      r1[k] = DIGITS_NONE;
      if (u < nunits) {
        if (ROWS == 2) {
          const longlong2 rr = kvm_ld2(reinterpret_cast<const int64_t*>(recs) + (u << 1));
          r0[k] = (uint64_t)rr.x;
          r1[k] = (uint64_t)rr.y;
        } else {
          r0[k] = kvm_ld1(recs + u);
        }
      }
    }
#pragma unroll
    for (int k = 0; k < REC_K; ++k) {
      const int64_t u = base + k * 32 + lane;
      const int64_t i = ROWS == 2 ? (u << 1) : u;
      if (!__any_sync(KVM_FULL, (r0[k] & r1[k] & DIGITS_NONE) != DIGITS_NONE)) continue;
      row(i, r0[k]);
      if (ROWS == 2) row(i + 1, r1[k]);
    }
  }
  drain();
  if (ROWS == 2 && (n & 1) && blockIdx.x == 0 && threadIdx.x < 32) {  // odd tail row, one warp
    const int64_t i = n - 1;
    row(i, lane == 0 ? recs[i] : DIGITS_NONE);
    drain();
  }
  if (PHASE == 1) return;
  __syncthreads();
  // flush the stages: one global cursor bump per non-empty target, then a cooperative copy
  if (threadIdx.x < MAXQ) {
    const uint32_t c = S.stage_n[threadIdx.x] < (uint32_t)STAGE_SLOTS ? S.stage_n[threadIdx.x] : (uint32_t)STAGE_SLOTS;
    S.stage_n[threadIdx.x] = c;
    if (c) S.stage_pos[threadIdx.x] =
        atomicAdd(reinterpret_cast<unsigned long long*>(&ws->bucket_counts[threadIdx.x]), (unsigned long long)c);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < MAXQ * STAGE_SLOTS; i += SEL_THREADS) {
    const int j = i / STAGE_SLOTS, k = i % STAGE_SLOTS;
    if ((uint32_t)k < S.stage_n[j]) {
      const int64_t pos = (int64_t)S.stage_pos[j] + k;
      if (pos < S.tcount[j]) arena[S.bk_off[j] + pos] = S.stage[j][k];
    }
  }
  for (int i = threadIdx.x; i < NH * NB_MAX; i += SEL_THREADS) {
    const uint32_t c = (&S.edge[0][0])[i];
    if (c) atomicAdd(&ws->edge20[0][0] + i, (unsigned long long)c);
  }
}

__global__ void __launch_bounds__(SEL_THREADS, 2)
column_compact_kernel(const double* __restrict__ values, int64_t n, int mode,
                      const void* __restrict__ member, bool vec, SelWs* ws, uint64_t* arena) {
  compact_body<1>(ws, arena, [&](auto&& f) { for_each_column_key(values, n, mode, member, vec, f); });
}

// ------------------------------------------------------------------ finish
constexpr int FIN_THREADS = 512;
constexpr int FIN_BITS = 11;
constexpr int FIN_BINS = 1 << FIN_BITS;
constexpr int FIN_SMALL = 128;   // candidates ranked by counting once this few remain (counting is quadratic)
constexpr int FIN_SURV = 2048;   // candidates gathered into a list of their own once this few remain
constexpr int HIST_CTAS = 16;    // extra CTAs that assemble the value histograms
constexpr int FIN_CACHE = 14336;  // bucket keys kept in shared memory (112 KB) by the finish: fits beside one CTA of the one-pass kernel

// CTAs 0..MAXQ-1: one query each - radix select inside the query's bucket, FIN_BITS per pass,
// switching to rank-by-counting in shared memory once <= FIN_SMALL candidates remain.
// CTAs MAXQ..: assemble the value histograms (edge counters + fine bins without an edge).
__global__ void __launch_bounds__(FIN_THREADS, 2)
select_finish_kernel(SelWs* ws, SrcPack src, const int64_t* __restrict__ src_counts,
                     kvm_query_result* __restrict__ results, int64_t* __restrict__ hist_out) {
  __shared__ uint32_t hist[FIN_BINS];
  __shared__ uint32_t chunk_sum[FIN_THREADS];
  __shared__ uint32_t sel_digit, sel_count, small_n;
  __shared__ int64_t sel_below;
  __shared__ uint64_t answer;

  if (blockIdx.x >= MAXQ) {
    // ---- value histograms (hist_out is zero-initialised by the caller): every CTA takes a
    // slice of the fine bins; bins without an edge go wholesale into the bin of their lowest key
    const kvm_hist_spec hs = ws->hspec;
    if (hist_out == nullptr || hs.nhist == 0 || ws->status != KVM_OK) return;
    const int nbins = hs.num_bins;
    const int part = blockIdx.x - MAXQ;
    unsigned long long* out = reinterpret_cast<unsigned long long*>(hist_out);
    const int nsets = ws->info.nsets;
    for (int t = part * FIN_THREADS + threadIdx.x; t < hs.nhist * BINS; t += HIST_CTAS * FIN_THREADS) {
      const int h = t / BINS, d = t % BINS;
      if (!ws->active20[h] || ((ws->flag20[h][d >> 5] >> (d & 31)) & 1u)) continue;
      int64_t c = 0;
      for (int b = 0; b < nsets; ++b)
        if ((hs.set_mask[h] >> b) & 1u) c += (int64_t)ws->hist[b][d];
      if (c) {
        const ColParam cp = ws->col[hs.col[h]];
        const uint64_t lowkey = cp.kbase + ((uint64_t)d << cp.shift);
        atomicAdd(out + h * nbins + bin_of(lowkey, ws->thr20[h], nbins), (unsigned long long)c);
      }
    }
    if (part == 0)
      for (int t = threadIdx.x; t < hs.nhist * nbins; t += FIN_THREADS) {
        const int h = t / nbins, b = t % nbins;
        const unsigned long long e = ws->edge20[h][b];
        if (e) atomicAdd(out + t, e);
      }
    return;
  }

  const int q = blockIdx.x;
  if (q >= ws->nq) return;
  const QState s = ws->q[q];
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  if (ws->status != KVM_OK || s.status != KVM_Q_OK || s.resolved) {
    if (threadIdx.x == 0) {
      kvm_query_result r;
      r.value = (ws->status == KVM_OK && s.status == KVM_Q_OK) ? kvm_unkey(s.key) : nan;
      r.n = s.n;
      r.rank = s.rank;
      // a failed select (arena overflow) is visible in every result, for callers that hold no
      // workspace pointer to read the status word from (kvm_column_select_static)
      r.status = ws->status != KVM_OK ? KVM_Q_FAILED : s.status;
      r.pad = 0;
      results[q] = r;
    }
    return;
  }
  const Bucket b = ws->bucket[s.bucket];
  const int64_t off = ws->bucket_offsets[s.bucket];
  // The bucket's keys (all sources) are brought into shared memory once - every load of a thread
  // in flight together - when they fit; the narrowing passes below then never go back to global
  // memory (each pass used to cost several dependent round trips to L2).
  extern __shared__ uint64_t kcache[];   // [FIN_CACHE]
  __shared__ int cached_n;
  {
    int64_t tot = 0;
    for (int sidx = 0; sidx < src.nsrc; ++sidx) {
      const int64_t c = src_counts[(size_t)sidx * MAXQ + s.bucket];
      tot += c < b.count ? c : b.count;
    }
    if (threadIdx.x == 0) cached_n = tot <= FIN_CACHE ? (int)tot : -1;
    if (tot <= FIN_CACHE) {
      int64_t base = 0;
      for (int sidx = 0; sidx < src.nsrc; ++sidx) {
        int64_t cnt = src_counts[(size_t)sidx * MAXQ + s.bucket];
        cnt = cnt < b.count ? cnt : b.count;
        const uint64_t* __restrict__ keys = src.arena[sidx] + off;
        for (int64_t i = threadIdx.x; i < cnt; i += FIN_THREADS) kcache[base + i] = keys[i];
        base += cnt;
      }
    }
    __syncthreads();
  }
  const int ncache = cached_n;
  int rem = b.shift;        // unresolved low bits
  uint64_t prefix = 0;      // resolved part of (key - b.base) >> rem
  int64_t resid = s.resid;
  int64_t live = b.count;   // candidates matching the prefix
  bool done = false;
  // Once few enough candidates match the prefix they are gathered into a list of their own (two
  // lists, used in turn): every later pass touches the survivors only, and a key repeated many
  // times costs its count per pass instead of the whole bucket.
  // (list 0 is an array of its own, list 1 reuses the start of the key cache, which is dead once
  // list 0 has been gathered from it)
  __shared__ uint64_t surv0[FIN_SURV];
  uint64_t* const surv[2] = {surv0, kcache};
  int cur = -1;             // list holding exactly the candidates that match the prefix (-1: none yet)
  int curn = 0;

  while (rem > 0 && !done) {
    if (live <= FIN_SURV && (cur < 0 || live < curn)) {
      const int nxt = cur < 0 ? 0 : cur ^ 1;
      if (threadIdx.x == 0) small_n = 0;
      __syncthreads();
      auto take = [&](uint64_t k) {
        if (shr64(k - b.base, rem) == prefix) {
          const uint32_t p = atomicAdd(&small_n, 1u);
          if (p < (uint32_t)FIN_SURV) surv[nxt][p] = k;
        }
      };
      if (cur >= 0) {
        for (int i = threadIdx.x; i < curn; i += FIN_THREADS) take(surv[cur][i]);
      } else if (ncache >= 0) {
        for (int i = threadIdx.x; i < ncache; i += FIN_THREADS) take(kcache[i]);
      } else {
        for (int sidx = 0; sidx < src.nsrc; ++sidx) {
          const int64_t cnt = src_counts[(size_t)sidx * MAXQ + s.bucket];
          const uint64_t* __restrict__ keys = src.arena[sidx] + off;
          for (int64_t i = threadIdx.x; i < cnt; i += FIN_THREADS) take(keys[i]);
        }
      }
      __syncthreads();
      cur = nxt;
      curn = (int)(small_n < (uint32_t)FIN_SURV ? small_n : (uint32_t)FIN_SURV);
      __syncthreads();
    }
    if (cur >= 0 && curn <= FIN_SMALL) {
      // rank by counting among the survivors
      const uint32_t m = (uint32_t)curn;
      for (uint32_t i = threadIdx.x; i < m; i += FIN_THREADS) {
        const uint64_t k = surv[cur][i];
        uint32_t less = 0, leq = 0;
        for (uint32_t j = 0; j < m; ++j) {
          const uint64_t o = surv[cur][j];
          less += o < k;
          leq += o <= k;
        }
        if ((int64_t)less <= resid && resid < (int64_t)leq) answer = k;  // all writers agree
      }
      __syncthreads();
      done = true;
      break;
    }
    const int take_bits = rem < FIN_BITS ? rem : FIN_BITS;
    const int sh = rem - take_bits;
    const uint32_t dmask = (1u << take_bits) - 1u;
    for (int i = threadIdx.x; i < FIN_BINS; i += FIN_THREADS) hist[i] = 0;
    __syncthreads();
    if (cur >= 0) {
      for (int i = threadIdx.x; i < curn; i += FIN_THREADS)
        atomicAdd(&hist[(uint32_t)((surv[cur][i] - b.base) >> sh) & dmask], 1u);
    } else if (ncache >= 0) {
      for (int i = threadIdx.x; i < ncache; i += FIN_THREADS) {
        const uint64_t low = kcache[i] - b.base;
        if (shr64(low, rem) == prefix) atomicAdd(&hist[(uint32_t)(low >> sh) & dmask], 1u);
      }
    } else
    for (int sidx = 0; sidx < src.nsrc; ++sidx) {
      const int64_t cnt = src_counts[(size_t)sidx * MAXQ + s.bucket];
      const uint64_t* __restrict__ keys = src.arena[sidx] + off;
      for (int64_t i0 = threadIdx.x; i0 < cnt; i0 += 4 * FIN_THREADS) {
        uint64_t kk[4];  // four independent loads in flight before the shared-memory atomics
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int64_t i = i0 + (int64_t)u * FIN_THREADS;
          kk[u] = i < cnt ? keys[i] : b.base + ~0ull;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int64_t i = i0 + (int64_t)u * FIN_THREADS;
          const uint64_t low = kk[u] - b.base;
          if (i < cnt && shr64(low, rem) == prefix) atomicAdd(&hist[(uint32_t)(low >> sh) & dmask], 1u);
        }
      }
    }
    __syncthreads();
    // locate the digit holding `resid`: each thread owns FIN_BINS/FIN_THREADS consecutive bins
    constexpr int PER = FIN_BINS / FIN_THREADS;
    uint32_t local = 0;
#pragma unroll
    for (int k = 0; k < PER; ++k) local += hist[threadIdx.x * PER + k];
    chunk_sum[threadIdx.x] = local;
    __syncthreads();
    if (threadIdx.x < 32) {
      constexpr int CH = FIN_THREADS / 32;
      uint32_t mine = 0;
      for (int k = 0; k < CH; ++k) mine += chunk_sum[threadIdx.x * CH + k];
      uint32_t incl = mine;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        uint32_t o = __shfl_up_sync(KVM_FULL, incl, d);
        if ((threadIdx.x & 31) >= d) incl += o;
      }
      const unsigned hit = __ballot_sync(KVM_FULL, (int64_t)incl > resid);
      if (hit && (int)(__ffs(hit) - 1) == (int)threadIdx.x) {
        int64_t below = (int64_t)(incl - mine);
        int bin = threadIdx.x * CH * PER;
        for (;; ++bin) {
          const int64_t c = hist[bin];
          if (below + c > resid) break;
          below += c;
        }
        sel_digit = (uint32_t)bin;
        sel_below = below;
        sel_count = hist[bin];
      }
      if (!hit && threadIdx.x == 0) {  // cannot happen with consistent counts
        sel_digit = 0;
        sel_below = 0;
        sel_count = 0;
      }
    }
    __syncthreads();
    prefix = (prefix << take_bits) | sel_digit;
    resid -= sel_below;
    live = sel_count;
    rem = sh;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    kvm_query_result r;
    r.value = kvm_unkey(done ? answer : b.base + prefix);
    r.n = s.n;
    r.rank = s.rank;
    r.status = KVM_Q_OK;
    r.pad = 0;
    results[q] = r;
  }
}

// ------------------------------------------------------------------ host helpers
static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

static bool log_vec_ok(const kvm_request_log* log, const uint8_t* flags) {
  return aligned16(log->latency_ms) && aligned16(log->ttfb_ms) && aligned16(log->tllt_ms) &&
         aligned16(log->completion_tokens) && (reinterpret_cast<uintptr_t>(flags) & 1u) == 0;
}

static bool column_vec_ok(const double* v, int mode, const void* member) {
  return aligned16(v) && (mode != KVM_MEMBER_STATUS || aligned16(member));
}

static int ensure_smem(const void* fn, size_t bytes) {
  if (bytes > 48 * 1024) {
    KVM_CUDA_CHECK(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  }
  return KVM_OK;
}

static uint64_t* arena_of(void* workspace) {
  return reinterpret_cast<uint64_t*>(reinterpret_cast<unsigned char*>(workspace) + WS_HEADER);
}

static const kvm_hist_spec kNoHists = {0, 0, {0, 0, 0, 0}, {0, 0, 0, 0}};

static kvm_hist_spec log_hist_spec(int32_t num_bins) {
  kvm_hist_spec hs;
  hs.nhist = 4;
  hs.num_bins = num_bins;
  hs.col[0] = KVM_COL_TTFB;
  hs.col[1] = KVM_COL_TLLT;
  hs.col[2] = KVM_COL_GEN;
  hs.col[3] = KVM_COL_PERTOK;
  hs.set_mask[0] = (1u << KVM_SET_TTFB_COLD) | (1u << KVM_SET_TTFB_WARM);
  hs.set_mask[1] = 1u << KVM_SET_TLLT;
  hs.set_mask[2] = 1u << KVM_SET_GEN;
  hs.set_mask[3] = 1u << KVM_SET_PERTOK;
  return hs;
}

}  // namespace

extern "C" size_t kvm_select_workspace_bytes(int64_t arena_capacity) {
  if (arena_capacity < 0) arena_capacity = 0;
  return WS_HEADER + kvm_align_up((size_t)arena_capacity * sizeof(uint64_t), 256) + 256;
}

extern "C" int64_t kvm_select_default_arena(int64_t n) {
  // buckets hold ~1/2048 of the populated key range; n/16 + 1M leaves a wide margin for
  // up to KVM_MAX_QUERIES buckets of concentrated data.
  return n / 16 + (1 << 20);
}

extern "C" int32_t kvm_select_bins(void) { return BINS; }
extern "C" uint64_t* kvm_select_hist_ptr(void* ws) { return &reinterpret_cast<SelWs*>(ws)->hist[0][0]; }
extern "C" uint64_t* kvm_select_edge_ptr(void* ws) {
  return reinterpret_cast<uint64_t*>(&reinterpret_cast<SelWs*>(ws)->edge20[0][0]);
}
extern "C" uint64_t* kvm_select_arena_ptr(void* ws) { return arena_of(ws); }
extern "C" int64_t* kvm_select_bucket_counts_ptr(void* ws) { return reinterpret_cast<SelWs*>(ws)->bucket_counts; }
extern "C" int64_t* kvm_select_bucket_offsets_ptr(void* ws) { return reinterpret_cast<SelWs*>(ws)->bucket_offsets; }
extern "C" kvm_set_info* kvm_select_info_ptr(void* ws) { return &reinterpret_cast<SelWs*>(ws)->info; }

extern "C" int kvm_select_plan(const kvm_set_info* info, const kvm_query* queries, int32_t nq,
                               const kvm_hist_spec* hists, void* workspace, size_t workspace_bytes,
                               int64_t arena_capacity, void* stream) {
  return kvm_select_plan_refine(info, queries, nq, hists, workspace, workspace_bytes, arena_capacity, 0, 0, stream);
}

extern "C" int kvm_select_plan_refine(const kvm_set_info* info, const kvm_query* queries, int32_t nq,
                                      const kvm_hist_spec* hists, void* workspace, size_t workspace_bytes,
                                      int64_t arena_capacity, int32_t refine_rounds, int64_t refine_min_keys,
                                      void* stream) {
  return kvm_select_plan_internal(info, queries, nq, hists, workspace, workspace_bytes, arena_capacity,
                                  refine_rounds, refine_min_keys, 0, stream);
}

// preset != 0: the digit parameters (ws->col) and the histograms are already in the workspace -
// written by the fused pass (csrc/kvm_fused.cu) from a sample; the plan leaves both alone.
int kvm_select_plan_internal(const kvm_set_info* info, const kvm_query* queries, int32_t nq,
                             const kvm_hist_spec* hists, void* workspace, size_t workspace_bytes,
                             int64_t arena_capacity, int32_t refine_rounds, int64_t refine_min_keys,
                             int32_t preset, void* stream) {
  KVM_REQUIRE(info && queries && workspace, "kvm_select_plan: null argument");
  KVM_REQUIRE(refine_rounds >= 0 && refine_rounds <= 4 && refine_min_keys >= 0,
              "kvm_select_plan: refine_rounds=%d outside [0,4] or negative refine_min_keys", refine_rounds);
  KVM_REQUIRE(nq >= 1 && nq <= MAXQ, "kvm_select_plan: nq=%d outside [1,%d]", nq, MAXQ);
  const kvm_hist_spec hs = hists ? *hists : kNoHists;
  KVM_REQUIRE(hs.nhist >= 0 && hs.nhist <= NH, "kvm_select_plan: nhist=%d outside [0,%d]", hs.nhist, NH);
  KVM_REQUIRE(hs.nhist == 0 || (hs.num_bins >= 1 && hs.num_bins <= NB_MAX),
              "kvm_select_plan: num_bins=%d outside [1,%d]", hs.num_bins, NB_MAX);
  if (workspace_bytes < kvm_select_workspace_bytes(arena_capacity)) {
    kvm_set_error("kvm_select_plan: workspace %zu < %zu bytes", workspace_bytes,
                  kvm_select_workspace_bytes(arena_capacity));
    return KVM_ERR_WORKSPACE;
  }
  SelWs* ws = reinterpret_cast<SelWs*>(workspace);
  QueryPack qp;
  for (int i = 0; i < MAXQ; ++i) qp.q[i] = i < nq ? queries[i] : kvm_query{0, 0, 0.0, 0};
  qp.nq = nq;
  select_plan_kernel<<<1, PLAN_THREADS, 0, (cudaStream_t)stream>>>(ws, info, qp, hs, arena_capacity,
                                                                  refine_rounds, refine_min_keys, preset ? 1 : 0);
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_log_select_hist(const kvm_request_log* log, const uint8_t* flags,
                                   const kvm_log_stats* stats, const kvm_log_extras* extras,
                                   void* workspace, void* stream) {
  KVM_REQUIRE(log && flags && workspace, "kvm_log_select_hist: null argument");
  const bool series = extras && extras->series_window_ms > 0.0;
  if (series)
    KVM_REQUIRE(stats && extras->series && extras->series_dropped && extras->series_max_bins >= 1,
                "kvm_log_select_hist: series requested without stats / buffers");
  cudaStream_t st = (cudaStream_t)stream;
  SeriesArgs sa{};
  if (series) {
    sa.start_ms = log->start_ms;
    sa.stats = stats;
    sa.window_ms = extras->series_window_ms;
    sa.max_bins = extras->series_max_bins;
    sa.series = reinterpret_cast<unsigned long long*>(extras->series);
    sa.dropped = reinterpret_cast<unsigned long long*>(extras->series_dropped);
    KVM_CUDA_CHECK(cudaMemsetAsync(extras->series, 0, sizeof(int64_t) * 3 * (size_t)extras->series_max_bins, st));
    KVM_CUDA_CHECK(cudaMemsetAsync(extras->series_dropped, 0, sizeof(int64_t), st));
  }
  const size_t smem = (size_t)KVM_LOG_NSETS * BINS * sizeof(uint32_t);
  uint64_t* digits = extras ? reinterpret_cast<uint64_t*>(extras->digit_records) : nullptr;
  const bool vec = log_vec_ok(log, flags) && (!series || aligned16(log->start_ms)) &&
                   (digits == nullptr || aligned16(digits));
  const int grid = kvm_grid_for(log->n, SEL_THREADS * 2 * 4, 2);
  SelWs* ws = reinterpret_cast<SelWs*>(workspace);
#define KVM_LAUNCH_HIST(R, S)                                                    \
  do {                                                                           \
    int rc = ensure_smem((const void*)log_hist_kernel<R, S>, smem);              \
    if (rc) return rc;                                                           \
    log_hist_kernel<R, S><<<grid, SEL_THREADS, smem, st>>>(*log, flags, ws, sa, digits); \
  } while (0)
  if (vec && series) KVM_LAUNCH_HIST(2, true);
  else if (vec) KVM_LAUNCH_HIST(2, false);
  else if (series) KVM_LAUNCH_HIST(1, true);
  else KVM_LAUNCH_HIST(1, false);
#undef KVM_LAUNCH_HIST
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_column_select_hist(const double* values, int64_t n, int32_t mode,
                                      const void* member, void* workspace, void* stream) {
  KVM_REQUIRE(workspace && (values || n == 0), "kvm_column_select_hist: null argument");
  KVM_REQUIRE(mode == KVM_MEMBER_ALL || member, "kvm_column_select_hist: member array is null");
  const int nsets = mode == KVM_MEMBER_SET_IDS ? KVM_MAX_SETS : 1;
  const size_t smem = (size_t)nsets * BINS * sizeof(uint32_t);
  int rc = ensure_smem((const void*)column_hist_kernel, smem);
  if (rc) return rc;
  const int grid = kvm_grid_for(n, SEL_THREADS * 2 * 4, 2);
  column_hist_kernel<<<grid, SEL_THREADS, smem, (cudaStream_t)stream>>>(
      values, n, mode, member, column_vec_ok(values, mode, member), nsets,
      reinterpret_cast<SelWs*>(workspace));
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_select_resolve(void* workspace, void* stream) {
  KVM_REQUIRE(workspace, "kvm_select_resolve: null workspace");
  const size_t smem = (size_t)KVM_MAX_SETS * BINS * sizeof(uint64_t);
  int rc = ensure_smem((const void*)select_resolve_kernel, smem);
  if (rc) return rc;
  select_resolve_kernel<<<1, RES_THREADS, smem, (cudaStream_t)stream>>>(reinterpret_cast<SelWs*>(workspace));
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

namespace {
template <int PHASE>
int launch_log_rec(const kvm_request_log* log, const uint8_t* flags, const uint64_t* digit_records,
                   void* workspace, cudaStream_t st) {
  const size_t smem = ((sizeof(CompactSmem) + 15) / 16) * 16 + sizeof(RecQueue);
  const int grid = kvm_stream_grid(log->n, SEL_THREADS * 2 * 8);
  int rc;
  if (aligned16(digit_records) && (reinterpret_cast<uintptr_t>(flags) & 1u) == 0) {
    if ((rc = ensure_smem((const void*)log_rec_kernel<2, PHASE>, smem))) return rc;
    log_rec_kernel<2, PHASE><<<grid, SEL_THREADS, smem, st>>>(
        *log, flags, digit_records, reinterpret_cast<SelWs*>(workspace), arena_of(workspace));
  } else {
    if ((rc = ensure_smem((const void*)log_rec_kernel<1, PHASE>, smem))) return rc;
    log_rec_kernel<1, PHASE><<<grid, SEL_THREADS, smem, st>>>(
        *log, flags, digit_records, reinterpret_cast<SelWs*>(workspace), arena_of(workspace));
  }
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}
}  // namespace

extern "C" uint64_t* kvm_select_hist2_ptr(void* ws) { return &reinterpret_cast<SelWs*>(ws)->hist2[0][0]; }
extern "C" int64_t kvm_select_hist2_words(void) { return (int64_t)MAXQ * BINS; }

extern "C" int kvm_log_select_refine_hist(const kvm_request_log* log, const uint8_t* flags,
                                          const uint64_t* digit_records, void* workspace, void* stream) {
  KVM_REQUIRE(log && flags && workspace && digit_records,
              "kvm_log_select_refine_hist: null argument (a refine round needs the digit records)");
  SelWs* ws = reinterpret_cast<SelWs*>(workspace);
  KVM_CUDA_CHECK(cudaMemsetAsync(&ws->hist2[0][0], 0, sizeof(ws->hist2), (cudaStream_t)stream));
  return launch_log_rec<1>(log, flags, digit_records, workspace, (cudaStream_t)stream);
}

extern "C" int kvm_select_refine(void* workspace, void* stream) {
  KVM_REQUIRE(workspace, "kvm_select_refine: null workspace");
  select_refine_kernel<<<1, REF_THREADS, 0, (cudaStream_t)stream>>>(reinterpret_cast<SelWs*>(workspace));
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_log_select_compact(const kvm_request_log* log, const uint8_t* flags,
                                      const uint64_t* digit_records, void* workspace, void* stream) {
  KVM_REQUIRE(log && flags && workspace, "kvm_log_select_compact: null argument");
  if (digit_records) return launch_log_rec<0>(log, flags, digit_records, workspace, (cudaStream_t)stream);
  const size_t smem = ((sizeof(CompactSmem) + 15) / 16) * 16 + sizeof(CompactQueue);
  const int grid = kvm_grid_for(log->n, SEL_THREADS * 2 * 4, 2);
  int rc;
  if (log_vec_ok(log, flags)) {
    if ((rc = ensure_smem((const void*)log_compact_kernel<2>, smem))) return rc;
    log_compact_kernel<2><<<grid, SEL_THREADS, smem, (cudaStream_t)stream>>>(
        *log, flags, reinterpret_cast<SelWs*>(workspace), arena_of(workspace));
  } else {
    if ((rc = ensure_smem((const void*)log_compact_kernel<1>, smem))) return rc;
    log_compact_kernel<1><<<grid, SEL_THREADS, smem, (cudaStream_t)stream>>>(
        *log, flags, reinterpret_cast<SelWs*>(workspace), arena_of(workspace));
  }
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_column_select_compact(const double* values, int64_t n, int32_t mode,
                                         const void* member, void* workspace, void* stream) {
  KVM_REQUIRE(workspace && (values || n == 0), "kvm_column_select_compact: null argument");
  int rc = ensure_smem((const void*)column_compact_kernel, sizeof(CompactSmem));
  if (rc) return rc;
  const int grid = kvm_grid_for(n, SEL_THREADS * 2 * 4, 2);
  column_compact_kernel<<<grid, SEL_THREADS, sizeof(CompactSmem), (cudaStream_t)stream>>>(
      values, n, mode, member, column_vec_ok(values, mode, member),
      reinterpret_cast<SelWs*>(workspace), arena_of(workspace));
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_select_finish(void* workspace, const uint64_t* const* arenas, int32_t nsrc,
                                 const int64_t* src_counts, kvm_query_result* results,
                                 int64_t* hist_counts, void* stream) {
  KVM_REQUIRE(workspace && results, "kvm_select_finish: null argument");
  KVM_REQUIRE(nsrc >= 0 && nsrc <= MAX_SRC, "kvm_select_finish: nsrc=%d outside [0,%d]", nsrc, MAX_SRC);
  SelWs* ws = reinterpret_cast<SelWs*>(workspace);
  SrcPack sp;
  for (int i = 0; i < MAX_SRC; ++i) sp.arena[i] = nullptr;
  if (nsrc == 0 || arenas == nullptr) {
    sp.nsrc = 1;
    sp.arena[0] = arena_of(workspace);
    src_counts = ws->bucket_counts;
  } else {
    KVM_REQUIRE(src_counts, "kvm_select_finish: src_counts is null");
    sp.nsrc = nsrc;
    for (int i = 0; i < nsrc; ++i) sp.arena[i] = arenas[i];
  }
  if (hist_counts)  // the assembling CTAs accumulate into it
    KVM_CUDA_CHECK(cudaMemsetAsync(hist_counts, 0, sizeof(int64_t) * KVM_MAX_HISTS * KVM_HIST_MAX_BINS,
                                   (cudaStream_t)stream));
  {
    const int rc = ensure_smem((const void*)select_finish_kernel, sizeof(uint64_t) * FIN_CACHE);
    if (rc) return rc;
  }
  select_finish_kernel<<<MAXQ + HIST_CTAS, FIN_THREADS, sizeof(uint64_t) * FIN_CACHE, (cudaStream_t)stream>>>(
      ws, sp, src_counts, results, hist_counts);
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_select_status(const void* workspace, void* stream) {
  KVM_REQUIRE(workspace, "kvm_select_status: null workspace");
  int32_t status = 0;
  KVM_CUDA_CHECK(cudaMemcpyAsync(&status, workspace, sizeof(status), cudaMemcpyDeviceToHost,
                                 (cudaStream_t)stream));
  KVM_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
  if (status == KVM_ERR_ARENA) kvm_set_error("select: candidate arena too small; retry with a larger arena_capacity");
  return status;
}

extern "C" int kvm_log_select(const kvm_request_log* log, const uint8_t* flags,
                              const kvm_log_stats* stats, const kvm_query* queries, int32_t nq,
                              kvm_query_result* results, const kvm_log_extras* extras,
                              void* workspace, size_t workspace_bytes, int64_t arena_capacity,
                              void* stream) {
  KVM_REQUIRE(log && flags && stats && results, "kvm_log_select: null argument");
  const bool hists = extras && extras->num_bins > 0;
  KVM_REQUIRE(!hists || extras->hist_counts, "kvm_log_select: hist_counts is null");
  const kvm_hist_spec hs = hists ? log_hist_spec(extras->num_bins) : kNoHists;
  const uint64_t* recs = extras ? reinterpret_cast<const uint64_t*>(extras->digit_records) : nullptr;
  // large logs: one refine round keeps the candidate buckets (and the finish) small
  const int rounds = (recs && log->n >= KVM_REFINE_MIN_ROWS) ? 1 : 0;
  int rc = kvm_select_plan_refine(&stats->sets, queries, nq, &hs, workspace, workspace_bytes, arena_capacity,
                                  rounds, KVM_REFINE_MIN_KEYS, stream);
  if (rc) return rc;
  if ((rc = kvm_log_select_hist(log, flags, stats, extras, workspace, stream))) return rc;
  if ((rc = kvm_select_resolve(workspace, stream))) return rc;
  for (int r = 0; r < rounds; ++r) {
    if ((rc = kvm_log_select_refine_hist(log, flags, recs, workspace, stream))) return rc;
    if ((rc = kvm_select_refine(workspace, stream))) return rc;
  }
  if ((rc = kvm_log_select_compact(log, flags, recs, workspace, stream))) return rc;
  return kvm_select_finish(workspace, nullptr, 0, nullptr, results, hists ? extras->hist_counts : nullptr, stream);
}

extern "C" int kvm_column_select(const double* values, int64_t n, int32_t mode, const void* member,
                                 int32_t nsets, const kvm_query* queries, int32_t nq,
                                 kvm_query_result* results, void* workspace, size_t workspace_bytes,
                                 int64_t arena_capacity, void* stream) {
  KVM_REQUIRE(results && workspace, "kvm_column_select: null argument");
  KVM_REQUIRE(mode != KVM_MEMBER_SET_IDS ? nsets == 1 : (nsets >= 1 && nsets <= KVM_MAX_SETS),
              "kvm_column_select: nsets=%d invalid for member mode %d", nsets, mode);
  if (workspace_bytes < kvm_select_workspace_bytes(arena_capacity)) {
    kvm_set_error("kvm_column_select: workspace %zu < %zu bytes", workspace_bytes,
                  kvm_select_workspace_bytes(arena_capacity));
    return KVM_ERR_WORKSPACE;
  }
  SelWs* ws = reinterpret_cast<SelWs*>(workspace);
  // the stats pass writes straight into the workspace's info block
  int rc = kvm_column_stats(values, n, mode, member, nsets, &ws->info, ws->stats_scratch,
                            sizeof(ws->stats_scratch), stream);
  if (rc) return rc;
  if ((rc = kvm_select_plan(&ws->info, queries, nq, nullptr, workspace, workspace_bytes, arena_capacity, stream))) return rc;
  if ((rc = kvm_column_select_hist(values, n, mode, member, workspace, stream))) return rc;
  if ((rc = kvm_select_resolve(workspace, stream))) return rc;
  if ((rc = kvm_column_select_compact(values, n, mode, member, workspace, stream))) return rc;
  return kvm_select_finish(workspace, nullptr, 0, nullptr, results, nullptr, stream);
}

// ------------------------------------------------------------------ allocation-free variant
// kvm_column_select_static: the same select as kvm_column_select, for callers that have no
// device memory of their own to offer as a workspace.  Typical case: a latency monitor
// that shares the GPU with an inference server which has already claimed the HBM pool for
// weights and KV cache - a cudaMalloc from the monitor either fails or takes memory the
// server counted on - and host processes whose policy forbids runtime allocation once
// serving has started.  The scratch lives in the array below: it is part of the library
// image, so every GPU that loads libkvmini_b200.so holds these bytes until the process exits.
namespace {

constexpr int64_t STATIC_MAX_N = 1 << 20;  // values per call
constexpr size_t STATIC_SCRATCH_BYTES = WS_HEADER + (size_t)STATIC_MAX_N * sizeof(uint64_t) + 256;
constexpr int STATIC_MAX_DEVICES = 64;

static __device__ __align__(256) unsigned char g_select_scratch[STATIC_SCRATCH_BYTES];

// In host code `g_select_scratch` is only a handle for the runtime, not an address a
// kernel can use; the device address is looked up once per device (that device current).
static void* g_scratch_addr[STATIC_MAX_DEVICES];
static bool g_scratch_ready[STATIC_MAX_DEVICES];
static std::mutex g_scratch_mutex;   // two host threads may make a device's first call at once

static int static_scratch_for_current_device(void** out) {
  int dev = -1;
  KVM_CUDA_CHECK(cudaGetDevice(&dev));
  KVM_REQUIRE(dev >= 0 && dev < STATIC_MAX_DEVICES, "kvm_column_select_static: device %d unsupported", dev);
  std::lock_guard<std::mutex> guard(g_scratch_mutex);
  if (!g_scratch_ready[dev]) {
    void* ptr = nullptr;
    KVM_CUDA_CHECK(cudaGetSymbolAddress(&ptr, g_select_scratch));
    g_scratch_addr[dev] = ptr;
    g_scratch_ready[dev] = true;
  }
  *out = g_scratch_addr[dev];
  return KVM_OK;
}

}  // namespace

extern "C" int64_t kvm_column_select_static_max_n(void) { return STATIC_MAX_N; }

extern "C" int kvm_column_select_static(const double* values, int64_t n, int32_t mode,
                                        const void* member, const kvm_query* queries, int32_t nq,
                                        kvm_query_result* results, void* stream) {
  KVM_REQUIRE(mode == KVM_MEMBER_ALL || mode == KVM_MEMBER_STATUS,
              "kvm_column_select_static: member mode %d not supported (only ALL and STATUS: one set, "
              "so the candidates of all queries together never exceed n)", mode);
  KVM_REQUIRE(n >= 0 && n <= STATIC_MAX_N,
              "kvm_column_select_static: n=%lld exceeds the built-in scratch (max %lld values); use "
              "kvm_column_select with a caller-provided workspace", (long long)n, (long long)STATIC_MAX_N);
  KVM_REQUIRE(queries && nq >= 1 && nq <= MAXQ, "kvm_column_select_static: nq=%d outside [1,%d]", nq, MAXQ);
  void* scratch = nullptr;
  int rc = static_scratch_for_current_device(&scratch);
  if (rc) return rc;
  // There is one set (set 0).  Candidate buckets are shared between queries only when their
  // set masks are identical, so masks that differ in bits no set answers to (1 vs 3) would give
  // several queries their own copy of one bin and the copies together could outgrow the scratch
  // (n keys).  Reduce every mask to bit 0: the copies collapse and the candidates of all queries
  // together never exceed n.
  kvm_query q1[MAXQ];
  for (int i = 0; i < nq; ++i) {
    q1[i] = queries[i];
    q1[i].set_mask &= 1u;
  }
  // Everything that must start from zero is cleared inside the pipeline itself (stats
  // ticket, histograms, append cursors), so stale contents of the previous call are harmless.
  return kvm_column_select(values, n, mode, member, 1, q1, nq, results, scratch,
                           STATIC_SCRATCH_BYTES, STATIC_MAX_N, stream);
}

// ------------------------------------------------------------------ multi-GPU exchange helpers
// After the compact pass every rank exchanges one small record
//   [bucket cursors (KVM_MAX_QUERIES) | histogram edge counters (KVM_MAX_HISTS*KVM_HIST_MAX_BINS) | dropped]
// with one all-gather; these two kernels pack it and fold the gathered records back in
// (cursors stay per rank - the finish kernel reads them per source; edge counters and the
// dropped counter are summed over the ranks).
namespace {
constexpr int POST_WORDS = MAXQ + NH * NB_MAX + 1;

__global__ void post_pack_kernel(const SelWs* ws, const long long* dropped, long long* rec) {
  for (int i = threadIdx.x; i < POST_WORDS; i += blockDim.x) {
    long long v;
    if (i < MAXQ) v = ws->bucket_counts[i];
    else if (i < MAXQ + NH * NB_MAX) v = (long long)(&ws->edge20[0][0])[i - MAXQ];
    else v = dropped ? *dropped : 0;
    rec[i] = v;
  }
}

__global__ void post_unpack_kernel(SelWs* ws, const long long* all, int world, long long* counts_all,
                                   long long* dropped) {
  for (int i = threadIdx.x; i < POST_WORDS; i += blockDim.x) {
    if (i < MAXQ) {
      for (int r = 0; r < world; ++r) counts_all[(size_t)r * MAXQ + i] = all[(size_t)r * POST_WORDS + i];
    } else {
      long long s = 0;
      for (int r = 0; r < world; ++r) s += all[(size_t)r * POST_WORDS + i];
      if (i < MAXQ + NH * NB_MAX) (&ws->edge20[0][0])[i - MAXQ] = (unsigned long long)s;
      else if (dropped) *dropped = s;
    }
  }
}
}  // namespace

// The exchanged candidate slots hold `capacity` keys per rank; with more candidates than that
// (global bucket sizes - every rank agrees) the finish must not run over the slots.
__global__ void require_capacity_kernel(SelWs* ws, long long capacity) {
  if (ws->status == KVM_OK && ws->bucket_offsets[MAXQ] > capacity) ws->status = KVM_ERR_ARENA;
}

extern "C" int kvm_select_require_capacity(void* workspace, int64_t capacity_keys, void* stream) {
  KVM_REQUIRE(workspace && capacity_keys >= 0, "kvm_select_require_capacity: bad argument");
  require_capacity_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(reinterpret_cast<SelWs*>(workspace), capacity_keys);
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int32_t kvm_select_post_words(void) { return POST_WORDS; }

extern "C" int kvm_select_post_pack(const void* workspace, const int64_t* dropped, int64_t* record, void* stream) {
  KVM_REQUIRE(workspace && record, "kvm_select_post_pack: null argument");
  post_pack_kernel<<<1, 128, 0, (cudaStream_t)stream>>>(reinterpret_cast<const SelWs*>(workspace),
                                                       reinterpret_cast<const long long*>(dropped),
                                                       reinterpret_cast<long long*>(record));
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_select_post_unpack(void* workspace, const int64_t* gathered, int32_t world,
                                      int64_t* counts_all, int64_t* dropped, void* stream) {
  KVM_REQUIRE(workspace && gathered && counts_all && world >= 1, "kvm_select_post_unpack: bad argument");
  post_unpack_kernel<<<1, 128, 0, (cudaStream_t)stream>>>(reinterpret_cast<SelWs*>(workspace),
                                                         reinterpret_cast<const long long*>(gathered), world,
                                                         reinterpret_cast<long long*>(counts_all),
                                                         reinterpret_cast<long long*>(dropped));
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

// ---- peer-memory form of the candidate exchange ---------------------------------------------
// Every rank copies the filled part of each of its buckets straight into its slot of every
// rank's exchange buffer (same bucket offsets everywhere), so no fixed-size gather and no host
// round trip is needed: the device knows how many candidates there are.
namespace {
__global__ void __launch_bounds__(256)
peer_put_candidates_kernel(SelWs* ws, const uint64_t* __restrict__ arena, KvmPeerTable pt,
                           int64_t slot_offset, int64_t slot_capacity) {
  if (ws->status != KVM_OK) return;
  if (ws->bucket_offsets[MAXQ] > slot_capacity) {  // global bucket sizes: every rank takes this branch
    if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) ws->status = KVM_ERR_ARENA;
    return;
  }
  uint64_t* __restrict__ dst = reinterpret_cast<uint64_t*>(pt.base[blockIdx.y] + slot_offset);
  const int nb = ws->nbuckets;
  for (int j = 0; j < nb; ++j) {
    const int64_t off = ws->bucket_offsets[j];
    int64_t cnt = ws->bucket_counts[j];
    const int64_t cap = ws->bucket[j].count;
    cnt = cnt < cap ? cnt : cap;
    // 16-byte stores over the aligned middle of the bucket, single words at its ends
    const int64_t head = (off & 1) && cnt > 0 ? 1 : 0;
    const int64_t pairs = (cnt - head) >> 1;
    const uint64_t* s8 = arena + off;
    uint64_t* d8 = dst + off;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      if (head) d8[0] = s8[0];
      if ((cnt - head) & 1) d8[cnt - 1] = s8[cnt - 1];
    }
    const ulonglong2* s16 = reinterpret_cast<const ulonglong2*>(s8 + head);
    ulonglong2* d16 = reinterpret_cast<ulonglong2*>(d8 + head);
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < pairs; i += (int64_t)gridDim.x * 256)
      d16[i] = s16[i];
  }
}
}  // namespace

extern "C" int kvm_peer_put_candidates(void* workspace, int64_t slot_offset, int64_t slot_capacity,
                                       void* const* peers, int32_t world, int32_t rank, void* stream) {
  KvmPeerTable pt;
  KVM_REQUIRE(workspace && kvm_peer_table(&pt, peers, world, rank) == 0,
              "kvm_peer_put_candidates: bad argument (world=%d rank=%d)", world, rank);
  KVM_REQUIRE(slot_offset % 16 == 0 && slot_capacity >= 0, "kvm_peer_put_candidates: slot offset must be 16-byte aligned");
  peer_put_candidates_kernel<<<dim3(48, world), 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<SelWs*>(workspace), arena_of(workspace), pt, slot_offset, slot_capacity);
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}
