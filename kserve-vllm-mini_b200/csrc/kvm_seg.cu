// kvm_seg.cu - K4: segmented run analysis for sweep roll-ups (BASELINE.json configs[2]:
// 4096 runs x 250k requests).  One persistent CTA per run does the WHOLE analysis of that
// run with all intermediate state in shared memory - no global histograms, no workspace,
// no inter-CTA synchronisation:
//
//   P1 scan      window, cold/warm, counts, token sums, key ranges           (56 B/row + 1 B flags)
//   P2 plan      digit parameters per value column, rank of every query
//   P3 hist      7 x 2048-bin digit histograms in shared memory              (33 B/row)
//   P4 resolve   bin + residual rank per query, candidate buckets, 20-bin thresholds
//   P5 compact   candidates -> shared-memory arena; rows whose fine bin holds a 20-bin
//                edge are binned exactly                                      (33 B/row)
//   P6 finish    each query: exact rank inside its bucket by counting (candidates on chip)
//   P7 emit      stats, order statistics, 4 x num_bins histogram counts
//
// Replaces the per-run loop of grid-sweep.sh:103-174 (bench.sh -> analyze.py per run) for the
// numeric core of analyze.py; semantics per row are the shared device functions of
// kvm_log_device.cuh, so a run analysed here equals the same run through the flat kernels.
// The 20-bin histograms are derived from the fine digit histogram: a fine bin that contains no
// 20-bin edge lies inside one 20-bin (added wholesale in P7); only rows in the <= 4*(nb-1)
// edge bins are binned one by one in P5.
// Runs whose candidates do not fit the on-chip arena are flagged (status 1) and redone by the
// host through the flat path; results stay exact either way.
// HBM-bound: 123 B/row over three streaming phases.

#include "kvm_common.cuh"
#include "kvm_log_device.cuh"

namespace {

constexpr int SEG_THREADS = 512;
constexpr int SEG_WARPS = SEG_THREADS / 32;
constexpr int SEG_ARENA = 12288;      // candidate keys kept on chip per run
constexpr int SEG_BUCKET_MAX = 2048;  // largest bucket finished by counting
constexpr int SEG_COLD = 256;
constexpr int SEG_MAXQ = KVM_MAX_QUERIES;
constexpr int SEG_NB_MAX = 64;
constexpr int NH = 4;  // 20-bin histograms: ttfb, tllt, generation, per-token

struct SegBucket {
  uint64_t base;
  int32_t shift, col;
  uint32_t set_mask, bin;
  uint32_t count, offset;
};

struct SegQ {
  int64_t n, rank, resid;
  uint64_t key;
  int32_t status, bucket, col, resolved;
};

struct SegCol {
  uint64_t kbase;
  int32_t shift, width;
};

struct SegSmem {
  uint32_t hist[KVM_LOG_NSETS][BINS];
  uint64_t arena[SEG_ARENA];
  uint64_t red[SEG_WARPS][NSLOTS];
  uint64_t fin[NSLOTS];
  uint64_t thr[NH][SEG_NB_MAX];
  uint32_t edge[NH][SEG_NB_MAX];
  uint32_t cnt20[NH][SEG_NB_MAX];
  uint32_t flag20[NH][BINS / 32];
  uint32_t bmap[KVM_LOG_NCOLS][BINS / 32];
  double cold_lo[SEG_COLD];
  double cold_hi[SEG_COLD];
  SegBucket bk[SEG_MAXQ];
  SegQ q[SEG_MAXQ];
  SegCol col[KVM_MAX_COLS];
  uint32_t q_bin[SEG_MAXQ];
  uint32_t q_cnt[SEG_MAXQ];
  uint32_t cursor[SEG_MAXQ];
  int32_t active20[NH];
  int32_t nbuckets;
  int32_t overflow;
  kvm_log_stats stats;
};

__device__ __forceinline__ int hist_col(int h) {
  return h == 0 ? KVM_COL_TTFB : h == 1 ? KVM_COL_TLLT : h == 2 ? KVM_COL_GEN : KVM_COL_PERTOK;
}
__device__ __forceinline__ uint32_t hist_sets(int h) {
  return h == 0 ? ((1u << KVM_SET_TTFB_COLD) | (1u << KVM_SET_TTFB_WARM))
         : h == 1 ? (1u << KVM_SET_TLLT) : h == 2 ? (1u << KVM_SET_GEN) : (1u << KVM_SET_PERTOK);
}
__device__ __forceinline__ int col_hist(int col) {  // inverse of hist_col, -1 for latency
  return col == KVM_COL_TTFB ? 0 : col == KVM_COL_TLLT ? 1 : col == KVM_COL_GEN ? 2 : col == KVM_COL_PERTOK ? 3 : -1;
}

// rows [r0, r1) of the log: f(i, start, lat, ttfb, tllt, status, ctok, ttok)
template <class F>
__device__ __forceinline__ void seg_rows_full(const kvm_request_log& log, int64_t r0, int64_t r1, bool vec, F&& f) {
  int64_t a = r0;
  if (vec) {
    if ((a & 1) && a < r1) {
      if (threadIdx.x == 0)
        f(a, log.start_ms[a], log.latency_ms[a], log.ttfb_ms[a], log.tllt_ms[a], log.status[a],
          log.completion_tokens[a], log.total_tokens[a]);
      ++a;
    }
    const int64_t npairs = (r1 - a) >> 1;
    for (int64_t p = threadIdx.x; p < npairs; p += SEG_THREADS) {
      const int64_t i = a + (p << 1);
      const double2 st = kvm_ld2(log.start_ms + i), la = kvm_ld2(log.latency_ms + i);
      const double2 tf = kvm_ld2(log.ttfb_ms + i), tl = kvm_ld2(log.tllt_ms + i);
      const longlong2 ss = kvm_ld2(log.status + i), ct = kvm_ld2(log.completion_tokens + i);
      const longlong2 tt = kvm_ld2(log.total_tokens + i);
      f(i, st.x, la.x, tf.x, tl.x, ss.x, ct.x, tt.x);
      f(i + 1, st.y, la.y, tf.y, tl.y, ss.y, ct.y, tt.y);
    }
    a += npairs << 1;
    if (a < r1 && threadIdx.x == 0)
      f(a, log.start_ms[a], log.latency_ms[a], log.ttfb_ms[a], log.tllt_ms[a], log.status[a],
        log.completion_tokens[a], log.total_tokens[a]);
  } else {
    for (int64_t i = a + threadIdx.x; i < r1; i += SEG_THREADS)
      f(i, kvm_ld1(log.start_ms + i), kvm_ld1(log.latency_ms + i), kvm_ld1(log.ttfb_ms + i),
        kvm_ld1(log.tllt_ms + i), kvm_ld1(log.status + i), kvm_ld1(log.completion_tokens + i),
        kvm_ld1(log.total_tokens + i));
  }
}

// rows [r0, r1): f(flags, lat, ttfb, tllt, ctok) - the columns the select passes read
template <class F>
__device__ __forceinline__ void seg_rows_keys(const kvm_request_log& log, const uint8_t* __restrict__ flags,
                                              int64_t r0, int64_t r1, bool vec, F&& f) {
  int64_t a = r0;
  if (vec) {
    if ((a & 1) && a < r1) {
      if (threadIdx.x == 0) f(flags[a], log.latency_ms[a], log.ttfb_ms[a], log.tllt_ms[a], log.completion_tokens[a]);
      ++a;
    }
    const int64_t npairs = (r1 - a) >> 1;
    for (int64_t p = threadIdx.x; p < npairs; p += SEG_THREADS) {
      const int64_t i = a + (p << 1);
      const uint16_t ff = *reinterpret_cast<const uint16_t*>(flags + i);
      if (!((ff | (ff >> 8)) & KVM_F_SUCCESS)) continue;
      const double2 la = kvm_ld2(log.latency_ms + i), tf = kvm_ld2(log.ttfb_ms + i);
      const double2 tl = kvm_ld2(log.tllt_ms + i);
      const longlong2 ct = kvm_ld2(log.completion_tokens + i);
      f((uint8_t)(ff & 0xff), la.x, tf.x, tl.x, ct.x);
      f((uint8_t)(ff >> 8), la.y, tf.y, tl.y, ct.y);
    }
    a += npairs << 1;
    if (a < r1 && threadIdx.x == 0) f(flags[a], log.latency_ms[a], log.ttfb_ms[a], log.tllt_ms[a], log.completion_tokens[a]);
  } else {
    for (int64_t i = a + threadIdx.x; i < r1; i += SEG_THREADS) {
      const uint8_t fl = flags[i];
      if (!(fl & KVM_F_SUCCESS)) continue;
      f(fl, kvm_ld1(log.latency_ms + i), kvm_ld1(log.ttfb_ms + i), kvm_ld1(log.tllt_ms + i),
        kvm_ld1(log.completion_tokens + i));
    }
  }
}

struct SegQueries {
  kvm_query q[SEG_MAXQ];
  int32_t nq;
};

__global__ void __launch_bounds__(SEG_THREADS, 1)
seg_analyze_kernel(kvm_request_log log, const int64_t* __restrict__ run_offsets, int64_t runs,
                   const int64_t* __restrict__ cold_offsets, const double* __restrict__ cold_times,
                   double cold_window, SegQueries qp, int nb, bool base_aligned,
                   uint8_t* __restrict__ flags, kvm_log_stats* __restrict__ out_stats,
                   kvm_query_result* __restrict__ out_results, int64_t* __restrict__ out_hist,
                   int32_t* __restrict__ out_status) {
  extern __shared__ __align__(16) unsigned char seg_raw[];
  SegSmem& S = *reinterpret_cast<SegSmem*>(seg_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const double nan = __longlong_as_double(0x7ff8000000000000ll);

  for (int64_t run = blockIdx.x; run < runs; run += gridDim.x) {
    const int64_t r0 = run_offsets[run], r1 = run_offsets[run + 1];
    const bool vec = base_aligned;
    // ---------------- P1 scan
    ColdTable cold{S.cold_lo, S.cold_hi, 0};
    bool cold_too_many = false;
    if (cold_offsets) {
      const int64_t c0 = cold_offsets[run], c1 = cold_offsets[run + 1];
      const int nc = (int)(c1 - c0);
      if (nc > SEG_COLD) cold_too_many = true;   // host redoes this run through the flat path
      else {
        cold.n = nc;
        for (int i = tid; i < nc; i += SEG_THREADS)
          cold_interval(cold_times[c0 + i], cold_window, &S.cold_lo[i], &S.cold_hi[i]);
      }
    }
    for (int i = tid; i < KVM_LOG_NSETS * BINS; i += SEG_THREADS) (&S.hist[0][0])[i] = 0;
    for (int i = tid; i < NH * SEG_NB_MAX; i += SEG_THREADS) {
      (&S.edge[0][0])[i] = 0;
      (&S.cnt20[0][0])[i] = 0;
      (&S.thr[0][0])[i] = ~0ull;
    }
    for (int i = tid; i < NH * (BINS / 32); i += SEG_THREADS) (&S.flag20[0][0])[i] = 0;
    for (int i = tid; i < KVM_LOG_NCOLS * (BINS / 32); i += SEG_THREADS) (&S.bmap[0][0])[i] = 0;
    if (tid < SEG_MAXQ) S.cursor[tid] = 0;
    if (tid == 0) {
      S.nbuckets = 0;
      S.overflow = 0;
    }
    __syncthreads();

    LogAcc a;
    a.init();
    seg_rows_full(log, r0, r1, vec, [&](int64_t i, double st, double la, double tf, double tl, int64_t ss,
                                        int64_t ct, int64_t tt) {
      flags[i] = process_row(a, st, la, tf, tl, ss, ct, tt, cold, -1);
    });
    acc_warp_reduce(a, S.red[warp], lane);
    __syncthreads();
    if (tid < NMAX_END) {
      uint64_t acc = S.red[0][tid];
      for (int w = 1; w < SEG_WARPS; ++w) {
        const uint64_t v = S.red[w][tid];
        if (tid < NSUM) acc = (uint64_t)((int64_t)acc + (int64_t)v);
        else if (tid < NMIN_END) acc = v < acc ? v : acc;
        else acc = v > acc ? v : acc;
      }
      S.fin[tid] = acc;
    } else if (tid == NMAX_END || tid == NMAX_END + 1) {
      const int s = tid == NMAX_END ? D_TTFB : D_LAT;
      kvm_dd acc{__longlong_as_double((long long)S.red[0][s]), __longlong_as_double((long long)S.red[0][s + 1])};
      for (int w = 1; w < SEG_WARPS; ++w)
        acc = kvm_dd_add(acc, kvm_dd{__longlong_as_double((long long)S.red[w][s]),
                                     __longlong_as_double((long long)S.red[w][s + 1])});
      S.fin[s] = (uint64_t)__double_as_longlong(acc.hi);
      S.fin[s + 1] = (uint64_t)__double_as_longlong(acc.lo);
    }
    __syncthreads();
    // ---------------- P2 plan
    if (tid == 0) {
      stats_from_slots(S.fin, r1 - r0, S.stats);
      if (cold_too_many) S.overflow = 1;
    }
    __syncthreads();
    if (tid < KVM_MAX_COLS) {
      SegCol cp{0, 0, 0};
      if (tid < KVM_LOG_NCOLS && S.stats.sets.kmin[tid] <= S.stats.sets.kmax[tid]) {
        const uint64_t range = S.stats.sets.kmax[tid] - S.stats.sets.kmin[tid];
        cp.kbase = S.stats.sets.kmin[tid];
        cp.width = range ? 64 - __clzll((long long)range) : 0;
        cp.shift = cp.width > BINS_LOG2 ? cp.width - BINS_LOG2 : 0;
      }
      S.col[tid] = cp;
    }
    if (tid >= 32 && tid < 32 + SEG_MAXQ) {
      const int qi = tid - 32;
      SegQ s{0, 0, 0, 0, KVM_Q_EMPTY, -1, 0, 1};
      if (qi < qp.nq) {
        const kvm_query q = qp.q[qi];
        int64_t n = 0;
        int col = -1;
        for (int b = 0; b < KVM_LOG_NSETS; ++b)
          if ((q.set_mask >> b) & 1u) {
            n += S.stats.sets.count[b];
            if (col < 0) col = S.stats.sets.set_col[b];
          }
        s.n = n;
        s.col = col < 0 ? 0 : col;
        if (n > 0) {
          s.rank = resolve_rank(q.rank_kind, q.p, n, q.rank);
          s.status = (s.rank < 0 || s.rank >= n) ? KVM_Q_OUT_OF_RANGE : KVM_Q_OK;
          s.resolved = s.status != KVM_Q_OK;
        }
      }
      S.q[qi] = s;
    }
    // 20-bin thresholds (bisection with the reference expression); one thread per edge
    if (tid >= 128 && tid < 128 + NH * (SEG_NB_MAX - 1)) {
      const int t = tid - 128, h = t / (SEG_NB_MAX - 1), b = 1 + t % (SEG_NB_MAX - 1);
      const int c = hist_col(h);
      const uint64_t klo = S.stats.sets.kmin[c], khi = S.stats.sets.kmax[c];
      if (b == 1) S.active20[h] = klo < khi ? 1 : 0;
      if (b < nb && klo < khi) S.thr[h][b - 1] = bin_threshold(klo, khi, nb, b);
    }
    __syncthreads();
    // edge bins of the fine histogram (those that contain a 20-bin threshold)
    if (tid < NH * (SEG_NB_MAX - 1)) {
      const int h = tid / (SEG_NB_MAX - 1), b = 1 + tid % (SEG_NB_MAX - 1);
      const uint64_t T = S.thr[h][b - 1];
      if (b < nb && S.active20[h] && T != ~0ull) {
        const SegCol cp = S.col[hist_col(h)];
        const uint32_t d = digit_of(T, cp.kbase, cp.shift);
        atomicOr(&S.flag20[h][d >> 5], 1u << (d & 31));
      }
    }
    __syncthreads();
    // ---------------- P3 hist
    seg_rows_keys(log, flags, r0, r1, vec, [&](uint8_t fl, double la, double tf, double tl, int64_t ct) {
      log_row_keys(fl, la, tf, tl, ct, [&](int set, int col, uint64_t key) {
        atomicAdd(&S.hist[set][digit_of(key, S.col[col].kbase, S.col[col].shift)], 1u);
      });
    });
    __syncthreads();
    // ---------------- P4 resolve (one warp per query)
    for (int qi = warp; qi < qp.nq; qi += SEG_WARPS) {
      SegQ s = S.q[qi];
      if (s.status != KVM_Q_OK) continue;
      const SegCol cp = S.col[s.col];
      const uint32_t mask = qp.q[qi].set_mask;
      if (cp.width == 0) {
        if (lane == 0) {
          s.key = cp.kbase;
          s.resolved = 1;
          S.q[qi] = s;
        }
        continue;
      }
      int64_t running = 0;
      bool found = false;
      for (int base = 0; base < BINS && !found; base += 32) {
        int64_t c = 0;
        for (int b = 0; b < KVM_LOG_NSETS; ++b)
          if ((mask >> b) & 1u) c += (int64_t)S.hist[b][base + lane];
        int64_t incl = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const int64_t o = __shfl_up_sync(KVM_FULL, incl, d);
          if (lane >= d) incl += o;
        }
        const unsigned hit = __ballot_sync(KVM_FULL, running + incl > s.rank);
        if (hit) {
          const int l = __ffs(hit) - 1;
          const int64_t excl = __shfl_sync(KVM_FULL, incl - c, l);
          const int64_t cnt = __shfl_sync(KVM_FULL, c, l);
          if (lane == 0) {
            S.q_bin[qi] = (uint32_t)(base + l);
            S.q_cnt[qi] = (uint32_t)cnt;
            s.resid = s.rank - (running + excl);
            if (cp.shift == 0) {
              s.key = cp.kbase + (uint64_t)(base + l);
              s.resolved = 1;
            } else {
              s.resolved = 0;
              s.bucket = -2;
            }
            S.q[qi] = s;
          }
          found = true;
        }
        running += __shfl_sync(KVM_FULL, incl, 31);
      }
      if (!found && lane == 0) {
        s.status = KVM_Q_OUT_OF_RANGE;
        s.resolved = 1;
        S.q[qi] = s;
      }
    }
    __syncthreads();
    if (tid == 0) {
      int nbk = 0;
      uint32_t off = 0;
      for (int qi = 0; qi < qp.nq; ++qi) {
        SegQ& s = S.q[qi];
        if (s.status != KVM_Q_OK || s.resolved) continue;
        const uint32_t mask = qp.q[qi].set_mask;
        int j = -1;
        for (int k = 0; k < nbk; ++k)
          if (S.bk[k].col == s.col && S.bk[k].set_mask == mask && S.bk[k].bin == S.q_bin[qi]) {
            j = k;
            break;
          }
        if (j < 0) {
          j = nbk++;
          const SegCol cp = S.col[s.col];
          SegBucket b;
          b.base = cp.kbase + ((uint64_t)S.q_bin[qi] << cp.shift);
          b.shift = cp.shift;
          b.col = s.col;
          b.set_mask = mask;
          b.bin = S.q_bin[qi];
          b.count = S.q_cnt[qi];
          b.offset = off;
          if (b.count > (uint32_t)SEG_BUCKET_MAX) S.overflow = 1;
          off += b.count;
          S.bk[j] = b;
          S.bmap[s.col][b.bin >> 5] |= 1u << (b.bin & 31);
        }
        s.bucket = j;
      }
      S.nbuckets = nbk;
      if (off > (uint32_t)SEG_ARENA) S.overflow = 1;
    }
    __syncthreads();
    if (S.overflow) {  // host redoes this run through the flat path
      if (tid == 0) {
        out_status[run] = 1;
        out_stats[run] = S.stats;
      }
      __syncthreads();
      continue;
    }
    // ---------------- P5 compact + exact binning of the rows in edge bins
    {
      const int nbk = S.nbuckets;
      seg_rows_keys(log, flags, r0, r1, vec, [&](uint8_t fl, double la, double tf, double tl, int64_t ct) {
        log_row_keys(fl, la, tf, tl, ct, [&](int set, int col, uint64_t key) {
          const uint32_t d = digit_of(key, S.col[col].kbase, S.col[col].shift);
          if ((S.bmap[col][d >> 5] >> (d & 31)) & 1u) {
            for (int j = 0; j < nbk; ++j)
              if (S.bk[j].col == col && S.bk[j].bin == d && ((S.bk[j].set_mask >> set) & 1u)) {
                const uint32_t pos = atomicAdd(&S.cursor[j], 1u);
                if (pos < S.bk[j].count) S.arena[S.bk[j].offset + pos] = key;
              }
          }
          const int h = col_hist(col);
          if (h >= 0 && ((S.flag20[h][d >> 5] >> (d & 31)) & 1u))
            atomicAdd(&S.edge[h][bin_of(key, S.thr[h], nb)], 1u);
        });
      });
    }
    __syncthreads();
    // ---------------- P6 finish: rank inside the bucket by counting (one warp per query)
    for (int qi = warp; qi < qp.nq; qi += SEG_WARPS) {
      SegQ s = S.q[qi];
      if (s.status != KVM_Q_OK || s.resolved) continue;
      const SegBucket b = S.bk[s.bucket];
      const uint64_t* keys = S.arena + b.offset;
      const uint32_t m = b.count;
      for (uint32_t i = lane; i < m; i += 32) {
        const uint64_t k = keys[i];
        uint32_t less = 0, leq = 0;
        for (uint32_t j = 0; j < m; ++j) {
          const uint64_t o = keys[j];
          less += o < k;
          leq += o <= k;
        }
        if ((int64_t)less <= s.resid && s.resid < (int64_t)leq) S.q[qi].key = k;  // all writers agree
      }
    }
    __syncthreads();
    // ---------------- P7 emit
    // fine bins without an edge lie inside a single 20-bin: add them wholesale
    for (int t = tid; t < NH * BINS; t += SEG_THREADS) {
      const int h = t / BINS, d = t % BINS;
      if (!S.active20[h] || ((S.flag20[h][d >> 5] >> (d & 31)) & 1u)) continue;
      const uint32_t sets = hist_sets(h);
      uint32_t c = 0;
      for (int b = 0; b < KVM_LOG_NSETS; ++b)
        if ((sets >> b) & 1u) c += S.hist[b][d];
      if (c) {
        const SegCol cp = S.col[hist_col(h)];
        const uint64_t lowkey = cp.kbase + ((uint64_t)d << cp.shift);
        atomicAdd(&S.cnt20[h][bin_of(lowkey, S.thr[h], nb)], c);
      }
    }
    __syncthreads();
    for (int t = tid; t < NH * nb; t += SEG_THREADS) {
      const int h = t / nb, b = t % nb;
      out_hist[(run * NH + h) * nb + b] = (int64_t)S.cnt20[h][b] + (int64_t)S.edge[h][b];
    }
    if (tid < qp.nq) {
      const SegQ s = S.q[tid];
      kvm_query_result r;
      r.value = s.status == KVM_Q_OK ? kvm_unkey(s.key) : nan;
      r.n = s.n;
      r.rank = s.rank;
      r.status = s.status;
      r.pad = 0;
      out_results[run * qp.nq + tid] = r;
    }
    if (tid == 0) {
      out_status[run] = 0;
      out_stats[run] = S.stats;
    }
    __syncthreads();
  }
}

}  // namespace

extern "C" int kvm_seg_analyze(const kvm_request_log* log, const int64_t* run_offsets, int64_t runs,
                               const int64_t* cold_offsets, const double* cold_times,
                               double cold_window_s, const kvm_query* queries, int32_t nq,
                               int32_t num_bins, uint8_t* flags, kvm_log_stats* stats,
                               kvm_query_result* results, int64_t* hist_counts, int32_t* run_status,
                               void* stream) {
  KVM_REQUIRE(log && run_offsets && queries && flags && stats && results && hist_counts && run_status,
              "kvm_seg_analyze: null argument");
  KVM_REQUIRE(runs >= 0, "kvm_seg_analyze: negative run count");
  KVM_REQUIRE(nq >= 1 && nq <= SEG_MAXQ, "kvm_seg_analyze: nq=%d outside [1,%d]", nq, SEG_MAXQ);
  KVM_REQUIRE(num_bins >= 1 && num_bins <= SEG_NB_MAX, "kvm_seg_analyze: num_bins=%d outside [1,%d]",
              num_bins, SEG_NB_MAX);
  KVM_REQUIRE(cold_offsets == nullptr || cold_times != nullptr, "kvm_seg_analyze: cold_times is null");
  if (runs == 0) return KVM_OK;
  SegQueries qp;
  for (int i = 0; i < SEG_MAXQ; ++i) qp.q[i] = i < nq ? queries[i] : kvm_query{0, 0, 0.0, 0};
  qp.nq = nq;
  const size_t smem = sizeof(SegSmem);
  KVM_CUDA_CHECK(cudaFuncSetAttribute((const void*)seg_analyze_kernel,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; };
  const bool aligned = al16(log->start_ms) && al16(log->latency_ms) && al16(log->ttfb_ms) &&
                       al16(log->tllt_ms) && al16(log->status) && al16(log->completion_tokens) &&
                       al16(log->total_tokens) && (reinterpret_cast<uintptr_t>(flags) & 1u) == 0;
  int grid = kvm_num_sms();
  if ((int64_t)grid > runs) grid = (int)runs;
  seg_analyze_kernel<<<grid, SEG_THREADS, smem, (cudaStream_t)stream>>>(
      *log, run_offsets, runs, cold_offsets, cold_times, cold_window_s, qp, num_bins, aligned, flags,
      stats, results, hist_counts, run_status);
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}
