// kvm_seg.cu - K4: segmented run analysis for sweep roll-ups (BASELINE.json configs[2]:
// 4096 runs x 250k requests).  One persistent CTA (two per SM) takes a run at a time and does the
// WHOLE analysis of that run with all intermediate state in shared memory - no global
// histograms, no inter-CTA synchronisation - reading the run's columns ONCE:
//
//   P0 sample    512 rows of the run -> digit parameters of every value column (kvm_fused_row.cuh)
//   P1 stream    the one-pass row of kvm_fused.cu: window, cold/warm, counts, token sums, exact key
//                ranges, 7 x 2048-bin digit histograms in shared memory; per row one flag byte
//                and one 8-byte digit record leave the SM          (56 B/row read, 9 B/row written)
//   P2 plan      statistics, rank of every query, 20-bin thresholds from the exact ranges
//   P4 resolve   bin + residual rank per query, candidate buckets
//   P5 20-bin    fine bins that hold no 20-bin edge are added wholesale; the shared-memory
//                histograms are then dead and their space becomes the candidate arena
//   P6 compact   second look at the run through its 9 B/row records only: rows whose digit
//                falls into a candidate bucket or a fine bin with a 20-bin edge fetch their value
//                (a few percent of the rows)
//   P7 finish    each query: exact rank inside its bucket by counting (candidates on chip)
//   P8 emit      stats, order statistics, 4 x num_bins histogram counts
//
// Replaces the per-run loop of grid-sweep.sh:103-174 (bench.sh -> analyze.py per run) for the
// numeric core of analyze.py; semantics per row are the shared device functions of
// kvm_log_device.cuh / kvm_fused_row.cuh, so a run analysed here equals the same run through the
// flat kernels.  The digit parameters come from a sample, so the two end bins of a column are
// catch-alls for keys outside the sampled hull: buckets are defined by DIGIT (not key range), and
// an end bin that holds such keys is binned row by row for the 20-bin histograms.
// Runs whose candidates do not fit the on-chip arena are flagged (status 1) and redone by the
// host through the flat path; results stay exact either way.
// HBM-bound by design: 56 B/row algorithmic; 74 B/row moved (records written and read back).

#include "kvm_common.cuh"
#include "kvm_log_device.cuh"
#include "kvm_fused_row.cuh"

namespace {

constexpr int SEG_THREADS = 384;      // x 2 CTAs per SM, like the flat one-pass kernel
constexpr int SEG_WARPS = SEG_THREADS / 32;
constexpr int SEG_ARENA = KVM_LOG_NSETS * BINS * 4 / 8;   // candidate keys per run: the space of the histograms
constexpr int SEG_BUCKET_MAX = SEG_ARENA;  // a bucket may fill the arena (the finish is linear in its size)
constexpr int SEG_COLD = 256;
constexpr int SEG_HITQ = 128;   // (row, column) pairs a warp collects before it fetches their values
constexpr int SEG_DIRECT = 96;  // more than this in one iteration (of up to 320): served where found
constexpr int SEG_RK = 4;       // row pairs per lane and iteration of the records pass
constexpr int SEG_MAXQ = KVM_MAX_QUERIES;
constexpr int SEG_NB_MAX = 64;
constexpr int NH = 4;  // 20-bin histograms: ttfb, tllt, generation, per-token

struct SegBucket {
  int32_t col;
  uint32_t set_mask, bin;
  uint32_t count, offset;
};

struct SegQ {
  int64_t n, rank, resid;
  uint64_t key;
  int32_t status, bucket, col, resolved;
};

struct SegSmem {
  uint32_t hist[KVM_LOG_NSETS * BINS];        // P1-P5; the candidate arena from P6 on
  FusedCols fc;
  uint64_t red[SEG_WARPS][NSLOTS];
  uint64_t fin[NSLOTS];
  unsigned long long s_inv[SEG_WARPS][NC], s_max[SEG_WARPS][NC];
  uint64_t thr[NH][SEG_NB_MAX];
  uint32_t edge[NH][SEG_NB_MAX];
  uint32_t cnt20[NH][SEG_NB_MAX];
  uint32_t flag20[NH][BINS / 32];
  uint32_t bmap[NC][BINS / 32];
  uint8_t hit8[NC][2 * BINS];                 // bmap | flag20 as bytes (1 << column): a 12-bit record field indexes it unchecked
  double cold_lo[SEG_COLD];
  double cold_hi[SEG_COLD];
  SegBucket bk[SEG_MAXQ];
  SegQ q[SEG_MAXQ];
  uint32_t q_bin[SEG_MAXQ];
  uint32_t q_cnt[SEG_MAXQ];
  uint32_t cursor[SEG_MAXQ];
  int32_t active20[NH];
  int32_t impure[NC];                         // bit 0: keys below kbase exist, bit 1: keys beyond the last bin exist
  int32_t nbuckets;
  int32_t overflow;
  uint32_t arena_used;
  uint64_t hitq[SEG_WARPS][SEG_HITQ];          // P6: rows waiting for their value, per warp
  uint32_t hitq_n[SEG_WARPS];
  uint32_t gsum[KVM_LOG_NSETS][BINS / 32];     // totals of every 32-bin group of every set
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

// Where the time of a run goes (debug aid): nanoseconds summed over all runs, per phase -
// [0] sample + set-up, [1] stream, [2] plan / resolve / wholesale bins, [3] records pass, [4] finish + emit.
__device__ unsigned long long seg_phase_ns[8];
__device__ __forceinline__ unsigned long long seg_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

static_assert(sizeof(SegSmem) <= 113 * 1024, "two CTAs per SM");

struct SegQueries {
  kvm_query q[SEG_MAXQ];
  int32_t nq;
};

// A row of P6 whose digit in column c is wanted: fetch the value ...
__device__ __forceinline__ double seg_value(const kvm_request_log& log, int64_t i, int c) {
  if (c == KVM_COL_LAT) return __ldg(log.latency_ms + i);
  if (c == KVM_COL_TTFB) return __ldg(log.ttfb_ms + i);
  if (c == KVM_COL_TLLT) return __ldg(log.tllt_ms + i);
  const double g = __ldg(log.tllt_ms + i) - __ldg(log.ttfb_ms + i);   // analyze.py:144
  if (c == KVM_COL_GEN) return g;
  return g / (double)__ldg(log.completion_tokens + i);                 // analyze.py:146
}
// ... then append it to the buckets it belongs to and / or bin it exactly.  Out of line: a few
// percent of the rows come here.
__device__ __noinline__ void seg_place(SegSmem& S, int c, uint32_t d, uint32_t fl, uint64_t key, int nb) {
  const bool cold_row = fl & KVM_F_COLD;
  const int set = c == KVM_COL_LAT ? (cold_row ? KVM_SET_LAT_COLD : KVM_SET_LAT_WARM)
                  : c == KVM_COL_TTFB ? (cold_row ? KVM_SET_TTFB_COLD : KVM_SET_TTFB_WARM)
                  : c == KVM_COL_TLLT ? KVM_SET_TLLT : c == KVM_COL_GEN ? KVM_SET_GEN : KVM_SET_PERTOK;
  uint64_t* arena = reinterpret_cast<uint64_t*>(S.hist);
  if ((S.bmap[c][d >> 5] >> (d & 31)) & 1u) {
    const int nbk = S.nbuckets;
    for (int j = 0; j < nbk; ++j)
      if (S.bk[j].col == c && S.bk[j].bin == d && ((S.bk[j].set_mask >> set) & 1u)) {
        const uint32_t pos = atomicAdd(&S.cursor[j], 1u);
        if (pos < S.bk[j].count) arena[S.bk[j].offset + pos] = key;
      }
  }
  const int h = col_hist(c);
  if (h >= 0 && ((S.flag20[h][d >> 5] >> (d & 31)) & 1u)) atomicAdd(&S.edge[h][bin_of(key, S.thr[h], nb)], 1u);
}
__device__ __forceinline__ void seg_hit(SegSmem& S, const kvm_request_log& log, int64_t i, int c, uint32_t d,
                                        uint32_t fl, int nb) {
  seg_place(S, c, d, fl, kvm_key(seg_value(log, i, c)), nb);
}

__global__ void __launch_bounds__(SEG_THREADS, 2)
seg_analyze_kernel(kvm_request_log log, const int64_t* __restrict__ run_offsets, int64_t runs,
                   const int64_t* __restrict__ cold_offsets, const double* __restrict__ cold_times,
                   double cold_window, SegQueries qp, int nb, bool base_aligned,
                   uint8_t* __restrict__ flags, uint64_t* __restrict__ recs, kvm_log_stats* __restrict__ out_stats,
                   kvm_query_result* __restrict__ out_results, int64_t* __restrict__ out_hist,
                   int32_t* __restrict__ out_status, unsigned int* __restrict__ ticket) {
  extern __shared__ __align__(16) unsigned char seg_raw[];
  SegSmem& S = *reinterpret_cast<SegSmem*>(seg_raw);
  uint32_t* sh = S.hist;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const double nan = __longlong_as_double(0x7ff8000000000000ll);

  // runs are handed out by ticket (the first one per CTA by its index): a CTA that finishes early
  // takes the next run instead of idling behind a static schedule
  __shared__ unsigned int next_run;
  for (int64_t run = blockIdx.x; run < runs;) {
    if (tid == 0) next_run = gridDim.x + atomicAdd(ticket, 1u);
    const int64_t r0 = run_offsets[run], r1 = run_offsets[run + 1];
    unsigned long long tp = tid == 0 ? seg_now() : 0ull;
    auto phase = [&](int k) {   // after a __syncthreads
      if (tid == 0) {
        const unsigned long long now = seg_now();
        atomicAdd(&seg_phase_ns[k], now - tp);
        tp = now;
      }
    };
    // ---------------- P0 sample + set-up
    kvm_request_log sub = log;   // the run as a log of its own (sampling)
    sub.start_ms += r0; sub.latency_ms += r0; sub.ttfb_ms += r0; sub.tllt_ms += r0;
    sub.status += r0; sub.completion_tokens += r0; sub.total_tokens += r0;
    sub.n = r1 - r0;
    uint64_t s_inv[NC], s_mx[NC];
    double s_st = 0.0;
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      s_inv[c] = 0ull;
      s_mx[c] = 0ull;
    }
    fused_sample_rows<512>(sub, tid, SEG_THREADS, s_inv, s_mx, s_st);
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
    for (int i = tid; i < KVM_LOG_NSETS * BINS / 4; i += SEG_THREADS) reinterpret_cast<uint4*>(sh)[i] = make_uint4(0, 0, 0, 0);
    for (int i = tid; i < NH * SEG_NB_MAX; i += SEG_THREADS) {
      (&S.edge[0][0])[i] = 0;
      (&S.cnt20[0][0])[i] = 0;
      (&S.thr[0][0])[i] = ~0ull;
    }
    for (int i = tid; i < NH * (BINS / 32); i += SEG_THREADS) (&S.flag20[0][0])[i] = 0;
    for (int i = tid; i < NC * (BINS / 32); i += SEG_THREADS) (&S.bmap[0][0])[i] = 0;
    for (int i = tid; i < NC * 2 * BINS / 4; i += SEG_THREADS) reinterpret_cast<uint32_t*>(&S.hit8[0][0])[i] = 0;
    if (tid < SEG_MAXQ) S.cursor[tid] = 0;
    if (tid == 0) {
      S.nbuckets = 0;
      S.overflow = 0;
    }
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      const uint64_t a = kvm_warp_max(s_inv[c]), b = kvm_warp_max(s_mx[c]);
      if (lane == 0) {
        S.s_inv[warp][c] = a;
        S.s_max[warp][c] = b;
      }
    }
    __syncthreads();
    if (tid < NC) {
      uint64_t inv = 0, mx = 0;
      for (int w = 0; w < SEG_WARPS; ++w) {
        inv = S.s_inv[w][tid] > inv ? S.s_inv[w][tid] : inv;
        mx = S.s_max[w][tid] > mx ? S.s_max[w][tid] : mx;
      }
      fused_cols_setup(S.fc, tid, inv, mx);
    }
    __syncthreads();
    if (tid == 0) fused_cols_fast_params(S.fc);
    __syncthreads();
    phase(0);

    // ---------------- P1 stream: two consecutive rows per thread, one 128-bit load per column
    const double cold_first = cold.n > 0 ? cold.lo[0] : 0.0, cold_last = cold.n > 0 ? cold.hi[cold.n - 1] : -1.0;
    FAcc a;
    a.init();
    auto one_row = [&](int64_t i) {
      uint32_t f;
      recs[i] = fused_row(a, S.fc, sh, cold, cold_first, cold_last, log.start_ms[i], log.latency_ms[i], log.ttfb_ms[i],
                          log.tllt_ms[i], log.status[i], log.completion_tokens[i], log.total_tokens[i], f);
      flags[i] = (uint8_t)f;
    };
    // columns are 16-byte aligned at even rows (base_aligned): an odd first / last row is peeled
    const int64_t a0 = base_aligned ? r0 + (r0 & 1) : r1;
    const int64_t npairs = a0 < r1 ? (r1 - a0) >> 1 : 0;
    if (base_aligned) {
      if (tid == 0 && a0 > r0 && r0 < r1) one_row(r0);
      if (tid == 32 && a0 < r1 && ((r1 - a0) & 1)) one_row(r1 - 1);
    } else {
      for (int64_t i = r0 + tid; i < r1; i += SEG_THREADS) {
        uint32_t f;
        recs[i] = fused_row(a, S.fc, sh, cold, cold_first, cold_last, kvm_ld1(log.start_ms + i), kvm_ld1(log.latency_ms + i),
                            kvm_ld1(log.ttfb_ms + i), kvm_ld1(log.tllt_ms + i), kvm_ld1(log.status + i),
                            kvm_ld1(log.completion_tokens + i), kvm_ld1(log.total_tokens + i), f);
        flags[i] = (uint8_t)f;
      }
    }
    for (int64_t p = tid; p < npairs; p += SEG_THREADS) {
      const int64_t i = a0 + (p << 1);
      // next iteration's lines on their way from DRAM to L2 while this one computes (one lane per 128-byte line)
      if ((lane & 7) == 0 && p + SEG_THREADS < npairs) {
        const int64_t j = i + 2 * SEG_THREADS;
        kvm_prefetch_l2(log.start_ms + j); kvm_prefetch_l2(log.latency_ms + j); kvm_prefetch_l2(log.ttfb_ms + j);
        kvm_prefetch_l2(log.tllt_ms + j); kvm_prefetch_l2(log.status + j); kvm_prefetch_l2(log.completion_tokens + j);
        kvm_prefetch_l2(log.total_tokens + j);
      }
      const double2 qs = kvm_ld2(log.start_ms + i), ql = kvm_ld2(log.latency_ms + i);
      const double2 qf = kvm_ld2(log.ttfb_ms + i), qt = kvm_ld2(log.tllt_ms + i);
      const longlong2 qx = kvm_ld2(log.status + i), qc = kvm_ld2(log.completion_tokens + i);
      const longlong2 qo = kvm_ld2(log.total_tokens + i);
      uint32_t f0, f1;
      const uint64_t q0 = fused_row(a, S.fc, sh, cold, cold_first, cold_last, qs.x, ql.x, qf.x, qt.x, qx.x, qc.x, qo.x, f0);
      const uint64_t q1 = fused_row(a, S.fc, sh, cold, cold_first, cold_last, qs.y, ql.y, qf.y, qt.y, qx.y, qc.y, qo.y, f1);
      *reinterpret_cast<ulonglong2*>(recs + i) = make_ulonglong2(q0, q1);
      *reinterpret_cast<uint16_t*>(flags + i) = (uint16_t)(f0 | (f1 << 8));
    }
    facc_warp_reduce(a, S.red[warp], lane);
    __syncthreads();
    phase(1);
    // totals of every 32-bin group of every set (one warp reduction each); a set's count is the sum of
    // its 64 group totals
    for (int g = warp; g < KVM_LOG_NSETS * (BINS / 32); g += SEG_WARPS) {
      const uint32_t v = __reduce_add_sync(KVM_FULL, sh[g * 32 + lane]);
      if (lane == 0) (&S.gsum[0][0])[g] = v;
    }
    __syncthreads();
    if (warp < KVM_LOG_NSETS) {
      const uint32_t v = __reduce_add_sync(KVM_FULL, S.gsum[warp][lane] + S.gsum[warp][32 + lane]);
      if (lane == 0)
        for (int w = 0; w < SEG_WARPS; ++w) S.red[w][S_CNT0 + warp] = w == 0 ? v : 0u;
    }
    __syncthreads();
    if (tid < NMAX_END) {
      const int s = tid;
      uint64_t acc;
      if (s >= M_KMIN0 && s < M_KMIN0 + NC) acc = S.fc.kmin[s - M_KMIN0];
      else if (s >= X_KMAX0 && s < X_KMAX0 + NC) acc = S.fc.kmax[s - X_KMAX0];
      else {
        acc = S.red[0][s];
        for (int w = 1; w < SEG_WARPS; ++w) acc = slot_combine(s, acc, S.red[w][s]);
      }
      S.fin[s] = acc;
    } else if (tid == NMAX_END || tid == NMAX_END + 1) {
      const int s = tid == NMAX_END ? D_TTFB : D_LAT;
      kvm_dd acc = dd_from_bits(S.red[0][s], S.red[0][s + 1]);
      for (int w = 1; w < SEG_WARPS; ++w) acc = kvm_dd_add(acc, dd_from_bits(S.red[w][s], S.red[w][s + 1]));
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
    if (tid < NC) {   // which end bins hold keys from outside the sampled hull
      int imp = 0;
      const uint64_t kmin = S.stats.sets.kmin[tid], kmax = S.stats.sets.kmax[tid], kb = S.fc.kbase[tid];
      if (kmin <= kmax) {
        if (kmin < kb) imp |= 1;
        if (kmax >= kb && ((kmax - kb) >> S.fc.shift[tid]) >= (uint64_t)BINS) imp |= 2;
      }
      S.impure[tid] = imp;
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
    if (tid >= 96 && tid < 96 + NH * (SEG_NB_MAX - 1)) {
      const int t = tid - 96, h = t / (SEG_NB_MAX - 1), b = 1 + t % (SEG_NB_MAX - 1);
      const int c = hist_col(h);
      const uint64_t klo = S.stats.sets.kmin[c], khi = S.stats.sets.kmax[c];
      if (b == 1) S.active20[h] = klo < khi ? 1 : 0;
      if (b < nb && klo < khi) S.thr[h][b - 1] = bin_threshold(klo, khi, nb, b);
    }
    __syncthreads();
    // fine bins that must be binned row by row: those that contain a 20-bin threshold, and the
    // end bins that hold keys from outside the sampled hull
    if (tid < NH * (SEG_NB_MAX - 1)) {
      const int h = tid / (SEG_NB_MAX - 1), b = 1 + tid % (SEG_NB_MAX - 1);
      const uint64_t T = S.thr[h][b - 1];
      const int c = hist_col(h);
      if (b < nb && S.active20[h] && T != ~0ull) {
        const uint32_t d = digit_of(T, S.fc.kbase[c], S.fc.shift[c]);
        atomicOr(&S.flag20[h][d >> 5], 1u << (d & 31));
      }
      if (b == 1 && S.active20[h]) {
        if (S.impure[c] & 1) atomicOr(&S.flag20[h][0], 1u);
        if (S.impure[c] & 2) atomicOr(&S.flag20[h][(BINS - 1) >> 5], 1u << ((BINS - 1) & 31));
      }
    }
    __syncthreads();
    // ---------------- P4 resolve (one warp per query)
    for (int qi = warp; qi < qp.nq; qi += SEG_WARPS) {
      SegQ s = S.q[qi];
      if (s.status != KVM_Q_OK) continue;
      const uint32_t mask = qp.q[qi].set_mask;
      const uint64_t ckmin = S.stats.sets.kmin[s.col], ckmax = S.stats.sets.kmax[s.col];
      if (ckmin == ckmax) {   // one distinct key in the column
        if (lane == 0) {
          s.key = ckmin;
          s.resolved = 1;
          S.q[qi] = s;
        }
        continue;
      }
      // two-level walk: 64 group totals (2 per lane) locate the 32-bin group, one scan finishes
      int64_t g0 = 0, g1 = 0;
      for (int b = 0; b < KVM_LOG_NSETS; ++b)
        if ((mask >> b) & 1u) {
          g0 += (int64_t)S.gsum[b][lane];
          g1 += (int64_t)S.gsum[b][32 + lane];
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
      i1 += __shfl_sync(KVM_FULL, i0, 31);
      const unsigned h0 = __ballot_sync(KVM_FULL, i0 > s.rank), h1 = __ballot_sync(KVM_FULL, i1 > s.rank);
      bool found = (h0 | h1) != 0;
      if (found) {
        const int gl = h0 ? __ffs(h0) - 1 : __ffs(h1) - 1;
        const int base = (h0 ? gl : 32 + gl) * 32;
        const int64_t running = h0 ? __shfl_sync(KVM_FULL, i0 - g0, gl) : __shfl_sync(KVM_FULL, i1 - g1, gl);
        int64_t c = 0;
        for (int b = 0; b < KVM_LOG_NSETS; ++b)
          if ((mask >> b) & 1u) c += (int64_t)sh[b * BINS + base + lane];
        int64_t incl = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const int64_t o = __shfl_up_sync(KVM_FULL, incl, d);
          if (lane >= d) incl += o;
        }
        const unsigned hit = __ballot_sync(KVM_FULL, running + incl > s.rank);
        found = hit != 0;
        if (hit) {
          const int l = __ffs(hit) - 1;
          const int64_t excl = __shfl_sync(KVM_FULL, incl - c, l);
          const int64_t cnt = __shfl_sync(KVM_FULL, c, l);
          if (lane == 0) {
            const int bin = base + l;
            S.q_bin[qi] = (uint32_t)bin;
            S.q_cnt[qi] = (uint32_t)cnt;
            s.resid = s.rank - (running + excl);
            const int imp = S.impure[s.col];
            const bool pure = !((bin == 0 && (imp & 1)) || (bin == BINS - 1 && (imp & 2)));
            if (S.fc.shift[s.col] == 0 && pure) {   // one key per bin
              s.key = S.fc.kbase[s.col] + (uint64_t)bin;
              s.resolved = 1;
            } else {
              s.resolved = 0;
              s.bucket = -2;
            }
            S.q[qi] = s;
          }
        }
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
          SegBucket b;
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
      S.arena_used = off;
      if (off > (uint32_t)SEG_ARENA || r1 - r0 >= (1ll << 32)) S.overflow = 1;
    }
    __syncthreads();
    if (S.overflow) {  // host redoes this run through the flat path
      if (tid == 0) {
        out_status[run] = 1;
        out_stats[run] = S.stats;
      }
      __syncthreads();
      run = next_run;
      __syncthreads();
      continue;
    }
    // ---------------- P5 fine bins without an edge lie inside a single 20-bin: add them wholesale
    for (int t = tid; t < NH * BINS; t += SEG_THREADS) {
      const int h = t / BINS, d = t % BINS;
      if (!S.active20[h] || ((S.flag20[h][d >> 5] >> (d & 31)) & 1u)) continue;
      const uint32_t sets = hist_sets(h);
      uint32_t c = 0;
      for (int b = 0; b < KVM_LOG_NSETS; ++b)
        if ((sets >> b) & 1u) c += sh[b * BINS + d];
      if (c) {
        const int col = hist_col(h);
        const uint64_t lowkey = S.fc.kbase[col] + ((uint64_t)d << S.fc.shift[col]);
        atomicAdd(&S.cnt20[h][bin_of(lowkey, S.thr[h], nb)], c);
      }
    }
    for (int t = tid; t < NC * BINS; t += SEG_THREADS) {
      const int c = t / BINS, d = t % BINS;
      const int h = col_hist(c);
      const uint32_t w = S.bmap[c][d >> 5] | (h >= 0 ? S.flag20[h][d >> 5] : 0u);
      S.hit8[c][d] = (uint8_t)(((w >> (d & 31)) & 1u) << c);
    }
    __syncthreads();   // the histograms are dead from here on: their space is the arena
    phase(2);
    // ---------------- P6 compact + exact binning of the rows in edge bins, from the records.
    // A wanted (row, column) is rare and costs a dependent fetch: the warp queues them and serves
    // 32 at a time with every lane busy, instead of stalling on each where it is found.
    {
      auto rec_row = [&](int64_t i, uint64_t r, uint32_t fl) {
#pragma unroll
        for (int c = 0; c < NC; ++c) {
          const uint32_t d = (uint32_t)(r >> (12 * c)) & 0xFFFu;
          if (S.hit8[c][d]) seg_hit(S, log, i, c, d, fl, nb);
        }
      };
      if (base_aligned) {
        if (tid == 0 && a0 > r0 && r0 < r1) rec_row(r0, recs[r0], flags[r0]);
        if (tid == 32 && a0 < r1 && ((r1 - a0) & 1)) rec_row(r1 - 1, recs[r1 - 1], flags[r1 - 1]);
      } else {
        for (int64_t i = r0 + tid; i < r1; i += SEG_THREADS) rec_row(i, recs[i], flags[i]);
      }
      uint64_t* hq = S.hitq[warp];
      int qn = 0;   // warp-uniform copy of hitq_n[warp]
      if (lane == 0) S.hitq_n[warp] = 0;
      __syncwarp();
      auto flush = [&]() {
        __syncwarp();
        for (int e0 = 0; e0 < qn; e0 += 64) {
          uint64_t w[2];
          double v[2];
#pragma unroll
          for (int u = 0; u < 2; ++u) {   // the fetches of both entries are in flight together
            const int e = e0 + u * 32 + lane;
            w[u] = e < qn ? hq[e] : ~0ull;
            if (e < qn) v[u] = seg_value(log, r0 + (int64_t)(uint32_t)w[u], (int)((w[u] >> 44) & 7u));
          }
#pragma unroll
          for (int u = 0; u < 2; ++u)
            if (e0 + u * 32 + lane < qn)
              seg_place(S, (int)((w[u] >> 44) & 7u), (uint32_t)(w[u] >> 32) & 0xFFFu, (uint32_t)(w[u] >> 48) & 0xFFu,
                        kvm_key(v[u]), nb);
        }
        if (lane == 0) S.hitq_n[warp] = 0;
        __syncwarp();
        qn = 0;
      };
      // four row pairs per lane and iteration (every load of a warp covers 512 consecutive bytes)
      for (int64_t pb = (int64_t)warp * (32 * SEG_RK); pb < npairs; pb += (int64_t)SEG_THREADS * SEG_RK) {
        ulonglong2 rr[SEG_RK];
        uint32_t ff[SEG_RK];
#pragma unroll
        for (int k = 0; k < SEG_RK; ++k) {
          const int64_t p = pb + k * 32 + lane;
          rr[k] = make_ulonglong2(~0ull, ~0ull);   // fields 0xFFF: wanted nowhere
          ff[k] = 0;
          if (p < npairs) {
            rr[k] = *reinterpret_cast<const ulonglong2*>(recs + a0 + (p << 1));
            ff[k] = *reinterpret_cast<const uint16_t*>(flags + a0 + (p << 1));
          }
        }
        // one byte per (column, digit): bit c set when the row's value in column c is wanted
        uint64_t hm = 0;
#pragma unroll
        for (int k = 0; k < SEG_RK; ++k) {
          uint32_t m0 = 0, m1 = 0;
#pragma unroll
          for (int c = 0; c < NC; ++c) {
            m0 |= S.hit8[c][(uint32_t)(rr[k].x >> (12 * c)) & 0xFFFu];
            m1 |= S.hit8[c][(uint32_t)(rr[k].y >> (12 * c)) & 0xFFFu];
          }
          hm |= (uint64_t)(m0 | (m1 << NC)) << (k * 2 * NC);
        }
        const int nh = __popcll(hm);
        const int total = __reduce_add_sync(KVM_FULL, nh);
        if (total == 0) continue;
        if (total > SEG_DIRECT || qn + total > SEG_HITQ) flush();
        const bool direct = total > SEG_DIRECT;   // a dense stretch: served where found
        int pos = 0;
        if (!direct && nh) pos = (int)atomicAdd(&S.hitq_n[warp], (uint32_t)nh);
        while (hm) {
          const int b = __ffsll((long long)hm) - 1;
          hm &= hm - 1;
          const int k = b / (2 * NC), second = (b % (2 * NC)) >= NC, c = b % NC;
          uint64_t r = rr[0].x;
          uint32_t f = ff[0];
#pragma unroll
          for (int kk = 0; kk < SEG_RK; ++kk)
            if (kk == k) {
              r = second ? rr[kk].y : rr[kk].x;
              f = ff[kk];
            }
          const uint32_t fl = second ? f >> 8 : f & 0xffu;
          const int64_t i = a0 + ((pb + k * 32 + lane) << 1) + second;
          const uint64_t d = (r >> (12 * c)) & 0xFFFull;
          if (direct) seg_hit(S, log, i, c, (uint32_t)d, fl, nb);
          else hq[pos++] = (uint64_t)(uint32_t)(i - r0) | (d << 32) | ((uint64_t)c << 44) | ((uint64_t)fl << 48);
        }
        if (!direct) qn += total;
      }
      flush();
    }
    __syncthreads();
    phase(3);
    // ---------------- P7 finish: a warp per query selects the wanted rank among its bucket's keys by
    // radix steps over the bits that differ inside the bucket, five bits (one lane per bin) a step
    {
      const uint64_t* arena = reinterpret_cast<const uint64_t*>(sh);
      uint32_t* cnt = reinterpret_cast<uint32_t*>(S.hitq[warp]);
      for (int qi = warp; qi < qp.nq; qi += SEG_WARPS) {
        const SegQ s = S.q[qi];
        if (s.status != KVM_Q_OK || s.resolved) continue;
        const SegBucket b = S.bk[s.bucket];
        const uint64_t* keys = arena + b.offset;
        const uint32_t m = b.count;
        const uint64_t k0 = keys[0];
        uint64_t diff = 0;
        for (uint32_t i = lane; i < m; i += 32) diff |= keys[i] ^ k0;
        diff = ((uint64_t)__reduce_or_sync(KVM_FULL, (uint32_t)(diff >> 32)) << 32) |
               (uint64_t)__reduce_or_sync(KVM_FULL, (uint32_t)diff);
        int top = bit_width64(diff);   // the low `top` bits vary inside the bucket
        uint64_t fixed = top >= 64 ? 0ull : ~0ull << top, value = k0 & fixed;
        int64_t r = s.resid;
        while (top > 0) {
          const int w = top >= 5 ? 5 : top, sft = top - w;
          cnt[lane] = 0;
          __syncwarp();
          for (uint32_t i = lane; i < m; i += 32) {
            const uint64_t k = keys[i];
            if ((k & fixed) == value) atomicAdd(&cnt[(uint32_t)(k >> sft) & ((1u << w) - 1u)], 1u);
          }
          __syncwarp();
          const uint32_t c = cnt[lane];
          uint32_t incl = c;
#pragma unroll
          for (int d = 1; d < 32; d <<= 1) {
            const uint32_t o = __shfl_up_sync(KVM_FULL, incl, d);
            if (lane >= d) incl += o;
          }
          const unsigned hit = __ballot_sync(KVM_FULL, (int64_t)incl > r);
          const int l = hit ? __ffs(hit) - 1 : 31;
          r -= (int64_t)__shfl_sync(KVM_FULL, incl - c, l);
          value |= (uint64_t)l << sft;
          fixed |= (uint64_t)((1u << w) - 1u) << sft;
          top = sft;
          __syncwarp();
        }
        if (lane == 0) S.q[qi].key = value;
      }
    }
    __syncthreads();
    // ---------------- P8 emit
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
    phase(4);
    run = next_run;
    __syncthreads();   // (next_run is rewritten at the top of the next run)
  }
}

}  // namespace

// debug aid: the per-phase nanoseconds summed over all runs since the last reset (synchronises)
extern "C" int kvm_seg_phase_times(uint64_t* out5, int32_t reset) {
  KVM_REQUIRE(out5, "kvm_seg_phase_times: null argument");
  unsigned long long h[8];
  KVM_CUDA_CHECK(cudaMemcpyFromSymbol(h, seg_phase_ns, sizeof(h)));
  for (int i = 0; i < 5; ++i) out5[i] = h[i];
  if (reset) {
    for (int i = 0; i < 8; ++i) h[i] = 0;
    KVM_CUDA_CHECK(cudaMemcpyToSymbol(seg_phase_ns, h, sizeof(h)));
  }
  return KVM_OK;
}

// 8 B/row digit records (+ 2 rows of slack) and one 16-byte slot for the run ticket
extern "C" size_t kvm_seg_analyze_scratch_bytes(int64_t n) { return sizeof(uint64_t) * (size_t)(n < 0 ? 0 : n + 2) + 16; }

extern "C" int kvm_seg_analyze(const kvm_request_log* log, const int64_t* run_offsets, int64_t runs,
                               const int64_t* cold_offsets, const double* cold_times,
                               double cold_window_s, const kvm_query* queries, int32_t nq,
                               int32_t num_bins, uint8_t* flags, void* scratch, size_t scratch_bytes,
                               kvm_log_stats* stats, kvm_query_result* results, int64_t* hist_counts,
                               int32_t* run_status, void* stream) {
  KVM_REQUIRE(log && run_offsets && queries && flags && scratch && stats && results && hist_counts && run_status,
              "kvm_seg_analyze: null argument");
  KVM_REQUIRE(runs >= 0, "kvm_seg_analyze: negative run count");
  KVM_REQUIRE(nq >= 1 && nq <= SEG_MAXQ, "kvm_seg_analyze: nq=%d outside [1,%d]", nq, SEG_MAXQ);
  KVM_REQUIRE(num_bins >= 1 && num_bins <= SEG_NB_MAX, "kvm_seg_analyze: num_bins=%d outside [1,%d]",
              num_bins, SEG_NB_MAX);
  KVM_REQUIRE(cold_offsets == nullptr || cold_times != nullptr, "kvm_seg_analyze: cold_times is null");
  if (scratch_bytes < kvm_seg_analyze_scratch_bytes(log->n)) {
    kvm_set_error("kvm_seg_analyze: scratch %zu < %zu bytes", scratch_bytes, kvm_seg_analyze_scratch_bytes(log->n));
    return KVM_ERR_WORKSPACE;
  }
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
                       al16(log->total_tokens) && (reinterpret_cast<uintptr_t>(flags) & 1u) == 0 && al16(scratch);
  int grid = 2 * kvm_num_sms();
  if ((int64_t)grid > runs) grid = (int)runs;
  unsigned int* ticket = reinterpret_cast<unsigned int*>(reinterpret_cast<uint64_t*>(scratch) + (log->n + 2));
  KVM_CUDA_CHECK(cudaMemsetAsync(ticket, 0, 16, (cudaStream_t)stream));
  seg_analyze_kernel<<<grid, SEG_THREADS, smem, (cudaStream_t)stream>>>(
      *log, run_offsets, runs, cold_offsets, cold_times, cold_window_s, qp, num_bins, aligned, flags,
      reinterpret_cast<uint64_t*>(scratch), stats, results, hist_counts, run_status, ticket);
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}
