// kvm_stats.cu - K1 (+K7): fused first pass over the request log, and the generic
// one-column count/min/max pass that feeds the select planner.
//
// Replaces (reference file:line)
//   analyze.py:183-189   window bounds            -> min_start_ms / max_end_ms
//   analyze.py:402-419   cold/warm classification -> KVM_F_COLD flag
//   analyze.py:496-514   success / token sums     -> int64 sums (exact)
//   analyze.py:68,89,143 NaN / tokens>0 filters   -> per-set counts + key ranges
//   analyze.py:581       statistics.mean(ttfs)    -> double-double sum
//   cost_estimator.py:117-124,293-308; cost_calculator.py:42-50
//
// Roofline: HBM.  Algorithmic bytes = 56 B/row read (4 fp64 + 3 int64) + 1 B/row written.
// Layout: each thread owns two consecutive rows per tile so that every column is read
// with one 128-bit load per thread; a warp reads 512 contiguous bytes per column.

#include "kvm_common.cuh"
#include "kvm_log_device.cuh"

namespace {

constexpr int STATS_THREADS = 256;
constexpr int STATS_WARPS = STATS_THREADS / 32;
constexpr int COLD_SMEM = 512;  // cold-start intervals staged in shared memory

struct StatsWs {
  unsigned int ticket;
  unsigned int pad[63];
  uint64_t partial[1];  // [grid][NSLOTS]
};

// One thread per cold start: the closed start_ms interval that the reference classifies cold.
__global__ void cold_table_kernel(const double* __restrict__ cold_times, int nc, double window,
                                  double* __restrict__ lo, double* __restrict__ hi) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nc) cold_interval(cold_times[i], window, lo + i, hi + i);
}

template <bool VEC, bool COLD_SMEM_TABLE>
__global__ void __launch_bounds__(STATS_THREADS, 2)
log_scan_kernel(kvm_request_log log, const double* __restrict__ cold_lo_g,
                const double* __restrict__ cold_hi_g, int nc,
                const uint8_t* __restrict__ cold_mask, uint8_t* __restrict__ flags,
                kvm_log_stats* __restrict__ out, StatsWs* ws) {
  __shared__ double cold_lo_s[COLD_SMEM], cold_hi_s[COLD_SMEM];
  __shared__ uint64_t red[STATS_WARPS][NSLOTS];
  __shared__ bool is_last;

  // the table pointer is known at compile time to be shared or global (no generic loads)
  ColdTable cold{COLD_SMEM_TABLE ? cold_lo_s : cold_lo_g, COLD_SMEM_TABLE ? cold_hi_s : cold_hi_g, nc};
  if (COLD_SMEM_TABLE) {
    for (int i = threadIdx.x; i < nc; i += STATS_THREADS) {
      cold_lo_s[i] = cold_lo_g[i];
      cold_hi_s[i] = cold_hi_g[i];
    }
    __syncthreads();
  }

  LogAcc a;
  a.init();
  const int64_t n = log.n;

  if (VEC) {
    // software-pipelined: the seven 128-bit loads of the next pair are in flight while the
    // current pair is processed (the per-row arithmetic is a long dependent chain)
    const int64_t npairs = n >> 1;
    const int64_t stride = (int64_t)gridDim.x * STATS_THREADS;
    int64_t p = (int64_t)blockIdx.x * STATS_THREADS + threadIdx.x;
    double2 st, la, tf, tl;
    longlong2 ss, ct, tt;
    uint16_t cm = 0;
    auto load = [&](int64_t q, double2& st_, double2& la_, double2& tf_, double2& tl_, longlong2& ss_,
                    longlong2& ct_, longlong2& tt_, uint16_t& cm_) {
      const int64_t i = q << 1;
      st_ = kvm_ld2(log.start_ms + i);
      la_ = kvm_ld2(log.latency_ms + i);
      tf_ = kvm_ld2(log.ttfb_ms + i);
      tl_ = kvm_ld2(log.tllt_ms + i);
      ss_ = kvm_ld2(log.status + i);
      ct_ = kvm_ld2(log.completion_tokens + i);
      tt_ = kvm_ld2(log.total_tokens + i);
      if (cold_mask) cm_ = kvm_ld_u8x2(cold_mask + i);
    };
    if (p < npairs) load(p, st, la, tf, tl, ss, ct, tt, cm);
    while (p < npairs) {
      const int64_t pn = p + stride;
      double2 st2 = st, la2 = la, tf2 = tf, tl2 = tl;
      longlong2 ss2 = ss, ct2 = ct, tt2 = tt;
      uint16_t cm2 = cm;
      if (pn < npairs) load(pn, st2, la2, tf2, tl2, ss2, ct2, tt2, cm2);
      const int64_t i = p << 1;
      const int c0 = cold_mask ? (int)(cm & 0xff) : -1, c1 = cold_mask ? (int)(cm >> 8) : -1;
      const uint8_t f0 = process_row(a, st.x, la.x, tf.x, tl.x, ss.x, ct.x, tt.x, cold, c0);
      const uint8_t f1 = process_row(a, st.y, la.y, tf.y, tl.y, ss.y, ct.y, tt.y, cold, c1);
      if (flags) *reinterpret_cast<uint16_t*>(flags + i) = (uint16_t)f0 | ((uint16_t)f1 << 8);
      st = st2; la = la2; tf = tf2; tl = tl2; ss = ss2; ct = ct2; tt = tt2; cm = cm2;
      p = pn;
    }
    if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
      const int64_t i = n - 1;
      const uint8_t f = process_row(a, log.start_ms[i], log.latency_ms[i], log.ttfb_ms[i],
                                    log.tllt_ms[i], log.status[i], log.completion_tokens[i],
                                    log.total_tokens[i], cold, cold_mask ? (int)cold_mask[i] : -1);
      if (flags) flags[i] = f;
    }
  } else {
    for (int64_t i = (int64_t)blockIdx.x * STATS_THREADS + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * STATS_THREADS) {
      const uint8_t f = process_row(a, kvm_ld1(log.start_ms + i), kvm_ld1(log.latency_ms + i),
                                    kvm_ld1(log.ttfb_ms + i), kvm_ld1(log.tllt_ms + i),
                                    kvm_ld1(log.status + i), kvm_ld1(log.completion_tokens + i),
                                    kvm_ld1(log.total_tokens + i), cold,
                                    cold_mask ? (int)kvm_ld_u8(cold_mask + i) : -1);
      if (flags) flags[i] = f;
    }
  }

  // ---- warp reduction, then fixed-order reduction across warps and across CTAs
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  acc_warp_reduce(a, red[warp], lane);
  __syncthreads();

  auto combine = [](int slot, uint64_t acc, uint64_t v) -> uint64_t {
    if (slot < NSUM) return (uint64_t)((int64_t)acc + (int64_t)v);
    if (slot < NMIN_END) return v < acc ? v : acc;
    return v > acc ? v : acc;
  };
  auto dd_of = [](uint64_t hi, uint64_t lo) -> kvm_dd {
    return kvm_dd{__longlong_as_double((long long)hi), __longlong_as_double((long long)lo)};
  };

  uint64_t* mine = ws->partial + (size_t)blockIdx.x * NSLOTS;
  if (threadIdx.x < NMAX_END) {
    const int s = threadIdx.x;
    uint64_t acc = red[0][s];
    for (int w = 1; w < STATS_WARPS; ++w) acc = combine(s, acc, red[w][s]);
    mine[s] = acc;
  } else if (threadIdx.x == NMAX_END || threadIdx.x == NMAX_END + 1) {
    const int s = threadIdx.x == NMAX_END ? D_TTFB : D_LAT;
    kvm_dd acc = dd_of(red[0][s], red[0][s + 1]);
    for (int w = 1; w < STATS_WARPS; ++w) acc = kvm_dd_add(acc, dd_of(red[w][s], red[w][s + 1]));
    mine[s] = (uint64_t)__double_as_longlong(acc.hi);
    mine[s + 1] = (uint64_t)__double_as_longlong(acc.lo);
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int t = atomicAdd(&ws->ticket, 1u);
    is_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();

  // last CTA: warp w reduces slots w, w+8, ... - its lanes stride over the CTAs' partials (the
  // loads are independent and go to L2), then a fixed-shape shuffle tree combines the lanes.
  // Integer sums / min / max are order-independent; the double-double sums always follow the
  // same tree for a given grid size, so the result is run-to-run deterministic.
  __shared__ uint64_t fin[NSLOTS];
  const uint64_t* part = ws->partial;
  for (int s = warp; s < NSLOTS; s += STATS_WARPS) {
    if (s < NMAX_END) {
      uint64_t acc = s < NSUM ? 0ull : (s < NMIN_END ? ~0ull : 0ull);
      for (unsigned g = lane; g < gridDim.x; g += 32) acc = combine(s, acc, __ldcg(part + (size_t)g * NSLOTS + s));
#pragma unroll
      for (int m = 16; m; m >>= 1) acc = combine(s, acc, kvm_shfl_xor(acc, m));
      if (lane == 0) fin[s] = acc;
    } else if (s == D_TTFB || s == D_LAT) {
      kvm_dd acc{0.0, 0.0};
      for (unsigned g = lane; g < gridDim.x; g += 32) {
        const size_t o = (size_t)g * NSLOTS + s;
        acc = kvm_dd_add(acc, dd_of(__ldcg(part + o), __ldcg(part + o + 1)));
      }
      acc = kvm_warp_sum(acc);
      if (lane == 0) {
        fin[s] = (uint64_t)__double_as_longlong(acc.hi);
        fin[s + 1] = (uint64_t)__double_as_longlong(acc.lo);
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    kvm_log_stats r;
    stats_from_slots(fin, n, r);
    *out = r;
    ws->ticket = 0;  // ready for the next launch on this workspace
  }
}

// ------------------------------------------------------------------ generic column stats
constexpr int CS_SLOTS = KVM_MAX_SETS + 2;  // counts, kmin, kmax

struct ColStatsWs {
  unsigned int ticket;
  unsigned int pad[63];
  uint64_t partial[1];  // [grid][CS_SLOTS]
};

__device__ __forceinline__ int member_set(int mode, const void* member, int64_t i) {
  if (mode == KVM_MEMBER_ALL) return 0;
  if (mode == KVM_MEMBER_SET_IDS) return (int)kvm_ld_u8(reinterpret_cast<const uint8_t*>(member) + i);
  return kvm_ld1(reinterpret_cast<const int64_t*>(member) + i) == 200 ? 0 : 255;
}

// MODE is a compile-time constant so the single-set modes (all / status==200) stream with
// 128-bit loads and one counter; only the set-id mode pays for per-set counters.
template <int MODE, bool VEC>
__global__ void __launch_bounds__(STATS_THREADS, 4)
column_stats_kernel(const double* __restrict__ values, int64_t n, const void* __restrict__ member,
                    int nsets, kvm_set_info* out, ColStatsWs* ws) {
  __shared__ uint64_t red[STATS_WARPS][CS_SLOTS];
  __shared__ bool is_last;
  int64_t cnt[KVM_MAX_SETS];
#pragma unroll
  for (int s = 0; s < KVM_MAX_SETS; ++s) cnt[s] = 0;
  const double inf = __longlong_as_double(0x7ff0000000000000ll);
  double vmin = inf, vmax = -inf;
  int32_t c0 = 0;
  if (MODE != KVM_MEMBER_SET_IDS && VEC) {
    const int64_t npairs = n >> 1;
    const int64_t* st = reinterpret_cast<const int64_t*>(member);
    for (int64_t p = (int64_t)blockIdx.x * STATS_THREADS + threadIdx.x; p < npairs;
         p += (int64_t)gridDim.x * STATS_THREADS) {
      const int64_t i = p << 1;
      double2 v = kvm_ld2(values + i);
      if (MODE == KVM_MEMBER_STATUS) {
        const longlong2 s2 = kvm_ld2(st + i);
        if (s2.x != 200) v.x = __longlong_as_double(0x7ff8000000000000ll);
        if (s2.y != 200) v.y = __longlong_as_double(0x7ff8000000000000ll);
      }
      c0 += (v.x == v.x) + (v.y == v.y);
      vmin = fmin(vmin, fmin(v.x, v.y));   // fmin / fmax skip NaN
      vmax = fmax(vmax, fmax(v.x, v.y));
    }
    if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
      const int64_t i = n - 1;
      double v = values[i];
      if (MODE == KVM_MEMBER_STATUS && st[i] != 200) v = __longlong_as_double(0x7ff8000000000000ll);
      c0 += (v == v);
      vmin = fmin(vmin, v);
      vmax = fmax(vmax, v);
    }
    cnt[0] = c0;
  } else {
    for (int64_t i = (int64_t)blockIdx.x * STATS_THREADS + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * STATS_THREADS) {
      const double v = kvm_ld1(values + i);
      const int s = member_set(MODE, member, i);
      if (s < nsets && !kvm_isnan(v)) {
#pragma unroll
        for (int k = 0; k < KVM_MAX_SETS; ++k) cnt[k] += (s == k);
        vmin = fmin(vmin, v);
        vmax = fmax(vmax, v);
      }
    }
  }
  // ordered keys; a zero bound covers both signed zeros; untouched range -> kmin > kmax
  uint64_t kmin = vmin == 0.0 ? kvm_key(-0.0) : kvm_key(vmin);
  uint64_t kmax = vmax == 0.0 ? kvm_key(0.0) : kvm_key(vmax);
  if (vmin > vmax) {
    kmin = ~0ull;
    kmax = 0ull;
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int s = 0; s < KVM_MAX_SETS; ++s) {
    int64_t v = kvm_warp_sum(cnt[s]);
    if (lane == 0) red[warp][s] = (uint64_t)v;
  }
  kmin = kvm_warp_min(kmin);
  kmax = kvm_warp_max(kmax);
  if (lane == 0) {
    red[warp][KVM_MAX_SETS] = kmin;
    red[warp][KVM_MAX_SETS + 1] = kmax;
  }
  __syncthreads();
  auto combine = [](int slot, uint64_t acc, uint64_t v) -> uint64_t {
    if (slot < KVM_MAX_SETS) return acc + v;
    if (slot == KVM_MAX_SETS) return v < acc ? v : acc;
    return v > acc ? v : acc;
  };
  if (threadIdx.x < CS_SLOTS) {
    const int s = threadIdx.x;
    uint64_t acc = red[0][s];
    for (int w = 1; w < STATS_WARPS; ++w) acc = combine(s, acc, red[w][s]);
    ws->partial[(size_t)blockIdx.x * CS_SLOTS + s] = acc;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = (atomicAdd(&ws->ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  __shared__ uint64_t fin[CS_SLOTS];
  const uint64_t* part = ws->partial;
  for (int s = warp; s < CS_SLOTS; s += STATS_WARPS) {
    uint64_t acc = s < KVM_MAX_SETS ? 0ull : (s == KVM_MAX_SETS ? ~0ull : 0ull);
    for (unsigned g = lane; g < gridDim.x; g += 32) acc = combine(s, acc, __ldcg(part + (size_t)g * CS_SLOTS + s));
#pragma unroll
    for (int m = 16; m; m >>= 1) acc = combine(s, acc, kvm_shfl_xor(acc, m));
    if (lane == 0) fin[s] = acc;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    kvm_set_info r;
    for (int s = 0; s < KVM_MAX_SETS; ++s) {
      r.count[s] = s < nsets ? (int64_t)fin[s] : 0;
      r.set_col[s] = 0;
    }
    for (int c = 0; c < KVM_MAX_COLS; ++c) {
      r.kmin[c] = ~0ull;
      r.kmax[c] = 0ull;
    }
    r.kmin[0] = fin[KVM_MAX_SETS];
    r.kmax[0] = fin[KVM_MAX_SETS + 1];
    r.nsets = nsets;
    r.ncols = 1;
    *out = r;
    ws->ticket = 0;
  }
}

constexpr int MAX_GRID = 148 * 8;  // workspace sizing bound (B200: 148 SMs)

}  // namespace

constexpr int MAX_COLD = 16384;  // cold starts per log the workspace has room for
static size_t scan_partials_bytes() { return kvm_align_up(256 + sizeof(uint64_t) * (size_t)NSLOTS * MAX_GRID, 256); }

extern "C" size_t kvm_log_scan_workspace_bytes(void) {
  return scan_partials_bytes() + 2 * sizeof(double) * (size_t)MAX_COLD;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

extern "C" int kvm_log_scan(const kvm_request_log* log, const double* cold_times, int32_t n_cold,
                            double cold_window_s, const uint8_t* cold_mask, uint8_t* flags,
                            kvm_log_stats* stats, void* workspace, size_t workspace_bytes,
                            void* stream) {
  KVM_REQUIRE(log && stats && workspace, "kvm_log_scan: null argument");
  KVM_REQUIRE(log->n >= 0 && n_cold >= 0, "kvm_log_scan: negative size");
  KVM_REQUIRE(n_cold == 0 || cold_times, "kvm_log_scan: cold_times is null");
  KVM_REQUIRE(n_cold <= MAX_COLD, "kvm_log_scan: %d cold starts exceed the supported %d", n_cold, MAX_COLD);
  if (workspace_bytes < kvm_log_scan_workspace_bytes()) {
    kvm_set_error("kvm_log_scan: workspace %zu < %zu bytes", workspace_bytes,
                  kvm_log_scan_workspace_bytes());
    return KVM_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  StatsWs* ws = reinterpret_cast<StatsWs*>(workspace);
  KVM_CUDA_CHECK(cudaMemsetAsync(ws, 0, 256, st));
  int grid = kvm_grid_for(log->n, STATS_THREADS * 2 * 4, 2);
  if (grid > MAX_GRID) grid = MAX_GRID;
  const bool vec = aligned16(log->start_ms) && aligned16(log->latency_ms) &&
                   aligned16(log->ttfb_ms) && aligned16(log->tllt_ms) && aligned16(log->status) &&
                   aligned16(log->completion_tokens) && aligned16(log->total_tokens) &&
                   (flags == nullptr || (reinterpret_cast<uintptr_t>(flags) & 1u) == 0) &&
                   (cold_mask == nullptr || (reinterpret_cast<uintptr_t>(cold_mask) & 1u) == 0);
  double* cold_lo = reinterpret_cast<double*>(reinterpret_cast<unsigned char*>(workspace) + scan_partials_bytes());
  double* cold_hi = cold_lo + MAX_COLD;
  if (cold_mask) n_cold = 0;  // the mask replaces the time-based classification
  if (n_cold > 0)
    cold_table_kernel<<<(n_cold + 127) / 128, 128, 0, st>>>(cold_times, n_cold, cold_window_s, cold_lo, cold_hi);
  const bool smem_table = n_cold <= COLD_SMEM;
#define KVM_LAUNCH_SCAN(V, T) \
  log_scan_kernel<V, T><<<grid, STATS_THREADS, 0, st>>>(*log, cold_lo, cold_hi, n_cold, cold_mask, flags, stats, ws)
  if (vec && smem_table) KVM_LAUNCH_SCAN(true, true);
  else if (vec) KVM_LAUNCH_SCAN(true, false);
  else if (smem_table) KVM_LAUNCH_SCAN(false, true);
  else KVM_LAUNCH_SCAN(false, false);
#undef KVM_LAUNCH_SCAN
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" size_t kvm_column_stats_workspace_bytes(void) {
  return kvm_align_up(256 + sizeof(uint64_t) * (size_t)CS_SLOTS * MAX_GRID, 256);
}

extern "C" int kvm_column_stats(const double* values, int64_t n, int32_t mode, const void* member,
                                int32_t nsets, kvm_set_info* info, void* workspace,
                                size_t workspace_bytes, void* stream) {
  KVM_REQUIRE(info && workspace && (values || n == 0), "kvm_column_stats: null argument");
  KVM_REQUIRE(n >= 0 && nsets >= 1 && nsets <= KVM_MAX_SETS, "kvm_column_stats: bad n/nsets");
  KVM_REQUIRE(mode == KVM_MEMBER_ALL || member, "kvm_column_stats: member array is null");
  const size_t need = kvm_column_stats_workspace_bytes();
  if (workspace_bytes < need) {
    kvm_set_error("kvm_column_stats: workspace %zu < %zu bytes", workspace_bytes, need);
    return KVM_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  ColStatsWs* ws = reinterpret_cast<ColStatsWs*>(workspace);
  KVM_CUDA_CHECK(cudaMemsetAsync(ws, 0, 256, st));
  int grid = kvm_grid_for(n, STATS_THREADS * 8, 4);
  if (grid > MAX_GRID) grid = MAX_GRID;   // the partials buffer holds MAX_GRID rows
  const bool vec = (reinterpret_cast<uintptr_t>(values) & 15u) == 0 &&
                   (mode != KVM_MEMBER_STATUS || (reinterpret_cast<uintptr_t>(member) & 15u) == 0);
#define KVM_LAUNCH_CS(M, V) \
  column_stats_kernel<M, V><<<grid, STATS_THREADS, 0, st>>>(values, n, member, nsets, info, ws)
  if (mode == KVM_MEMBER_ALL) { if (vec) KVM_LAUNCH_CS(KVM_MEMBER_ALL, true); else KVM_LAUNCH_CS(KVM_MEMBER_ALL, false); }
  else if (mode == KVM_MEMBER_STATUS) { if (vec) KVM_LAUNCH_CS(KVM_MEMBER_STATUS, true); else KVM_LAUNCH_CS(KVM_MEMBER_STATUS, false); }
  else KVM_LAUNCH_CS(KVM_MEMBER_SET_IDS, false);
#undef KVM_LAUNCH_CS
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

// ------------------------------------------------------------------ multi-GPU combine helpers
// A log sharded over several GPUs is scanned per shard; the shard statistics are combined with
// two NCCL all-reduces the host issues between these two kernels: `sums` (int64, SUM) and
// `mins` (int64, MIN).  Minima and maxima both travel through the MIN reduction: keys are
// mapped to order-preserving signed integers and maxima are complemented (~x reverses order).
namespace {
constexpr int PACK_SUMS = 8 + KVM_MAX_SETS;
constexpr int PACK_MINS = 2 * KVM_MAX_COLS + 2;
constexpr long long MIN64 = (long long)0x8000000000000000ull;

__device__ __forceinline__ long long ordered_of_double(double v) {  // NaN handled by the caller
  const long long b = __double_as_longlong(v);
  return b < 0 ? (b ^ 0x7fffffffffffffffll) : b;
}
__device__ __forceinline__ double double_of_ordered(long long o) {
  return __longlong_as_double(o < 0 ? (o ^ 0x7fffffffffffffffll) : o);
}

__global__ void stats_pack_kernel(const kvm_log_stats* st, long long* sums, long long* mins) {
  const int t = threadIdx.x;
  const double inf = __longlong_as_double(0x7ff0000000000000ll);
  if (t == 0) {
    sums[0] = st->n_total;
    sums[1] = st->n_success;
    sums[2] = st->n_cold;
    sums[3] = st->n_cold_success;
    sums[4] = st->sum_completion_tokens;
    sums[5] = st->sum_total_tokens;
    sums[6] = st->sum_total_tokens_cold;
    sums[7] = st->n_gen_rows;
    const double a = st->min_start_ms, b = st->max_end_ms;
    mins[2 * KVM_MAX_COLS] = ordered_of_double(a != a ? inf : a);
    mins[2 * KVM_MAX_COLS + 1] = ~ordered_of_double(b != b ? -inf : b);
  }
  if (t < KVM_MAX_SETS) sums[8 + t] = st->sets.count[t];
  if (t < KVM_MAX_COLS) {
    mins[t] = (long long)(st->sets.kmin[t]) ^ MIN64;
    mins[KVM_MAX_COLS + t] = ~((long long)(st->sets.kmax[t]) ^ MIN64);
  }
}

__global__ void stats_unpack_kernel(kvm_log_stats* st, const long long* sums, const long long* mins) {
  const int t = threadIdx.x;
  const double inf = __longlong_as_double(0x7ff0000000000000ll);
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  if (t == 0) {
    st->n_total = sums[0];
    st->n_success = sums[1];
    st->n_cold = sums[2];
    st->n_cold_success = sums[3];
    st->sum_completion_tokens = sums[4];
    st->sum_total_tokens = sums[5];
    st->sum_total_tokens_cold = sums[6];
    st->n_gen_rows = sums[7];
    const double a = double_of_ordered(mins[2 * KVM_MAX_COLS]);
    const double b = double_of_ordered(~mins[2 * KVM_MAX_COLS + 1]);
    st->min_start_ms = a == inf ? nan : a;
    st->max_end_ms = b == -inf ? nan : b;
  }
  if (t < KVM_MAX_SETS) st->sets.count[t] = sums[8 + t];
  if (t < KVM_MAX_COLS) {
    st->sets.kmin[t] = (uint64_t)(mins[t] ^ MIN64);
    st->sets.kmax[t] = (uint64_t)((~mins[KVM_MAX_COLS + t]) ^ MIN64);
  }
}
}  // namespace

extern "C" int32_t kvm_log_stats_pack_sums(void) { return PACK_SUMS; }
extern "C" int32_t kvm_log_stats_pack_mins(void) { return PACK_MINS; }

extern "C" int kvm_log_stats_pack(const kvm_log_stats* stats, int64_t* sums, int64_t* mins, void* stream) {
  KVM_REQUIRE(stats && sums && mins, "kvm_log_stats_pack: null argument");
  stats_pack_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(stats, reinterpret_cast<long long*>(sums),
                                                       reinterpret_cast<long long*>(mins));
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_log_stats_unpack(kvm_log_stats* stats, const int64_t* sums, const int64_t* mins, void* stream) {
  KVM_REQUIRE(stats && sums && mins, "kvm_log_stats_unpack: null argument");
  stats_unpack_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(stats, reinterpret_cast<const long long*>(sums),
                                                         reinterpret_cast<const long long*>(mins));
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

// ---- one-collective form: every rank all-gathers its packed partial
// [sums | mins | ttfb_hi, ttfb_lo, lat_hi, lat_lo] and combines all of them locally, the
// double-double sums in RANK ORDER (every rank computes bit-identical totals).
namespace {
constexpr int PACK_WORDS = PACK_SUMS + PACK_MINS + 4;

__global__ void stats_pack_words_kernel(const kvm_log_stats* st, long long* packed) {
  // same packing as stats_pack_kernel, into one contiguous record
  const int t = threadIdx.x;
  const double inf = __longlong_as_double(0x7ff0000000000000ll);
  long long* sums = packed;
  long long* mins = packed + PACK_SUMS;
  if (t == 0) {
    sums[0] = st->n_total;
    sums[1] = st->n_success;
    sums[2] = st->n_cold;
    sums[3] = st->n_cold_success;
    sums[4] = st->sum_completion_tokens;
    sums[5] = st->sum_total_tokens;
    sums[6] = st->sum_total_tokens_cold;
    sums[7] = st->n_gen_rows;
    const double a = st->min_start_ms, b = st->max_end_ms;
    mins[2 * KVM_MAX_COLS] = ordered_of_double(a != a ? inf : a);
    mins[2 * KVM_MAX_COLS + 1] = ~ordered_of_double(b != b ? -inf : b);
    long long* dd = packed + PACK_SUMS + PACK_MINS;
    dd[0] = __double_as_longlong(st->ttfb_sum_hi);
    dd[1] = __double_as_longlong(st->ttfb_sum_lo);
    dd[2] = __double_as_longlong(st->lat_sum_hi);
    dd[3] = __double_as_longlong(st->lat_sum_lo);
  }
  if (t < KVM_MAX_SETS) sums[8 + t] = st->sets.count[t];
  if (t < KVM_MAX_COLS) {
    mins[t] = (long long)(st->sets.kmin[t]) ^ MIN64;
    mins[KVM_MAX_COLS + t] = ~((long long)(st->sets.kmax[t]) ^ MIN64);
  }
}

__global__ void stats_combine_kernel(kvm_log_stats* st, const long long* gathered, int world) {
  __shared__ long long sums[PACK_SUMS];
  __shared__ long long mins[PACK_MINS];
  const int t = threadIdx.x;
  if (t < PACK_SUMS) {
    long long a = 0;
    for (int r = 0; r < world; ++r) a += gathered[(size_t)r * PACK_WORDS + t];
    sums[t] = a;
  } else if (t < PACK_SUMS + PACK_MINS) {
    const int k = t - PACK_SUMS;
    long long a = gathered[PACK_SUMS + k];
    for (int r = 1; r < world; ++r) {
      const long long v = gathered[(size_t)r * PACK_WORDS + PACK_SUMS + k];
      a = v < a ? v : a;
    }
    mins[k] = a;
  } else if (t == PACK_SUMS + PACK_MINS || t == PACK_SUMS + PACK_MINS + 1) {
    const int o = PACK_SUMS + PACK_MINS + 2 * (t - PACK_SUMS - PACK_MINS);
    kvm_dd acc{0.0, 0.0};
    for (int r = 0; r < world; ++r)   // rank order
      acc = kvm_dd_add(acc, kvm_dd{__longlong_as_double(gathered[(size_t)r * PACK_WORDS + o]),
                                   __longlong_as_double(gathered[(size_t)r * PACK_WORDS + o + 1])});
    if (o == PACK_SUMS + PACK_MINS) {
      st->ttfb_sum_hi = acc.hi;
      st->ttfb_sum_lo = acc.lo;
    } else {
      st->lat_sum_hi = acc.hi;
      st->lat_sum_lo = acc.lo;
    }
  }
  __syncthreads();
  const double inf = __longlong_as_double(0x7ff0000000000000ll);
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  if (t == 0) {
    st->n_total = sums[0];
    st->n_success = sums[1];
    st->n_cold = sums[2];
    st->n_cold_success = sums[3];
    st->sum_completion_tokens = sums[4];
    st->sum_total_tokens = sums[5];
    st->sum_total_tokens_cold = sums[6];
    st->n_gen_rows = sums[7];
    const double a = double_of_ordered(mins[2 * KVM_MAX_COLS]);
    const double b = double_of_ordered(~mins[2 * KVM_MAX_COLS + 1]);
    st->min_start_ms = a == inf ? nan : a;
    st->max_end_ms = b == -inf ? nan : b;
  }
  if (t < KVM_MAX_SETS) st->sets.count[t] = sums[8 + t];
  if (t < KVM_MAX_COLS) {
    st->sets.kmin[t] = (uint64_t)(mins[t] ^ MIN64);
    st->sets.kmax[t] = (uint64_t)((~mins[KVM_MAX_COLS + t]) ^ MIN64);
  }
}
}  // namespace

extern "C" int32_t kvm_log_stats_pack_words(void) { return PACK_WORDS; }

extern "C" int kvm_log_stats_pack_words_into(const kvm_log_stats* stats, int64_t* packed, void* stream) {
  KVM_REQUIRE(stats && packed, "kvm_log_stats_pack_words_into: null argument");
  stats_pack_words_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(stats, reinterpret_cast<long long*>(packed));
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_log_stats_combine(kvm_log_stats* stats, const int64_t* gathered, int32_t world, void* stream) {
  KVM_REQUIRE(stats && gathered && world >= 1, "kvm_log_stats_combine: bad argument");
  stats_combine_kernel<<<1, 64, 0, (cudaStream_t)stream>>>(stats, reinterpret_cast<const long long*>(gathered), world);
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}
