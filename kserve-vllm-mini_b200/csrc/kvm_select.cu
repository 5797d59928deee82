// kvm_select.cu - K2 (+K6): exact multi-rank radix select over fp64 keys.
//
// Replaces analyze.py:59-81 `percentile()` (25 calls / run over 9 multisets, each a
// fresh NaN filter + list.sort) and the upper-median of energy/collector.py:333.
// The kernels return exact ORDER STATISTICS; the host evaluates the reference's
// interpolation expression in Python floats, so the result is bit-identical.
//
// Algorithm (range-adaptive radix select; DESIGN.md "select"):
//   plan     per value column: kbase = min key, width = bits(max-min),
//            shift = max(0, width-11): digit(key) = (key-kbase) >> shift   (< 2048)
//            per query: n = sum of its base-set counts, rank from (kind, p, n)
//   hist     ONE pass over the data: per base set a 2048-bin histogram of digit(key),
//            privatised in shared memory, flushed with 64-bit global atomics
//   resolve  per query: bin holding the rank, residual rank, candidate bucket
//            (column, set mask, bin); buckets de-duplicated and laid out in the arena
//   compact  ONE pass over the data: keys whose digit hits a bucket are appended to it
//   finish   one CTA per query: 11-bit-per-pass radix select inside its bucket
// With adaptive digits a bucket holds ~n/2048 of the populated range, so everything
// after `compact` is tiny; the two streaming passes are HBM-bound.
//
// Multi-GPU: `hist` is all-reduced and the arenas all-gathered by the host between the
// stages; every rank then resolves / finishes identically.

#include "kvm_common.cuh"
#include "kvm_log_device.cuh"

namespace {

constexpr int SEL_THREADS = 512;
constexpr int MAXQ = KVM_MAX_QUERIES;
constexpr int MAX_SRC = 16;
constexpr int STATS_SCRATCH = 256 + 8 * (KVM_MAX_SETS + 2) * 148 * 8;

struct ColParam {
  uint64_t kbase;
  int32_t shift;  // digit = (key - kbase) >> shift
  int32_t width;  // significant bits of (kmax - kmin); 0 => column holds one distinct key
};

struct Bucket {
  uint64_t base;  // keys of the bucket lie in [base, base + 2^shift)
  int32_t shift;
  int32_t col;
  uint32_t set_mask;
  uint32_t bin;
  int64_t count;  // global number of keys in the bucket
};

struct QState {
  int64_t n;
  int64_t rank;
  int64_t resid;  // rank inside the bucket
  uint64_t key;   // valid when resolved
  int32_t status; // KVM_Q_*
  int32_t bucket; // -1 when resolved without candidates
  int32_t col;
  int32_t resolved;
};

struct SelWs {
  int32_t status;
  int32_t nq;
  int32_t nbuckets;
  int32_t pad0;
  int64_t arena_capacity;
  int64_t arena_used;
  kvm_set_info info;
  kvm_query queries[MAXQ];
  ColParam col[KVM_MAX_COLS];
  QState q[MAXQ];
  Bucket bucket[MAXQ];
  int64_t bucket_offsets[MAXQ + 1];
  int64_t bucket_counts[MAXQ];  // local append cursors (compact)
  uint32_t col_bitmap[KVM_MAX_COLS][BINS / 32];
  uint64_t hist[KVM_MAX_SETS][BINS];
  unsigned char stats_scratch[STATS_SCRATCH];
};

constexpr size_t WS_HEADER = (sizeof(SelWs) + 255) / 256 * 256;

struct QueryPack {
  kvm_query q[MAXQ];
  int32_t nq;
};

struct SrcPack {
  const uint64_t* arena[MAX_SRC];
  int32_t nsrc;
};

// ------------------------------------------------------------------ plan
__global__ void select_plan_kernel(SelWs* ws, const kvm_set_info* info_in, QueryPack qp,
                                   int64_t arena_capacity) {
  __shared__ kvm_set_info info;
  if (threadIdx.x == 0) {
    info = *info_in;
    ws->info = info;
    ws->status = KVM_OK;
    ws->nq = qp.nq;
    ws->nbuckets = 0;
    ws->arena_capacity = arena_capacity;
    ws->arena_used = 0;
  }
  __syncthreads();
  const int t = threadIdx.x;
  if (t < KVM_MAX_COLS) {
    ColParam cp{0, 0, 0};
    if (t < info.ncols && info.kmin[t] <= info.kmax[t]) {
      const uint64_t range = info.kmax[t] - info.kmin[t];
      cp.kbase = info.kmin[t];
      cp.width = range ? 64 - __clzll((long long)range) : 0;
      cp.shift = cp.width > BINS_LOG2 ? cp.width - BINS_LOG2 : 0;
    }
    ws->col[t] = cp;
  }
  if (t < MAXQ) {
    QState s{0, 0, 0, 0, KVM_Q_EMPTY, -1, 0, 1};
    ws->bucket_counts[t] = 0;
    if (t < qp.nq) {
      const kvm_query q = qp.q[t];
      ws->queries[t] = q;
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
    ws->q[t] = s;
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
template <class Stream>
__device__ __forceinline__ void hist_body(SelWs* ws, int nsets, Stream&& stream) {
  extern __shared__ uint32_t sh[];  // [nsets][BINS]
  __shared__ ColParams cp;
  for (int i = threadIdx.x; i < nsets * BINS; i += blockDim.x) sh[i] = 0;
  if (threadIdx.x < KVM_MAX_COLS) {
    cp.kbase[threadIdx.x] = ws->col[threadIdx.x].kbase;
    cp.shift[threadIdx.x] = ws->col[threadIdx.x].shift;
  }
  __syncthreads();
  stream([&](int set, int col, uint64_t key) {
    atomicAdd(&sh[set * BINS + digit_of(key, cp.kbase[col], cp.shift[col])], 1u);
  });
  __syncthreads();
  for (int i = threadIdx.x; i < nsets * BINS; i += blockDim.x) {
    const uint32_t c = sh[i];
    if (c) atomicAdd(reinterpret_cast<unsigned long long*>(&ws->hist[0][0]) + i, (unsigned long long)c);
  }
}

__global__ void __launch_bounds__(SEL_THREADS, 2)
log_hist_kernel(kvm_request_log log, const uint8_t* __restrict__ flags, bool vec, SelWs* ws) {
  hist_body(ws, KVM_LOG_NSETS, [&](auto&& f) { for_each_log_key(log, flags, vec, f); });
}

__global__ void __launch_bounds__(SEL_THREADS, 2)
column_hist_kernel(const double* __restrict__ values, int64_t n, int mode,
                   const void* __restrict__ member, bool vec, int nsets, SelWs* ws) {
  hist_body(ws, nsets, [&](auto&& f) { for_each_column_key(values, n, mode, member, vec, f); });
}

// ------------------------------------------------------------------ resolve
// One warp per query walks the (summed) histogram of its sets to the bin holding the rank.
__global__ void __launch_bounds__(1024, 1) select_resolve_kernel(SelWs* ws) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  __shared__ uint32_t q_bin[MAXQ];
  __shared__ int64_t q_cnt[MAXQ];
  const int nq = ws->nq;
  const int nsets = ws->info.nsets;

  for (int i = threadIdx.x; i < KVM_MAX_COLS * (BINS / 32); i += blockDim.x)
    (&ws->col_bitmap[0][0])[i] = 0;

  for (int q = warp; q < nq; q += nwarps) {
    QState s = ws->q[q];
    if (s.status != KVM_Q_OK) continue;
    const ColParam cp = ws->col[s.col];
    const uint32_t mask = ws->queries[q].set_mask;
    if (cp.width == 0) {  // one distinct key in the column
      if (lane == 0) {
        s.key = cp.kbase;
        s.resolved = 1;
        s.bucket = -1;
        ws->q[q] = s;
      }
      continue;
    }
    int64_t running = 0;
    bool found = false;
    for (int base = 0; base < BINS && !found; base += 32) {
      int64_t c = 0;
      for (int b = 0; b < nsets; ++b)
        if ((mask >> b) & 1u) c += (int64_t)ws->hist[b][base + lane];
      int64_t incl = c;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        int64_t o = __shfl_up_sync(KVM_FULL, incl, d);
        if (lane >= d) incl += o;
      }
      const unsigned hit = __ballot_sync(KVM_FULL, running + incl > s.rank);
      if (hit) {
        const int l = __ffs(hit) - 1;
        const int64_t excl = __shfl_sync(KVM_FULL, incl - c, l);
        const int64_t cnt = __shfl_sync(KVM_FULL, c, l);
        if (lane == 0) {
          q_bin[q] = (uint32_t)(base + l);
          q_cnt[q] = cnt;
          s.resid = s.rank - (running + excl);
          if (cp.shift == 0) {  // the digit is the whole key
            s.key = cp.kbase + (uint64_t)(base + l);
            s.resolved = 1;
            s.bucket = -1;
          } else {
            s.resolved = 0;
            s.bucket = -2;  // assigned below
          }
          ws->q[q] = s;
        }
        found = true;
      }
      running += __shfl_sync(KVM_FULL, incl, 31);
    }
    if (!found && lane == 0) {  // histogram inconsistent with counts (caller error)
      s.status = KVM_Q_OUT_OF_RANGE;
      s.resolved = 1;
      ws->q[q] = s;
    }
  }
  __syncthreads();
  __threadfence_block();
  if (threadIdx.x == 0) {
    int nb = 0;
    int64_t off = 0;
    for (int q = 0; q < nq; ++q) {
      QState s = ws->q[q];
      if (s.status != KVM_Q_OK || s.resolved) continue;
      const uint32_t mask = ws->queries[q].set_mask;
      int j = -1;
      for (int k = 0; k < nb; ++k)
        if (ws->bucket[k].col == s.col && ws->bucket[k].set_mask == mask &&
            ws->bucket[k].bin == q_bin[q]) {
          j = k;
          break;
        }
      if (j < 0) {
        j = nb++;
        const ColParam cp = ws->col[s.col];
        Bucket b;
        b.base = cp.kbase + ((uint64_t)q_bin[q] << cp.shift);
        b.shift = cp.shift;
        b.col = s.col;
        b.set_mask = mask;
        b.bin = q_bin[q];
        b.count = q_cnt[q];
        ws->bucket[j] = b;
        ws->bucket_offsets[j] = off;
        off += b.count;
        ws->col_bitmap[s.col][q_bin[q] >> 5] |= 1u << (q_bin[q] & 31);
      }
      ws->q[q].bucket = j;
    }
    ws->bucket_offsets[nb] = off;
    for (int k = nb + 1; k <= MAXQ; ++k) ws->bucket_offsets[k] = off;
    ws->nbuckets = nb;
    ws->arena_used = off;
    if (off > ws->arena_capacity) ws->status = KVM_ERR_ARENA;
  }
}

// ------------------------------------------------------------------ compact
template <class Stream>
__device__ __forceinline__ void compact_body(SelWs* ws, uint64_t* __restrict__ arena,
                                             Stream&& stream) {
  __shared__ ColParams cp;
  __shared__ uint32_t bitmap[KVM_MAX_COLS][BINS / 32];
  __shared__ Bucket bk[MAXQ];
  __shared__ int64_t bk_off[MAXQ];
  __shared__ int nb_s;
  if (ws->status != KVM_OK) return;
  if (threadIdx.x == 0) nb_s = ws->nbuckets;
  if (threadIdx.x < KVM_MAX_COLS) {
    cp.kbase[threadIdx.x] = ws->col[threadIdx.x].kbase;
    cp.shift[threadIdx.x] = ws->col[threadIdx.x].shift;
  }
  for (int i = threadIdx.x; i < KVM_MAX_COLS * (BINS / 32); i += blockDim.x)
    (&bitmap[0][0])[i] = (&ws->col_bitmap[0][0])[i];
  for (int i = threadIdx.x; i < MAXQ; i += blockDim.x) {
    bk[i] = ws->bucket[i];
    bk_off[i] = ws->bucket_offsets[i];
  }
  __syncthreads();
  const int nb = nb_s;
  if (nb == 0) return;
  stream([&](int set, int col, uint64_t key) {
    const uint32_t d = digit_of(key, cp.kbase[col], cp.shift[col]);
    if (!((bitmap[col][d >> 5] >> (d & 31)) & 1u)) return;
    for (int j = 0; j < nb; ++j) {
      if (bk[j].col == col && bk[j].bin == d && ((bk[j].set_mask >> set) & 1u)) {
        const unsigned long long pos =
            atomicAdd(reinterpret_cast<unsigned long long*>(&ws->bucket_counts[j]), 1ull);
        if ((int64_t)pos < bk[j].count) arena[bk_off[j] + (int64_t)pos] = key;
      }
    }
  });
}

__global__ void __launch_bounds__(SEL_THREADS, 2)
log_compact_kernel(kvm_request_log log, const uint8_t* __restrict__ flags, bool vec, SelWs* ws,
                   uint64_t* arena) {
  compact_body(ws, arena, [&](auto&& f) { for_each_log_key(log, flags, vec, f); });
}

__global__ void __launch_bounds__(SEL_THREADS, 2)
column_compact_kernel(const double* __restrict__ values, int64_t n, int mode,
                      const void* __restrict__ member, bool vec, SelWs* ws, uint64_t* arena) {
  compact_body(ws, arena, [&](auto&& f) { for_each_column_key(values, n, mode, member, vec, f); });
}

// ------------------------------------------------------------------ finish
constexpr int FIN_THREADS = 512;
constexpr int FIN_BITS = 11;
constexpr int FIN_BINS = 1 << FIN_BITS;

// One CTA per query: radix select inside the query's bucket, FIN_BITS per pass.
__global__ void __launch_bounds__(FIN_THREADS, 1)
select_finish_kernel(SelWs* ws, SrcPack src, const int64_t* __restrict__ src_counts,
                     kvm_query_result* __restrict__ results) {
  __shared__ uint32_t hist[FIN_BINS];
  __shared__ uint32_t chunk_sum[FIN_THREADS];
  __shared__ uint32_t sel_digit;
  __shared__ int64_t sel_below;
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
      r.status = s.status;
      r.pad = 0;
      results[q] = r;
    }
    return;
  }
  const Bucket b = ws->bucket[s.bucket];
  const int64_t off = ws->bucket_offsets[s.bucket];
  int rem = b.shift;        // unresolved low bits
  uint64_t prefix = 0;      // resolved part of (key - b.base) >> rem
  int64_t resid = s.resid;

  while (rem > 0) {
    const int take = rem < FIN_BITS ? rem : FIN_BITS;
    const int sh = rem - take;
    const uint32_t dmask = (1u << take) - 1u;
    for (int i = threadIdx.x; i < FIN_BINS; i += FIN_THREADS) hist[i] = 0;
    __syncthreads();
    for (int sidx = 0; sidx < src.nsrc; ++sidx) {
      const int64_t cnt = src_counts[(size_t)sidx * MAXQ + s.bucket];
      const uint64_t* __restrict__ keys = src.arena[sidx] + off;
      for (int64_t i = threadIdx.x; i < cnt; i += FIN_THREADS) {
        const uint64_t low = keys[i] - b.base;
        if ((rem >= 64 ? 0 : (low >> rem)) == prefix) atomicAdd(&hist[(uint32_t)(low >> sh) & dmask], 1u);
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
      // warp 0 scans the chunk sums (16 per lane)
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
      }
      if (!hit && threadIdx.x == 0) {  // cannot happen with consistent counts
        sel_digit = 0;
        sel_below = 0;
      }
    }
    __syncthreads();
    prefix = (prefix << take) | sel_digit;
    resid -= sel_below;
    rem = sh;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    kvm_query_result r;
    r.value = kvm_unkey(b.base + prefix);
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

}  // namespace

extern "C" size_t kvm_select_workspace_bytes(int64_t arena_capacity) {
  if (arena_capacity < 0) arena_capacity = 0;
  return WS_HEADER + kvm_align_up((size_t)arena_capacity * sizeof(uint64_t), 256) + 256;
}

extern "C" int64_t kvm_select_default_arena(int64_t n) {
  // buckets hold ~1/2048 of the populated key range; n/16 + 1M leaves a wide margin for
  // up to KVM_MAX_QUERIES buckets of concentrated data.
  int64_t cap = n / 16 + (1 << 20);
  return cap;
}

extern "C" int32_t kvm_select_bins(void) { return BINS; }
extern "C" uint64_t* kvm_select_hist_ptr(void* ws) { return &reinterpret_cast<SelWs*>(ws)->hist[0][0]; }
extern "C" uint64_t* kvm_select_arena_ptr(void* ws) { return arena_of(ws); }
extern "C" int64_t* kvm_select_bucket_counts_ptr(void* ws) { return reinterpret_cast<SelWs*>(ws)->bucket_counts; }
extern "C" int64_t* kvm_select_bucket_offsets_ptr(void* ws) { return reinterpret_cast<SelWs*>(ws)->bucket_offsets; }
extern "C" kvm_set_info* kvm_select_info_ptr(void* ws) { return &reinterpret_cast<SelWs*>(ws)->info; }

extern "C" int kvm_select_plan(const kvm_set_info* info, const kvm_query* queries, int32_t nq,
                               void* workspace, size_t workspace_bytes, int64_t arena_capacity,
                               void* stream) {
  KVM_REQUIRE(info && queries && workspace, "kvm_select_plan: null argument");
  KVM_REQUIRE(nq >= 1 && nq <= MAXQ, "kvm_select_plan: nq=%d outside [1,%d]", nq, MAXQ);
  if (workspace_bytes < kvm_select_workspace_bytes(arena_capacity)) {
    kvm_set_error("kvm_select_plan: workspace %zu < %zu bytes", workspace_bytes,
                  kvm_select_workspace_bytes(arena_capacity));
    return KVM_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  SelWs* ws = reinterpret_cast<SelWs*>(workspace);
  QueryPack qp;
  for (int i = 0; i < MAXQ; ++i) qp.q[i] = i < nq ? queries[i] : kvm_query{0, 0, 0.0, 0};
  qp.nq = nq;
  KVM_CUDA_CHECK(cudaMemsetAsync(&ws->hist[0][0], 0, sizeof(ws->hist), st));
  select_plan_kernel<<<1, 64, 0, st>>>(ws, info, qp, arena_capacity);
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_log_select_hist(const kvm_request_log* log, const uint8_t* flags,
                                   void* workspace, void* stream) {
  KVM_REQUIRE(log && flags && workspace, "kvm_log_select_hist: null argument");
  const size_t smem = (size_t)KVM_LOG_NSETS * BINS * sizeof(uint32_t);
  int rc = ensure_smem((const void*)log_hist_kernel, smem);
  if (rc) return rc;
  const int grid = kvm_grid_for(log->n, SEL_THREADS * 2 * 4, 2);
  log_hist_kernel<<<grid, SEL_THREADS, smem, (cudaStream_t)stream>>>(
      *log, flags, log_vec_ok(log, flags), reinterpret_cast<SelWs*>(workspace));
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
  select_resolve_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(reinterpret_cast<SelWs*>(workspace));
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_log_select_compact(const kvm_request_log* log, const uint8_t* flags,
                                      void* workspace, void* stream) {
  KVM_REQUIRE(log && flags && workspace, "kvm_log_select_compact: null argument");
  const int grid = kvm_grid_for(log->n, SEL_THREADS * 2 * 4, 2);
  log_compact_kernel<<<grid, SEL_THREADS, 0, (cudaStream_t)stream>>>(
      *log, flags, log_vec_ok(log, flags), reinterpret_cast<SelWs*>(workspace), arena_of(workspace));
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_column_select_compact(const double* values, int64_t n, int32_t mode,
                                         const void* member, void* workspace, void* stream) {
  KVM_REQUIRE(workspace && (values || n == 0), "kvm_column_select_compact: null argument");
  const int grid = kvm_grid_for(n, SEL_THREADS * 2 * 4, 2);
  column_compact_kernel<<<grid, SEL_THREADS, 0, (cudaStream_t)stream>>>(
      values, n, mode, member, column_vec_ok(values, mode, member),
      reinterpret_cast<SelWs*>(workspace), arena_of(workspace));
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_select_finish(void* workspace, const uint64_t* const* arenas, int32_t nsrc,
                                 const int64_t* src_counts, kvm_query_result* results,
                                 void* stream) {
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
  select_finish_kernel<<<MAXQ, FIN_THREADS, 0, (cudaStream_t)stream>>>(ws, sp, src_counts, results);
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
                              kvm_query_result* results, void* workspace, size_t workspace_bytes,
                              int64_t arena_capacity, void* stream) {
  KVM_REQUIRE(log && flags && stats && results, "kvm_log_select: null argument");
  int rc = kvm_select_plan(&stats->sets, queries, nq, workspace, workspace_bytes, arena_capacity, stream);
  if (rc) return rc;
  if ((rc = kvm_log_select_hist(log, flags, workspace, stream))) return rc;
  if ((rc = kvm_select_resolve(workspace, stream))) return rc;
  if ((rc = kvm_log_select_compact(log, flags, workspace, stream))) return rc;
  return kvm_select_finish(workspace, nullptr, 0, nullptr, results, stream);
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
  if ((rc = kvm_select_plan(&ws->info, queries, nq, workspace, workspace_bytes, arena_capacity, stream))) return rc;
  if ((rc = kvm_column_select_hist(values, n, mode, member, workspace, stream))) return rc;
  if ((rc = kvm_select_resolve(workspace, stream))) return rc;
  if ((rc = kvm_column_select_compact(values, n, mode, member, workspace, stream))) return rc;
  return kvm_select_finish(workspace, nullptr, 0, nullptr, results, stream);
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

static int static_scratch_for_current_device(void** out) {
  int dev = -1;
  KVM_CUDA_CHECK(cudaGetDevice(&dev));
  KVM_REQUIRE(dev >= 0 && dev < STATIC_MAX_DEVICES, "kvm_column_select_static: device %d unsupported", dev);
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
  void* scratch = nullptr;
  int rc = static_scratch_for_current_device(&scratch);
  if (rc) return rc;
  // Everything that must start from zero is cleared inside the pipeline itself (stats
  // ticket, histograms, append cursors), so stale contents of the previous call are harmless.
  return kvm_column_select(values, n, mode, member, 1, queries, nq, results, scratch,
                           STATIC_SCRATCH_BYTES, STATIC_MAX_N, stream);
}
