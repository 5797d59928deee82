// kvm_quantile.cu - the 16 B/record percentile path in ONE pass over the data
// (latency p50/p95/p99 over the status-200 rows; analyze.py:59-81).
//
// The general select (kvm_select.cu) streams the column three times (range, histogram, compact).
// For a handful of quantiles of one multiset a sample tells where to look:
//   sample   4096 stratified pseudo-random members are binned in shared memory (one CTA, key-space
//            histogram); for every requested p the sample ranks p*m +- (5 sigma + 6), widened to
//            bin edges, give a key bracket [lo, hi] that holds the wanted order statistic except
//            with probability ~1e-6 (then the verification below fails and the call declines)
//   stream   ONE pass (16 B/record): per bracket the exact count of member keys below it, a
//            2048-bin histogram of the keys inside it, and the inside keys themselves appended to
//            a per-CTA region of the arena (no global atomics in the loop)
//   collect  every CTA derives from the exact counts the fine bin that holds each rank and
//            gathers the arena keys of those bins (~n/20000 per query)
//   rank     one CTA per query ranks its few hundred keys by counting
// The answer is EXACT whenever it is produced: the bracket is verified with the exact counts
// (count_below <= rank < count_below + count_inside).  If the verification fails, a per-CTA
// region or a list overflows (heavy duplicates, adversarial order), the status word says
// KVM_ERR_UNSUPPORTED and the caller runs the general select - nothing approximate is returned.

#include "kvm_common.cuh"
#include "kvm_log_device.cuh"

namespace {

constexpr int QS_MAXQ = 8;        // queries per call
constexpr int QS_NB = 4;          // brackets (distinct p)
constexpr int QS_SAMPLE = 4096;
constexpr int QS_BINS = 2048;
constexpr int QS_BITS = 11;
constexpr int QS_THREADS = 512;
constexpr int QS_MAX_GRID = 148 * 2 * 2;
constexpr int QS_LIST = 4096;     // keys one query may gather for the final ranking
constexpr uint64_t QS_NONE = ~0ull;

struct QsBracket {
  uint64_t lo, hi;  // member key is inside iff lo <= key <= hi; below iff key < lo
  uint64_t base;    // digit = clamp((key - base) >> shift, 0, 2047); base = lo unless the bracket is open below
  int32_t shift;
  int32_t active;
};

__device__ __forceinline__ uint32_t qs_digit(uint64_t key, uint64_t base, int shift) {
  if (key < base) return 0u;
  const uint64_t d = (key - base) >> shift;
  return d < (uint64_t)2047 ? (uint32_t)d : 2047u;
}

struct QsPart {       // one per CTA of the streaming pass
  long long members;
  long long below[QS_NB];
  long long inside[QS_NB];  // appended (or attempted) keys
};

struct QsQuery {
  int32_t bracket, kind;
  double p;
  int32_t valid, bin;   // filled by collect
  long long resid, rank, n;
  int32_t qstatus, pad;
};

struct QsWs {
  int32_t status;       // KVM_OK / KVM_ERR_UNSUPPORTED
  int32_t nq, nb, grid;
  long long region_cap; // keys per (CTA, bracket) region
  QsBracket br[QS_NB];
  QsQuery q[QS_MAXQ];
  unsigned long long list_n[QS_MAXQ];
  unsigned long long hist[QS_NB][QS_BINS];
  QsPart part[QS_MAX_GRID];
  uint64_t list[QS_MAXQ][QS_LIST];
  uint64_t rank_scratch[QS_MAXQ][QS_LIST];
};

constexpr size_t QS_HEADER = (sizeof(QsWs) + 255) / 256 * 256;

__device__ __forceinline__ uint32_t qs_hash(uint32_t x) {
  x ^= x >> 16;
  x *= 0x7feb352du;
  x ^= x >> 15;
  x *= 0x846ca68bu;
  x ^= x >> 16;
  return x;
}

template <int MODE>
__device__ __forceinline__ bool qs_member(const void* member, int64_t i) {
  if (MODE == KVM_MEMBER_STATUS) return reinterpret_cast<const int64_t*>(member)[i] == 200;
  return true;
}

struct QsArgs {
  double p[QS_NB];
  int32_t nb;
  int32_t nq;
  int32_t q_bracket[QS_MAXQ];
  int32_t q_kind[QS_MAXQ];
  double q_p[QS_MAXQ];
};

// ------------------------------------------------------------------ sample
template <int MODE>
__global__ void __launch_bounds__(1024, 1)
qs_sample_kernel(const double* __restrict__ values, int64_t n, const void* __restrict__ member, QsArgs a,
                 QsWs* ws, int grid, long long region_cap) {
  __shared__ uint64_t keys[QS_SAMPLE];
  __shared__ uint32_t sh[QS_BINS];   // sample histogram, then (in place) its exclusive prefix
  __shared__ uint64_t wmin[32], wmax[32];
  __shared__ uint32_t wtot[32];
  __shared__ int m_eff_s;
  const int t = threadIdx.x;
  // clear what the streaming pass accumulates into
  for (int i = t; i < QS_NB * QS_BINS; i += 1024) (&ws->hist[0][0])[i] = 0;
  if (t < QS_MAXQ) ws->list_n[t] = 0;
  if (t == 0) {
    ws->status = KVM_OK;
    ws->nq = a.nq;
    ws->nb = a.nb;
    ws->grid = grid;
    ws->region_cap = region_cap;
    m_eff_s = 0;
  }
  // stratified pseudo-random sample: one element from each of 4096 equal strata
  const int64_t width = n / QS_SAMPLE;
#pragma unroll
  for (int k = 0; k < QS_SAMPLE / 1024; ++k) {
    const int s = t + 1024 * k;
    const int64_t i = (int64_t)s * width + (int64_t)(qs_hash((uint32_t)s * 2654435761u + 12345u) % (uint64_t)width);
    const double v = values[i];
    keys[s] = (qs_member<MODE>(member, i) && v == v) ? kvm_key(v) : QS_NONE;
  }
  __syncthreads();
  // No sort: a 2048-bin histogram of the sample in key space (range-adaptive between the sample's
  // extremes) locates every wanted sample rank to within one bin; the bracket takes the outer
  // edges of those bins, so it is a superset of the order-statistic bracket.
  uint64_t mn = QS_NONE, mx = 0;
  int cnt = 0;
#pragma unroll
  for (int k = 0; k < QS_SAMPLE / 1024; ++k) {
    const uint64_t key = keys[t + 1024 * k];
    if (key != QS_NONE) {
      mn = key < mn ? key : mn;
      mx = key > mx ? key : mx;
      ++cnt;
    }
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    const uint64_t a0 = __shfl_xor_sync(KVM_FULL, mn, o), a1 = __shfl_xor_sync(KVM_FULL, mx, o);
    mn = a0 < mn ? a0 : mn;
    mx = a1 > mx ? a1 : mx;
    cnt += __shfl_xor_sync(KVM_FULL, cnt, o);
  }
  if ((t & 31) == 0) {
    wmin[t >> 5] = mn;
    wmax[t >> 5] = mx;
    atomicAdd(&m_eff_s, cnt);
  }
  sh[t] = 0;
  sh[t + 1024] = 0;
  __syncthreads();
  if (t < 32) {
    mn = wmin[t];
    mx = wmax[t];
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      const uint64_t a0 = __shfl_xor_sync(KVM_FULL, mn, o), a1 = __shfl_xor_sync(KVM_FULL, mx, o);
      mn = a0 < mn ? a0 : mn;
      mx = a1 > mx ? a1 : mx;
    }
    if (t == 0) {
      wmin[0] = mn;
      wmax[0] = mx;
    }
  }
  __syncthreads();
  const int m = m_eff_s;
  const uint64_t kmin = wmin[0], kmax = wmax[0];
  const uint64_t krange = m > 0 ? kmax - kmin : 0;
  const int kbits = krange ? 64 - __clzll((long long)krange) : 0;
  const int kshift = kbits > QS_BITS ? kbits - QS_BITS : 0;
#pragma unroll
  for (int k = 0; k < QS_SAMPLE / 1024; ++k) {
    const uint64_t key = keys[t + 1024 * k];
    if (key != QS_NONE) atomicAdd(&sh[(uint32_t)((key - kmin) >> kshift)], 1u);
  }
  __syncthreads();
  // exclusive prefix over the 2048 bins: two bins per thread, warp scan, scan of the warp totals
  {
    const uint32_t c0 = sh[2 * t], c1 = sh[2 * t + 1];
    uint32_t incl = c0 + c1;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(KVM_FULL, incl, d);
      if ((t & 31) >= d) incl += o;
    }
    if ((t & 31) == 31) wtot[t >> 5] = incl;
    __syncthreads();
    if (t < 32) {
      uint32_t w = wtot[t];
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(KVM_FULL, w, d);
        if (t >= d) w += o;
      }
      wtot[t] = w;  // inclusive
    }
    __syncthreads();
    const uint32_t before = incl - (c0 + c1) + ((t >> 5) ? wtot[(t >> 5) - 1] : 0u);
    sh[2 * t] = before;          // every thread read its two bins before the barriers above
    sh[2 * t + 1] = before + c0;
  }
  __syncthreads();
  if (t < QS_NB) {
    QsBracket b{QS_NONE, 0, 0, 0, 0};
    if (t < a.nb && m >= 512) {
      const double p = fmax(0.0, fmin(a.p[t], 1.0));
      const double pos = p * (double)(m - 1);
      const double w = 5.0 * sqrt(p * (1.0 - p) * (double)m) + 6.0;
      const int li = (int)floor(pos - w), hi = (int)ceil(pos + w);
      const bool open_lo = li <= 0, open_hi = hi >= m - 1;
      // bin holding a sample rank r: the last bin whose exclusive prefix is <= r
      auto bin_of_rank = [&](int r) {
        int lo_b = 0, hi_b = QS_BINS - 1;
        while (lo_b < hi_b) {
          const int mid = (lo_b + hi_b + 1) >> 1;
          if ((int)sh[mid] <= r) lo_b = mid;
          else hi_b = mid - 1;
        }
        return lo_b;
      };
      b.lo = open_lo ? 0ull : kmin + ((uint64_t)bin_of_rank(li) << kshift);
      b.hi = open_hi ? QS_NONE - 1 : kmin + (((uint64_t)bin_of_rank(hi) + 1) << kshift) - 1;
      // digits are scaled to the part of the bracket the sample has seen; an open end keeps the
      // outermost bin for whatever lies beyond the sample's extreme
      b.base = open_lo ? kmin : b.lo;
      const uint64_t top = open_hi ? kmax : b.hi;
      const uint64_t range = top > b.base ? top - b.base : 0;
      const int bits = range ? 64 - __clzll((long long)range) : 0;
      const int span = (open_lo || open_hi) ? QS_BITS - 1 : QS_BITS;   // leave half of the bins beyond the sample
      b.shift = bits > span ? bits - span : 0;
      b.active = 1;
    }
    ws->br[t] = b;
  }
  if (t < QS_MAXQ) {
    QsQuery q{};
    q.bracket = a.q_bracket[t];
    q.kind = a.q_kind[t];
    q.p = a.q_p[t];
    ws->q[t] = q;
  }
  if (t == 0 && m < 512) ws->status = KVM_ERR_UNSUPPORTED;  // too few members to bracket anything
}

// ------------------------------------------------------------------ stream
struct QsSmem {
  uint32_t hist[QS_NB][QS_BINS];
  uint32_t cursor[QS_NB];
};

template <int MODE, bool VEC, int NBT>   // NBT = active brackets (the loops below are unrolled over them)
__global__ void __launch_bounds__(QS_THREADS, 2)
qs_stream_kernel(const double* __restrict__ values, int64_t n, const void* __restrict__ member, QsWs* ws,
                 uint64_t* __restrict__ arena) {
  __shared__ QsSmem S;
  __shared__ long long red[QS_THREADS / 32][1 + QS_NB];
  if (ws->status != KVM_OK) return;
  for (int i = threadIdx.x; i < NBT * QS_BINS; i += QS_THREADS) (&S.hist[0][0])[i] = 0;
  if (threadIdx.x < QS_NB) S.cursor[threadIdx.x] = 0;
  uint64_t lo[NBT], hi[NBT], bs[NBT];
  int sh[NBT];
#pragma unroll
  for (int b = 0; b < NBT; ++b) {
    lo[b] = ws->br[b].lo;
    hi[b] = ws->br[b].hi;
    bs[b] = ws->br[b].base;
    sh[b] = ws->br[b].shift;
  }
  const long long cap = ws->region_cap;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const unsigned lt = (1u << lane) - 1u;
  int32_t members = 0, below[NBT];
#pragma unroll
  for (int b = 0; b < NBT; ++b) below[b] = 0;

  // two records per lane: their membership ballots share one cursor bump per bracket
  auto two = [&](double v0, bool m0, double v1, bool m1) {
    m0 = m0 && (v0 == v0);
    m1 = m1 && (v1 == v1);
    const uint64_t k0 = kvm_key(v0), k1 = kvm_key(v1);
    members += (int)m0 + (int)m1;
#pragma unroll
    for (int b = 0; b < NBT; ++b) {
      const bool l0 = k0 < lo[b], l1 = k1 < lo[b];
      below[b] += (int)(m0 & l0) + (int)(m1 & l1);
      const bool in0 = m0 & !l0 & (k0 <= hi[b]);
      const bool in1 = m1 & !l1 & (k1 <= hi[b]);
      const unsigned mk0 = __ballot_sync(KVM_FULL, in0), mk1 = __ballot_sync(KVM_FULL, in1);
      if (mk0 | mk1) {
        const int c0 = __popc(mk0);
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(&S.cursor[b], (uint32_t)(c0 + __popc(mk1)));
        base = __shfl_sync(KVM_FULL, base, 0);
        uint64_t* __restrict__ region = arena + ((size_t)blockIdx.x * QS_NB + b) * (size_t)cap;
        if (in0) {
          const long long pos = (long long)base + __popc(mk0 & lt);
          if (pos < cap) region[pos] = k0;
          atomicAdd(&S.hist[b][qs_digit(k0, bs[b], sh[b])], 1u);
        }
        if (in1) {
          const long long pos = (long long)base + c0 + __popc(mk1 & lt);
          if (pos < cap) region[pos] = k1;
          atomicAdd(&S.hist[b][qs_digit(k1, bs[b], sh[b])], 1u);
        }
      }
    }
  };
  auto one = [&](double v, bool mem) { two(v, mem, 0.0, false); };

  if (VEC) {
    const int64_t npairs = n >> 1;
    // warp-uniform trip count (the ballots need all lanes)
    const int64_t stride = (int64_t)gridDim.x * QS_THREADS;
    for (int64_t p0 = (int64_t)blockIdx.x * QS_THREADS + (threadIdx.x & ~31); p0 < npairs; p0 += stride) {
      const int64_t p = p0 + lane;
      const bool live = p < npairs;
      const int64_t i = p << 1;
      double2 v = make_double2(0.0, 0.0);
      bool m0 = false, m1 = false;
      if (live) {
        v = kvm_ld2(values + i);
        if (MODE == KVM_MEMBER_STATUS) {
          const longlong2 s2 = kvm_ld2(reinterpret_cast<const int64_t*>(member) + i);
          m0 = s2.x == 200;
          m1 = s2.y == 200;
        } else {
          m0 = m1 = true;
        }
      }
      two(v.x, m0, v.y, m1);
    }
    if ((n & 1) && blockIdx.x == 0 && threadIdx.x < 32) {
      const int64_t i = n - 1;
      const bool l0 = lane == 0;
      one(l0 ? values[i] : 0.0, l0 && qs_member<MODE>(member, i));
    }
  } else {
    const int64_t stride = (int64_t)gridDim.x * QS_THREADS;
    for (int64_t i0 = (int64_t)blockIdx.x * QS_THREADS + (threadIdx.x & ~31); i0 < n; i0 += stride) {
      const int64_t i = i0 + lane;
      const bool live = i < n;
      one(live ? kvm_ld1(values + i) : 0.0, live && qs_member<MODE>(member, i));
    }
  }
  __syncthreads();
  // flush: histogram, per-CTA partial counts
  for (int i = threadIdx.x; i < NBT * QS_BINS; i += QS_THREADS) {
    const uint32_t c = (&S.hist[0][0])[i];
    if (c) atomicAdd(&ws->hist[0][0] + i, (unsigned long long)c);
  }
  const int warp = threadIdx.x >> 5;
  {
    const long long m = kvm_warp_sum((int64_t)members);
    if (lane == 0) red[warp][0] = m;
#pragma unroll
    for (int b = 0; b < QS_NB; ++b) {
      const long long v = kvm_warp_sum((int64_t)(b < NBT ? below[b < NBT ? b : 0] : 0));
      if (lane == 0) red[warp][1 + b] = v;
    }
  }
  __syncthreads();
  if (threadIdx.x < 1 + QS_NB) {
    long long s = 0;
    for (int w = 0; w < QS_THREADS / 32; ++w) s += red[w][threadIdx.x];
    QsPart& P = ws->part[blockIdx.x];
    if (threadIdx.x == 0) P.members = s;
    else P.below[threadIdx.x - 1] = s;
  }
  if (threadIdx.x < QS_NB) ws->part[blockIdx.x].inside[threadIdx.x] = (long long)S.cursor[threadIdx.x];
}

// ------------------------------------------------------------------ collect
constexpr int QS_COLLECT_CTAS = 296;

__global__ void __launch_bounds__(QS_THREADS, 2)
qs_collect_kernel(QsWs* ws, const uint64_t* __restrict__ arena) {
  __shared__ uint32_t hist[QS_NB][QS_BINS];  // n < 2^31 on this path (checked by the host)
  __shared__ long long tot[1 + 2 * QS_NB];  // members, below[], inside[]
  __shared__ int overflow_s;
  __shared__ QsQuery qs[QS_MAXQ];
  __shared__ QsBracket br[QS_NB];
  if (ws->status != KVM_OK) return;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int grid1 = ws->grid, nq = ws->nq;
  const long long cap = ws->region_cap;
  if (t < 1 + 2 * QS_NB) tot[t] = 0;
  if (t == 0) overflow_s = 0;
  for (int i = t; i < QS_NB * QS_BINS; i += QS_THREADS) (&hist[0][0])[i] = (uint32_t)(&ws->hist[0][0])[i];
  if (t < QS_MAXQ) qs[t] = ws->q[t];
  if (t < QS_NB) br[t] = ws->br[t];
  __syncthreads();
  // exact totals from the per-CTA partials (1 + 2*NB columns, one warp each)
  for (int col = warp; col < 1 + 2 * QS_NB; col += QS_THREADS / 32) {
    long long s = 0;
    int ovf = 0;
    for (int c = lane; c < grid1; c += 32) {
      const QsPart& P = ws->part[c];
      const long long v = col == 0 ? P.members : (col <= QS_NB ? P.below[col - 1] : P.inside[col - 1 - QS_NB]);
      if (col > QS_NB && v > cap) ovf = 1;
      s += v;
    }
    s = kvm_warp_sum(s);
    ovf = __any_sync(KVM_FULL, ovf);
    if (lane == 0) {
      tot[col] = s;
      if (ovf) overflow_s = 1;
    }
  }
  __syncthreads();
  // every query: rank -> bracket check -> fine bin + residual (one warp per query)
  for (int q = warp; q < nq; q += QS_THREADS / 32) {
    QsQuery s = qs[q];
    const long long n = tot[0];
    s.n = n;
    s.valid = 0;
    s.bin = -1;
    if (n == 0) {
      s.qstatus = KVM_Q_EMPTY;
      s.valid = 1;
    } else {
      s.rank = resolve_rank(s.kind, s.p, n, 0);
      if (s.rank < 0 || s.rank >= n) {
        s.qstatus = KVM_Q_OUT_OF_RANGE;
        s.valid = 1;
      } else {
        s.qstatus = KVM_Q_OK;
        const int b = s.bracket;
        const long long r = s.rank - tot[1 + b];
        if (!overflow_s && br[b].active && r >= 0 && r < tot[1 + QS_NB + b]) {
          // walk the bracket's histogram, 32 bins per step
          long long running = 0;
          for (int base = 0; base < QS_BINS; base += 32) {
            const long long c = (long long)hist[b][base + lane];
            long long incl = c;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
              const long long o = __shfl_up_sync(KVM_FULL, incl, d);
              if (lane >= d) incl += o;
            }
            const unsigned hit = __ballot_sync(KVM_FULL, running + incl > r);
            if (hit) {
              const int l = __ffs(hit) - 1;
              s.bin = base + l;
              s.resid = r - (running + __shfl_sync(KVM_FULL, incl - c, l));
              s.valid = 1;
              break;
            }
            running += __shfl_sync(KVM_FULL, incl, 31);
          }
        }
      }
    }
    if (lane == 0) qs[q] = s;
  }
  __syncthreads();
  if (blockIdx.x == 0 && t < QS_MAXQ) ws->q[t] = qs[t];
  // gather the arena keys of the wanted bins: regions are dealt round-robin to the CTAs
  const int nb = ws->nb;
  for (int reg = blockIdx.x; reg < grid1 * nb; reg += gridDim.x) {
    const int cta = reg / nb, b = reg % nb;
    long long cnt = ws->part[cta].inside[b];
    cnt = cnt < cap ? cnt : cap;
    const uint64_t* __restrict__ keys = arena + ((size_t)cta * QS_NB + b) * (size_t)cap;
    const uint64_t base = br[b].base;
    const int sh = br[b].shift;
    int want[QS_MAXQ];   // bins this bracket's queries need (-1: not this bracket)
#pragma unroll
    for (int q = 0; q < QS_MAXQ; ++q)
      want[q] = (q < nq && qs[q].bracket == b && qs[q].qstatus == KVM_Q_OK) ? qs[q].bin : -1;
    for (long long i = t; i < cnt; i += QS_THREADS) {
      const uint64_t k = keys[i];
      const int d = (int)qs_digit(k, base, sh);
#pragma unroll
      for (int q = 0; q < QS_MAXQ; ++q)
        if (want[q] == d) {
          const unsigned long long pos = atomicAdd(&ws->list_n[q], 1ull);
          if (pos < (unsigned long long)QS_LIST) ws->list[q][pos] = k;
        }
    }
  }
}

// ------------------------------------------------------------------ rank
__global__ void __launch_bounds__(1024, 1)
qs_rank_kernel(QsWs* ws, kvm_query_result* __restrict__ results) {
  __shared__ uint64_t keys[QS_LIST];
  __shared__ uint32_t hist[QS_BINS];
  __shared__ uint64_t wmin[32], wmax[32];
  __shared__ uint32_t wtot[32];
  __shared__ uint32_t sel_bin, sel_resid, small_n;
  __shared__ uint64_t answer;
  uint64_t* small = reinterpret_cast<uint64_t*>(ws->rank_scratch[blockIdx.x]);  // m2 is tiny: L2-resident
  const int q = blockIdx.x;
  if (q >= ws->nq) return;
  const QsQuery s = ws->q[q];
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  const unsigned long long m = ws->list_n[q];
  const bool need = ws->status == KVM_OK && s.valid && s.qstatus == KVM_Q_OK;
  bool ok = ws->status == KVM_OK && s.valid;
  if (need && (m > (unsigned long long)QS_LIST || (long long)m <= s.resid)) ok = false;
  if (!ok) {
    if (threadIdx.x == 0) {
      ws->status = KVM_ERR_UNSUPPORTED;  // all writers agree
      results[q] = kvm_query_result{nan, s.n, s.rank, s.qstatus, 0};
    }
    return;
  }
  if (!need) {  // empty multiset / rank out of range: nothing to look up
    if (threadIdx.x == 0) results[q] = kvm_query_result{nan, s.n, s.rank, s.qstatus, 0};
    return;
  }
  // narrow once with a range-adaptive 2048-bin histogram of the gathered keys, then rank the
  // handful of keys in the bin that holds the residual rank by counting
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  uint64_t mn = QS_NONE, mx = 0;
  for (unsigned i = t; i < m; i += 1024) {
    const uint64_t k = ws->list[q][i];
    keys[i] = k;
    mn = k < mn ? k : mn;
    mx = k > mx ? k : mx;
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    const uint64_t a0 = __shfl_xor_sync(KVM_FULL, mn, o), a1 = __shfl_xor_sync(KVM_FULL, mx, o);
    mn = a0 < mn ? a0 : mn;
    mx = a1 > mx ? a1 : mx;
  }
  if (lane == 0) {
    wmin[warp] = mn;
    wmax[warp] = mx;
  }
  hist[t] = 0;
  hist[t + 1024] = 0;
  if (t == 0) small_n = 0;
  __syncthreads();
  mn = wmin[0];
  mx = wmax[0];
  for (int w = 1; w < 32; ++w) {
    mn = wmin[w] < mn ? wmin[w] : mn;
    mx = wmax[w] > mx ? wmax[w] : mx;
  }
  const uint64_t range = mx - mn;
  const int bits = range ? 64 - __clzll((long long)range) : 0;
  const int sh2 = bits > QS_BITS ? bits - QS_BITS : 0;
  for (unsigned i = t; i < m; i += 1024) atomicAdd(&hist[(uint32_t)((keys[i] - mn) >> sh2)], 1u);
  __syncthreads();
  {
    const uint32_t c0 = hist[2 * t], c1 = hist[2 * t + 1];
    uint32_t incl = c0 + c1;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(KVM_FULL, incl, d);
      if (lane >= d) incl += o;
    }
    if (lane == 31) wtot[warp] = incl;
    __syncthreads();
    uint32_t before = incl - (c0 + c1);
    for (int w = 0; w < warp; ++w) before += wtot[w];
    const long long r = s.resid;
    if ((long long)before <= r && r < (long long)before + c0) {
      sel_bin = 2 * t;
      sel_resid = (uint32_t)(r - before);
    } else if ((long long)before + c0 <= r && r < (long long)before + c0 + c1) {
      sel_bin = 2 * t + 1;
      sel_resid = (uint32_t)(r - before - c0);
    }
  }
  __syncthreads();
  const uint32_t sb = sel_bin, sr = sel_resid;
  for (unsigned i = t; i < m; i += 1024) {
    const uint64_t k = keys[i];
    if ((uint32_t)((k - mn) >> sh2) == sb) {
      const uint32_t pos = atomicAdd(&small_n, 1u);
      small[pos] = k;   // at most m <= QS_LIST entries
    }
  }
  __syncthreads();
  const uint32_t m2 = small_n;
  for (unsigned i = t; i < m2; i += 1024) {
    const uint64_t k = small[i];
    uint32_t less = 0, leq = 0;
    for (unsigned j = 0; j < m2; ++j) {
      const uint64_t o = small[j];
      less += o < k;
      leq += o <= k;
    }
    if (less <= sr && sr < leq) answer = k;  // all writers agree
  }
  __syncthreads();
  if (threadIdx.x == 0) results[q] = kvm_query_result{kvm_unkey(answer), s.n, s.rank, KVM_Q_OK, 0};
}

long long qs_region_cap(int64_t n, int grid) {
  // brackets hold ~2-10 % of the members each; a CTA sees n/grid records: room for 40 % of them per bracket
  long long cap = (long long)((n / grid) * 2 / 5) + 1024;
  return (cap + 1) & ~1ll;
}

int qs_grid(int64_t n) { return kvm_grid_for(n, QS_THREADS * 2 * 4, 2); }

}  // namespace

extern "C" int64_t kvm_column_select_sampled_min_n(void) { return 1 << 18; }

extern "C" size_t kvm_column_select_sampled_workspace_bytes(int64_t n) {
  if (n < 0) n = 0;
  const int grid = qs_grid(n);
  return QS_HEADER + (size_t)grid * QS_NB * (size_t)qs_region_cap(n, grid) * sizeof(uint64_t) + 256;
}

extern "C" int kvm_column_select_sampled(const double* values, int64_t n, int32_t mode, const void* member,
                                         const kvm_query* queries, int32_t nq, kvm_query_result* results,
                                         void* workspace, size_t workspace_bytes, void* stream) {
  KVM_REQUIRE(values && queries && results && workspace, "kvm_column_select_sampled: null argument");
  if (!(mode == KVM_MEMBER_ALL || (mode == KVM_MEMBER_STATUS && member)) || nq < 1 || nq > QS_MAXQ ||
      n < kvm_column_select_sampled_min_n() || n >= (1ll << 31)) {
    kvm_set_error("kvm_column_select_sampled: needs member mode ALL/STATUS, 1..%d queries, %lld <= n < 2^31; use kvm_column_select",
                  QS_MAXQ, (long long)kvm_column_select_sampled_min_n());
    return KVM_ERR_UNSUPPORTED;
  }
  QsArgs a{};
  for (int q = 0; q < nq; ++q) {
    const kvm_query& Q = queries[q];
    double p = Q.p;
    if (Q.rank_kind == KVM_RANK_UPPER_MEDIAN) p = 0.5;
    const bool has_p = Q.rank_kind == KVM_RANK_LERP_LO || Q.rank_kind == KVM_RANK_LERP_HI ||
                       Q.rank_kind == KVM_RANK_FLOOR_PN || Q.rank_kind == KVM_RANK_FLOOR_PNM1 ||
                       Q.rank_kind == KVM_RANK_UPPER_MEDIAN;
    if (!has_p || Q.set_mask != 1u || !(p == p)) {
      kvm_set_error("kvm_column_select_sampled: query %d is not a quantile of set 0; use kvm_column_select", q);
      return KVM_ERR_UNSUPPORTED;
    }
    int b = -1;
    for (int k = 0; k < a.nb; ++k)
      if (a.p[k] == p) b = k;
    if (b < 0) {
      if (a.nb == QS_NB) {
        kvm_set_error("kvm_column_select_sampled: more than %d distinct quantiles; use kvm_column_select", QS_NB);
        return KVM_ERR_UNSUPPORTED;
      }
      b = a.nb++;
      a.p[b] = p;
    }
    a.q_bracket[q] = b;
    a.q_kind[q] = Q.rank_kind;
    a.q_p[q] = Q.p;
  }
  a.nq = nq;
  if (workspace_bytes < kvm_column_select_sampled_workspace_bytes(n)) {
    kvm_set_error("kvm_column_select_sampled: workspace %zu < %zu bytes", workspace_bytes,
                  kvm_column_select_sampled_workspace_bytes(n));
    return KVM_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  QsWs* ws = reinterpret_cast<QsWs*>(workspace);
  uint64_t* arena = reinterpret_cast<uint64_t*>(reinterpret_cast<unsigned char*>(workspace) + QS_HEADER);
  const int grid = qs_grid(n);
  const long long cap = qs_region_cap(n, grid);
  const bool vec = (reinterpret_cast<uintptr_t>(values) & 15u) == 0 &&
                   (mode != KVM_MEMBER_STATUS || (reinterpret_cast<uintptr_t>(member) & 15u) == 0);
#define KVM_QS_STREAM(M, V, NBT) \
  qs_stream_kernel<M, V, NBT><<<grid, QS_THREADS, 0, st>>>(values, n, member, ws, arena)
#define KVM_QS_NB(M, V)                 \
  do {                                  \
    if (a.nb == 1) KVM_QS_STREAM(M, V, 1);      \
    else if (a.nb == 2) KVM_QS_STREAM(M, V, 2); \
    else if (a.nb == 3) KVM_QS_STREAM(M, V, 3); \
    else KVM_QS_STREAM(M, V, 4);                \
  } while (0)
  if (mode == KVM_MEMBER_STATUS) {
    qs_sample_kernel<KVM_MEMBER_STATUS><<<1, 1024, 0, st>>>(values, n, member, a, ws, grid, cap);
    if (vec) KVM_QS_NB(KVM_MEMBER_STATUS, true);
    else KVM_QS_NB(KVM_MEMBER_STATUS, false);
  } else {
    qs_sample_kernel<KVM_MEMBER_ALL><<<1, 1024, 0, st>>>(values, n, member, a, ws, grid, cap);
    if (vec) KVM_QS_NB(KVM_MEMBER_ALL, true);
    else KVM_QS_NB(KVM_MEMBER_ALL, false);
  }
#undef KVM_QS_NB
#undef KVM_QS_STREAM
  qs_collect_kernel<<<QS_COLLECT_CTAS, QS_THREADS, 0, st>>>(ws, arena);
  qs_rank_kernel<<<nq, 1024, 0, st>>>(ws, results);
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}
