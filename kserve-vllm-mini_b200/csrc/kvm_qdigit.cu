// kvm_qdigit.cu - one-and-an-eighth-pass exact quantiles of one multiset (the 16 B/record
// percentile path; analyze.py:59-81).
//
//   sample   4096 stratified pseudo-random members give an ESTIMATE of the key range (one CTA)
//   hist     ONE pass over the data (16 B/record): 13-bit digit of every member key relative to
//            the estimated range (bins 0 and 8191 catch what lies far outside it), histogram in
//            shared memory, and the digit written as a 2-byte record (0xFFFF = not a member)
//   resolve  one CTA: exact member count = sum of the histogram, rank of every query, the bin
//            holding it and the residual rank
//   collect  pass over the 2 B/record digit records; the value of a record is fetched only when
//            its digit is a wanted bin (~0.1 % of the records)
//   rank     one CTA per query: one narrowing histogram over the gathered keys, then counting
// Exact whenever it answers.  It declines (status word KVM_ERR_UNSUPPORTED, the caller runs the
// general select) when a rank falls into a catch-all bin (the sample missed the range), when a
// list overflows (heavy duplicates) or when the sample holds too few members.

#include <cstddef>

#include "kvm_common.cuh"
#include "kvm_log_device.cuh"

namespace {

constexpr int QD_MAXQ = 8;
constexpr int QD_SAMPLE = 4096;
constexpr int QD_BINS = 8192;
constexpr int QD_SPAN_BITS = 12;   // the widened sample range spans at most 4096 bins; the rest lies beyond it
constexpr int QD_THREADS = 512;
constexpr int QD_SMALL = 4096;     // keys ranked by counting at the very end
constexpr int QD_QUEUE = 2048;     // wanted rows a CTA queues for the drain after its streaming loop
constexpr int QD_STAGE = 384;      // keys a CTA stages per wanted bin before going to global memory
constexpr uint64_t QD_NONE = ~0ull;

struct QdQuery {
  int32_t kind, slot;   // slot: index of the wanted bin this query reads its keys from
  double p;
  int32_t valid, bin;
  long long resid, rank, n;
  int32_t qstatus, pad;
};

struct QdWs {
  int32_t status;
  int32_t nq, nslots, pad;
  uint64_t base;        // digit = key < base ? 0 : min(1 + ((key - base) >> shift), 8191)
  int32_t shift, pad2;
  long long list_cap;   // keys per slot list
  QdQuery q[QD_MAXQ];
  int32_t slot_bin[QD_MAXQ];
  // ---- cleared by one memset before every call (QD_CLEAR_* below)
  unsigned long long s_inv_min, s_max;   // sample extremes (the minimum travels complemented)
  unsigned int s_count, s_ticket;
  unsigned long long list_n[QD_MAXQ];
  unsigned long long hist[QD_BINS];
  // ----
  uint64_t small[QD_MAXQ][QD_SMALL];
};

constexpr size_t QD_HEADER = (sizeof(QdWs) + 255) / 256 * 256;
constexpr size_t QD_CLEAR_BEGIN = offsetof(QdWs, s_inv_min), QD_CLEAR_END = offsetof(QdWs, small);

struct QdArgs {
  int32_t nq;
  int32_t kind[QD_MAXQ];
  double p[QD_MAXQ];
};

__device__ __forceinline__ uint32_t qd_hash(uint32_t x) {
  x ^= x >> 16;
  x *= 0x7feb352du;
  x ^= x >> 15;
  x *= 0x846ca68bu;
  x ^= x >> 16;
  return x;
}

__device__ __forceinline__ uint32_t qd_digit(uint64_t key, uint64_t base, int shift) {
  if (key < base) return 0u;
  const uint64_t d = ((key - base) >> shift) + 1ull;
  return d < (uint64_t)(QD_BINS - 1) ? (uint32_t)d : (uint32_t)(QD_BINS - 1);
}

template <int MODE>
__device__ __forceinline__ bool qd_member(const void* member, int64_t i) {
  if (MODE == KVM_MEMBER_STATUS) return reinterpret_cast<const int64_t*>(member)[i] == 200;
  return true;
}

// ------------------------------------------------------------------ sample
// 16 CTAs x 256 threads, one sample each (one CTA would serialise 4096 random DRAM accesses);
// the last CTA to finish turns the sample's extremes into the digit parameters.  The words the
// CTAs accumulate into are cleared by the host-side memset in front of the launch.
constexpr int QD_SAMPLE_CTAS = 16;

template <int MODE>
__global__ void __launch_bounds__(QD_SAMPLE / QD_SAMPLE_CTAS)
qd_sample_kernel(const double* __restrict__ values, int64_t n, const void* __restrict__ member, QdArgs a,
                 QdWs* ws, long long list_cap) {
  __shared__ bool is_last;
  const int t = threadIdx.x;
  const int s = blockIdx.x * blockDim.x + t;
  const int64_t width = n / QD_SAMPLE;
  const int64_t i = (int64_t)s * width + (int64_t)(qd_hash((uint32_t)s * 2654435761u + 12345u) % (uint64_t)width);
  const double v = values[i];
  const bool mem = qd_member<MODE>(member, i) && v == v;
  uint64_t mx = mem ? kvm_key(v) : 0ull, inv_mn = mem ? ~kvm_key(v) : 0ull;
  int cnt = mem;
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    const uint64_t a0 = __shfl_xor_sync(KVM_FULL, inv_mn, o), a1 = __shfl_xor_sync(KVM_FULL, mx, o);
    inv_mn = a0 > inv_mn ? a0 : inv_mn;
    mx = a1 > mx ? a1 : mx;
    cnt += __shfl_xor_sync(KVM_FULL, cnt, o);
  }
  if ((t & 31) == 0 && cnt) {
    atomicMax(&ws->s_inv_min, (unsigned long long)inv_mn);
    atomicMax(&ws->s_max, (unsigned long long)mx);
    atomicAdd(&ws->s_count, (unsigned int)cnt);
  }
  __threadfence();
  __syncthreads();
  if (t == 0) is_last = atomicAdd(&ws->s_ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  if (t == 0) {
    const uint64_t mn = ~__ldcg(&ws->s_inv_min), mxk = __ldcg(&ws->s_max);
    const int m = (int)__ldcg(&ws->s_count);
    ws->nq = a.nq;
    ws->list_cap = list_cap;
    if (m < 256) {
      ws->status = KVM_ERR_UNSUPPORTED;  // too few members in the sample to trust its range
    } else {
      // widen the sampled range by an eighth on both sides (saturating), 4096 in-range bins at most
      const uint64_t range = mxk - mn, margin = (range >> 3) + 1;
      const uint64_t lo = mn > margin ? mn - margin : 0ull;
      const uint64_t hi = mxk < QD_NONE - 1 - margin ? mxk + margin : QD_NONE - 1;
      const uint64_t r2 = hi - lo;
      const int bits = r2 ? 64 - __clzll((long long)r2) : 0;
      ws->base = lo;
      ws->shift = bits > QD_SPAN_BITS ? bits - QD_SPAN_BITS : 0;
      ws->status = KVM_OK;
    }
  }
  if (t < QD_MAXQ) {
    QdQuery q{};
    q.kind = a.kind[t];
    q.p = a.p[t];
    ws->q[t] = q;
  }
}

// ------------------------------------------------------------------ hist (+ digit records)
template <int MODE, bool VEC>
__global__ void __launch_bounds__(QD_THREADS, 2)
qd_hist_kernel(const double* __restrict__ values, int64_t n, const void* __restrict__ member, QdWs* ws,
               uint16_t* __restrict__ recs) {
  __shared__ uint32_t hist[QD_BINS];
  if (ws->status != KVM_OK) return;
  for (int i = threadIdx.x; i < QD_BINS; i += QD_THREADS) hist[i] = 0;
  const uint64_t base = ws->base;
  const int shift = ws->shift;
  __syncthreads();
  auto digit_of_value = [&](double v, bool mem) -> uint32_t {
    if (!(mem && v == v)) return 0xFFFFu;
    const uint32_t d = qd_digit(kvm_key(v), base, shift);
    atomicAdd(&hist[d], 1u);
    return d;
  };
  if (VEC) {
    // four records per thread and iteration: two 16-byte value loads, one 8-byte record store
    const int64_t nquads = n >> 2;
    for (int64_t qd = (int64_t)blockIdx.x * QD_THREADS + threadIdx.x; qd < nquads;
         qd += (int64_t)gridDim.x * QD_THREADS) {
      const int64_t i = qd << 2;
      const double2 v0 = kvm_ld2(values + i), v1 = kvm_ld2(values + i + 2);
      bool m0 = true, m1 = true, m2 = true, m3 = true;
      if (MODE == KVM_MEMBER_STATUS) {
        const int64_t* st = reinterpret_cast<const int64_t*>(member);
        const longlong2 s0 = kvm_ld2(st + i), s1 = kvm_ld2(st + i + 2);
        m0 = s0.x == 200;
        m1 = s0.y == 200;
        m2 = s1.x == 200;
        m3 = s1.y == 200;
      }
      const uint32_t d0 = digit_of_value(v0.x, m0), d1 = digit_of_value(v0.y, m1);
      const uint32_t d2 = digit_of_value(v1.x, m2), d3 = digit_of_value(v1.y, m3);
      uint2 out;
      out.x = d0 | (d1 << 16);
      out.y = d2 | (d3 << 16);
      *reinterpret_cast<uint2*>(recs + i) = out;
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
      const int64_t i = (n & ~3ll) + threadIdx.x;
      recs[i] = (uint16_t)digit_of_value(values[i], qd_member<MODE>(member, i));
    }
  } else {
    for (int64_t i = (int64_t)blockIdx.x * QD_THREADS + threadIdx.x; i < n; i += (int64_t)gridDim.x * QD_THREADS)
      recs[i] = (uint16_t)digit_of_value(kvm_ld1(values + i), qd_member<MODE>(member, i));
  }
  __syncthreads();
  for (int i = threadIdx.x; i < QD_BINS; i += QD_THREADS) {
    const uint32_t c = hist[i];
    if (c) atomicAdd(&ws->hist[i], (unsigned long long)c);
  }
}

// ------------------------------------------------------------------ resolve
__global__ void __launch_bounds__(1024, 1) qd_resolve_kernel(QdWs* ws) {
  __shared__ uint32_t cum[QD_BINS];   // exclusive prefix of the histogram (n < 2^31 on this path)
  __shared__ unsigned long long wtot[32];
  __shared__ QdQuery qs[QD_MAXQ];
  if (ws->status != KVM_OK) return;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  constexpr int PER = QD_BINS / 1024;
  unsigned long long c[PER], local = 0;
#pragma unroll
  for (int k = 0; k < PER; ++k) {
    c[k] = ws->hist[t * PER + k];
    local += c[k];
  }
  unsigned long long incl = local;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned long long o = __shfl_up_sync(KVM_FULL, incl, d);
    if (lane >= d) incl += o;
  }
  if (lane == 31) wtot[warp] = incl;
  if (t < QD_MAXQ) qs[t] = ws->q[t];
  __syncthreads();
  unsigned long long before = incl - local;
  for (int w = 0; w < warp; ++w) before += wtot[w];
#pragma unroll
  for (int k = 0; k < PER; ++k) {
    cum[t * PER + k] = (uint32_t)before;
    before += c[k];
  }
  __syncthreads();
  unsigned long long total = 0;
  for (int w = 0; w < 32; ++w) total += wtot[w];
  const int nq = ws->nq;
  if (t < nq) {
    QdQuery s = qs[t];
    s.n = (long long)total;
    s.valid = 1;
    s.bin = -1;
    s.slot = -1;
    if (total == 0) {
      s.qstatus = KVM_Q_EMPTY;
    } else {
      s.rank = resolve_rank(s.kind, s.p, s.n, 0);
      if (s.rank < 0 || s.rank >= s.n) {
        s.qstatus = KVM_Q_OUT_OF_RANGE;
      } else {
        s.qstatus = KVM_Q_OK;
        int lo = 0, hi = QD_BINS - 1;   // last bin whose exclusive prefix is <= rank
        while (lo < hi) {
          const int mid = (lo + hi + 1) >> 1;
          if ((long long)cum[mid] <= s.rank) lo = mid;
          else hi = mid - 1;
        }
        s.bin = lo;
        s.resid = s.rank - (long long)cum[lo];
        if (lo == 0 || lo == QD_BINS - 1) s.valid = 0;   // catch-all bins: the sample missed the range
      }
    }
    qs[t] = s;
  }
  __syncthreads();
  if (t == 0) {
    int nslots = 0;
    bool ok = true;
    for (int q = 0; q < nq; ++q) {
      QdQuery& s = qs[q];
      if (!s.valid) ok = false;
      if (s.qstatus != KVM_Q_OK || !s.valid) continue;
      int slot = -1;
      for (int k = 0; k < nslots; ++k)
        if (ws->slot_bin[k] == s.bin) slot = k;
      if (slot < 0) {
        slot = nslots++;
        ws->slot_bin[slot] = s.bin;
      }
      s.slot = slot;
    }
    for (int k = nslots; k < QD_MAXQ; ++k) ws->slot_bin[k] = -1;
    ws->nslots = nslots;
    if (!ok) ws->status = KVM_ERR_UNSUPPORTED;
    for (int q = 0; q < QD_MAXQ; ++q) ws->q[q] = qs[q];
  }
}

// ------------------------------------------------------------------ collect
// Streams the 2 B/record digit records; a record whose digit is a wanted bin has its value
// fetched and its key appended to that bin's list (staged per CTA, one cursor bump per flush).
template <bool VEC>
__global__ void __launch_bounds__(QD_THREADS, 2)
qd_collect_kernel(const double* __restrict__ values, int64_t n, const uint16_t* __restrict__ recs, QdWs* ws,
                  uint64_t* __restrict__ lists) {
  __shared__ uint64_t stage[QD_MAXQ][QD_STAGE];
  __shared__ uint32_t stage_n[QD_MAXQ];
  __shared__ unsigned long long stage_pos[QD_MAXQ];
  __shared__ uint32_t queue[QD_QUEUE];
  __shared__ uint32_t queue_n;
  if (ws->status != KVM_OK) return;
  const int nslots = ws->nslots;
  const long long cap = ws->list_cap;
  uint32_t want[QD_MAXQ];
#pragma unroll
  for (int k = 0; k < QD_MAXQ; ++k) want[k] = k < nslots ? (uint32_t)ws->slot_bin[k] : 0xFFFFFFFFu;
  if (threadIdx.x < QD_MAXQ) stage_n[threadIdx.x] = 0;
  if (threadIdx.x == 0) queue_n = 0;
  __syncthreads();
  // a wanted record only queues its row index: the value fetch (a random access) happens after
  // the streaming loop, all lanes at once, instead of stalling the warp in the middle of it
  auto hit = [&](int64_t i, uint32_t d) {
    const uint32_t qp = atomicAdd(&queue_n, 1u);
    if (qp < (uint32_t)QD_QUEUE) {
      queue[qp] = (uint32_t)i;   // n < 2^31 on this path
      return;
    }
    // queue full (tens of thousands of wanted records in one CTA): handle it in place
#pragma unroll
    for (int k = 0; k < QD_MAXQ; ++k)
      if (d == want[k]) {
        const uint64_t key = kvm_key(values[i]);
        const uint32_t sp = atomicAdd(&stage_n[k], 1u);
        if (sp < (uint32_t)QD_STAGE) {
          stage[k][sp] = key;
        } else {  // the stage is full: straight to the list
          const unsigned long long pos = atomicAdd(&ws->list_n[k], 1ull);
          if (pos < (unsigned long long)cap) lists[(size_t)k * (size_t)cap + pos] = key;
        }
      }
  };
  // branch-free membership test in registers: a record's digit is compared with the wanted bins
  // (4 of them when that is all there are - the usual case: three percentiles whose lo and hi
  // ranks share a bin); only a real hit leaves the straight-line code
  // both 16-bit records of a word are tested at once: x = w ^ (bin | bin << 16) has a zero half
  // where a record equals the bin, and (x - 0x00010001) & ~x & 0x80008000 flags zero halves (a
  // borrow can only raise a false flag above a true one; the hit path re-checks exactly)
  const bool few = nslots <= 4;
  uint32_t want2[QD_MAXQ];
#pragma unroll
  for (int k = 0; k < QD_MAXQ; ++k) want2[k] = k < nslots ? (want[k] & 0xFFFFu) * 0x00010001u : 0xFFFEFFFEu;  // 0xFFFE: no record has it
  auto word = [&](int64_t i, uint32_t w) {
    uint32_t z = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint32_t x = w ^ want2[k];
      z |= (x - 0x00010001u) & ~x;
    }
    if (!few) {
#pragma unroll
      for (int k = 4; k < QD_MAXQ; ++k) {
        const uint32_t x = w ^ want2[k];
        z |= (x - 0x00010001u) & ~x;
      }
    }
    z &= 0x80008000u;
    if (z) {
      const uint32_t d0 = w & 0xFFFFu, d1 = w >> 16;
      if ((z & 0x8000u) && d0 != 0xFFFFu) hit(i, d0);
      if ((z >> 16) && d1 != 0xFFFFu) hit(i + 1, d1);
    }
  };
  if (VEC) {
    const int64_t noct = n >> 3;
    // 32 contiguous bytes (16 records) per thread and iteration, two loads in flight
    const int64_t npair = noct >> 1;
    const uint4* __restrict__ r16 = reinterpret_cast<const uint4*>(recs);
    for (int64_t o = (int64_t)blockIdx.x * QD_THREADS + threadIdx.x; o < npair; o += (int64_t)gridDim.x * QD_THREADS) {
      const uint4 r0 = __ldg(r16 + 2 * o), r1 = __ldg(r16 + 2 * o + 1);
      const int64_t i = o << 4;
      word(i, r0.x);
      word(i + 2, r0.y);
      word(i + 4, r0.z);
      word(i + 6, r0.w);
      word(i + 8, r1.x);
      word(i + 10, r1.y);
      word(i + 12, r1.z);
      word(i + 14, r1.w);
    }
    if ((noct & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
      const uint4 r0 = __ldg(r16 + noct - 1);
      const int64_t i = (noct - 1) << 3;
      word(i, r0.x);
      word(i + 2, r0.y);
      word(i + 4, r0.z);
      word(i + 6, r0.w);
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 7)) {
      const int64_t i = (n & ~7ll) + threadIdx.x;
      word(i, (uint32_t)recs[i] | 0xFFFF0000u);   // upper half: "not a member"
    }
  } else {
    for (int64_t i = (int64_t)blockIdx.x * QD_THREADS + threadIdx.x; i < n; i += (int64_t)gridDim.x * QD_THREADS)
      word(i, (uint32_t)recs[i] | 0xFFFF0000u);
  }
  __syncthreads();
  {  // drain the queue: every thread takes queued rows, fetches the value, appends the key
    const uint32_t qn = queue_n < (uint32_t)QD_QUEUE ? queue_n : (uint32_t)QD_QUEUE;
    for (uint32_t j = threadIdx.x; j < qn; j += QD_THREADS) {
      const int64_t i = (int64_t)queue[j];
      const uint32_t d = recs[i];
      const uint64_t key = kvm_key(values[i]);
#pragma unroll
      for (int k = 0; k < QD_MAXQ; ++k)
        if (d == want[k]) {
          const uint32_t sp = atomicAdd(&stage_n[k], 1u);
          if (sp < (uint32_t)QD_STAGE) {
            stage[k][sp] = key;
          } else {
            const unsigned long long pos = atomicAdd(&ws->list_n[k], 1ull);
            if (pos < (unsigned long long)cap) lists[(size_t)k * (size_t)cap + pos] = key;
          }
        }
    }
  }
  __syncthreads();
  if (threadIdx.x < QD_MAXQ) {
    const uint32_t c = stage_n[threadIdx.x] < (uint32_t)QD_STAGE ? stage_n[threadIdx.x] : (uint32_t)QD_STAGE;
    stage_n[threadIdx.x] = c;
    if (c) stage_pos[threadIdx.x] = atomicAdd(&ws->list_n[threadIdx.x], (unsigned long long)c);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < QD_MAXQ * QD_STAGE; i += QD_THREADS) {
    const int k = i / QD_STAGE, j = i % QD_STAGE;
    if ((uint32_t)j < stage_n[k]) {
      const unsigned long long pos = stage_pos[k] + j;
      if (pos < (unsigned long long)cap) lists[(size_t)k * (size_t)cap + pos] = stage[k][j];
    }
  }
}

// ------------------------------------------------------------------ rank
__global__ void __launch_bounds__(1024, 1)
qd_rank_kernel(QdWs* ws, const uint64_t* __restrict__ lists, kvm_query_result* __restrict__ results) {
  __shared__ uint32_t hist[2048];
  __shared__ uint64_t wmin[32], wmax[32];
  __shared__ uint32_t wtot[32];
  __shared__ uint32_t sel_bin, sel_resid, small_n;
  __shared__ uint64_t answer;
  const int q = blockIdx.x;
  if (q >= ws->nq) return;
  const QdQuery s = ws->q[q];
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const bool need = ws->status == KVM_OK && s.valid && s.qstatus == KVM_Q_OK;
  if (!need) {
    // empty multiset / rank out of range: nothing to look up; a declined call reports NaN
    if (t == 0) results[q] = kvm_query_result{nan, s.n, s.rank, s.qstatus, 0};
    return;
  }
  const long long cap = ws->list_cap;
  const unsigned long long m = ws->list_n[s.slot];
  if (m > (unsigned long long)cap || (long long)m <= s.resid) {
    if (t == 0) {
      ws->status = KVM_ERR_UNSUPPORTED;  // all writers agree
      results[q] = kvm_query_result{nan, s.n, s.rank, s.qstatus, 0};
    }
    return;
  }
  const uint64_t* __restrict__ keys = lists + (size_t)s.slot * (size_t)cap;
  uint64_t* small = ws->small[q];
  // the keys of bin d lie in [base + (d-1) << shift, base + d << shift): no min/max pass needed
  (void)wmin;
  (void)wmax;
  hist[t] = 0;
  hist[t + 1024] = 0;
  if (t == 0) small_n = 0;
  const int bshift = ws->shift;
  const uint64_t mn = ws->base + ((uint64_t)(s.bin - 1) << bshift);
  const int sh2 = bshift > 11 ? bshift - 11 : 0;
  __syncthreads();
  for (unsigned long long i0 = t; i0 < m; i0 += 8 * 1024) {
    uint64_t kk[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) kk[u] = i0 + u * 1024 < m ? __ldcg(keys + i0 + u * 1024) : QD_NONE;
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (kk[u] != QD_NONE) atomicAdd(&hist[(uint32_t)((kk[u] - mn) >> sh2) & 2047u], 1u);
  }
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
    long long before = (long long)(incl - (c0 + c1));
    for (int w = 0; w < warp; ++w) before += wtot[w];
    const long long r = s.resid;
    if (before <= r && r < before + c0) {
      sel_bin = 2 * t;
      sel_resid = (uint32_t)(r - before);
    } else if (before + c0 <= r && r < before + c0 + c1) {
      sel_bin = 2 * t + 1;
      sel_resid = (uint32_t)(r - before - c0);
    }
  }
  __syncthreads();
  const uint32_t sb = sel_bin, sr = sel_resid;
  for (unsigned long long i0 = t; i0 < m; i0 += 8 * 1024) {
    uint64_t kk[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) kk[u] = i0 + u * 1024 < m ? __ldcg(keys + i0 + u * 1024) : QD_NONE;
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (kk[u] != QD_NONE && ((uint32_t)((kk[u] - mn) >> sh2) & 2047u) == sb) {
        const uint32_t pos = atomicAdd(&small_n, 1u);
        if (pos < (uint32_t)QD_SMALL) small[pos] = kk[u];
      }
  }
  __syncthreads();
  const uint32_t m2 = small_n;
  if (m2 > (uint32_t)QD_SMALL) {  // one value repeated thousands of times: the general select's job
    if (t == 0) {
      ws->status = KVM_ERR_UNSUPPORTED;
      results[q] = kvm_query_result{nan, s.n, s.rank, s.qstatus, 0};
    }
    return;
  }
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
  if (t == 0) results[q] = kvm_query_result{kvm_unkey(answer), s.n, s.rank, KVM_Q_OK, 0};
}

int qd_grid(int64_t n) { return kvm_grid_for(n, QD_THREADS * 4 * 4, 2); }
long long qd_list_cap(int64_t n) {
  const long long c = (long long)(n / 64) + 16384;   // a bin holds ~n/1500 of smooth data
  return (c + 1) & ~1ll;
}

}  // namespace

extern "C" int64_t kvm_column_select_sampled_min_n(void) { return 1 << 18; }

extern "C" size_t kvm_column_select_sampled_workspace_bytes(int64_t n) {
  if (n < 0) n = 0;
  return QD_HEADER + kvm_align_up((size_t)n * sizeof(uint16_t) + 16, 256) +
         (size_t)QD_MAXQ * (size_t)qd_list_cap(n) * sizeof(uint64_t) + 256;
}

extern "C" int kvm_column_select_sampled(const double* values, int64_t n, int32_t mode, const void* member,
                                       const kvm_query* queries, int32_t nq, kvm_query_result* results,
                                       void* workspace, size_t workspace_bytes, void* stream) {
  KVM_REQUIRE(values && queries && results && workspace, "kvm_column_select_sampled: null argument");
  if (!(mode == KVM_MEMBER_ALL || (mode == KVM_MEMBER_STATUS && member)) || nq < 1 || nq > QD_MAXQ ||
      n < (1 << 18) || n >= (1ll << 31)) {
    kvm_set_error("kvm_column_select_sampled: needs member mode ALL/STATUS, 1..%d queries, 2^18 <= n < 2^31; "
                  "use kvm_column_select", QD_MAXQ);
    return KVM_ERR_UNSUPPORTED;
  }
  QdArgs a{};
  a.nq = nq;
  for (int q = 0; q < nq; ++q) {
    const kvm_query& Q = queries[q];
    if (Q.rank_kind == KVM_RANK_ABSOLUTE || Q.rank_kind < 0 || Q.rank_kind > KVM_RANK_FLOOR_PNM1 || Q.set_mask != 1u) {
      kvm_set_error("kvm_column_select_sampled: query %d is not a quantile of set 0; use kvm_column_select", q);
      return KVM_ERR_UNSUPPORTED;
    }
    a.kind[q] = Q.rank_kind;
    a.p[q] = Q.p;
  }
  if (workspace_bytes < kvm_column_select_sampled_workspace_bytes(n)) {
    kvm_set_error("kvm_column_select_sampled: workspace %zu < %zu bytes", workspace_bytes,
                  kvm_column_select_sampled_workspace_bytes(n));
    return KVM_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  unsigned char* wsb = reinterpret_cast<unsigned char*>(workspace);
  QdWs* ws = reinterpret_cast<QdWs*>(wsb);
  uint16_t* recs = reinterpret_cast<uint16_t*>(wsb + QD_HEADER);
  uint64_t* lists = reinterpret_cast<uint64_t*>(wsb + QD_HEADER + kvm_align_up((size_t)n * sizeof(uint16_t) + 16, 256));
  const long long cap = qd_list_cap(n);
  const int grid = qd_grid(n);
  const bool vec = (reinterpret_cast<uintptr_t>(values) & 15u) == 0 &&
                   (mode != KVM_MEMBER_STATUS || (reinterpret_cast<uintptr_t>(member) & 15u) == 0);
  KVM_CUDA_CHECK(cudaMemsetAsync(wsb + QD_CLEAR_BEGIN, 0, QD_CLEAR_END - QD_CLEAR_BEGIN, st));
  constexpr int STH = QD_SAMPLE / QD_SAMPLE_CTAS;
  if (mode == KVM_MEMBER_STATUS) {
    qd_sample_kernel<KVM_MEMBER_STATUS><<<QD_SAMPLE_CTAS, STH, 0, st>>>(values, n, member, a, ws, cap);
    if (vec) qd_hist_kernel<KVM_MEMBER_STATUS, true><<<grid, QD_THREADS, 0, st>>>(values, n, member, ws, recs);
    else qd_hist_kernel<KVM_MEMBER_STATUS, false><<<grid, QD_THREADS, 0, st>>>(values, n, member, ws, recs);
  } else {
    qd_sample_kernel<KVM_MEMBER_ALL><<<QD_SAMPLE_CTAS, STH, 0, st>>>(values, n, member, a, ws, cap);
    if (vec) qd_hist_kernel<KVM_MEMBER_ALL, true><<<grid, QD_THREADS, 0, st>>>(values, n, member, ws, recs);
    else qd_hist_kernel<KVM_MEMBER_ALL, false><<<grid, QD_THREADS, 0, st>>>(values, n, member, ws, recs);
  }
  qd_resolve_kernel<<<1, 1024, 0, st>>>(ws);
  qd_collect_kernel<true><<<kvm_grid_for(n, QD_THREADS * 8 * 4, 2), QD_THREADS, 0, st>>>(values, n, recs, ws, lists);
  qd_rank_kernel<<<nq, 1024, 0, st>>>(ws, lists, results);
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}
