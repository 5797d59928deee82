// kvm_qdigit.cu - one-and-an-eighth-pass exact quantiles of one multiset (the 16 B/record
// percentile path; analyze.py:59-81).
//
// Two launches, no memset; the second is a programmatic dependent launch (its CTAs become
// resident while the first kernel drains, so there is no launch gap between the two):
//   hist     prologue: every CTA draws the SAME 512 stratified pseudo-random members (one load per
//            thread, L2 hits for all but the first CTA) and derives an ESTIMATE of the key range
//            from them - no separate sampling launch, no grid-wide dependency; CTA 0 also clears
//            the accumulators and publishes a per-call token that the flushes wait for.
//            body: ONE pass over the data (16 B/record, 256-bit evict-first loads): 13-bit digit
//            of every member key relative to the estimated range (bins 0 and 8191 catch what lies
//            far outside it), histogram in shared memory, and the digit written as a 2-byte record
//            (0xFFFF = not a member; stored with an L2 keep policy - the next kernel reads them).
//            epilogue (last CTA to finish, ticket): exact member count = sum of the histogram,
//            rank of every query, the bin holding it and the residual rank.
//   collect  pass over the 2 B/record digit records; the value of a record is fetched only when
//            its digit is a wanted bin (~0.1 % of the records): its key goes to that bin's list and
//            its next 11 bits into that bin's sub-histogram.
//            epilogue (last CTA, ticket): one 64-thread group per query walks the sub-histogram to
//            the sub-bin holding the rank; all threads then stream the lists once and the handful
//            of keys of each query's sub-bin is ranked.
// Exact whenever it answers.  It declines (status word KVM_ERR_UNSUPPORTED, the caller runs the
// general select) when a rank falls into a catch-all bin (the sample missed the range), when a
// list overflows (heavy duplicates spread over many keys) or when the sample holds too few members.

#include <atomic>
#include <cstddef>
#include <cstdlib>

#include "kvm_common.cuh"
#include "kvm_log_device.cuh"

namespace {

constexpr int QD_MAXQ = 8;
constexpr int QD_THREADS = 512;
constexpr int QD_SAMPLE = QD_THREADS;  // one stratified sample per thread, the same in every CTA
constexpr int QD_BINS = 8192;
constexpr int QD_SPAN_BITS = 12;   // the widened sample range spans at most 4096 bins; the rest lies beyond it
constexpr int QD_SUB_BITS = 11;    // sub-histogram of a wanted bin
constexpr int QD_SUB = 1 << QD_SUB_BITS;
constexpr int QD_SMALL = 4096;     // keys ranked by counting at the very end
constexpr int QD_QUEUE = 2048;     // wanted rows a CTA queues for the drain after its streaming loop
constexpr int QD_STAGE = 384;      // keys a CTA stages per wanted bin before going to global memory
constexpr int QD_GROUP = QD_THREADS / QD_MAXQ;   // threads that finish one query in the last CTA
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
  unsigned long long ready;   // == the call's token once CTA 0 of the hist kernel has cleared the block below
  // ---- cleared by CTA 0 of the hist kernel at the start of every call (QD_CLEAR_* below)
  alignas(16) unsigned int s_ticket;
  unsigned int c_ticket, pad4[2];
  unsigned long long list_n[QD_MAXQ];
  unsigned long long hist[QD_BINS];
  unsigned int hist2[QD_MAXQ][QD_SUB];
  // ----
  alignas(16) uint64_t small[QD_MAXQ][QD_SMALL];
};

constexpr size_t QD_HEADER = (sizeof(QdWs) + 255) / 256 * 256;
constexpr size_t QD_CLEAR_BEGIN = offsetof(QdWs, s_ticket), QD_CLEAR_END = offsetof(QdWs, small);
static_assert(QD_CLEAR_BEGIN % 16 == 0 && QD_CLEAR_END % 16 == 0, "cleared with 16-byte stores");

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

__device__ __forceinline__ unsigned long long qd_ld_acquire(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void qd_st_release(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// ------------------------------------------------------------------ resolve (last CTA of the hist kernel)
// `cum` (shared, QD_BINS words) receives the exclusive prefix of the global histogram; every
// query finds its bin by bisection.
__device__ __forceinline__ void qd_resolve(QdWs* ws, uint32_t* cum, const QdArgs& a) {
  __shared__ unsigned long long wtot[QD_THREADS / 32];
  __shared__ QdQuery qs[QD_MAXQ];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  constexpr int PER = QD_BINS / QD_THREADS;
  unsigned long long c[PER], local = 0;
#pragma unroll
  for (int k = 0; k < PER; ++k) {
    c[k] = __ldcg(&ws->hist[t * PER + k]);
    local += c[k];
  }
  unsigned long long incl = local;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned long long o = __shfl_up_sync(KVM_FULL, incl, d);
    if (lane >= d) incl += o;
  }
  if (lane == 31) wtot[warp] = incl;
  __syncthreads();
  unsigned long long before = incl - local;
  for (int w = 0; w < warp; ++w) before += wtot[w];
#pragma unroll
  for (int k = 0; k < PER; ++k) {
    cum[t * PER + k] = (uint32_t)before;   // n < 2^31 on this path
    before += c[k];
  }
  __syncthreads();
  unsigned long long total = 0;
  for (int w = 0; w < QD_THREADS / 32; ++w) total += wtot[w];
  const int nq = a.nq;
  if (t < QD_MAXQ) {
    QdQuery s{};
    if (t < nq) {
      int kind = 0;
      double p = 0.0;
#pragma unroll
      for (int k = 0; k < QD_MAXQ; ++k)   // (no dynamic indexing of the kernel parameter)
        if (k == t) {
          kind = a.kind[k];
          p = a.p[k];
        }
      s.kind = kind;
      s.p = p;
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
    }
    qs[t] = s;
  }
  __syncthreads();
  if (t == 0) {
    int nslots = 0;
    bool ok = true;
    int sb[QD_MAXQ];
    for (int q = 0; q < nq; ++q) {
      QdQuery& s = qs[q];
      if (!s.valid) ok = false;
      if (s.qstatus != KVM_Q_OK || !s.valid) continue;
      int slot = -1;
      for (int k = 0; k < nslots; ++k)
        if (sb[k] == s.bin) slot = k;
      if (slot < 0) {
        slot = nslots++;
        sb[slot] = s.bin;
      }
      s.slot = slot;
    }
    for (int k = 0; k < QD_MAXQ; ++k) ws->slot_bin[k] = k < nslots ? sb[k] : -1;
    ws->nslots = nslots;
    if (!ok) ws->status = KVM_ERR_UNSUPPORTED;
    for (int q = 0; q < QD_MAXQ; ++q) ws->q[q] = qs[q];
  }
}

// ------------------------------------------------------------------ hist (sample + digit records + resolve)
// VEC: both columns are 32-byte aligned -> eight records per thread and iteration (two 256-bit
// loads per column, one 16-byte record store); CTA b owns the contiguous range of row octets
// [b*chunk, (b+1)*chunk), so every CTA streams the same number of bytes.
template <int MODE, bool VEC>
__global__ void __launch_bounds__(QD_THREADS, 2)
qd_hist_kernel(const double* __restrict__ values, int64_t n, const void* __restrict__ member, QdWs* ws,
               uint16_t* __restrict__ recs, QdArgs a, long long list_cap, unsigned long long token) {
  __shared__ uint32_t hist[QD_BINS];
  __shared__ unsigned long long red_inv[QD_THREADS / 32], red_mx[QD_THREADS / 32];
  __shared__ int red_cnt[QD_THREADS / 32];
  __shared__ unsigned long long s_base;
  __shared__ int s_shift, s_ok;
  __shared__ bool is_last;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  kvm_pdl_launch_dependents();   // the collect kernel may take the SM slots this grid frees as it drains

  // ---- CTA 0: clear what the call accumulates into, then publish the call's token.  No other
  // CTA touches those words before it has seen the token (the wait in front of the flush below),
  // and CTA 0 waits for nobody, so there is no ordering between CTAs to rely on.
  if (blockIdx.x == 0) {
    uint4* z = reinterpret_cast<uint4*>(reinterpret_cast<unsigned char*>(ws) + QD_CLEAR_BEGIN);
    constexpr int NZ = (int)((QD_CLEAR_END - QD_CLEAR_BEGIN) / 16);
    for (int i = t; i < NZ; i += QD_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
    __threadfence();
    __syncthreads();
    if (t == 0) qd_st_release(&ws->ready, token);
  }
  for (int i = t; i < QD_BINS; i += QD_THREADS) hist[i] = 0;

  // ---- the first tile of the stream is requested before the sample is reduced (its latency
  // hides the sample's)
  const int64_t nocts = n >> 3;
  const int64_t chunk = (nocts + gridDim.x - 1) / gridDim.x;
  const int64_t o_end = min(nocts, ((int64_t)blockIdx.x + 1) * chunk);
  int64_t o = (int64_t)blockIdx.x * chunk + t;
  const int64_t* st = reinterpret_cast<const int64_t*>(member);
  kvm_u64x4 v0{}, v1{}, s0{}, s1{};
  auto load8 = [&](int64_t oc, kvm_u64x4& a0, kvm_u64x4& a1, kvm_u64x4& b0, kvm_u64x4& b1) {
    const int64_t i = oc << 3;
    a0 = kvm_ld4_stream(values + i);
    a1 = kvm_ld4_stream(values + i + 4);
    if (MODE == KVM_MEMBER_STATUS) {
      b0 = kvm_ld4_stream(st + i);
      b1 = kvm_ld4_stream(st + i + 4);
    }
  };
  bool have = VEC && o < o_end;
  if (have) load8(o, v0, v1, s0, s1);

  // ---- sample: thread t draws one element of stratum t - the same element in every CTA, so all
  // CTAs derive identical digit parameters without talking to each other
  {
    const int64_t width = n / QD_SAMPLE;
    const int64_t i = (int64_t)t * width + (int64_t)(qd_hash((uint32_t)t * 2654435761u + 12345u) % (uint64_t)width);
    const double v = __ldg(values + i);
    bool mem = v == v;
    if (MODE == KVM_MEMBER_STATUS) mem = mem && __ldg(st + i) == 200;
    unsigned long long mx = mem ? kvm_key(v) : 0ull, inv_mn = mem ? ~kvm_key(v) : 0ull;
    int cnt = mem;
#pragma unroll
    for (int d = 16; d; d >>= 1) {
      const unsigned long long a0 = __shfl_xor_sync(KVM_FULL, inv_mn, d), a1 = __shfl_xor_sync(KVM_FULL, mx, d);
      inv_mn = a0 > inv_mn ? a0 : inv_mn;
      mx = a1 > mx ? a1 : mx;
      cnt += __shfl_xor_sync(KVM_FULL, cnt, d);
    }
    if (lane == 0) {
      red_inv[warp] = inv_mn;
      red_mx[warp] = mx;
      red_cnt[warp] = cnt;
    }
    __syncthreads();
    if (t == 0) {
      unsigned long long inv = 0, mxk = 0;
      int m = 0;
      for (int w = 0; w < QD_THREADS / 32; ++w) {
        inv = red_inv[w] > inv ? red_inv[w] : inv;
        mxk = red_mx[w] > mxk ? red_mx[w] : mxk;
        m += red_cnt[w];
      }
      const uint64_t mn = ~inv;
      const int ok = m >= QD_SAMPLE / 8;   // too few members in the sample to trust its range otherwise
      uint64_t lo = 0;
      int shift = 0;
      if (ok) {
        // widen the sampled range by an eighth on both sides (saturating), 4096 in-range bins at most
        const uint64_t range = mxk - mn, margin = (range >> 3) + 1;
        lo = mn > margin ? mn - margin : 0ull;
        const uint64_t hi = mxk < QD_NONE - 1 - margin ? mxk + margin : QD_NONE - 1;
        const uint64_t r2 = hi - lo;
        const int bits = r2 ? 64 - __clzll((long long)r2) : 0;
        shift = bits > QD_SPAN_BITS ? bits - QD_SPAN_BITS : 0;
      }
      s_base = lo;
      s_shift = shift;
      s_ok = ok;
      if (blockIdx.x == 0) {
        ws->nq = a.nq;
        ws->list_cap = list_cap;
        ws->base = lo;
        ws->shift = shift;
        ws->nslots = 0;
        ws->status = ok ? KVM_OK : KVM_ERR_UNSUPPORTED;
        if (!ok)
          for (int q = 0; q < QD_MAXQ; ++q) ws->q[q] = QdQuery{};
      }
    }
    __syncthreads();
  }
  if (!s_ok) return;   // every CTA takes the same decision; the status word says "declined"
  const uint64_t base = s_base;
  const int shift = s_shift;
  auto digit_of_value = [&](double v, bool mem) -> uint32_t {
    if (!(mem && v == v)) return 0xFFFFu;
    const uint32_t d = qd_digit(kvm_key(v), base, shift);
    atomicAdd(&hist[d], 1u);
    return d;
  };
  auto dig = [&](unsigned long long vb, unsigned long long sb) -> uint32_t {
    return digit_of_value(__longlong_as_double((long long)vb), MODE != KVM_MEMBER_STATUS || sb == 200ull);
  };
  if (VEC) {
    const unsigned long long keep = kvm_policy_evict_last();
    while (have) {
      uint4 out;
      out.x = dig(v0.x, s0.x) | (dig(v0.y, s0.y) << 16);
      out.y = dig(v0.z, s0.z) | (dig(v0.w, s0.w) << 16);
      out.z = dig(v1.x, s1.x) | (dig(v1.y, s1.y) << 16);
      out.w = dig(v1.z, s1.z) | (dig(v1.w, s1.w) << 16);
      kvm_st4_u32_keep(recs + (o << 3), out, keep);
      // (no register double-buffering: 128 B in flight per thread x 1024 threads per SM is
      // several times what the memory system needs; the other warps cover this one's wait)
      o += QD_THREADS;
      have = o < o_end;
      if (have) load8(o, v0, v1, s0, s1);
    }
    if (blockIdx.x == 0 && t < (n & 7)) {
      const int64_t i = (n & ~7ll) + t;
      recs[i] = (uint16_t)digit_of_value(values[i], qd_member<MODE>(member, i));
    }
  } else {
    for (int64_t i = (int64_t)blockIdx.x * QD_THREADS + t; i < n; i += (int64_t)gridDim.x * QD_THREADS)
      recs[i] = (uint16_t)digit_of_value(kvm_ld1(values + i), qd_member<MODE>(member, i));
  }
  __syncthreads();
  if (t == 0) {
    while (qd_ld_acquire(&ws->ready) != token) {}   // CTA 0 has cleared the accumulators (long ago)
  }
  __syncthreads();
  for (int i = t; i < QD_BINS; i += QD_THREADS) {
    const uint32_t c = hist[i];
    if (c) atomicAdd(&ws->hist[i], (unsigned long long)c);
  }
  __threadfence();
  __syncthreads();
  if (t == 0) is_last = atomicAdd(&ws->s_ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  qd_resolve(ws, hist, a);
}

// ------------------------------------------------------------------ collect (+ rank in the last CTA)
struct QdCollectSmem {
  uint64_t stage[QD_MAXQ][QD_STAGE];
  uint32_t queue[QD_QUEUE];
  uint32_t stage_n[QD_MAXQ];
  unsigned long long stage_pos[QD_MAXQ];
  uint32_t queue_n;
  int is_last;
};

// Finish of all queries by the last CTA of the collect kernel.  Three phases:
//   A  one group of QD_GROUP threads per query walks the wanted bin's sub-histogram to the sub-bin
//      holding the residual rank;
//   B  ALL threads stream the slot lists once (every load of a thread is in flight at the same
//      time - a 64-thread group would need dozens of dependent round trips to L2); a key that
//      lies in the sub-bin of a query updates that query's min / max and is gathered;
//   C  per query: min == max settles it (thousands of copies of one key - values rounded to three
//      decimals - need no ranking), otherwise the handful of gathered keys is ranked by counting.
struct QdFinSmem {
  uint32_t gsum[QD_MAXQ][QD_GROUP];
  uint32_t sb[QD_MAXQ], sr[QD_MAXQ], scount[QD_MAXQ], cnt[QD_MAXQ];
  int32_t state[QD_MAXQ];   // 0: nothing to look up, 1: answered, 2: needs its sub-bin's keys, 3: decline
  int32_t slot[QD_MAXQ];
  unsigned long long kmin[QD_MAXQ], kmax[QD_MAXQ];
  unsigned long long slot_m[QD_MAXQ], slot_base[QD_MAXQ], slot_begin[QD_MAXQ + 1];
};

__device__ __noinline__ void qd_finish(QdWs* ws, const uint64_t* __restrict__ lists, QdFinSmem& F,
                                          kvm_query_result* __restrict__ results) {
  const int t = threadIdx.x, g = t / QD_GROUP, gt = t % QD_GROUP;
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  const int nq = ws->nq, nslots = ws->nslots;
  const long long cap = ws->list_cap;
  const int bshift = ws->shift;
  const int sh2 = bshift > QD_SUB_BITS ? bshift - QD_SUB_BITS : 0;
  const uint64_t base = ws->base;
  if (t < QD_MAXQ) {
    const unsigned long long m = t < nslots ? __ldcg(&ws->list_n[t]) : 0ull;
    F.slot_m[t] = m;
    F.slot_base[t] = t < nslots ? base + ((uint64_t)(ws->slot_bin[t] - 1) << bshift) : 0ull;
    F.kmin[t] = QD_NONE;
    F.kmax[t] = 0ull;
    F.cnt[t] = 0;
    F.state[t] = 0;
  }
  __syncthreads();
  if (t == 0) {
    unsigned long long acc = 0;
    for (int k = 0; k < QD_MAXQ; ++k) {
      F.slot_begin[k] = acc;
      acc += F.slot_m[k] <= (unsigned long long)cap ? F.slot_m[k] : 0ull;   // an overflowed list is not scanned
    }
    F.slot_begin[QD_MAXQ] = acc;
  }
  // ---- A: sub-bin of every query
  QdQuery s{};
  bool need = false;
  if (g < nq) {
    s = ws->q[g];
    need = s.valid && s.qstatus == KVM_Q_OK;
    if (gt == 0) F.slot[g] = need ? s.slot : -1;
  }
  constexpr int PER = QD_SUB / QD_GROUP;
  uint32_t c[PER], local = 0;
  bool overflow = false;
  if (need) {
    const unsigned long long m = F.slot_m[s.slot];
    overflow = m > (unsigned long long)cap || (long long)m <= s.resid;
    if (!overflow) {
      const unsigned int* __restrict__ h2 = ws->hist2[s.slot];
#pragma unroll
      for (int k = 0; k < PER; k += 4) {
        const uint4 x = __ldcg(reinterpret_cast<const uint4*>(h2 + gt * PER + k));
        c[k] = x.x; c[k + 1] = x.y; c[k + 2] = x.z; c[k + 3] = x.w;
        local += x.x + x.y + x.z + x.w;
      }
    }
    F.gsum[g][gt] = local;
  }
  __syncthreads();
  if (need && !overflow) {
    uint32_t before = 0;
    for (int k = 0; k < gt; ++k) before += F.gsum[g][k];
    const uint32_t r = (uint32_t)s.resid;
    if (before <= r && r < before + local) {
      uint32_t b = before;
      int k = 0;
      for (; k < PER - 1; ++k) {
        if (r < b + c[k]) break;
        b += c[k];
      }
      F.sb[g] = (uint32_t)(gt * PER + k);   // the sub-bin
      F.sr[g] = r - b;                      // rank inside it
      F.scount[g] = c[k];                   // its population
      F.state[g] = sh2 == 0 ? 1 : (c[k] <= (uint32_t)QD_SMALL ? 2 : 4);   // 4: too many to gather - min/max only
    }
  } else if (need && gt == 0) {
    F.state[g] = 3;
  }
  __syncthreads();
  // ---- B: one pass of all threads over all slot lists
  {
    const unsigned long long total = F.slot_begin[QD_MAXQ];
    constexpr int U = 8;
    for (unsigned long long i0 = t; i0 < total; i0 += (unsigned long long)U * QD_THREADS) {
      uint64_t kk[U];
      int ks[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const unsigned long long i = i0 + (unsigned long long)u * QD_THREADS;
        ks[u] = -1;
        kk[u] = 0;
        if (i < total) {
          int k = 0;
#pragma unroll
          for (int j = 1; j < QD_MAXQ; ++j) k += i >= F.slot_begin[j];
          ks[u] = k;
          kk[u] = __ldcg(lists + (size_t)k * (size_t)cap + (size_t)(i - F.slot_begin[k]));
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (ks[u] < 0) continue;
        const uint32_t sub = (uint32_t)((kk[u] - F.slot_base[ks[u]]) >> sh2);
        for (int q = 0; q < nq; ++q) {
          const int stq = F.state[q];
          if ((stq == 2 || stq == 4) && F.slot[q] == ks[u] && F.sb[q] == sub) {
            atomicMin(&F.kmin[q], (unsigned long long)kk[u]);
            atomicMax(&F.kmax[q], (unsigned long long)kk[u]);
            if (stq == 2) {
              const uint32_t pos = atomicAdd(&F.cnt[q], 1u);
              if (pos < (uint32_t)QD_SMALL) ws->small[q][pos] = kk[u];
            }
          }
        }
      }
    }
  }
  __threadfence_block();
  __syncthreads();
  // ---- C: answers
  if (g >= nq) return;
  const int stq = F.state[g];
  if (stq == 0) {   // empty multiset / rank out of range: nothing to look up
    if (gt == 0) results[g] = kvm_query_result{nan, s.n, s.rank, s.qstatus, 0};
    return;
  }
  auto decline = [&]() {
    if (gt == 0) {
      ws->status = KVM_ERR_UNSUPPORTED;  // all writers agree
      results[g] = kvm_query_result{nan, s.n, s.rank, s.qstatus, 0};
    }
  };
  if (stq == 3) {
    decline();
    return;
  }
  if (stq == 1) {   // the sub-bin is one key
    if (gt == 0) results[g] = kvm_query_result{kvm_unkey(F.slot_base[s.slot] + F.sb[g]), s.n, s.rank, KVM_Q_OK, 0};
    return;
  }
  if (F.kmin[g] == F.kmax[g]) {   // one distinct key (any number of copies)
    if (gt == 0) results[g] = kvm_query_result{kvm_unkey(F.kmin[g]), s.n, s.rank, KVM_Q_OK, 0};
    return;
  }
  const uint32_t m2 = F.cnt[g], sr = F.sr[g];
  if (stq == 4 || m2 != F.scount[g]) {   // too many distinct candidates for the scratch: the general select's job
    decline();
    return;
  }
  const uint64_t* small = ws->small[g];
  for (uint32_t i = gt; i < m2; i += QD_GROUP) {
    const uint64_t k = __ldcg(small + i);
    uint32_t less = 0, leq = 0;
    for (uint32_t j = 0; j < m2; ++j) {
      const uint64_t x = __ldcg(small + j);
      less += x < k;
      leq += x <= k;
    }
    if (less <= sr && sr < leq) results[g] = kvm_query_result{kvm_unkey(k), s.n, s.rank, KVM_Q_OK, 0};  // all writers agree
  }
}

// Streams the 2 B/record digit records; a record whose digit is a wanted bin has its value
// fetched, its key appended to that bin's list (staged per CTA, one cursor bump per flush) and
// the key's next 11 bits counted into the bin's sub-histogram.
template <bool VEC>
__global__ void __launch_bounds__(QD_THREADS, 2)
qd_collect_kernel(const double* __restrict__ values, int64_t n, const uint16_t* __restrict__ recs, QdWs* ws,
                  uint64_t* __restrict__ lists, kvm_query_result* __restrict__ results) {
  __shared__ QdCollectSmem S;
  const int t = threadIdx.x;
  if (t < QD_MAXQ) S.stage_n[t] = 0;
  if (t == 0) S.queue_n = 0;
  kvm_pdl_wait();   // everything above overlapped the hist kernel's tail; its results are visible from here on
  if (ws->status != KVM_OK) {   // declined by the sample or by the resolve: report and re-arm the handshake
    if (blockIdx.x == 0) {
      const double nan = __longlong_as_double(0x7ff8000000000000ll);
      if (t < ws->nq) results[t] = kvm_query_result{nan, ws->q[t].n, ws->q[t].rank, ws->q[t].qstatus, 0};
      if (t == 0) ws->ready = 0ull;
    }
    return;
  }
  const int nslots = ws->nslots;
  const long long cap = ws->list_cap;
  const int bshift = ws->shift;
  const uint64_t base = ws->base;
  const int sh2 = bshift > QD_SUB_BITS ? bshift - QD_SUB_BITS : 0;
  uint32_t want[QD_MAXQ];
#pragma unroll
  for (int k = 0; k < QD_MAXQ; ++k) want[k] = k < nslots ? (uint32_t)ws->slot_bin[k] : 0xFFFFFFFFu;
  __syncthreads();
  // key of a wanted record -> sub-histogram + staged list entry
  auto take = [&](int64_t i, uint32_t d) {
    const uint64_t key = kvm_key(values[i]);
#pragma unroll
    for (int k = 0; k < QD_MAXQ; ++k)
      if (d == want[k]) {
        const uint64_t bin_base = base + ((uint64_t)(d - 1) << bshift);
        atomicAdd(&ws->hist2[k][(uint32_t)((key - bin_base) >> sh2) & (QD_SUB - 1)], 1u);
        const uint32_t sp = atomicAdd(&S.stage_n[k], 1u);
        if (sp < (uint32_t)QD_STAGE) {
          S.stage[k][sp] = key;
        } else {  // the stage is full: straight to the list
          const unsigned long long pos = atomicAdd(&ws->list_n[k], 1ull);
          if (pos < (unsigned long long)cap) lists[(size_t)k * (size_t)cap + pos] = key;
        }
      }
  };
  // a wanted record only queues its row index: the value fetch (a random access) happens after
  // the streaming loop, all lanes at once, instead of stalling the warp in the middle of it
  auto hit = [&](int64_t i, uint32_t d) {
    const uint32_t qp = atomicAdd(&S.queue_n, 1u);
    if (qp < (uint32_t)QD_QUEUE) S.queue[qp] = (uint32_t)i;   // n < 2^31 on this path
    else take(i, d);                                            // queue full: handle it in place
  };
  // branch-free membership test in registers: both 16-bit records of a word are tested at once:
  // x = w ^ (bin | bin << 16) has a zero half where a record equals the bin, and
  // (x - 0x00010001) & ~x & 0x80008000 flags zero halves (a borrow can only raise a false flag
  // above a true one; the hit path re-checks exactly)
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
      if ((z & 0x8000u) && d0 != 0xFFFFu) {
        bool w0 = false;
#pragma unroll
        for (int k = 0; k < QD_MAXQ; ++k) w0 |= d0 == want[k];
        if (w0) hit(i, d0);
      }
      if ((z >> 16) && d1 != 0xFFFFu) {
        bool w1 = false;
#pragma unroll
        for (int k = 0; k < QD_MAXQ; ++k) w1 |= d1 == want[k];
        if (w1) hit(i + 1, d1);
      }
    }
  };
  if (VEC) {
    // 32 contiguous bytes (16 records) per thread and iteration; CTA b owns a contiguous range
    const int64_t nunits = n >> 4;
    const int64_t chunk = (nunits + gridDim.x - 1) / gridDim.x;
    const int64_t u_end = min(nunits, ((int64_t)blockIdx.x + 1) * chunk);
    for (int64_t u0 = (int64_t)blockIdx.x * chunk + t; u0 < u_end; u0 += 2 * QD_THREADS) {
      const int64_t u1 = u0 + QD_THREADS;
      const kvm_u64x4 r0 = kvm_ld4(recs + (u0 << 4));
      kvm_u64x4 r1{~0ull, ~0ull, ~0ull, ~0ull};
      if (u1 < u_end) r1 = kvm_ld4(recs + (u1 << 4));
      const int64_t i = u0 << 4, j = u1 << 4;
      word(i, (uint32_t)r0.x); word(i + 2, (uint32_t)(r0.x >> 32));
      word(i + 4, (uint32_t)r0.y); word(i + 6, (uint32_t)(r0.y >> 32));
      word(i + 8, (uint32_t)r0.z); word(i + 10, (uint32_t)(r0.z >> 32));
      word(i + 12, (uint32_t)r0.w); word(i + 14, (uint32_t)(r0.w >> 32));
      word(j, (uint32_t)r1.x); word(j + 2, (uint32_t)(r1.x >> 32));
      word(j + 4, (uint32_t)r1.y); word(j + 6, (uint32_t)(r1.y >> 32));
      word(j + 8, (uint32_t)r1.z); word(j + 10, (uint32_t)(r1.z >> 32));
      word(j + 12, (uint32_t)r1.w); word(j + 14, (uint32_t)(r1.w >> 32));
    }
    if (blockIdx.x == 0 && t < (n & 15)) {
      const int64_t i = (n & ~15ll) + t;
      word(i, (uint32_t)recs[i] | 0xFFFF0000u);   // upper half: "not a member"
    }
  } else {
    for (int64_t i = (int64_t)blockIdx.x * QD_THREADS + t; i < n; i += (int64_t)gridDim.x * QD_THREADS)
      word(i, (uint32_t)recs[i] | 0xFFFF0000u);
  }
  __syncthreads();
  {  // drain the queue: every thread takes queued rows, fetches the value, appends the key
    const uint32_t qn = S.queue_n < (uint32_t)QD_QUEUE ? S.queue_n : (uint32_t)QD_QUEUE;
    for (uint32_t j = t; j < qn; j += QD_THREADS) {
      const int64_t i = (int64_t)S.queue[j];
      take(i, (uint32_t)recs[i]);
    }
  }
  __syncthreads();
  if (t < QD_MAXQ) {
    const uint32_t c = S.stage_n[t] < (uint32_t)QD_STAGE ? S.stage_n[t] : (uint32_t)QD_STAGE;
    S.stage_n[t] = c;
    if (c) S.stage_pos[t] = atomicAdd(&ws->list_n[t], (unsigned long long)c);
  }
  __syncthreads();
  for (int i = t; i < QD_MAXQ * QD_STAGE; i += QD_THREADS) {
    const int k = i / QD_STAGE, j = i % QD_STAGE;
    if ((uint32_t)j < S.stage_n[k]) {
      const unsigned long long pos = S.stage_pos[k] + j;
      if (pos < (unsigned long long)cap) lists[(size_t)k * (size_t)cap + pos] = S.stage[k][j];
    }
  }
  // ---- last CTA to get here finishes the queries, one group of threads per query
  __threadfence();
  __syncthreads();
  if (t == 0) S.is_last = atomicAdd(&ws->c_ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!S.is_last) return;
  __threadfence();
  // (the stages are free now: the finish uses their shared memory)
  static_assert(sizeof(QdFinSmem) <= sizeof(S.stage), "finish state lives in the stage area");
  qd_finish(ws, lists, *reinterpret_cast<QdFinSmem*>(&S.stage[0][0]), results);
  __syncthreads();
  if (t == 0) ws->ready = 0ull;   // a replay of the same launch sequence (same token) waits for its own clearing
}

// two resident CTAs per SM, each with a contiguous share of the rows
int qd_grid(int64_t n, int rows_per_thread_iter) {
  return kvm_grid_for(n, QD_THREADS * rows_per_thread_iter, 2);
}
long long qd_list_cap(int64_t n) {
  const long long c = (long long)(n / 64) + 16384;   // a bin holds ~n/1500 of smooth data
  return (c + 1) & ~1ll;
}

}  // namespace

extern "C" int64_t kvm_column_select_sampled_min_n(void) { return 1 << 18; }

extern "C" size_t kvm_column_select_sampled_workspace_bytes(int64_t n) {
  if (n < 0) n = 0;
  return QD_HEADER + kvm_align_up((size_t)n * sizeof(uint16_t) + 32, 256) +
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
  KVM_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255u) == 0, "kvm_column_select_sampled: workspace must be 256-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  unsigned char* wsb = reinterpret_cast<unsigned char*>(workspace);
  QdWs* ws = reinterpret_cast<QdWs*>(wsb);
  uint16_t* recs = reinterpret_cast<uint16_t*>(wsb + QD_HEADER);
  uint64_t* lists = reinterpret_cast<uint64_t*>(wsb + QD_HEADER + kvm_align_up((size_t)n * sizeof(uint16_t) + 32, 256));
  const long long cap = qd_list_cap(n);
  const int grid = qd_grid(n, 8 * 4);
  const bool vec = (reinterpret_cast<uintptr_t>(values) & 31u) == 0 &&
                   (mode != KVM_MEMBER_STATUS || (reinterpret_cast<uintptr_t>(member) & 31u) == 0);
  // per-call token for the clear-then-flush handshake inside the hist kernel (never 0; the collect
  // kernel resets the word at its end, so a graph replay with the same token works)
  static std::atomic<unsigned long long> call_counter{0x9e3779b97f4a7c15ull};
  const unsigned long long token = call_counter.fetch_add(0x9e3779b97f4a7c15ull) | 1ull;
  if (mode == KVM_MEMBER_STATUS) {
    if (vec) qd_hist_kernel<KVM_MEMBER_STATUS, true><<<grid, QD_THREADS, 0, st>>>(values, n, member, ws, recs, a, cap, token);
    else qd_hist_kernel<KVM_MEMBER_STATUS, false><<<grid, QD_THREADS, 0, st>>>(values, n, member, ws, recs, a, cap, token);
  } else {
    if (vec) qd_hist_kernel<KVM_MEMBER_ALL, true><<<grid, QD_THREADS, 0, st>>>(values, n, member, ws, recs, a, cap, token);
    else qd_hist_kernel<KVM_MEMBER_ALL, false><<<grid, QD_THREADS, 0, st>>>(values, n, member, ws, recs, a, cap, token);
  }
  KVM_CUDA_CHECK(cudaGetLastError());
  // the collect kernel is a programmatic dependent launch: it may become resident while the hist
  // kernel drains and blocks in griddepcontrol.wait until that kernel has completed
  static const bool no_pdl = getenv("KVM_NO_PDL") != nullptr;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)qd_grid(n, 16 * 4));
  cfg.blockDim = dim3(QD_THREADS);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = no_pdl ? 0 : 1;
  const uint16_t* recs_c = recs;
  KVM_CUDA_CHECK(cudaLaunchKernelEx(&cfg, qd_collect_kernel<true>, values, n, recs_c, ws, lists, results));
  return KVM_OK;
}
