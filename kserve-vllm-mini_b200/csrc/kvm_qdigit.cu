// kvm_qdigit.cu - one-and-an-eighth-pass exact quantiles of one multiset (the 16 B/record
// percentile path; analyze.py:59-81).
//
// Three launches, no memset; the second and third are programmatic dependent launches (their
// CTAs become resident while the kernel in front drains):
//   hist     prologue: every CTA draws the SAME 512 stratified pseudo-random members (one load per
//            thread, L2 hits for all but the first CTA) and derives an ESTIMATE of the key range
//            from them - no separate sampling launch, no grid-wide dependency; CTA 0 also clears
//            the accumulators and publishes a per-call token that the flushes wait for.
//            body: ONE pass over the data (16 B/record, 256-bit loads): 13-bit digit
//            of every member key relative to the estimated range (bins 0 and 8191 catch what lies
//            far outside it), histogram in shared memory, and the digit written as a 2-byte record
//            (0xFFFF = not a member).
//            epilogue (last CTA to finish, ticket): exact member count = sum of the histogram,
//            rank of every query, the bin holding it and the residual rank.
//   collect  pass over the 2 B/record digit records; the value of a record is fetched only when
//            its digit is a wanted bin (~0.1 % of the records): its key goes to that bin's list and
//            its next 11 bits into that bin's sub-histogram.
//   rank     one CTA per query: the sub-histogram gives the sub-bin holding the rank, the handful
//            of keys of that sub-bin is ranked by counting.
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
  unsigned long long dbg[16];   // phase time stamps of the last call (globaltimer ns / SM clocks), for profiling
  unsigned long long ready;   // == the call's token once CTA 0 of the hist kernel has cleared the block below
  // ---- cleared by CTA 0 of the hist kernel at the start of every call (QD_CLEAR_* below)
  alignas(16) unsigned int s_ticket;
  unsigned int pad4[3];
  unsigned long long list_n[QD_MAXQ];
  unsigned int hist[QD_BINS];   // n < 2^31 on this path
  // ----
  alignas(16) unsigned int hist2[QD_MAXQ][QD_SUB];   // rows of the wanted bins are cleared by the resolve
  alignas(16) uint64_t small[QD_MAXQ][QD_SMALL];
};

constexpr size_t QD_HEADER = (sizeof(QdWs) + 255) / 256 * 256;
constexpr size_t QD_CLEAR_BEGIN = offsetof(QdWs, s_ticket), QD_CLEAR_END = offsetof(QdWs, hist2);
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

__device__ __forceinline__ unsigned long long qd_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
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
  __shared__ unsigned long long wtot[QD_THREADS / 32 + 1];
  __shared__ QdQuery qs[QD_MAXQ];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  // warp w owns bins [w*SEG, (w+1)*SEG): lane-contiguous (coalesced) loads, all in flight at once,
  // then one warp scan per 32 bins; `cum` gets the exclusive prefix inside the warp's segment and
  // wtot the segment totals (a bin's global prefix = cum[bin] + sum of the segments before it)
  constexpr int NW = QD_THREADS / 32, SEG = QD_BINS / NW, IT = SEG / 32;
  uint32_t c[IT];
#pragma unroll
  for (int k = 0; k < IT; ++k) c[k] = __ldcg(&ws->hist[warp * SEG + k * 32 + lane]);
  uint32_t carry = 0;
#pragma unroll
  for (int k = 0; k < IT; ++k) {
    const uint32_t ck = c[k];
    uint32_t incl = ck;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(KVM_FULL, incl, d);
      if (lane >= d) incl += o;
    }
    cum[warp * SEG + k * 32 + lane] = carry + incl - ck;
    carry += __shfl_sync(KVM_FULL, incl, 31);
  }
  if (lane == 0) wtot[warp] = carry;
  __syncthreads();
  if (t == 0) {   // segment totals -> exclusive offsets, grand total in wtot[NW]
    unsigned long long acc = 0;
    for (int w = 0; w < NW; ++w) {
      const unsigned long long x = wtot[w];
      wtot[w] = acc;
      acc += x;
    }
    wtot[NW] = acc;
  }
  __syncthreads();
  const unsigned long long total = wtot[NW];
  auto prefix = [&](int bin) -> long long { return (long long)cum[bin] + (long long)wtot[bin / SEG]; };
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
            if (prefix(mid) <= s.rank) lo = mid;
            else hi = mid - 1;
          }
          s.bin = lo;
          s.resid = s.rank - prefix(lo);
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
    qs[0].pad = nslots;
  }
  __syncthreads();
  if (t < QD_MAXQ) {
    QdQuery o = qs[t];
    o.pad = 0;
    ws->q[t] = o;
  }
  {   // the sub-histograms the collect kernel counts into
    const int nslots = qs[0].pad;
    uint4* z = reinterpret_cast<uint4*>(&ws->hist2[0][0]);
    for (int i = t; i < nslots * (QD_SUB / 4); i += QD_THREADS) z[i] = make_uint4(0u, 0u, 0u, 0u);
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
  if (blockIdx.x == 0 && t == 0) ws->dbg[0] = qd_now();

  // ---- sample loads first (they are the longest latency of the prologue: cold TLB): thread t
  // draws one element - the same element in every CTA, so all CTAs derive identical digit
  // parameters without talking to each other.  16 regions spread evenly over the column (the
  // first at its start, the last at its end: a sorted column shows its extremes), 32
  // pseudo-random elements inside a 4096-element window of each: 512 values from ~32 pages.
  const int64_t* st = reinterpret_cast<const int64_t*>(member);
  double sample_v;
  long long sample_s = 200;
  {
    const int64_t region = t >> 5;
    const int64_t win = 4096;
    const int64_t off = region * ((n - win) / (QD_SAMPLE / 32 - 1));
    const int64_t i = off + (int64_t)(qd_hash((uint32_t)t * 2654435761u + 12345u) & (uint32_t)(win - 1));
    sample_v = __ldg(values + i);
    if (MODE == KVM_MEMBER_STATUS) sample_s = __ldg(st + i);
  }
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
  const int64_t o_end = nocts, o_step = (int64_t)gridDim.x * QD_THREADS;   // grid-stride (measured faster than per-CTA ranges)
  int64_t o = (int64_t)blockIdx.x * QD_THREADS + t;
  kvm_u64x4 v0{}, v1{}, s0{}, s1{};
  auto load8 = [&](int64_t oc, kvm_u64x4& a0, kvm_u64x4& a1, kvm_u64x4& b0, kvm_u64x4& b1) {
    const int64_t i = oc << 3;
    a0 = kvm_ld4_nc(values + i);
    a1 = kvm_ld4_nc(values + i + 4);
    if (MODE == KVM_MEMBER_STATUS) {
      b0 = kvm_ld4_nc(st + i);
      b1 = kvm_ld4_nc(st + i + 4);
    }
  };
  bool have = VEC && o < o_end;
  if (have) load8(o, v0, v1, s0, s1);

  // ---- reduce the sample
  {
    const double v = sample_v;
    const bool mem = v == v && sample_s == 200;
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
  if (blockIdx.x == 0 && t == 0) ws->dbg[1] = qd_now();
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
    while (have) {
      uint4 out;
      out.x = dig(v0.x, s0.x) | (dig(v0.y, s0.y) << 16);
      out.y = dig(v0.z, s0.z) | (dig(v0.w, s0.w) << 16);
      out.z = dig(v1.x, s1.x) | (dig(v1.y, s1.y) << 16);
      out.w = dig(v1.z, s1.z) | (dig(v1.w, s1.w) << 16);
      *reinterpret_cast<uint4*>(recs + (o << 3)) = out;
      // (no register double-buffering: 128 B in flight per thread x 1024 threads per SM is
      // several times what the memory system needs; the other warps cover this one's wait)
      o += o_step;
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
    if (c) atomicAdd(&ws->hist[i], c);
  }
  __threadfence();
  __syncthreads();
  if (t == 0) is_last = atomicAdd(&ws->s_ticket, 1u) == gridDim.x - 1;
  __syncthreads();
  if (blockIdx.x == 0 && t == 0) ws->dbg[2] = qd_now();
  if (!is_last) return;
  __threadfence();
  if (t == 0) ws->dbg[3] = qd_now();
  qd_resolve(ws, hist, a);
  if (t == 0) ws->dbg[4] = qd_now();
}

// ------------------------------------------------------------------ collect (+ rank in the last CTA)
struct QdCollectSmem {
  uint64_t stage[QD_MAXQ][QD_STAGE];
  uint32_t queue[QD_QUEUE];
  uint32_t stage_n[QD_MAXQ];
  unsigned long long stage_pos[QD_MAXQ];
  uint32_t queue_n;
  int is_last;
  // parameters of the rare paths (kept out of line so that the streaming loop stays small)
  uint32_t want[QD_MAXQ];
  uint64_t base;
  long long cap;
  int32_t bshift, sh2;
};

// key of a wanted record -> sub-histogram + staged list entry
__device__ __noinline__ void qd_take(QdCollectSmem& S, QdWs* ws, const double* __restrict__ values,
                                     uint64_t* __restrict__ lists, int64_t i, uint32_t d) {
  const uint64_t key = kvm_key(values[i]);
  for (int k = 0; k < QD_MAXQ; ++k)
    if (d == S.want[k]) {
      const uint64_t bin_base = S.base + ((uint64_t)(d - 1) << S.bshift);
      atomicAdd(&ws->hist2[k][(uint32_t)((key - bin_base) >> S.sh2) & (QD_SUB - 1)], 1u);
      const uint32_t sp = atomicAdd(&S.stage_n[k], 1u);
      if (sp < (uint32_t)QD_STAGE) {
        S.stage[k][sp] = key;
      } else {  // the stage is full: straight to the list
        const unsigned long long pos = atomicAdd(&ws->list_n[k], 1ull);
        if (pos < (unsigned long long)S.cap) lists[(size_t)k * (size_t)S.cap + pos] = key;
      }
    }
}

// a word of two records raised the (conservative) flag: re-check exactly; a wanted record only
// queues its row index - the value fetch (a random access) happens after the streaming loop, all
// lanes at once, instead of stalling the warp in the middle of it
__device__ __noinline__ void qd_hit_word(QdCollectSmem& S, QdWs* ws, const double* __restrict__ values,
                                         uint64_t* __restrict__ lists, int64_t i, uint32_t w) {
  for (int h = 0; h < 2; ++h) {
    const uint32_t d = h ? (w >> 16) : (w & 0xFFFFu);
    if (d == 0xFFFFu) continue;
    bool wanted = false;
    for (int k = 0; k < QD_MAXQ; ++k) wanted |= d == S.want[k];
    if (!wanted) continue;
    const uint32_t qp = atomicAdd(&S.queue_n, 1u);
    if (qp < (uint32_t)QD_QUEUE) S.queue[qp] = (uint32_t)(i + h);   // n < 2^31 on this path
    else qd_take(S, ws, values, lists, i + h, d);                    // queue full: handle it in place
  }
}

// Streams the 2 B/record digit records; a record whose digit is a wanted bin has its value
// fetched, its key appended to that bin's list (staged per CTA, one cursor bump per flush) and
// the key's next 11 bits counted into the bin's sub-histogram.
template <bool VEC>
__global__ void __launch_bounds__(QD_THREADS, 2)
qd_collect_kernel(const double* __restrict__ values, int64_t n, const uint16_t* __restrict__ recs, QdWs* ws,
                  uint64_t* __restrict__ lists) {
  __shared__ QdCollectSmem S;
  const int t = threadIdx.x;
  if (t < QD_MAXQ) S.stage_n[t] = 0;
  if (t == 0) S.queue_n = 0;
  kvm_pdl_launch_dependents();   // the rank kernel may become resident as this grid drains
  kvm_pdl_wait();   // everything above overlapped the hist kernel's tail; its results are visible from here on
  if (blockIdx.x == 0 && t == 0) ws->dbg[5] = qd_now();
  if (ws->status != KVM_OK) return;   // declined by the sample or by the resolve: the rank kernel reports
  const int nslots = ws->nslots;
  const long long cap = ws->list_cap;
  const int bshift = ws->shift;
  const uint64_t base = ws->base;
  const int sh2 = bshift > QD_SUB_BITS ? bshift - QD_SUB_BITS : 0;
  uint32_t want[QD_MAXQ];
#pragma unroll
  for (int k = 0; k < QD_MAXQ; ++k) want[k] = k < nslots ? (uint32_t)ws->slot_bin[k] : 0xFFFFFFFFu;
  if (t < QD_MAXQ) S.want[t] = t < nslots ? (uint32_t)ws->slot_bin[t] : 0xFFFFFFFFu;
  if (t == 0) {
    S.base = base;
    S.cap = cap;
    S.bshift = bshift;
    S.sh2 = sh2;
  }
  __syncthreads();
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
    if (z) qd_hit_word(S, ws, values, lists, i, w);
  };
  if (VEC) {
    // 32 contiguous bytes (16 records) per thread and iteration; CTA b owns a contiguous range
    const int64_t u_end = n >> 4, u_step = (int64_t)gridDim.x * QD_THREADS;
    for (int64_t u0 = (int64_t)blockIdx.x * QD_THREADS + t; u0 < u_end; u0 += 2 * u_step) {
      const int64_t u1 = u0 + u_step;
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
  if (blockIdx.x == 0 && t == 0) ws->dbg[6] = qd_now();
  {  // drain the queue: every thread takes queued rows, fetches the value, appends the key
    const uint32_t qn = S.queue_n < (uint32_t)QD_QUEUE ? S.queue_n : (uint32_t)QD_QUEUE;
    for (uint32_t j = t; j < qn; j += QD_THREADS) {
      const int64_t i = (int64_t)S.queue[j];
      qd_take(S, ws, values, lists, i, (uint32_t)recs[i]);
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
  if (blockIdx.x == 0 && t == 0) ws->dbg[7] = qd_now();
}

// ------------------------------------------------------------------ rank
// One CTA per query (a programmatic dependent launch behind the collect kernel).  The wanted
// bin's sub-histogram, its list length and the first 8 keys of every thread are requested
// together (one round trip); the sub-histogram gives the sub-bin holding the residual rank, the
// keys of that sub-bin - usually a handful, or thousands of copies of ONE key (values rounded to
// three decimals), which min == max settles without ranking anything - are gathered in shared
// memory and ranked by counting.
constexpr int QR_THREADS = 512;
constexpr int QR_U = 8;

__global__ void __launch_bounds__(QR_THREADS, 1)
qd_rank_kernel(QdWs* ws, const uint64_t* __restrict__ lists, kvm_query_result* __restrict__ results) {
  __shared__ uint32_t h2[QD_SUB];
  __shared__ uint32_t wsum[QR_THREADS / 32];
  __shared__ uint64_t small[QD_SMALL];
  __shared__ unsigned long long kmin_s, kmax_s;
  __shared__ uint32_t sb_s, sr_s, scount_s, cnt_s;
  __shared__ int found_s;
  const int q = blockIdx.x, t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  if (t == 0) {
    kmin_s = QD_NONE;
    kmax_s = 0ull;
    cnt_s = 0;
    found_s = 0;
  }
  kvm_pdl_wait();   // the collect kernel has completed; its lists and sub-histograms are visible
  if (q == 0 && t == 0) ws->dbg[8] = qd_now();
  const QdQuery s = ws->q[q];
  const int status = ws->status;
  const long long cap = ws->list_cap;
  const int bshift = ws->shift;
  const uint64_t base = ws->base;
  // last kernel of the call: a replay of the same launch sequence (same token) must wait for its
  // own clearing again (the hist kernel of the next call starts after this grid has completed)
  if (q == 0 && t == 0) ws->ready = 0ull;
  const bool need = status == KVM_OK && s.valid && s.qstatus == KVM_Q_OK;
  if (!need) {   // empty multiset / rank out of range: nothing to look up; a declined call reports NaN
    if (t == 0) results[q] = kvm_query_result{nan, s.n, s.rank, s.qstatus, 0};
    return;
  }
  const int sh2 = bshift > QD_SUB_BITS ? bshift - QD_SUB_BITS : 0;
  const uint64_t bin_base = base + ((uint64_t)(s.bin - 1) << bshift);
  const uint64_t* __restrict__ keys = lists + (size_t)s.slot * (size_t)cap;
  // ---- one round trip: list length, sub-histogram (coalesced), first keys of every thread
  const unsigned long long m = __ldcg(&ws->list_n[s.slot]);
  const uint4 hv = __ldcg(reinterpret_cast<const uint4*>(ws->hist2[s.slot]) + t);   // bins 4t .. 4t+3
  uint64_t kk[QR_U];
#pragma unroll
  for (int u = 0; u < QR_U; ++u) kk[u] = __ldcg(keys + (size_t)u * QR_THREADS + t);   // inside the list area (cap >= 16384)
  auto decline = [&]() {
    if (t == 0) {
      ws->status = KVM_ERR_UNSUPPORTED;  // all writers agree
      results[q] = kvm_query_result{nan, s.n, s.rank, s.qstatus, 0};
    }
  };
  if (m > (unsigned long long)cap || (long long)m <= s.resid) {
    decline();
    return;
  }
  // ---- sub-bin holding the residual rank
  const uint32_t lsum = hv.x + hv.y + hv.z + hv.w;
  uint32_t incl = lsum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t o = __shfl_up_sync(KVM_FULL, incl, d);
    if (lane >= d) incl += o;
  }
  if (lane == 31) wsum[warp] = incl;
  __syncthreads();
  uint32_t before = incl - lsum;
  for (int w = 0; w < warp; ++w) before += wsum[w];
  const uint32_t r = (uint32_t)s.resid;
  if (before <= r && r < before + lsum) {   // exactly one thread
    uint32_t b = before;
    int k = 0;
    const uint32_t cc[4] = {hv.x, hv.y, hv.z, hv.w};
    for (; k < 3; ++k) {
      if (r < b + cc[k]) break;
      b += cc[k];
    }
    sb_s = (uint32_t)(4 * t + k);
    sr_s = r - b;
    scount_s = cc[k];
    found_s = 1;
  }
  __syncthreads();
  if (!found_s) {   // sub-histogram inconsistent with the list length
    decline();
    return;
  }
  const uint32_t sb = sb_s, sr = sr_s, scount = scount_s;
  if (sh2 == 0) {   // the sub-bin is one key
    if (t == 0) results[q] = kvm_query_result{kvm_unkey(bin_base + sb), s.n, s.rank, KVM_Q_OK, 0};
    return;
  }
  // ---- keys of the sub-bin
  const bool gather = scount <= (uint32_t)QD_SMALL;
  uint64_t mn = QD_NONE, mx = 0;
  auto look = [&](uint64_t key) {
    if ((uint32_t)((key - bin_base) >> sh2) == sb) {
      mn = key < mn ? key : mn;
      mx = key > mx ? key : mx;
      if (gather) {
        const uint32_t pos = atomicAdd(&cnt_s, 1u);
        if (pos < (uint32_t)QD_SMALL) small[pos] = key;
      }
    }
  };
#pragma unroll
  for (int u = 0; u < QR_U; ++u)
    if ((unsigned long long)u * QR_THREADS + t < m) look(kk[u]);
  for (unsigned long long i0 = (unsigned long long)QR_U * QR_THREADS + t; i0 < m; i0 += (unsigned long long)QR_U * QR_THREADS) {
    uint64_t k2[QR_U];
#pragma unroll
    for (int u = 0; u < QR_U; ++u) k2[u] = i0 + (unsigned long long)u * QR_THREADS < m ? __ldcg(keys + i0 + (size_t)u * QR_THREADS) : QD_NONE;
#pragma unroll
    for (int u = 0; u < QR_U; ++u)
      if (k2[u] != QD_NONE) look(k2[u]);
  }
  mn = kvm_warp_min(mn);
  mx = kvm_warp_max(mx);
  if (lane == 0 && mn != QD_NONE) {
    atomicMin(&kmin_s, (unsigned long long)mn);
    atomicMax(&kmax_s, (unsigned long long)mx);
  }
  __syncthreads();
  if (kmin_s == kmax_s) {   // one distinct key (any number of copies)
    if (t == 0) results[q] = kvm_query_result{kvm_unkey(kmin_s), s.n, s.rank, KVM_Q_OK, 0};
    return;
  }
  const uint32_t m2 = cnt_s;
  if (!gather || m2 != scount) {   // too many distinct candidates for the scratch: the general select's job
    decline();
    return;
  }
  for (uint32_t i = t; i < m2; i += QR_THREADS) {
    const uint64_t k = small[i];
    uint32_t less = 0, leq = 0;
    for (uint32_t j = 0; j < m2; ++j) {
      const uint64_t x = small[j];
      less += x < k;
      leq += x <= k;
    }
    if (less <= sr && sr < leq) results[q] = kvm_query_result{kvm_unkey(k), s.n, s.rank, KVM_Q_OK, 0};  // all writers agree
  }
}

// two resident CTAs per SM, each with a contiguous share of the rows
int qd_grid(int64_t n, int rows_per_thread_iter) {
  // (a machine-filling grid leaves kvm_tail_slots() CTAs out: room for the rank kernel of another call in flight)
  return kvm_stream_grid(n, QD_THREADS * rows_per_thread_iter);
}
long long qd_list_cap(int64_t n) {
  const long long c = (long long)(n / 64) + 16384;   // a bin holds ~n/1500 of smooth data
  return (c + 1) & ~1ll;
}

}  // namespace

extern "C" int64_t kvm_column_select_sampled_min_n(void) { return 1 << 18; }
// byte offset of the 16 phase time stamps (ns, %globaltimer) of the last call inside the workspace
extern "C" int64_t kvm_column_select_sampled_dbg_offset(void) { return (int64_t)offsetof(QdWs, dbg); }

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
  KVM_CUDA_CHECK(cudaLaunchKernelEx(&cfg, qd_collect_kernel<true>, values, n, recs_c, ws, lists));
  cfg.gridDim = dim3((unsigned)nq);
  cfg.blockDim = dim3(QR_THREADS);
  const uint64_t* lists_c = lists;
  KVM_CUDA_CHECK(cudaLaunchKernelEx(&cfg, qd_rank_kernel, ws, lists_c, results));
  return KVM_OK;
}
