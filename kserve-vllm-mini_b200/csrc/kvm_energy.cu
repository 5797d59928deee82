// kvm_energy.cu - K9/K10/K11 front end: power-trace scan, trapezoidal integration,
// double-double prefix of segment energies and the per-interval energy join.
//
// Replaces (reference file:line)
//   energy/collector.py:133-149  trapezoidal_wh: keep non-null samples inside [t0,t1], stable
//                                sort by ts, sum ((p0+p1)/2.0)*((ts1-ts0)/3600.0) where dt>0
//   energy/collector.py:296-347  window from the sample span (--include-warmup), "series" idle
//                                tax (two more windows), "baseline" idle tax (counts outside the
//                                window, mean sample step; the upper median goes through kvm_select)
//   extension A16 (no reference function): E_i = trapezoidal_wh(samples, s_i, e_i) for every
//   request interval, as a difference of prefix sums of the segment energies.
//
// Trace layout: ts fp64[m] ascending (kvm_sort_pairs_f64 first if not), watts fp64[m], NaN = null.
// A "segment" ends at a non-null sample i and starts at the previous non-null sample j; it
// belongs to window [t0,t1] iff t0 <= ts[j] and ts[i] <= t1 (membership clip, no edge
// interpolation).  Nulls are rare, so j is found by stepping back from i-1.
// All sums are double-double with a fixed reduction order (deterministic); HBM-bound:
// 16 B/sample per pass, prefix pass +16 B/sample written.

#include "kvm_common.cuh"

namespace {

constexpr int E_THREADS = 256;
constexpr int E_WARPS = E_THREADS / 32;
constexpr int MAX_WINDOWS = 8;
constexpr int E_MAX_GRID = 148 * 8;

struct Windows {
  double t0[MAX_WINDOWS];
  double t1[MAX_WINDOWS];
  int k;
};

// x / 3600.0 without the 30-instruction fp64 division: q = RN(x * RN(1/3600)) corrected by the
// exactly computed residual (fma), then PROVEN: with r2 = x - 3600*q2 (exact, fma) the candidate
// q2 is the correctly rounded quotient iff |r2| / 3600 < ulp(q2) / 2 (a quotient is never a
// rounding tie); the check uses a slightly smaller bound and anything it does not accept -
// including non-finite, tiny and huge operands - takes the hardware division.  Bit-identical to
// `x / 3600.0` (collector.py:146) by construction.
__device__ __forceinline__ double div3600(double x) {
  const double d = 3600.0, rcp = 1.0 / 3600.0;   // RN(1/3600), folded at compile time
  const double q = x * rcp;
  const double q2 = __fma_rn(__fma_rn(-d, q, x), rcp, q);
  const double r2 = __fma_rn(-d, q2, x);
  const int hi = __double2hiint(q2) & 0x7ff00000;
  // half an ulp of q2 times 3600 = 2^(e-53) * 3600 with e the unbiased exponent of q2
  const double bound = __hiloint2double(hi - (53 << 20), 0) * (d * 0.999);
  const double ax = fabs(x);
  // (a power of two has a finer grid below it: left to the hardware division)
  const bool pow2 = ((__double2hiint(q2) & 0x000fffff) | __double2loint(q2)) == 0;
  if (ax > 1e-250 && ax < 1e250 && fabs(r2) < bound && !pow2) return q2;
  return x / d;
}

// energy of the segment ending at sample i (0 if i is null, has no non-null predecessor, or dt<=0)
__device__ __forceinline__ double segment_wh(const double* __restrict__ ts, const double* __restrict__ watts,
                                             int64_t i, double ti, double wi, double* t_prev) {
  if (kvm_isnan(wi)) return 0.0;
  int64_t j = i - 1;
  double wj = 0.0;
  while (j >= 0) {
    wj = watts[j];
    if (!kvm_isnan(wj)) break;
    --j;
  }
  if (j < 0) {
    *t_prev = __longlong_as_double(0x7ff8000000000000ll);
    return 0.0;
  }
  const double tj = ts[j];
  *t_prev = tj;
  const double dt_h = div3600(ti - tj);          // collector.py:146: (ts1 - ts0) / 3600.0, exactly
  if (!(dt_h > 0)) return 0.0;                   // collector.py:147
  return ((wj + wi) / 2.0) * dt_h;               // collector.py:148
}

// Eight consecutive samples of one thread (two 256-bit loads per column when the tile is full and
// the columns are 32-byte aligned).  The segment ending at sample i starts at the previous
// non-null sample; inside the thread that sample is carried in registers, only the first of the
// eight looks back into memory.  f(r, i, ti, tj, area) is called for every sample (area = 0 for a
// null sample, a sample without a non-null predecessor, or dt <= 0; tj = NaN without predecessor).
constexpr int SEG_ITEMS = 8;
template <class F>
__device__ __forceinline__ void for_eight_samples(const double* __restrict__ ts, const double* __restrict__ watts,
                                                  int64_t m, int64_t base, bool vec, F&& f) {
  if (base >= m) return;
  double t[SEG_ITEMS], w[SEG_ITEMS];
  if (vec && base + SEG_ITEMS <= m) {
    const kvm_u64x4 a0 = kvm_ld4_nc(ts + base), a1 = kvm_ld4_nc(ts + base + 4);
    const kvm_u64x4 b0 = kvm_ld4_nc(watts + base), b1 = kvm_ld4_nc(watts + base + 4);
#define KVM_D(x) __longlong_as_double((long long)(x))
    t[0] = KVM_D(a0.x); t[1] = KVM_D(a0.y); t[2] = KVM_D(a0.z); t[3] = KVM_D(a0.w);
    t[4] = KVM_D(a1.x); t[5] = KVM_D(a1.y); t[6] = KVM_D(a1.z); t[7] = KVM_D(a1.w);
    w[0] = KVM_D(b0.x); w[1] = KVM_D(b0.y); w[2] = KVM_D(b0.z); w[3] = KVM_D(b0.w);
    w[4] = KVM_D(b1.x); w[5] = KVM_D(b1.y); w[6] = KVM_D(b1.z); w[7] = KVM_D(b1.w);
#undef KVM_D
  } else {
#pragma unroll
    for (int r = 0; r < SEG_ITEMS; ++r) {
      const bool in = base + r < m;
      t[r] = in ? ts[base + r] : 0.0;
      w[r] = in ? watts[base + r] : __longlong_as_double(0x7ff8000000000000ll);
    }
  }
  // previous non-null sample before `base`
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  double tp = nan, wp = nan;
  for (int64_t j = base - 1; j >= 0; --j) {
    const double wj = watts[j];
    if (!kvm_isnan(wj)) {
      wp = wj;
      tp = ts[j];
      break;
    }
  }
#pragma unroll
  for (int r = 0; r < SEG_ITEMS; ++r) {
    if (base + r >= m) break;
    double area = 0.0, tj = nan;
    if (!kvm_isnan(w[r])) {
      tj = tp;
      if (!kvm_isnan(wp)) {
        const double dt_h = div3600(t[r] - tp);          // collector.py:146
        if (dt_h > 0) area = ((wp + w[r]) / 2.0) * dt_h;  // collector.py:147-148
      }
      tp = t[r];
      wp = w[r];
    }
    f(r, base + r, t[r], tj, area);
  }
}

// fixed-order block reduction of `cnt` double-double values per thread -> out[cnt] (thread 0 valid)
template <int CNT>
__device__ __forceinline__ void block_reduce_dd(kvm_dd (&v)[CNT], kvm_dd (*red)[CNT], kvm_dd (&out)[CNT]) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int c = 0; c < CNT; ++c) {
    kvm_dd r = kvm_warp_sum(v[c]);
    if (lane == 0) red[warp][c] = r;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int c = 0; c < CNT; ++c) {
      kvm_dd acc = red[0][c];
      for (int w = 1; w < E_WARPS; ++w) acc = kvm_dd_add(acc, red[w][c]);
      out[c] = acc;
    }
  }
}

struct PartialWs {
  unsigned int ticket;
  unsigned int pad[63];
  double partial[1];
};

// ------------------------------------------------------------------ power scan (stats)
enum { PS_VALID = 0, PS_OUT, PS_OUT_VALID, PS_UNSORTED, PS_NSUM };

__global__ void __launch_bounds__(E_THREADS)
power_scan_kernel(const double* __restrict__ ts, const double* __restrict__ watts, int64_t m,
                  double t0, double t1, uint8_t* __restrict__ outside_ids, kvm_power_stats* out,
                  PartialWs* ws) {
  __shared__ int64_t red_i[E_WARPS][PS_NSUM];
  __shared__ double red_mm[E_WARPS][2];
  __shared__ kvm_dd red_dd[E_WARPS][1];
  __shared__ bool is_last;
  int64_t cnt[PS_NSUM] = {0, 0, 0, 0};
  const double inf = __longlong_as_double(0x7ff0000000000000ll);
  double tmin = inf, tmax = -inf;
  kvm_dd step[1] = {{0.0, 0.0}};
  for (int64_t i = (int64_t)blockIdx.x * E_THREADS + threadIdx.x; i < m;
       i += (int64_t)gridDim.x * E_THREADS) {
    const double t = kvm_ld1(ts + i), w = kvm_ld1(watts + i);
    const bool valid = !kvm_isnan(w);
    const bool outside = (t < t0) || (t > t1);   // collector.py:330,345
    cnt[PS_VALID] += valid;
    cnt[PS_OUT] += outside;
    cnt[PS_OUT_VALID] += (outside && valid);
    tmin = fmin(tmin, t);
    tmax = fmax(tmax, t);
    if (i + 1 < m) {
      const double tn = ts[i + 1];
      cnt[PS_UNSORTED] += (tn < t);
      step[0] = kvm_dd_add(step[0], tn - t);     // collector.py:338-340, load order
    }
    if (outside_ids) outside_ids[i] = (outside && valid) ? 0 : 255;
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int c = 0; c < PS_NSUM; ++c) {
    int64_t v = kvm_warp_sum(cnt[c]);
    if (lane == 0) red_i[warp][c] = v;
  }
#pragma unroll
  for (int s = 16; s; s >>= 1) {
    tmin = fmin(tmin, kvm_shfl_xor(tmin, s));
    tmax = fmax(tmax, kvm_shfl_xor(tmax, s));
  }
  if (lane == 0) {
    red_mm[warp][0] = tmin;
    red_mm[warp][1] = tmax;
  }
  kvm_dd tot[1];
  block_reduce_dd<1>(step, red_dd, tot);   // contains a __syncthreads
  constexpr int SLOTS = PS_NSUM + 4;
  double* mine = ws->partial + (size_t)blockIdx.x * SLOTS;
  if (threadIdx.x == 0) {
    for (int c = 0; c < PS_NSUM; ++c) {
      int64_t a = 0;
      for (int w = 0; w < E_WARPS; ++w) a += red_i[w][c];
      mine[c] = __longlong_as_double(a);
    }
    double a = inf, b = -inf;
    for (int w = 0; w < E_WARPS; ++w) {
      a = fmin(a, red_mm[w][0]);
      b = fmax(b, red_mm[w][1]);
    }
    mine[PS_NSUM] = a;
    mine[PS_NSUM + 1] = b;
    mine[PS_NSUM + 2] = tot[0].hi;
    mine[PS_NSUM + 3] = tot[0].lo;
    __threadfence();
    is_last = (atomicAdd(&ws->ticket, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last || threadIdx.x >= 32) return;
  __threadfence();
  // last CTA, warp 0: lanes stride over the CTAs' partials (independent L2 loads), then a
  // fixed-shape shuffle tree - deterministic for a given grid size
  const double* part = ws->partial;
  const int ln = threadIdx.x;
  int64_t sums[PS_NSUM] = {0, 0, 0, 0};
  double a = inf, b = -inf;
  kvm_dd st{0.0, 0.0};
  for (unsigned g = ln; g < gridDim.x; g += 32) {
    const double* p = part + (size_t)g * SLOTS;
    for (int c = 0; c < PS_NSUM; ++c) sums[c] += __double_as_longlong(__ldcg(p + c));
    a = fmin(a, __ldcg(p + PS_NSUM));
    b = fmax(b, __ldcg(p + PS_NSUM + 1));
    st = kvm_dd_add(st, kvm_dd{__ldcg(p + PS_NSUM + 2), __ldcg(p + PS_NSUM + 3)});
  }
  for (int c = 0; c < PS_NSUM; ++c) sums[c] = kvm_warp_sum(sums[c]);
#pragma unroll
  for (int sft = 16; sft; sft >>= 1) {
    a = fmin(a, kvm_shfl_xor(a, sft));
    b = fmax(b, kvm_shfl_xor(b, sft));
  }
  st = kvm_warp_sum(st);
  if (ln != 0) return;
  kvm_power_stats r;
  r.m = m;
  r.n_valid = sums[PS_VALID];
  r.n_outside = sums[PS_OUT];
  r.n_outside_valid = sums[PS_OUT_VALID];
  r.n_unsorted = sums[PS_UNSORTED];
  r.ts_min = a;
  r.ts_max = b;
  r.step_sum_hi = st.hi;
  r.step_sum_lo = st.lo;
  *out = r;
  ws->ticket = 0;
}

// ------------------------------------------------------------------ K9 trapezoid over k windows
template <int K>   // accumulators for K windows live in registers (K = 1, 3 or 8)
__global__ void __launch_bounds__(E_THREADS)
trapezoid_kernel(const double* __restrict__ ts, const double* __restrict__ watts, int64_t m,
                 Windows win, double* __restrict__ out /* [k][2] */, PartialWs* ws) {
  __shared__ kvm_dd red[E_WARPS][K];
  __shared__ bool is_last;
  kvm_dd acc[K];
#pragma unroll
  for (int w = 0; w < K; ++w) acc[w] = {0.0, 0.0};
  const bool vec = ((reinterpret_cast<uintptr_t>(ts) | reinterpret_cast<uintptr_t>(watts)) & 31u) == 0;
  const int64_t ntiles = (m + E_THREADS * SEG_ITEMS - 1) / (E_THREADS * SEG_ITEMS);
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t base = (tile * E_THREADS + threadIdx.x) * SEG_ITEMS;
    for_eight_samples(ts, watts, m, base, vec, [&](int, int64_t, double ti, double tj, double area) {
      if (area != 0.0) {
#pragma unroll
        for (int w = 0; w < K; ++w)
          if (w < win.k && win.t1[w] > win.t0[w] && win.t0[w] <= tj && ti <= win.t1[w]) {
            // sum + accumulated rounding error (two_sum): 7 flops instead of a double-double add
            const double t = acc[w].hi + area;
            const double bb = t - acc[w].hi;
            acc[w].lo += (acc[w].hi - (t - bb)) + (area - bb);
            acc[w].hi = t;
          }
      }
    });
  }
#pragma unroll
  for (int w = 0; w < K; ++w) {   // (sum, error) -> normalised double-double
    const double hi = acc[w].hi + acc[w].lo;
    acc[w].lo = acc[w].lo - (hi - acc[w].hi);
    acc[w].hi = hi;
  }
  kvm_dd tot[K];
  block_reduce_dd<K>(acc, red, tot);
  double* mine = ws->partial + (size_t)blockIdx.x * (2 * MAX_WINDOWS);
  if (threadIdx.x == 0) {
    for (int w = 0; w < K; ++w) {
      mine[2 * w] = tot[w].hi;
      mine[2 * w + 1] = tot[w].lo;
    }
    __threadfence();
    is_last = (atomicAdd(&ws->ticket, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  {
    // warp w reduces window w: lanes stride over the partials, fixed-shape tree at the end
    const double* part = ws->partial;
    const int w = threadIdx.x >> 5, ln = threadIdx.x & 31;
    if (w < win.k) {
      kvm_dd a{0.0, 0.0};
      for (unsigned g = ln; g < gridDim.x; g += 32)
        a = kvm_dd_add(a, kvm_dd{__ldcg(part + (size_t)g * 2 * MAX_WINDOWS + 2 * w),
                                 __ldcg(part + (size_t)g * 2 * MAX_WINDOWS + 2 * w + 1)});
      a = kvm_warp_sum(a);
      if (ln == 0) {
        out[2 * w] = a.hi;
        out[2 * w + 1] = a.lo;
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) ws->ticket = 0;
}

// ------------------------------------------------------------------ K10 prefix (reduce, scan, apply)
constexpr int PFX_ITEMS = 8;
constexpr int PFX_TILE = E_THREADS * PFX_ITEMS;  // 2048 samples per tile

__global__ void __launch_bounds__(E_THREADS)
prefix_tile_sum_kernel(const double* __restrict__ ts, const double* __restrict__ watts, int64_t m,
                       int64_t ntiles, double* __restrict__ tile_sum /* [ntiles][2] */) {
  __shared__ kvm_dd red[E_WARPS][1];
  const bool vec = ((reinterpret_cast<uintptr_t>(ts) | reinterpret_cast<uintptr_t>(watts)) & 31u) == 0;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    // eight consecutive segments per thread are added in plain fp64 (their sum is ~1e-3 Wh: the
    // rounding error, ~1e-18 Wh, is far below the 1e-12 relative bound of any interval energy);
    // everything ACROSS threads and tiles is double-double
    double local = 0.0;
    const int64_t base = tile * PFX_TILE + (int64_t)threadIdx.x * PFX_ITEMS;
    for_eight_samples(ts, watts, m, base, vec, [&](int, int64_t, double, double, double area) { local += area; });
    kvm_dd acc[1] = {{local, 0.0}};
    kvm_dd tot[1];
    block_reduce_dd<1>(acc, red, tot);
    if (threadIdx.x == 0) {
      tile_sum[2 * tile] = tot[0].hi;
      tile_sum[2 * tile + 1] = tot[0].lo;
    }
    __syncthreads();
  }
}

// single CTA: exclusive scan of the tile sums (sequential per thread chunk + carry)
__global__ void __launch_bounds__(E_THREADS)
prefix_tile_scan_kernel(int64_t ntiles, double* __restrict__ tile_sum) {
  __shared__ kvm_dd chunk_tot[E_THREADS];
  const int64_t per = (ntiles + E_THREADS - 1) / E_THREADS;
  const int64_t b = (int64_t)threadIdx.x * per;
  const int64_t e = b + per < ntiles ? b + per : ntiles;
  kvm_dd acc{0.0, 0.0};
  for (int64_t t = b; t < e; ++t) acc = kvm_dd_add(acc, kvm_dd{tile_sum[2 * t], tile_sum[2 * t + 1]});
  chunk_tot[threadIdx.x] = acc;
  __syncthreads();
  kvm_dd run{0.0, 0.0};
  for (int k = 0; k < (int)threadIdx.x; ++k) run = kvm_dd_add(run, chunk_tot[k]);
  for (int64_t t = b; t < e; ++t) {
    const kvm_dd v{tile_sum[2 * t], tile_sum[2 * t + 1]};
    tile_sum[2 * t] = run.hi;
    tile_sum[2 * t + 1] = run.lo;
    run = kvm_dd_add(run, v);
  }
}

__global__ void __launch_bounds__(E_THREADS)
prefix_apply_kernel(const double* __restrict__ ts, const double* __restrict__ watts, int64_t m,
                    int64_t ntiles, const double* __restrict__ tile_excl, double* __restrict__ p_hi,
                    double* __restrict__ p_lo) {
  __shared__ kvm_dd thread_tot[E_THREADS];
  const bool vec = ((reinterpret_cast<uintptr_t>(ts) | reinterpret_cast<uintptr_t>(watts)) & 31u) == 0;
  const bool vec_out = ((reinterpret_cast<uintptr_t>(p_hi) | reinterpret_cast<uintptr_t>(p_lo)) & 31u) == 0;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t base = tile * PFX_TILE + (int64_t)threadIdx.x * PFX_ITEMS;
    double a[PFX_ITEMS];   // inclusive plain sums inside the thread (see prefix_tile_sum_kernel)
    double local = 0.0;
#pragma unroll
    for (int r = 0; r < PFX_ITEMS; ++r) a[r] = 0.0;
    for_eight_samples(ts, watts, m, base, vec, [&](int r, int64_t, double, double, double area) {
      local += area;
      a[r] = local;
    });
    kvm_dd acc{local, 0.0};
    // exclusive prefix over the threads of the tile: inclusive shuffle scan inside each warp,
    // then the warp totals are folded in warp order (fixed shape -> deterministic)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    kvm_dd incl = acc;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const kvm_dd o{__shfl_up_sync(KVM_FULL, incl.hi, d), __shfl_up_sync(KVM_FULL, incl.lo, d)};
      if (lane >= d) incl = kvm_dd_add(o, incl);
    }
    if (lane == 31) thread_tot[warp] = incl;
    __syncthreads();
    kvm_dd run{tile_excl[2 * tile], tile_excl[2 * tile + 1]};
    for (int w = 0; w < warp; ++w) run = kvm_dd_add(run, thread_tot[w]);
    {
      kvm_dd excl{__shfl_up_sync(KVM_FULL, incl.hi, 1), __shfl_up_sync(KVM_FULL, incl.lo, 1)};
      if (lane == 0) excl = kvm_dd{0.0, 0.0};
      run = kvm_dd_add(run, excl);   // + exclusive prefix inside the warp
    }
    double ph[PFX_ITEMS], pl[PFX_ITEMS];
#pragma unroll
    for (int r = 0; r < PFX_ITEMS; ++r) {
      const kvm_dd p = kvm_dd_add(run, a[r]);   // prefix before the thread + plain sum inside it
      ph[r] = p.hi;
      pl[r] = p.lo;
    }
    if (vec_out && base + PFX_ITEMS <= m) {
#define KVM_B(x) (unsigned long long)__double_as_longlong(x)
      asm volatile("st.global.v4.b64 [%0], {%1,%2,%3,%4};" ::"l"(p_hi + base), "l"(KVM_B(ph[0])), "l"(KVM_B(ph[1])), "l"(KVM_B(ph[2])), "l"(KVM_B(ph[3])) : "memory");
      asm volatile("st.global.v4.b64 [%0], {%1,%2,%3,%4};" ::"l"(p_hi + base + 4), "l"(KVM_B(ph[4])), "l"(KVM_B(ph[5])), "l"(KVM_B(ph[6])), "l"(KVM_B(ph[7])) : "memory");
      asm volatile("st.global.v4.b64 [%0], {%1,%2,%3,%4};" ::"l"(p_lo + base), "l"(KVM_B(pl[0])), "l"(KVM_B(pl[1])), "l"(KVM_B(pl[2])), "l"(KVM_B(pl[3])) : "memory");
      asm volatile("st.global.v4.b64 [%0], {%1,%2,%3,%4};" ::"l"(p_lo + base + 4), "l"(KVM_B(pl[4])), "l"(KVM_B(pl[5])), "l"(KVM_B(pl[6])), "l"(KVM_B(pl[7])) : "memory");
#undef KVM_B
    } else {
#pragma unroll
      for (int r = 0; r < PFX_ITEMS; ++r)
        if (base + r < m) {
          p_hi[base + r] = ph[r];
          p_lo[base + r] = pl[r];
        }
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------ K10 interval join
__device__ __forceinline__ int64_t lower_bound_ts(const double* __restrict__ ts, int64_t m, double x) {
  int64_t lo = 0, hi = m;  // first index with ts >= x
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (ts[mid] < x) lo = mid + 1; else hi = mid;
  }
  return lo;
}
__device__ __forceinline__ int64_t upper_bound_ts(const double* __restrict__ ts, int64_t m, double x) {
  int64_t lo = 0, hi = m;  // first index with ts > x
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (ts[mid] <= x) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// ---- search from a guess.  A power trace is sampled (nearly) regularly, so the index of time x
// is close to (x - ts[0]) * (m - 1) / (ts[m-1] - ts[0]); the search gallops outward from that
// guess and bisects the bracket - two or three probes instead of ~27 dependent loads, and (the
// intervals arriving in time order) neighbouring threads probe neighbouring cache lines: the
// join streams through ts and the prefix arrays instead of hopping over them.  Any trace, any
// interval order: the guess only decides how many probes are needed.
__device__ __forceinline__ int64_t lower_bound_from(const double* __restrict__ ts, int64_t m, double x, int64_t g) {
  g = g < 0 ? 0 : (g > m - 1 ? m - 1 : g);
  int64_t lo, hi;   // answer (first index with ts >= x) in [lo, hi]
  if (ts[g] < x) {
    lo = g + 1;
    for (int64_t step = 1;; step <<= 1) {
      hi = lo + step - 1;
      if (hi >= m) { hi = m; break; }
      if (!(ts[hi] < x)) break;
      lo = hi + 1;
    }
  } else {
    hi = g;
    for (int64_t step = 1;; step <<= 1) {
      const int64_t p = hi - step;
      if (p < 0) { lo = 0; break; }
      if (ts[p] < x) { lo = p + 1; break; }
      hi = p;
    }
  }
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (ts[mid] < x) lo = mid + 1; else hi = mid;
  }
  return lo;
}
__device__ __forceinline__ int64_t upper_bound_from(const double* __restrict__ ts, int64_t m, double x, int64_t g) {
  g = g < 0 ? 0 : (g > m - 1 ? m - 1 : g);
  int64_t lo, hi;   // answer (first index with ts > x) in [lo, hi]
  if (ts[g] <= x) {
    lo = g + 1;
    for (int64_t step = 1;; step <<= 1) {
      hi = lo + step - 1;
      if (hi >= m) { hi = m; break; }
      if (!(ts[hi] <= x)) break;
      lo = hi + 1;
    }
  } else {
    hi = g;
    for (int64_t step = 1;; step <<= 1) {
      const int64_t p = hi - step;
      if (p < 0) { lo = 0; break; }
      if (ts[p] <= x) { lo = p + 1; break; }
      hi = p;
    }
  }
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (ts[mid] <= x) lo = mid + 1; else hi = mid;
  }
  return lo;
}

struct TraceTable {
  kvm_power_trace tr[KVM_MAX_TRACES];
  int32_t n;
};

// Every interval against every trace in ONE pass: the interval columns are read once, the
// per-trace energies are added in trace order (PromQL `sum`, collector.py:45) in a register and
// written once.  accumulate != 0 adds to what energy[] holds.
__global__ void __launch_bounds__(E_THREADS)
interval_energy_kernel(TraceTable tt, const double* __restrict__ start_ms, const double* __restrict__ latency_ms,
                       int64_t n, double* __restrict__ energy, int accumulate) {
  for (int64_t r = (int64_t)blockIdx.x * E_THREADS + threadIdx.x; r < n;
       r += (int64_t)gridDim.x * E_THREADS) {
    const double st = kvm_ld1(start_ms + r), la = kvm_ld1(latency_ms + r);
    const double s = st / 1000.0, e = (st + la) / 1000.0;
    double total = accumulate ? energy[r] : 0.0;
    bool first_term = !accumulate;
    for (int k = 0; k < tt.n; ++k) {
      const kvm_power_trace T = tt.tr[k];
      const int64_t m = T.m;
      double wh = 0.0;
      if (e > s && m > 1) {
        const double t0 = T.ts[0], tn = T.ts[m - 1];
        const double rate = tn > t0 ? (double)(m - 1) / (tn - t0) : 0.0;
        const double gs = (s - t0) * rate;
        const int64_t g0 = gs > 0.0 ? (gs < (double)m ? (int64_t)gs : m - 1) : 0;
        int64_t first = lower_bound_from(T.ts, m, s, g0);
        const double span = (e - s) * rate;
        const int64_t g1 = first + (span < (double)m ? (int64_t)span : m);
        const int64_t last = upper_bound_from(T.ts, m, e, g1) - 1;
        while (first <= last && kvm_isnan(T.watts[first])) ++first;  // first kept sample of the interval
        if (first < last) {
          const kvm_dd d = kvm_dd_sub(kvm_dd{T.p_hi[last], T.p_lo[last]}, kvm_dd{T.p_hi[first], T.p_lo[first]});
          wh = d.hi + d.lo;
        }
      }
      total = first_term ? wh : total + wh;
      first_term = false;
    }
    energy[r] = total;
  }
}

int grid_for(int64_t items, int per_cta) {
  int g = kvm_grid_for(items, per_cta, 8);
  return g > E_MAX_GRID ? E_MAX_GRID : g;
}

}  // namespace

namespace {
__global__ void div3600_check_kernel(const double* __restrict__ x, double* __restrict__ out, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = div3600(x[i]);
}
}  // namespace

extern "C" int kvm_selfcheck_div3600(const double* x, double* out, int64_t n, void* stream) {
  KVM_REQUIRE(n >= 0 && (n == 0 || (x && out)), "kvm_selfcheck_div3600: bad argument");
  if (n == 0) return KVM_OK;
  div3600_check_kernel<<<grid_for(n, 1024), 256, 0, (cudaStream_t)stream>>>(x, out, n);
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" size_t kvm_energy_workspace_bytes(int64_t m) {
  const int64_t ntiles = (m + PFX_TILE - 1) / PFX_TILE + 1;
  size_t partial = 256 + sizeof(double) * (size_t)E_MAX_GRID * 2 * MAX_WINDOWS;
  size_t tiles = sizeof(double) * 2 * (size_t)ntiles;
  return kvm_align_up(partial, 256) + kvm_align_up(tiles, 256);
}

extern "C" int kvm_power_scan(const double* ts, const double* watts, int64_t m, double t0, double t1,
                              uint8_t* outside_ids, kvm_power_stats* stats, void* workspace,
                              size_t workspace_bytes, void* stream) {
  KVM_REQUIRE(stats && workspace && m >= 0 && ((ts && watts) || m == 0), "kvm_power_scan: bad argument");
  if (workspace_bytes < kvm_energy_workspace_bytes(m)) {
    kvm_set_error("kvm_power_scan: workspace %zu < %zu bytes", workspace_bytes, kvm_energy_workspace_bytes(m));
    return KVM_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  KVM_CUDA_CHECK(cudaMemsetAsync(workspace, 0, 256, st));
  power_scan_kernel<<<grid_for(m, E_THREADS * 8), E_THREADS, 0, st>>>(
      ts, watts, m, t0, t1, outside_ids, stats, reinterpret_cast<PartialWs*>(workspace));
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_trapezoid_wh(const double* ts, const double* watts, int64_t m,
                                const double* windows_host, int32_t k, double* out_dd,
                                void* workspace, size_t workspace_bytes, void* stream) {
  KVM_REQUIRE(out_dd && workspace && windows_host && m >= 0 && ((ts && watts) || m == 0),
              "kvm_trapezoid_wh: bad argument");
  KVM_REQUIRE(k >= 1 && k <= MAX_WINDOWS, "kvm_trapezoid_wh: k=%d outside [1,%d]", k, MAX_WINDOWS);
  if (workspace_bytes < kvm_energy_workspace_bytes(m)) {
    kvm_set_error("kvm_trapezoid_wh: workspace %zu < %zu bytes", workspace_bytes, kvm_energy_workspace_bytes(m));
    return KVM_ERR_WORKSPACE;
  }
  Windows win;
  for (int w = 0; w < MAX_WINDOWS; ++w) {
    win.t0[w] = w < k ? windows_host[2 * w] : 0.0;
    win.t1[w] = w < k ? windows_host[2 * w + 1] : 0.0;
  }
  win.k = k;
  cudaStream_t st = (cudaStream_t)stream;
  KVM_CUDA_CHECK(cudaMemsetAsync(workspace, 0, 256, st));
  const int grid = grid_for(m, E_THREADS * 8);
  PartialWs* pws = reinterpret_cast<PartialWs*>(workspace);
  if (k == 1) trapezoid_kernel<1><<<grid, E_THREADS, 0, st>>>(ts, watts, m, win, out_dd, pws);
  else if (k <= 3) trapezoid_kernel<3><<<grid, E_THREADS, 0, st>>>(ts, watts, m, win, out_dd, pws);
  else trapezoid_kernel<8><<<grid, E_THREADS, 0, st>>>(ts, watts, m, win, out_dd, pws);
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_energy_prefix(const double* ts, const double* watts, int64_t m, double* p_hi,
                                 double* p_lo, void* workspace, size_t workspace_bytes, void* stream) {
  KVM_REQUIRE(workspace && m >= 0 && ((ts && watts && p_hi && p_lo) || m == 0), "kvm_energy_prefix: bad argument");
  if (m == 0) return KVM_OK;
  if (workspace_bytes < kvm_energy_workspace_bytes(m)) {
    kvm_set_error("kvm_energy_prefix: workspace %zu < %zu bytes", workspace_bytes, kvm_energy_workspace_bytes(m));
    return KVM_ERR_WORKSPACE;
  }
  const int64_t ntiles = (m + PFX_TILE - 1) / PFX_TILE;
  const size_t partial = kvm_align_up(256 + sizeof(double) * (size_t)E_MAX_GRID * 2 * MAX_WINDOWS, 256);
  double* tile_sum = reinterpret_cast<double*>(reinterpret_cast<unsigned char*>(workspace) + partial);
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = grid_for(m, PFX_TILE);
  prefix_tile_sum_kernel<<<grid, E_THREADS, 0, st>>>(ts, watts, m, ntiles, tile_sum);
  prefix_tile_scan_kernel<<<1, E_THREADS, 0, st>>>(ntiles, tile_sum);
  prefix_apply_kernel<<<grid, E_THREADS, 0, st>>>(ts, watts, m, ntiles, tile_sum, p_hi, p_lo);
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

namespace {
// out[i] = ((s0[i] + s1[i]) + s2[i]) + ...: the contributions of the ranks added in rank order
__global__ void __launch_bounds__(256)
ordered_sum_kernel(const double* __restrict__ slices, int world, int64_t n, int64_t stride, double* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) {
    double acc = kvm_ld1(slices + i);
    for (int r = 1; r < world; ++r) acc = acc + kvm_ld1(slices + (size_t)r * stride + i);
    out[i] = acc;
  }
}
}  // namespace

extern "C" int kvm_ordered_sum_f64(const double* slices, int32_t world, int64_t n, int64_t stride, double* out,
                                   void* stream) {
  KVM_REQUIRE(world >= 1 && n >= 0 && stride >= n, "kvm_ordered_sum_f64: bad sizes");
  if (n == 0) return KVM_OK;
  KVM_REQUIRE(slices && out, "kvm_ordered_sum_f64: null argument");
  ordered_sum_kernel<<<grid_for(n, 256 * 8), 256, 0, (cudaStream_t)stream>>>(slices, world, n, stride, out);
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_interval_energy_multi(const kvm_power_trace* traces, int32_t ntraces, const double* start_ms,
                                         const double* latency_ms, int64_t n, double* energy_wh,
                                         int32_t accumulate, void* stream) {
  KVM_REQUIRE(traces && ntraces >= 1 && ntraces <= KVM_MAX_TRACES, "kvm_interval_energy_multi: 1..%d traces", KVM_MAX_TRACES);
  KVM_REQUIRE(n >= 0, "kvm_interval_energy_multi: negative size");
  if (n == 0) return KVM_OK;
  KVM_REQUIRE(start_ms && latency_ms && energy_wh, "kvm_interval_energy_multi: null argument");
  TraceTable tt{};
  tt.n = ntraces;
  for (int k = 0; k < ntraces; ++k) {
    KVM_REQUIRE(traces[k].m >= 0 && (traces[k].m == 0 || (traces[k].ts && traces[k].watts && traces[k].p_hi && traces[k].p_lo)),
                "kvm_interval_energy_multi: trace %d has null columns", k);
    tt.tr[k] = traces[k];
  }
  interval_energy_kernel<<<grid_for(n, E_THREADS * 4), E_THREADS, 0, (cudaStream_t)stream>>>(
      tt, start_ms, latency_ms, n, energy_wh, accumulate);
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_interval_energy(const double* ts, const double* watts, const double* p_hi,
                                   const double* p_lo, int64_t m, const double* start_ms,
                                   const double* latency_ms, int64_t n, double* energy_wh,
                                   int32_t accumulate, void* stream) {
  const kvm_power_trace one{ts, watts, p_hi, p_lo, m};
  return kvm_interval_energy_multi(&one, 1, start_ms, latency_ms, n, energy_wh, accumulate, stream);
}
