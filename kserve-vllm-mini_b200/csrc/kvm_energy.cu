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
  const double dt_h = (ti - tj) / 3600.0;        // collector.py:146
  if (!(dt_h > 0)) return 0.0;                   // collector.py:147
  return ((wj + wi) / 2.0) * dt_h;               // collector.py:148
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
  for (int64_t i = (int64_t)blockIdx.x * E_THREADS + threadIdx.x; i < m;
       i += (int64_t)gridDim.x * E_THREADS) {
    const double ti = kvm_ld1(ts + i), wi = kvm_ld1(watts + i);
    double tj;
    const double area = segment_wh(ts, watts, i, ti, wi, &tj);
    if (area != 0.0) {
#pragma unroll
      for (int w = 0; w < K; ++w)
        if (w < win.k && win.t1[w] > win.t0[w] && win.t0[w] <= tj && ti <= win.t1[w])
          acc[w] = kvm_dd_add(acc[w], area);
    }
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
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    kvm_dd acc[1] = {{0.0, 0.0}};
    const int64_t base = tile * PFX_TILE + (int64_t)threadIdx.x * PFX_ITEMS;
#pragma unroll
    for (int r = 0; r < PFX_ITEMS; ++r) {
      const int64_t i = base + r;
      if (i < m) {
        double tj;
        acc[0] = kvm_dd_add(acc[0], segment_wh(ts, watts, i, ts[i], watts[i], &tj));
      }
    }
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
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t base = tile * PFX_TILE + (int64_t)threadIdx.x * PFX_ITEMS;
    double a[PFX_ITEMS];
    kvm_dd acc{0.0, 0.0};
#pragma unroll
    for (int r = 0; r < PFX_ITEMS; ++r) {
      const int64_t i = base + r;
      double tj;
      a[r] = i < m ? segment_wh(ts, watts, i, ts[i], watts[i], &tj) : 0.0;
      acc = kvm_dd_add(acc, a[r]);
    }
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
#pragma unroll
    for (int r = 0; r < PFX_ITEMS; ++r) {
      const int64_t i = base + r;
      run = kvm_dd_add(run, a[r]);
      if (i < m) {
        p_hi[i] = run.hi;
        p_lo[i] = run.lo;
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

__global__ void __launch_bounds__(E_THREADS)
interval_energy_kernel(const double* __restrict__ ts, const double* __restrict__ watts,
                       const double* __restrict__ p_hi, const double* __restrict__ p_lo, int64_t m,
                       const double* __restrict__ start_ms, const double* __restrict__ latency_ms,
                       int64_t n, double* __restrict__ energy, int accumulate) {
  for (int64_t r = (int64_t)blockIdx.x * E_THREADS + threadIdx.x; r < n;
       r += (int64_t)gridDim.x * E_THREADS) {
    const double st = kvm_ld1(start_ms + r), la = kvm_ld1(latency_ms + r);
    const double s = st / 1000.0, e = (st + la) / 1000.0;
    double wh = 0.0;
    if (e > s && m > 1) {
      int64_t first = lower_bound_ts(ts, m, s);
      const int64_t last = upper_bound_ts(ts, m, e) - 1;
      while (first <= last && kvm_isnan(watts[first])) ++first;  // first kept sample of the interval
      if (first < last) {
        const kvm_dd d = kvm_dd_sub(kvm_dd{p_hi[last], p_lo[last]}, kvm_dd{p_hi[first], p_lo[first]});
        wh = d.hi + d.lo;
      }
    }
    energy[r] = accumulate ? energy[r] + wh : wh;
  }
}

int grid_for(int64_t items, int per_cta) {
  int g = kvm_grid_for(items, per_cta, 8);
  return g > E_MAX_GRID ? E_MAX_GRID : g;
}

}  // namespace

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

extern "C" int kvm_interval_energy(const double* ts, const double* watts, const double* p_hi,
                                   const double* p_lo, int64_t m, const double* start_ms,
                                   const double* latency_ms, int64_t n, double* energy_wh,
                                   int32_t accumulate, void* stream) {
  KVM_REQUIRE(n >= 0 && m >= 0, "kvm_interval_energy: negative size");
  if (n == 0) return KVM_OK;
  KVM_REQUIRE(start_ms && latency_ms && energy_wh && ((ts && watts && p_hi && p_lo) || m == 0),
              "kvm_interval_energy: null argument");
  interval_energy_kernel<<<grid_for(n, E_THREADS * 4), E_THREADS, 0, (cudaStream_t)stream>>>(
      ts, watts, p_hi, p_lo, m, start_ms, latency_ms, n, energy_wh, accumulate);
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}
