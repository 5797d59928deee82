// kvm_hist.cu - K5 uniform value histograms and K8 windowed series.
//
// K5 replaces analyze.py:84-122 `compute_histograms` (4 calls per run, :149-152):
//   idx = min(nb-1, int((v-min)/w)),  w = (max-min)/nb
// The bin index is a monotone function of v, so instead of dividing per element the
// planner finds, by bisection over the ordered-key space, the smallest key whose index
// reaches b (b = 1..nb-1) evaluating the reference expression itself; the streaming
// kernel then ranks each key against those nb-1 thresholds.  Bit-exact by construction,
// no fp64 division in the hot loop.
//
// K8 (extension A12, no reference function): 1-window series of request count, success
// count and completion tokens; bin = floor((start_ms - min_start_ms)/window_ms).
//
// Both are HBM-bound streaming passes; shared-memory histograms are privatised per CTA
// and lanes that hit the same bin are merged with __match_any_sync before the atomic.

#include "kvm_common.cuh"
#include "kvm_log_device.cuh"

namespace {

constexpr int H_THREADS = 512;
constexpr int MAX_BINS = 256;
constexpr int LOG_NHIST = 4;  // ttfb, tllt, generation, per-token

struct HistPlan {
  uint64_t thr[LOG_NHIST][MAX_BINS];  // thr[h][b-1] = smallest key with index >= b
  int32_t active[LOG_NHIST];          // 0: empty or min==max (host resolves from stats)
};

// thr for one histogram: thread b-1 bisects for bin b
__device__ void plan_thresholds(uint64_t klo, uint64_t khi, int nb, uint64_t* thr, int32_t* active) {
  const bool ok = klo < khi;
  if (threadIdx.x == 0) *active = ok ? 1 : 0;
  if (!ok) return;
  for (int b = 1 + threadIdx.x; b < nb; b += blockDim.x) thr[b - 1] = bin_threshold(klo, khi, nb, b);
}

__global__ void log_hist_plan_kernel(const kvm_log_stats* stats, int nb, HistPlan* plan) {
  const int h = blockIdx.x;
  const int col = h == 0 ? KVM_COL_TTFB : h == 1 ? KVM_COL_TLLT : h == 2 ? KVM_COL_GEN : KVM_COL_PERTOK;
  plan_thresholds(stats->sets.kmin[col], stats->sets.kmax[col], nb, plan->thr[h], &plan->active[h]);
}

__global__ void column_hist_plan_kernel(double lo, double hi, int nb, HistPlan* plan) {
  plan_thresholds(kvm_key(lo), kvm_key(hi), nb, plan->thr[0], &plan->active[0]);
}

// Warp-merged shared-memory increment; bin < 0 = lane has nothing to add.  Must be
// called by all 32 lanes.
__device__ __forceinline__ void warp_hist_add(uint32_t* h, int bin) {
  const unsigned peers = __match_any_sync(KVM_FULL, bin);
  if (bin >= 0 && (int)(__ffs(peers) - 1) == (int)(threadIdx.x & 31)) atomicAdd(&h[bin], (uint32_t)__popc(peers));
}

__global__ void __launch_bounds__(H_THREADS, 2)
log_hist20_kernel(kvm_request_log log, const uint8_t* __restrict__ flags,
                  const HistPlan* __restrict__ plan, int nb, int64_t* __restrict__ counts) {
  __shared__ uint64_t thr[LOG_NHIST][MAX_BINS];
  __shared__ uint32_t h[LOG_NHIST][MAX_BINS];
  __shared__ int32_t active[LOG_NHIST];
  for (int i = threadIdx.x; i < LOG_NHIST * MAX_BINS; i += H_THREADS) {
    (&h[0][0])[i] = 0;
    (&thr[0][0])[i] = (&plan->thr[0][0])[i];
  }
  if (threadIdx.x < LOG_NHIST) active[threadIdx.x] = plan->active[threadIdx.x];
  __syncthreads();
  const int64_t n = log.n;
  // warp-uniform trip count: every lane iterates while the warp's first row is in range
  const int64_t stride = (int64_t)gridDim.x * H_THREADS;
  for (int64_t base = (int64_t)blockIdx.x * H_THREADS + (threadIdx.x & ~31); base < n; base += stride) {
    const int64_t i = base + (threadIdx.x & 31);
    int b0 = -1, b1 = -1, b2 = -1, b3 = -1;
    if (i < n) {
      const uint8_t fl = kvm_ld_u8(flags + i);
      if (fl & KVM_F_SUCCESS) {
        const double tf = kvm_ld1(log.ttfb_ms + i);
        const double tl = kvm_ld1(log.tllt_ms + i);
        if (active[0] && !kvm_isnan(tf)) b0 = bin_of(kvm_key(tf), thr[0], nb);
        if (active[1] && !kvm_isnan(tl)) b1 = bin_of(kvm_key(tl), thr[1], nb);
        if (fl & KVM_F_GEN) {
          const double g = tl - tf;
          if (!kvm_isnan(g)) {
            if (active[2]) b2 = bin_of(kvm_key(g), thr[2], nb);
            const double pt = g / (double)kvm_ld1(log.completion_tokens + i);
            if (active[3] && !kvm_isnan(pt)) b3 = bin_of(kvm_key(pt), thr[3], nb);
          }
        }
      }
    }
    warp_hist_add(h[0], b0);
    warp_hist_add(h[1], b1);
    warp_hist_add(h[2], b2);
    warp_hist_add(h[3], b3);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < LOG_NHIST * nb; i += H_THREADS) {
    const int hh = i / nb, b = i % nb;
    if (h[hh][b]) atomicAdd(reinterpret_cast<unsigned long long*>(counts) + i, (unsigned long long)h[hh][b]);
  }
}

__global__ void __launch_bounds__(H_THREADS, 2)
column_hist_kernel2(const double* __restrict__ values, int64_t n, const HistPlan* __restrict__ plan,
                    int nb, int64_t* __restrict__ counts) {
  __shared__ uint64_t thr[MAX_BINS];
  __shared__ uint32_t h[MAX_BINS];
  for (int i = threadIdx.x; i < MAX_BINS; i += H_THREADS) {
    h[i] = 0;
    thr[i] = plan->thr[0][i];
  }
  __syncthreads();
  if (!plan->active[0]) return;
  const int64_t stride = (int64_t)gridDim.x * H_THREADS;
  for (int64_t base = (int64_t)blockIdx.x * H_THREADS + (threadIdx.x & ~31); base < n; base += stride) {
    const int64_t i = base + (threadIdx.x & 31);
    int b = -1;
    if (i < n) {
      const double v = kvm_ld1(values + i);
      if (!kvm_isnan(v)) b = bin_of(kvm_key(v), thr, nb);
    }
    warp_hist_add(h, b);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < nb; i += H_THREADS)
    if (h[i]) atomicAdd(reinterpret_cast<unsigned long long*>(counts) + i, (unsigned long long)h[i]);
}

// ------------------------------------------------------------------ K8 series
__global__ void __launch_bounds__(H_THREADS, 2)
log_series_kernel(kvm_request_log log, const uint8_t* __restrict__ flags,
                  const kvm_log_stats* __restrict__ stats, double window_ms, int64_t max_bins,
                  int64_t* __restrict__ series, int64_t* __restrict__ dropped) {
  const double s0 = stats->min_start_ms;
  const int64_t n = log.n;
  const int lane = threadIdx.x & 31;
  unsigned long long* req = reinterpret_cast<unsigned long long*>(series);
  unsigned long long* suc = req + max_bins;
  unsigned long long* tok = suc + max_bins;
  const int64_t stride = (int64_t)gridDim.x * H_THREADS;
  for (int64_t base = (int64_t)blockIdx.x * H_THREADS + (threadIdx.x & ~31); base < n; base += stride) {
    const int64_t i = base + lane;
    long long bin = -1;
    unsigned ok = 0;
    long long ct = 0;
    if (i < n) {
      const double st = kvm_ld1(log.start_ms + i);
      const double b = floor((st - s0) / window_ms);
      const uint8_t fl = kvm_ld_u8(flags + i);
      ok = fl & KVM_F_SUCCESS;
      if (ok) ct = kvm_ld1(log.completion_tokens + i);
      if (b >= 0.0 && b < (double)max_bins) bin = (long long)b;
      else if (b == b) bin = -2;  // out of range (NaN start: ignored)
    }
    // rows are time-ordered, so a warp usually spans one or two bins: merge per bin
    const unsigned peers = __match_any_sync(KVM_FULL, bin);
    const int leader = __ffs(peers) - 1;
    const unsigned okmask = __ballot_sync(KVM_FULL, ok != 0);
    // token sum over the peer group (segmented butterfly is overkill: loop over set bits in leader)
    long long tsum = 0;
    for (unsigned m = peers; m; m &= m - 1) {
      const int src = __ffs(m) - 1;
      const long long v = __shfl_sync(peers, ct, src);
      tsum += v;
    }
    if (lane == leader) {
      if (bin >= 0) {
        atomicAdd(&req[bin], (unsigned long long)__popc(peers));
        const int nok = __popc(peers & okmask);
        if (nok) {
          atomicAdd(&suc[bin], (unsigned long long)nok);
          atomicAdd(&tok[bin], (unsigned long long)tsum);
        }
      } else if (bin == -2) {
        atomicAdd(reinterpret_cast<unsigned long long*>(dropped), (unsigned long long)__popc(peers));
      }
    }
  }
}

}  // namespace

extern "C" size_t kvm_log_histograms_workspace_bytes(int32_t num_bins) {
  (void)num_bins;
  return kvm_align_up(sizeof(HistPlan), 256);
}

extern "C" int kvm_log_histograms(const kvm_request_log* log, const uint8_t* flags,
                                  const kvm_log_stats* stats, int32_t num_bins, int64_t* counts,
                                  void* workspace, size_t workspace_bytes, void* stream) {
  KVM_REQUIRE(log && flags && stats && counts && workspace, "kvm_log_histograms: null argument");
  KVM_REQUIRE(num_bins >= 1 && num_bins <= MAX_BINS, "kvm_log_histograms: num_bins=%d outside [1,%d]",
              num_bins, MAX_BINS);
  if (workspace_bytes < sizeof(HistPlan)) {
    kvm_set_error("kvm_log_histograms: workspace %zu < %zu bytes", workspace_bytes, sizeof(HistPlan));
    return KVM_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  HistPlan* plan = reinterpret_cast<HistPlan*>(workspace);
  KVM_CUDA_CHECK(cudaMemsetAsync(counts, 0, sizeof(int64_t) * LOG_NHIST * num_bins, st));
  KVM_CUDA_CHECK(cudaMemsetAsync(plan, 0xff, sizeof(HistPlan), st));
  log_hist_plan_kernel<<<LOG_NHIST, 64, 0, st>>>(stats, num_bins, plan);
  const int grid = kvm_grid_for(log->n, H_THREADS * 8, 2);
  log_hist20_kernel<<<grid, H_THREADS, 0, st>>>(*log, flags, plan, num_bins, counts);
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_column_histogram(const double* values, int64_t n, double lo, double hi,
                                    int32_t num_bins, int64_t* counts, void* workspace,
                                    size_t workspace_bytes, void* stream) {
  KVM_REQUIRE(counts && workspace && (values || n == 0), "kvm_column_histogram: null argument");
  KVM_REQUIRE(num_bins >= 1 && num_bins <= MAX_BINS, "kvm_column_histogram: num_bins=%d outside [1,%d]",
              num_bins, MAX_BINS);
  KVM_REQUIRE(lo < hi, "kvm_column_histogram: needs lo < hi (the caller resolves min==max)");
  if (workspace_bytes < sizeof(HistPlan)) {
    kvm_set_error("kvm_column_histogram: workspace %zu < %zu bytes", workspace_bytes, sizeof(HistPlan));
    return KVM_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  HistPlan* plan = reinterpret_cast<HistPlan*>(workspace);
  KVM_CUDA_CHECK(cudaMemsetAsync(counts, 0, sizeof(int64_t) * num_bins, st));
  KVM_CUDA_CHECK(cudaMemsetAsync(plan, 0xff, sizeof(HistPlan), st));
  column_hist_plan_kernel<<<1, 64, 0, st>>>(lo, hi, num_bins, plan);
  const int grid = kvm_grid_for(n, H_THREADS * 8, 2);
  column_hist_kernel2<<<grid, H_THREADS, 0, st>>>(values, n, plan, num_bins, counts);
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_log_series(const kvm_request_log* log, const uint8_t* flags,
                              const kvm_log_stats* stats, double window_ms, int64_t max_bins,
                              int64_t* series, int64_t* dropped, void* stream) {
  KVM_REQUIRE(log && flags && stats && series && dropped, "kvm_log_series: null argument");
  KVM_REQUIRE(window_ms > 0.0 && max_bins >= 1, "kvm_log_series: window_ms and max_bins must be positive");
  cudaStream_t st = (cudaStream_t)stream;
  KVM_CUDA_CHECK(cudaMemsetAsync(series, 0, sizeof(int64_t) * 3 * (size_t)max_bins, st));
  KVM_CUDA_CHECK(cudaMemsetAsync(dropped, 0, sizeof(int64_t), st));
  const int grid = kvm_grid_for(log->n, H_THREADS * 8, 2);
  log_series_kernel<<<grid, H_THREADS, 0, st>>>(*log, flags, stats, window_ms, max_bins, series, dropped);
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}
