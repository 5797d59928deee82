// kvm_sort.cu - K3: stable LSD radix sort of (fp64 key, fp64 payload) pairs, "onesweep" form.
//
// Replaces the list.sort calls that need a FULL order: energy/collector.py:141
// `ss.sort(key=lambda s: s.ts)` (power samples may arrive unsorted; the sort is stable, and
// stability decides which neighbours pair up when timestamps repeat).  Percentiles do not
// use this kernel - they go through the radix select (kvm_select.cu).
//
// 8-bit digits over the order-preserving u64 image of the key.
//   hist     ONE pass over the keys builds all eight digit histograms (8 B/pair);
//   plan     one small CTA: exclusive scan of every histogram, and which digits are constant over
//            all keys (for epoch timestamps most of the high bytes) - those passes do nothing.
//            Decided ON THE DEVICE: no host synchronisation anywhere in the sort.  The plan also
//            fixes the ping-pong so that the last pass that runs writes the caller's output.
//   pass x8  ONE kernel per digit (16 B read + 16 B written per pair, nothing read twice): a tile
//            of 4096 pairs per CTA, taken in ticket order.  Keys are ranked inside the tile with
//            __match_any_sync (stable), the tile's digit counts are published and the counts of
//            the tiles in front are collected by decoupled look-back (no separate count / scan
//            kernels), the pairs are reordered in shared memory and leave as runs of consecutive
//            addresses per digit (coalesced scatter).
// HBM-bound: 32 B/pair per executed pass + 8 B/pair for the histograms.

#include "kvm_common.cuh"

namespace {

constexpr int RADIX_BITS = 8;
constexpr int RADIX = 1 << RADIX_BITS;
constexpr int NPASS = 64 / RADIX_BITS;
constexpr int SORT_THREADS = 512;
constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr int ITEMS = 8;
constexpr int TILE = SORT_THREADS * ITEMS;  // 4096 pairs
constexpr uint32_t FLAG_AGG = 1u << 30, FLAG_PREFIX = 2u << 30, VALUE_MASK = (1u << 30) - 1u;

struct SortPlan {
  unsigned long long ghist[NPASS][RADIX];   // digit histograms, then exclusive offsets
  int32_t run[NPASS];                       // 1: the digit varies, the pass runs
  int32_t order[NPASS];                     // index of the pass among the executed ones
  int32_t executed;
  unsigned int ticket[NPASS];
};

__device__ __forceinline__ uint64_t sort_key(double v) {
  // NaN keys sort last (after +inf) like any value with all exponent bits set
  return kvm_key(v);
}

__global__ void __launch_bounds__(SORT_THREADS)
sort_digit_hist_kernel(const double* __restrict__ keys, int64_t m, SortPlan* plan) {
  __shared__ uint32_t h[NPASS][RADIX];
  for (int i = threadIdx.x; i < NPASS * RADIX; i += SORT_THREADS) (&h[0][0])[i] = 0;
  __syncthreads();
  const bool vec = (reinterpret_cast<uintptr_t>(keys) & 15u) == 0;
  auto count = [&](double v) {
    const uint64_t k = sort_key(v);
#pragma unroll
    for (int p = 0; p < NPASS; ++p) atomicAdd(&h[p][(k >> (p * RADIX_BITS)) & (RADIX - 1)], 1u);
  };
  if (vec) {
    const int64_t npairs = m >> 1;
    for (int64_t i = (int64_t)blockIdx.x * SORT_THREADS + threadIdx.x; i < npairs; i += (int64_t)gridDim.x * SORT_THREADS) {
      const double2 v = kvm_ld2(keys + 2 * i);
      count(v.x);
      count(v.y);
    }
    if ((m & 1) && blockIdx.x == 0 && threadIdx.x == 0) count(keys[m - 1]);
  } else {
    for (int64_t i = (int64_t)blockIdx.x * SORT_THREADS + threadIdx.x; i < m; i += (int64_t)gridDim.x * SORT_THREADS)
      count(kvm_ld1(keys + i));
  }
  __syncthreads();
  for (int i = threadIdx.x; i < NPASS * RADIX; i += SORT_THREADS)
    if ((&h[0][0])[i]) atomicAdd(&plan->ghist[0][0] + i, (unsigned long long)(&h[0][0])[i]);
}

// one warp per digit position: exclusive scan of its histogram, constant-digit test
__global__ void __launch_bounds__(32 * NPASS) sort_plan_kernel(SortPlan* plan, int64_t m) {
  __shared__ int run_s[NPASS];
  const int p = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned long long c[RADIX / 32], local = 0;
  bool constant = false;
#pragma unroll
  for (int k = 0; k < RADIX / 32; ++k) {
    c[k] = plan->ghist[p][lane * (RADIX / 32) + k];
    constant |= c[k] == (unsigned long long)m;
    local += c[k];
  }
  unsigned long long incl = local;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned long long o = __shfl_up_sync(KVM_FULL, incl, d);
    if (lane >= d) incl += o;
  }
  unsigned long long run = incl - local;
#pragma unroll
  for (int k = 0; k < RADIX / 32; ++k) {
    plan->ghist[p][lane * (RADIX / 32) + k] = run;
    run += c[k];
  }
  const bool any_const = __any_sync(KVM_FULL, constant);
  if (lane == 0) run_s[p] = any_const ? 0 : 1;
  __syncthreads();
  if (threadIdx.x == 0) {
    int k = 0;
    for (int q = 0; q < NPASS; ++q) {
      plan->run[q] = run_s[q];
      plan->order[q] = k;
      k += run_s[q];
      plan->ticket[q] = 0;
    }
    plan->executed = k;
  }
}

__device__ __forceinline__ uint32_t ld_status(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_status(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

struct SortSmem {
  uint64_t keys[TILE];
  double vals[TILE];
  uint32_t warp_hist[SORT_WARPS][RADIX];   // per warp digit counts, then exclusive prefix over the warps
  uint32_t tile_hist[RADIX];
  uint32_t digit_start[RADIX];              // exclusive scan of tile_hist
  unsigned long long global_base[RADIX];    // where the tile's run of every digit starts in the output
  unsigned int tile_id;
};

// One digit pass.  src / dst are chosen from the plan: the first executed pass reads the caller's
// input, the last one writes the caller's output, the ones in between ping-pong.
__global__ void __launch_bounds__(SORT_THREADS)
sort_onesweep_kernel(int pass, const double* __restrict__ in_k, const double* __restrict__ in_v,
                     double* __restrict__ out_k, double* __restrict__ out_v, double* __restrict__ tmp_k,
                     double* __restrict__ tmp_v, int64_t m, SortPlan* plan, uint32_t* __restrict__ status /* [ntiles][RADIX] */) {
  extern __shared__ __align__(16) unsigned char sraw[];
  SortSmem& S = *reinterpret_cast<SortSmem*>(sraw);
  if (!plan->run[pass]) return;   // constant digit: nothing to do
  const int ord = plan->order[pass], executed = plan->executed;
  // destination: out when an even number of executed passes follow, else tmp
  const bool dst_is_out = ((executed - 1 - ord) & 1) == 0;
  const double* __restrict__ src_k = ord == 0 ? in_k : (dst_is_out ? tmp_k : out_k);
  const double* __restrict__ src_v = ord == 0 ? in_v : (dst_is_out ? tmp_v : out_v);
  double* __restrict__ dst_k = dst_is_out ? out_k : tmp_k;
  double* __restrict__ dst_v = dst_is_out ? out_v : tmp_v;
  const int shift = pass * RADIX_BITS;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const unsigned lt = (1u << lane) - 1u;
  const int64_t ntiles = (m + TILE - 1) / TILE;
  for (;;) {
    __syncthreads();
    if (t == 0) S.tile_id = atomicAdd(&plan->ticket[pass], 1u);   // ticket order: a tile only waits for started tiles
    for (int i = t; i < SORT_WARPS * RADIX; i += SORT_THREADS) (&S.warp_hist[0][0])[i] = 0;
    __syncthreads();
    const int64_t tile = S.tile_id;
    if (tile >= ntiles) return;
    const int64_t base = tile * TILE + (int64_t)warp * (32 * ITEMS);
    // ---- load (striped inside the warp: item r of lane l is element r*32 + l of the warp's slice)
    uint64_t key[ITEMS];
    uint32_t rank[ITEMS];
#pragma unroll
    for (int r = 0; r < ITEMS; ++r) {
      const int64_t i = base + r * 32 + lane;
      key[r] = i < m ? sort_key(kvm_ld1(src_k + i)) : ~0ull;
    }
    // ---- stable rank inside the warp: items in order, lanes in order inside an item
#pragma unroll
    for (int r = 0; r < ITEMS; ++r) {
      const int64_t i = base + r * 32 + lane;
      const uint32_t d = (uint32_t)(key[r] >> shift) & (RADIX - 1);
      const bool live = i < m;
      const unsigned peers = __match_any_sync(KVM_FULL, live ? d : 0xffffffffu);
      uint32_t before = 0;
      if (live) {
        before = S.warp_hist[warp][d];   // every peer reads the same counter before the leader bumps it
      }
      __syncwarp();
      if (live && lane == (__ffs(peers) - 1)) S.warp_hist[warp][d] = before + (uint32_t)__popc(peers);
      __syncwarp();
      rank[r] = before + (uint32_t)__popc(peers & lt);
    }
    __syncthreads();
    // ---- digit totals of the tile, exclusive prefix over the warps (thread d owns digit d)
    if (t < RADIX) {
      uint32_t run = 0;
#pragma unroll
      for (int w = 0; w < SORT_WARPS; ++w) {
        const uint32_t c = S.warp_hist[w][t];
        S.warp_hist[w][t] = run;
        run += c;
      }
      S.tile_hist[t] = run;
      // publish, then look back over the tiles in front
      uint32_t* mine = status + (size_t)tile * RADIX + t;
      unsigned long long excl = 0;
      if (tile == 0) {
        st_status(mine, FLAG_PREFIX | run);
      } else {
        st_status(mine, FLAG_AGG | run);
        for (int64_t q = tile - 1; q >= 0; --q) {
          uint32_t v;
          do {   // (a volatile load: a plain one lets the compiler assume the loop ends and read once)
            v = ld_status(status + (size_t)q * RADIX + t);
          } while ((v >> 30) == 0u);
          excl += v & VALUE_MASK;
          if ((v >> 30) == 2u) break;
        }
        // the inclusive prefix may exceed 30 bits only when m >= 2^30: the entry point refuses that
        st_status(mine, FLAG_PREFIX | (uint32_t)(excl + run));
      }
      S.global_base[t] = plan->ghist[pass][t] + excl;
    }
    __syncthreads();
    // exclusive scan of the tile's digit counts (one warp, 8 digits per lane)
    if (warp == 0) {
      uint32_t c[RADIX / 32], local = 0;
#pragma unroll
      for (int k = 0; k < RADIX / 32; ++k) {
        c[k] = S.tile_hist[lane * (RADIX / 32) + k];
        local += c[k];
      }
      uint32_t incl = local;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(KVM_FULL, incl, d);
        if (lane >= d) incl += o;
      }
      uint32_t run = incl - local;
#pragma unroll
      for (int k = 0; k < RADIX / 32; ++k) {
        S.digit_start[lane * (RADIX / 32) + k] = run;
        run += c[k];
      }
    }
    __syncthreads();
    // ---- reorder in shared memory: slot = start of the digit in the tile + rank inside the digit
#pragma unroll
    for (int r = 0; r < ITEMS; ++r) {
      const int64_t i = base + r * 32 + lane;
      if (i < m) {
        const uint32_t d = (uint32_t)(key[r] >> shift) & (RADIX - 1);
        const uint32_t slot = S.digit_start[d] + S.warp_hist[warp][d] + rank[r];
        S.keys[slot] = key[r];
        S.vals[slot] = kvm_ld1(src_v + i);
      }
    }
    __syncthreads();
    // ---- scatter: consecutive threads write consecutive addresses inside every digit's run
    const int count = (int)min((int64_t)TILE, m - tile * TILE);
    for (int s = t; s < count; s += SORT_THREADS) {
      const uint64_t k = S.keys[s];
      const uint32_t d = (uint32_t)(k >> shift) & (RADIX - 1);
      const unsigned long long pos = S.global_base[d] + (unsigned long long)(s - S.digit_start[d]);
      dst_k[pos] = kvm_unkey(k);
      dst_v[pos] = S.vals[s];
    }
  }
}

// nothing ran (all keys equal, or one pair): the output is the input
__global__ void sort_copy_if_idle_kernel(const double* __restrict__ in_k, const double* __restrict__ in_v,
                                         double* __restrict__ out_k, double* __restrict__ out_v, int64_t m,
                                         const SortPlan* plan) {
  if (plan->executed != 0) return;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (int64_t)gridDim.x * blockDim.x) {
    out_k[i] = in_k[i];
    out_v[i] = in_v[i];
  }
}

struct SortLayout {
  size_t plan, status, tmp_k, tmp_v, total;
};
SortLayout sort_layout(int64_t m) {
  const int64_t ntiles = (m + TILE - 1) / TILE;
  SortLayout L;
  L.plan = 0;
  L.status = kvm_align_up(sizeof(SortPlan), 256);
  L.tmp_k = L.status + kvm_align_up(sizeof(uint32_t) * (size_t)RADIX * (size_t)(ntiles > 0 ? ntiles : 1), 256);
  L.tmp_v = L.tmp_k + kvm_align_up(sizeof(double) * (size_t)m, 256);
  L.total = L.tmp_v + kvm_align_up(sizeof(double) * (size_t)m, 256);
  return L;
}

}  // namespace

extern "C" size_t kvm_sort_pairs_workspace_bytes(int64_t m) { return sort_layout(m < 0 ? 0 : m).total; }

// Stream-ordered, no host synchronisation.
extern "C" int kvm_sort_pairs_f64(const double* keys_in, const double* vals_in, double* keys_out,
                                  double* vals_out, int64_t m, void* workspace,
                                  size_t workspace_bytes, void* stream) {
  KVM_REQUIRE(m >= 0, "kvm_sort_pairs_f64: negative size");
  if (m == 0) return KVM_OK;
  KVM_REQUIRE(m < (1ll << 30), "kvm_sort_pairs_f64: m=%lld exceeds the supported 2^30 - 1 pairs per call", (long long)m);
  KVM_REQUIRE(keys_in && vals_in && keys_out && vals_out && workspace, "kvm_sort_pairs_f64: null argument");
  KVM_REQUIRE(keys_in != keys_out && vals_in != vals_out, "kvm_sort_pairs_f64: in-place sort is not supported");
  const SortLayout L = sort_layout(m);
  if (workspace_bytes < L.total) {
    kvm_set_error("kvm_sort_pairs_f64: workspace %zu < %zu bytes", workspace_bytes, L.total);
    return KVM_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  unsigned char* wsb = reinterpret_cast<unsigned char*>(workspace);
  SortPlan* plan = reinterpret_cast<SortPlan*>(wsb + L.plan);
  uint32_t* status = reinterpret_cast<uint32_t*>(wsb + L.status);
  double* tmp_k = reinterpret_cast<double*>(wsb + L.tmp_k);
  double* tmp_v = reinterpret_cast<double*>(wsb + L.tmp_v);
  const int64_t ntiles = (m + TILE - 1) / TILE;
  KVM_CUDA_CHECK(cudaMemsetAsync(plan, 0, sizeof(SortPlan), st));
  sort_digit_hist_kernel<<<kvm_grid_for(m, SORT_THREADS * 16, 2), SORT_THREADS, 0, st>>>(keys_in, m, plan);
  sort_plan_kernel<<<1, 32 * NPASS, 0, st>>>(plan, m);
  KVM_CUDA_CHECK(cudaFuncSetAttribute((const void*)sort_onesweep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)sizeof(SortSmem)));
  const int grid = (int)(ntiles < 2 * kvm_num_sms() ? ntiles : 2 * kvm_num_sms());
  for (int p = 0; p < NPASS; ++p) {
    // (a pass whose digit is constant returns at once and never looks at the status words)
    KVM_CUDA_CHECK(cudaMemsetAsync(status, 0, sizeof(uint32_t) * (size_t)RADIX * (size_t)ntiles, st));
    sort_onesweep_kernel<<<grid, SORT_THREADS, sizeof(SortSmem), st>>>(p, keys_in, vals_in, keys_out, vals_out, tmp_k,
                                                                       tmp_v, m, plan, status);
  }
  sort_copy_if_idle_kernel<<<kvm_grid_for(m, 256 * 8, 2), 256, 0, st>>>(keys_in, vals_in, keys_out, vals_out, m, plan);
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}
