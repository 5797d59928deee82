// kvm_sort.cu - K3: stable LSD radix sort of (fp64 key, fp64 payload) pairs.
//
// Replaces the list.sort calls that need a FULL order: energy/collector.py:141
// `ss.sort(key=lambda s: s.ts)` (power samples may arrive unsorted; the sort is stable, and
// stability decides which neighbours pair up when timestamps repeat).  Percentiles do not
// use this kernel - they go through the radix select (kvm_select.cu).
//
// 8-bit digits over the order-preserving u64 image of the key.  One up-front pass builds
// the 8 digit histograms (passes whose digit is constant over all keys are skipped - for
// epoch timestamps that is most of the high bytes).  Each remaining pass:
//   count    per-tile digit counts              (tile = 2048 keys)
//   scan     exclusive offsets in (digit, tile) order
//   scatter  stable placement: each warp owns a contiguous slice of the tile and ranks its
//            keys round by round with __match_any_sync + a per-warp digit counter
// HBM-bound: 40 B/pair/pass (keys read twice, payload once, both written).

#include "kvm_common.cuh"

namespace {

constexpr int RADIX_BITS = 8;
constexpr int RADIX = 1 << RADIX_BITS;
constexpr int SORT_THREADS = 256;
constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr int ITEMS = 8;
constexpr int TILE = SORT_THREADS * ITEMS;  // 2048
constexpr int NPASS = 64 / RADIX_BITS;

__device__ __forceinline__ uint64_t sort_key(double v) {
  // NaN keys sort last (after +inf) like any value with all exponent bits set
  return kvm_key(v);
}

__global__ void __launch_bounds__(SORT_THREADS)
sort_digit_hist_kernel(const double* __restrict__ keys, int64_t m, unsigned long long* __restrict__ ghist) {
  __shared__ uint32_t h[NPASS][RADIX];
  for (int i = threadIdx.x; i < NPASS * RADIX; i += SORT_THREADS) (&h[0][0])[i] = 0;
  __syncthreads();
  for (int64_t i = (int64_t)blockIdx.x * SORT_THREADS + threadIdx.x; i < m;
       i += (int64_t)gridDim.x * SORT_THREADS) {
    const uint64_t k = sort_key(kvm_ld1(keys + i));
#pragma unroll
    for (int p = 0; p < NPASS; ++p) atomicAdd(&h[p][(k >> (p * RADIX_BITS)) & (RADIX - 1)], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < NPASS * RADIX; i += SORT_THREADS)
    if ((&h[0][0])[i]) atomicAdd(&ghist[i], (unsigned long long)(&h[0][0])[i]);
}

// tile_counts[d * ntiles + tile]
__global__ void __launch_bounds__(SORT_THREADS)
sort_count_kernel(const double* __restrict__ keys, int64_t m, int shift, int64_t ntiles,
                  uint32_t* __restrict__ tile_counts) {
  __shared__ uint32_t h[RADIX];
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    h[threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = tile * TILE;
#pragma unroll
    for (int r = 0; r < ITEMS; ++r) {
      const int64_t i = base + r * SORT_THREADS + threadIdx.x;
      if (i < m) atomicAdd(&h[(sort_key(kvm_ld1(keys + i)) >> shift) & (RADIX - 1)], 1u);
    }
    __syncthreads();
    tile_counts[(int64_t)threadIdx.x * ntiles + tile] = h[threadIdx.x];
    __syncthreads();
  }
}

// One CTA per digit: exclusive scan of that digit's tile counts, offset by the number of
// keys with a smaller digit (from the up-front histogram of this pass).
__global__ void __launch_bounds__(SORT_THREADS)
sort_scan_kernel(const unsigned long long* __restrict__ pass_hist, int64_t ntiles,
                 const uint32_t* __restrict__ tile_counts, int64_t* __restrict__ tile_offsets) {
  __shared__ int64_t warp_tot[SORT_WARPS];
  __shared__ int64_t carry;
  const int d = blockIdx.x;
  if (threadIdx.x == 0) {
    int64_t b = 0;
    for (int k = 0; k < d; ++k) b += (int64_t)pass_hist[k];
    carry = b;
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int64_t t0 = 0; t0 < ntiles; t0 += SORT_THREADS) {
    const int64_t t = t0 + threadIdx.x;
    const int64_t c = t < ntiles ? (int64_t)tile_counts[(int64_t)d * ntiles + t] : 0;
    int64_t incl = c;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      int64_t o = __shfl_up_sync(KVM_FULL, incl, s);
      if (lane >= s) incl += o;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    int64_t before = carry;
    for (int w = 0; w < warp; ++w) before += warp_tot[w];
    if (t < ntiles) tile_offsets[(int64_t)d * ntiles + t] = before + incl - c;
    __syncthreads();
    if (threadIdx.x == SORT_THREADS - 1) carry = before + incl;
    __syncthreads();
  }
}

__global__ void __launch_bounds__(SORT_THREADS)
sort_scatter_kernel(const double* __restrict__ keys_in, const double* __restrict__ vals_in,
                    double* __restrict__ keys_out, double* __restrict__ vals_out, int64_t m,
                    int shift, int64_t ntiles, const int64_t* __restrict__ tile_offsets) {
  __shared__ uint32_t warp_cnt[SORT_WARPS][RADIX];
  __shared__ int64_t digit_base[RADIX];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned lt_mask = (1u << lane) - 1u;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    for (int i = threadIdx.x; i < SORT_WARPS * RADIX; i += SORT_THREADS) (&warp_cnt[0][0])[i] = 0;
    __syncthreads();
    // warp w owns keys [base + w*256, base + (w+1)*256): 8 rounds of 32 consecutive keys
    const int64_t wbase = tile * TILE + (int64_t)warp * (32 * ITEMS);
    double k[ITEMS];
    uint32_t rank[ITEMS];
    int dig[ITEMS];
#pragma unroll
    for (int r = 0; r < ITEMS; ++r) {
      const int64_t i = wbase + r * 32 + lane;
      const bool live = i < m;
      k[r] = live ? kvm_ld1(keys_in + i) : 0.0;
      dig[r] = live ? (int)((sort_key(k[r]) >> shift) & (RADIX - 1)) : -1;
      const unsigned peers = __match_any_sync(KVM_FULL, dig[r]);
      uint32_t before = 0;
      if (live) before = warp_cnt[warp][dig[r]];
      __syncwarp();
      rank[r] = before + __popc(peers & lt_mask);
      if (live && (int)(__ffs(peers) - 1) == lane) warp_cnt[warp][dig[r]] = before + __popc(peers);
      __syncwarp();
    }
    __syncthreads();
    {
      // thread d: exclusive prefix of digit d over the warps, on top of the tile's global offset
      const int d = threadIdx.x;
      int64_t run = tile_offsets[(int64_t)d * ntiles + tile];
      digit_base[d] = run;
      uint32_t acc = 0;
      for (int w = 0; w < SORT_WARPS; ++w) {
        const uint32_t c = warp_cnt[w][d];
        warp_cnt[w][d] = acc;
        acc += c;
      }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < ITEMS; ++r) {
      const int64_t i = wbase + r * 32 + lane;
      if (i < m) {
        const int64_t pos = digit_base[dig[r]] + warp_cnt[warp][dig[r]] + rank[r];
        keys_out[pos] = k[r];
        vals_out[pos] = kvm_ld1(vals_in + i);
      }
    }
    __syncthreads();
  }
}

struct SortLayout {
  size_t hist_off, counts_off, offsets_off, tmpk_off, tmpv_off, total;
  int64_t ntiles;
};

SortLayout sort_layout(int64_t m) {
  SortLayout L;
  L.ntiles = (m + TILE - 1) / TILE;
  if (L.ntiles < 1) L.ntiles = 1;
  size_t off = 0;
  L.hist_off = off;
  off += kvm_align_up(sizeof(unsigned long long) * NPASS * RADIX, 256);
  L.counts_off = off;
  off += kvm_align_up(sizeof(uint32_t) * RADIX * (size_t)L.ntiles, 256);
  L.offsets_off = off;
  off += kvm_align_up(sizeof(int64_t) * RADIX * (size_t)L.ntiles, 256);
  L.tmpk_off = off;
  off += kvm_align_up(sizeof(double) * (size_t)(m > 0 ? m : 1), 256);
  L.tmpv_off = off;
  off += kvm_align_up(sizeof(double) * (size_t)(m > 0 ? m : 1), 256);
  L.total = off;
  return L;
}

}  // namespace

extern "C" size_t kvm_sort_pairs_workspace_bytes(int64_t m) { return sort_layout(m < 0 ? 0 : m).total; }

// Synchronises the stream once (to read the up-front digit histograms and skip constant digits).
extern "C" int kvm_sort_pairs_f64(const double* keys_in, const double* vals_in, double* keys_out,
                                  double* vals_out, int64_t m, void* workspace,
                                  size_t workspace_bytes, void* stream) {
  KVM_REQUIRE(m >= 0, "kvm_sort_pairs_f64: negative size");
  if (m == 0) return KVM_OK;
  KVM_REQUIRE(keys_in && vals_in && keys_out && vals_out && workspace, "kvm_sort_pairs_f64: null argument");
  KVM_REQUIRE(keys_in != keys_out && vals_in != vals_out, "kvm_sort_pairs_f64: in-place sort is not supported");
  const SortLayout L = sort_layout(m);
  if (workspace_bytes < L.total) {
    kvm_set_error("kvm_sort_pairs_f64: workspace %zu < %zu bytes", workspace_bytes, L.total);
    return KVM_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  unsigned char* ws = reinterpret_cast<unsigned char*>(workspace);
  unsigned long long* ghist = reinterpret_cast<unsigned long long*>(ws + L.hist_off);
  uint32_t* counts = reinterpret_cast<uint32_t*>(ws + L.counts_off);
  int64_t* offsets = reinterpret_cast<int64_t*>(ws + L.offsets_off);
  double* tmpk = reinterpret_cast<double*>(ws + L.tmpk_off);
  double* tmpv = reinterpret_cast<double*>(ws + L.tmpv_off);

  KVM_CUDA_CHECK(cudaMemsetAsync(ghist, 0, sizeof(unsigned long long) * NPASS * RADIX, st));
  sort_digit_hist_kernel<<<kvm_grid_for(m, SORT_THREADS * 8, 4), SORT_THREADS, 0, st>>>(keys_in, m, ghist);
  KVM_CUDA_CHECK(cudaGetLastError());
  static thread_local unsigned long long host_hist[NPASS * RADIX];
  KVM_CUDA_CHECK(cudaMemcpyAsync(host_hist, ghist, sizeof(host_hist), cudaMemcpyDeviceToHost, st));
  KVM_CUDA_CHECK(cudaStreamSynchronize(st));

  int active[NPASS], nactive = 0;
  for (int p = 0; p < NPASS; ++p) {
    bool constant = false;
    for (int d = 0; d < RADIX; ++d)
      if (host_hist[p * RADIX + d] == (unsigned long long)m) constant = true;
    if (!constant) active[nactive++] = p;
  }
  if (nactive == 0) {  // all keys equal: stable sort == copy
    KVM_CUDA_CHECK(cudaMemcpyAsync(keys_out, keys_in, sizeof(double) * m, cudaMemcpyDeviceToDevice, st));
    KVM_CUDA_CHECK(cudaMemcpyAsync(vals_out, vals_in, sizeof(double) * m, cudaMemcpyDeviceToDevice, st));
    return KVM_OK;
  }
  // ping-pong so that the last active pass lands in (keys_out, vals_out)
  const double* src_k = keys_in;
  const double* src_v = vals_in;
  const int grid = kvm_grid_for(m, TILE, 8);
  for (int a = 0; a < nactive; ++a) {
    const bool to_out = ((nactive - 1 - a) % 2) == 0;
    double* dst_k = to_out ? keys_out : tmpk;
    double* dst_v = to_out ? vals_out : tmpv;
    const int shift = active[a] * RADIX_BITS;
    sort_count_kernel<<<grid, SORT_THREADS, 0, st>>>(src_k, m, shift, L.ntiles, counts);
    sort_scan_kernel<<<RADIX, SORT_THREADS, 0, st>>>(ghist + active[a] * RADIX, L.ntiles, counts, offsets);
    sort_scatter_kernel<<<grid, SORT_THREADS, 0, st>>>(src_k, src_v, dst_k, dst_v, m, shift, L.ntiles, offsets);
    KVM_CUDA_CHECK(cudaGetLastError());
    src_k = dst_k;
    src_v = dst_v;
  }
  return KVM_OK;
}
