// kvm_peer.cu - the exchange steps of the sharded (multi-GPU) analysis as our own kernels over
// NVLink peer memory (no counterpart in the reference, which is one
// process).  Three primitives, all stream-ordered and CUDA-graph capturable:
//   put      copy a local record into the same offset of every rank's exchange buffer
//   add      add the non-zero words of a local u64 array into every rank's accumulator
//            (system-scope atomics: the histograms of all ranks sum up in place)
//   barrier  arrival counters: "everything this rank pushed before is visible; wait until
//            every rank has said the same"
// The epoch lives on the device, so a captured graph replays the barrier unchanged.

#include <cstdlib>

#include "kvm_common.cuh"
#include "kvm_peer.cuh"

namespace {

constexpr int PEER_THREADS = 256;
// flag block: [0, MAX_WORLD) arrival counters (written by the peers), then epoch and error (local)
constexpr int FLAG_EPOCH = KVM_PEER_MAX_WORLD;
constexpr int FLAG_ERROR = KVM_PEER_MAX_WORLD + 1;
// a dead peer must not hang the GPU for ever; ordinary host skew (a GC pause, a first-call module
// load, uneven H2D) must not trip it either: default 60 s, KVM_PEER_TIMEOUT_S overrides
constexpr double DEFAULT_TIMEOUT_S = 60.0;

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// grid (chunks, world): CTA column y writes to rank y (8-byte words: the records are small)
__global__ void __launch_bounds__(PEER_THREADS)
peer_put_kernel(KvmPeerTable pt, const unsigned long long* __restrict__ src, int64_t words, int64_t dst_offset) {
  unsigned long long* __restrict__ dst = reinterpret_cast<unsigned long long*>(pt.base[blockIdx.y] + dst_offset);
  for (int64_t i = (int64_t)blockIdx.x * PEER_THREADS + threadIdx.x; i < words;
       i += (int64_t)gridDim.x * PEER_THREADS)
    dst[i] = src[i];
}

__global__ void __launch_bounds__(PEER_THREADS)
peer_add_kernel(KvmPeerTable pt, const unsigned long long* __restrict__ src, int64_t words,
                int64_t dst_offset) {
  unsigned long long* __restrict__ dst = reinterpret_cast<unsigned long long*>(pt.base[blockIdx.y] + dst_offset);
  for (int64_t i = (int64_t)blockIdx.x * PEER_THREADS + threadIdx.x; i < words;
       i += (int64_t)gridDim.x * PEER_THREADS) {
    const unsigned long long v = src[i];
    if (v) atomicAdd_system(dst + i, v);
  }
}

// One CTA, one thread per rank: release everything this GPU wrote so far (the earlier kernels of
// the stream included - fences are cumulative), bump my counter in rank p's flag block, then
// wait for rank p's bump in mine.
//
// Time-out: a rank that waits longer than `spin_limit` cycles for peer p writes 1+p into the error
// word of EVERY rank's flag block (system-scope stores), so the late rank learns about it too -
// it would otherwise sail through its own waits (the early rank has already bumped every counter)
// and report results computed from half-exchanged data.  The error word is sticky: once set, every
// later barrier on that buffer returns at once (the data of the step is void anyway) and the host
// raises on every rank (dist_api.ShardedAnalyzer) until the fabric is rebuilt.
__device__ __forceinline__ void st_relaxed_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__global__ void peer_barrier_kernel(KvmPeerTable pt, int64_t flags_offset, long long spin_limit) {
  __shared__ unsigned long long want;
  unsigned long long* local = reinterpret_cast<unsigned long long*>(pt.base[pt.rank] + flags_offset);
  if (threadIdx.x == 0) {
    want = local[FLAG_EPOCH] + 1ull;
    local[FLAG_EPOCH] = want;
  }
  __syncthreads();
  const int p = threadIdx.x;
  if (p < pt.world) {
    __threadfence_system();
    unsigned long long* remote = reinterpret_cast<unsigned long long*>(pt.base[p] + flags_offset);
    atomicAdd_system(remote + pt.rank, 1ull);
    const long long t0 = clock64();
    while (ld_acquire_sys(local + p) < want) {
      if (ld_acquire_sys(local + FLAG_ERROR) != 0ull) break;   // somebody already gave up on this step
      if (clock64() - t0 > spin_limit) {
        for (int r = 0; r < pt.world; ++r)
          st_relaxed_sys(reinterpret_cast<unsigned long long*>(pt.base[r] + flags_offset) + FLAG_ERROR,
                         1ull + (unsigned long long)p);
        break;
      }
    }
    __threadfence_system();
  }
}

}  // namespace

// spin limit in SM cycles: KVM_PEER_TIMEOUT_S (seconds, > 0) or the default, at the device's clock
static double g_peer_timeout_s = 0.0;   // 0: not decided yet

static long long peer_spin_limit_cycles() {
  if (g_peer_timeout_s <= 0.0) {
    double secs = DEFAULT_TIMEOUT_S;
    if (const char* e = getenv("KVM_PEER_TIMEOUT_S")) {
      const double v = atof(e);
      if (v > 0.0) secs = v;
    }
    g_peer_timeout_s = secs;
  }
  int dev = 0, khz = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev) != cudaSuccess || khz <= 0)
    khz = 2000000;
  return (long long)(g_peer_timeout_s * 1000.0 * (double)khz);
}

extern "C" double kvm_peer_set_timeout_s(double seconds) {
  peer_spin_limit_cycles();
  const double old = g_peer_timeout_s;
  if (seconds > 0.0) g_peer_timeout_s = seconds;
  return old;
}

extern "C" int64_t kvm_peer_flag_bytes(void) { return (int64_t)sizeof(unsigned long long) * (KVM_PEER_MAX_WORLD + 2); }

extern "C" int kvm_peer_put(const void* src, int64_t bytes, int64_t dst_offset, void* const* peers,
                            int32_t world, int32_t rank, void* stream) {
  KvmPeerTable pt;
  KVM_REQUIRE(kvm_peer_table(&pt, peers, world, rank) == 0, "kvm_peer_put: bad peer table (world=%d rank=%d)", world, rank);
  KVM_REQUIRE(src && bytes >= 0 && bytes % 8 == 0 && dst_offset % 8 == 0 &&
                  (reinterpret_cast<uintptr_t>(src) & 7u) == 0,
              "kvm_peer_put: source, size and offset must be 8-byte aligned");
  if (bytes == 0) return KVM_OK;
  const int64_t words = bytes / 8;
  const int64_t want = (words + PEER_THREADS - 1) / PEER_THREADS;
  const int gx = (int)(want < 64 ? want : 64);
  peer_put_kernel<<<dim3(gx, world), PEER_THREADS, 0, (cudaStream_t)stream>>>(
      pt, reinterpret_cast<const unsigned long long*>(src), words, dst_offset);
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_peer_add_u64(const uint64_t* src, int64_t words, int64_t dst_offset, void* const* peers,
                                int32_t world, int32_t rank, void* stream) {
  KvmPeerTable pt;
  KVM_REQUIRE(kvm_peer_table(&pt, peers, world, rank) == 0, "kvm_peer_add_u64: bad peer table (world=%d rank=%d)", world, rank);
  KVM_REQUIRE(src && words >= 0 && dst_offset % 8 == 0, "kvm_peer_add_u64: bad argument");
  if (words == 0) return KVM_OK;
  const int64_t want = (words + PEER_THREADS - 1) / PEER_THREADS;
  const int gx = (int)(want < 128 ? want : 128);
  peer_add_kernel<<<dim3(gx, world), PEER_THREADS, 0, (cudaStream_t)stream>>>(
      pt, reinterpret_cast<const unsigned long long*>(src), words, dst_offset);
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}

extern "C" int kvm_peer_barrier(int64_t flags_offset, void* const* peers, int32_t world, int32_t rank,
                                void* stream) {
  KvmPeerTable pt;
  KVM_REQUIRE(kvm_peer_table(&pt, peers, world, rank) == 0, "kvm_peer_barrier: bad peer table (world=%d rank=%d)", world, rank);
  KVM_REQUIRE(flags_offset % 8 == 0 && flags_offset >= 0, "kvm_peer_barrier: bad flag offset");
  peer_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(pt, flags_offset, peer_spin_limit_cycles());
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}
