// kvm_nearest.cu - index of the member element closest to a target value.
//
// Replaces the exemplar search of report_generator.py:437-454: over the rows with status 200 and
// a latency, `sorted((abs(latency - p95), trace_id))[0]` - i.e. the FIRST row (stable sort) with
// the smallest |latency - p95| - which costs a CSV re-parse and an O(n log n) sort of tuples.
// Here: one streaming pass (16 B/row with the status membership), the result is the row index;
// the host looks the trace id up.  Distances are non-negative doubles, so their bit patterns
// order like the values; ties go to the lower index, NaN distances never win.

#include "kvm_common.cuh"

namespace {

constexpr int NEAR_THREADS = 256;

struct Best {
  unsigned long long dist;  // bits of |v - target|; ~0 = none
  long long idx;
};

__device__ __forceinline__ void take(Best& b, double v, bool member, double target, long long i) {
  const double d = fabs(v - target);
  if (member && d == d) {  // NaN (v NaN, or inf - inf) never wins
    const unsigned long long k = (unsigned long long)__double_as_longlong(d);
    if (k < b.dist || (k == b.dist && i < b.idx)) {
      b.dist = k;
      b.idx = i;
    }
  }
}

__device__ __forceinline__ Best better(Best a, Best b) {
  return (b.dist < a.dist || (b.dist == a.dist && b.idx < a.idx)) ? b : a;
}

__device__ __forceinline__ Best warp_best(Best b) {
#pragma unroll
  for (int m = 16; m; m >>= 1) {
    Best o;
    o.dist = __shfl_xor_sync(KVM_FULL, b.dist, m);
    o.idx = __shfl_xor_sync(KVM_FULL, b.idx, m);
    b = better(b, o);
  }
  return b;
}

struct NearWs {
  unsigned int arrived;
  unsigned int pad[3];
  Best part[1];  // [grid]
};

template <int MODE, bool VEC>
__global__ void __launch_bounds__(NEAR_THREADS)
column_nearest_kernel(const double* __restrict__ values, int64_t n, const void* __restrict__ member,
                      double target, kvm_nearest_result* out, NearWs* ws) {
  __shared__ Best red[NEAR_THREADS / 32];
  __shared__ bool is_last;
  Best b{~0ull, 0x7fffffffffffffffll};
  const int64_t* st = reinterpret_cast<const int64_t*>(member);
  const uint8_t* ids = reinterpret_cast<const uint8_t*>(member);
  if (VEC) {
    const int64_t npairs = n >> 1;
    for (int64_t p = (int64_t)blockIdx.x * NEAR_THREADS + threadIdx.x; p < npairs;
         p += (int64_t)gridDim.x * NEAR_THREADS) {
      const int64_t i = p << 1;
      const double2 v = kvm_ld2(values + i);
      bool m0 = true, m1 = true;
      if (MODE == KVM_MEMBER_STATUS) {
        const longlong2 s2 = kvm_ld2(st + i);
        m0 = s2.x == 200;
        m1 = s2.y == 200;
      }
      take(b, v.x, m0, target, i);
      take(b, v.y, m1, target, i + 1);
    }
    if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0)
      take(b, values[n - 1], MODE != KVM_MEMBER_STATUS || st[n - 1] == 200, target, n - 1);
  } else {
    for (int64_t i = (int64_t)blockIdx.x * NEAR_THREADS + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * NEAR_THREADS) {
      bool m = true;
      if (MODE == KVM_MEMBER_STATUS) m = st[i] == 200;
      if (MODE == KVM_MEMBER_SET_IDS) m = ids[i] != 255;
      take(b, values[i], m, target, i);
    }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  b = warp_best(b);
  if (lane == 0) red[warp] = b;
  __syncthreads();
  if (warp == 0) {
    b = lane < NEAR_THREADS / 32 ? red[lane] : Best{~0ull, 0x7fffffffffffffffll};
    b = warp_best(b);
    if (lane == 0) {
      ws->part[blockIdx.x] = b;
      __threadfence();
      is_last = atomicAdd(&ws->arrived, 1u) == gridDim.x - 1;
    }
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  Best g{~0ull, 0x7fffffffffffffffll};
  for (unsigned int k = threadIdx.x; k < gridDim.x; k += NEAR_THREADS) {
    Best o;
    o.dist = __ldcg(&ws->part[k].dist);
    o.idx = __ldcg(&ws->part[k].idx);
    g = better(g, o);
  }
  g = warp_best(g);
  if (lane == 0) red[warp] = g;
  __syncthreads();
  if (warp == 0) {
    g = lane < NEAR_THREADS / 32 ? red[lane] : Best{~0ull, 0x7fffffffffffffffll};
    g = warp_best(g);
    if (lane == 0) {
      kvm_nearest_result r;
      r.index = g.dist == ~0ull ? -1 : g.idx;
      r.distance = g.dist == ~0ull ? __longlong_as_double(0x7ff8000000000000ll)
                                   : __longlong_as_double((long long)g.dist);
      *out = r;
      ws->arrived = 0;  // ready for the next call
    }
  }
}

}  // namespace

extern "C" size_t kvm_column_nearest_workspace_bytes(void) {
  return sizeof(NearWs) + sizeof(Best) * (size_t)(4 * 160 * 8);
}

extern "C" int kvm_column_nearest(const double* values, int64_t n, int32_t mode, const void* member,
                                  double target, kvm_nearest_result* result, void* workspace,
                                  size_t workspace_bytes, void* stream) {
  KVM_REQUIRE(result && workspace && (values || n == 0), "kvm_column_nearest: null argument");
  KVM_REQUIRE(mode == KVM_MEMBER_ALL || mode == KVM_MEMBER_SET_IDS || mode == KVM_MEMBER_STATUS,
              "kvm_column_nearest: member mode %d", mode);
  KVM_REQUIRE(mode == KVM_MEMBER_ALL || member, "kvm_column_nearest: member array is null");
  if (workspace_bytes < kvm_column_nearest_workspace_bytes()) {
    kvm_set_error("kvm_column_nearest: workspace %zu < %zu bytes", workspace_bytes, kvm_column_nearest_workspace_bytes());
    return KVM_ERR_WORKSPACE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  NearWs* ws = reinterpret_cast<NearWs*>(workspace);
  KVM_CUDA_CHECK(cudaMemsetAsync(&ws->arrived, 0, sizeof(unsigned int), st));
  int grid = kvm_grid_for(n, NEAR_THREADS * 2 * 4, 8);
  const int cap = 4 * 160 * 8;
  if (grid > cap) grid = cap;
  const bool vec = mode != KVM_MEMBER_SET_IDS && (reinterpret_cast<uintptr_t>(values) & 15u) == 0 &&
                   (mode != KVM_MEMBER_STATUS || (reinterpret_cast<uintptr_t>(member) & 15u) == 0);
#define KVM_LAUNCH_NEAR(M, V) \
  column_nearest_kernel<M, V><<<grid, NEAR_THREADS, 0, st>>>(values, n, member, target, result, ws)
  if (mode == KVM_MEMBER_ALL) { if (vec) KVM_LAUNCH_NEAR(KVM_MEMBER_ALL, true); else KVM_LAUNCH_NEAR(KVM_MEMBER_ALL, false); }
  else if (mode == KVM_MEMBER_STATUS) { if (vec) KVM_LAUNCH_NEAR(KVM_MEMBER_STATUS, true); else KVM_LAUNCH_NEAR(KVM_MEMBER_STATUS, false); }
  else KVM_LAUNCH_NEAR(KVM_MEMBER_SET_IDS, false);
#undef KVM_LAUNCH_NEAR
  KVM_CUDA_CHECK(cudaGetLastError());
  return KVM_OK;
}
