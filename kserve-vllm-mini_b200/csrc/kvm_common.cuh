// kvm_common.cuh - shared device/host helpers for libkvmini_b200 (sm_100a).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/kvmini_b200.h"

// ------------------------------------------------------------------ host error plumbing
void kvm_set_error(const char* fmt, ...);

#define KVM_CUDA_CHECK(expr)                                                        \
  do {                                                                              \
    cudaError_t _e = (expr);                                                        \
    if (_e != cudaSuccess) {                                                        \
      kvm_set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                    __LINE__);                                                      \
      return KVM_ERR_CUDA;                                                          \
    }                                                                               \
  } while (0)

#define KVM_REQUIRE(cond, ...)    \
  do {                            \
    if (!(cond)) {                \
      kvm_set_error(__VA_ARGS__); \
      return KVM_ERR_INVALID;     \
    }                             \
  } while (0)

int kvm_num_sms();  // cached per process for the current device

static inline size_t kvm_align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ------------------------------------------------------------------ constants
constexpr int KVM_WARP = 32;
constexpr unsigned KVM_FULL = 0xffffffffu;

// ------------------------------------------------------------------ keys
// Order-preserving image of an fp64 value: flip the sign bit of non-negatives, all bits
// of negatives.  Unsigned comparison of keys == numeric comparison of the doubles
// (-0.0 sorts just below +0.0; the reference's sort treats them as equal - DESIGN.md).
__host__ __device__ __forceinline__ uint64_t kvm_key_from_bits(uint64_t b) {
  return b ^ ((b >> 63) ? 0xffffffffffffffffull : 0x8000000000000000ull);
}
__host__ __device__ __forceinline__ uint64_t kvm_bits_from_key(uint64_t k) {
  return k ^ ((k >> 63) ? 0x8000000000000000ull : 0xffffffffffffffffull);
}
__device__ __forceinline__ uint64_t kvm_key(double v) {
  return kvm_key_from_bits((uint64_t)__double_as_longlong(v));
}
__device__ __forceinline__ double kvm_unkey(uint64_t k) {
  return __longlong_as_double((long long)kvm_bits_from_key(k));
}
__device__ __forceinline__ bool kvm_isnan(double v) { return v != v; }

// ------------------------------------------------------------------ streaming loads
// 128-bit read-only loads that do not allocate in L1 (every column is read once per pass).
__device__ __forceinline__ double2 kvm_ld2(const double* p) {
  double2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];"
               : "=d"(r.x), "=d"(r.y)
               : "l"(p));
  return r;
}
__device__ __forceinline__ longlong2 kvm_ld2(const int64_t* p) {
  longlong2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.s64 {%0,%1}, [%2];"
               : "=l"(r.x), "=l"(r.y)
               : "l"(p));
  return r;
}
__device__ __forceinline__ double kvm_ld1(const double* p) {
  double r;
  asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ int64_t kvm_ld1(const int64_t* p) {
  long long r;
  asm volatile("ld.global.nc.L1::no_allocate.s64 %0, [%1];" : "=l"(r) : "l"(p));
  return (int64_t)r;
}
__device__ __forceinline__ uint64_t kvm_ld1(const uint64_t* p) {
  unsigned long long r;
  asm volatile("ld.global.nc.L1::no_allocate.u64 %0, [%1];" : "=l"(r) : "l"(p));
  return (uint64_t)r;
}
__device__ __forceinline__ uint16_t kvm_ld_u8x2(const uint8_t* p) {
  unsigned short r;
  asm volatile("ld.global.nc.L1::no_allocate.u16 %0, [%1];" : "=h"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ uint8_t kvm_ld_u8(const uint8_t* p) {
  unsigned short r;
  asm volatile("ld.global.nc.L1::no_allocate.u8 %0, [%1];" : "=h"(r) : "l"(p));
  return (uint8_t)r;
}

// 256-bit streaming loads (sm_100: LDG.256): four 8-byte elements per thread, not allocated in
// L1.  32-byte aligned addresses only.  Measured on B200 (tools/probe/bw_probe.cu, two 8-byte
// columns + 2 B/row stores, 1e8 rows): plain 256-bit loads 5.8 TB/s; the same loads marked
// L2::evict_first 3.9 TB/s - so the evict-first form below is NOT used by the streaming kernels.
struct kvm_u64x4 {
  unsigned long long x, y, z, w;
};
__device__ __forceinline__ kvm_u64x4 kvm_ld4_stream(const void* p) {
  kvm_u64x4 r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v4.b64 {%0,%1,%2,%3}, [%4];"
               : "=l"(r.x), "=l"(r.y), "=l"(r.z), "=l"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ kvm_u64x4 kvm_ld4_nc(const void* p) {
  kvm_u64x4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.b64 {%0,%1,%2,%3}, [%4];"
               : "=l"(r.x), "=l"(r.y), "=l"(r.z), "=l"(r.w)
               : "l"(p));
  return r;
}
// plain 256-bit load (data another kernel has just written and that should be served from L2)
__device__ __forceinline__ kvm_u64x4 kvm_ld4(const void* p) {
  kvm_u64x4 r;
  asm volatile("ld.global.v4.b64 {%0,%1,%2,%3}, [%4];"
               : "=l"(r.x), "=l"(r.y), "=l"(r.z), "=l"(r.w)
               : "l"(p));
  return r;
}
// L2 "keep" policy for small by-products that the next kernel reads back (evict-last)
__device__ __forceinline__ unsigned long long kvm_policy_evict_last() {
  unsigned long long pol;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void kvm_st4_u32_keep(void* p, uint4 v, unsigned long long pol) {
  asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w), "l"(pol)
               : "memory");
}
// programmatic dependent launch (PDL): the dependent kernel may be made resident while this one
// still runs / wait until every kernel this one depends on has completed and its writes are visible
__device__ __forceinline__ void kvm_prefetch_l2(const void* p) {
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}
__device__ __forceinline__ void kvm_prefetch_l1(const void* p) {
  asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
}
__device__ __forceinline__ void kvm_pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;"); }
__device__ __forceinline__ void kvm_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// ------------------------------------------------------------------ warp reductions (64-bit)
__device__ __forceinline__ int64_t kvm_shfl_xor(int64_t v, int m) {
  return (int64_t)__shfl_xor_sync(KVM_FULL, (long long)v, m);
}
__device__ __forceinline__ uint64_t kvm_shfl_xor(uint64_t v, int m) {
  return (uint64_t)__shfl_xor_sync(KVM_FULL, (unsigned long long)v, m);
}
__device__ __forceinline__ double kvm_shfl_xor(double v, int m) {
  return __shfl_xor_sync(KVM_FULL, v, m);
}
__device__ __forceinline__ int64_t kvm_warp_sum(int64_t v) {
#pragma unroll
  for (int m = 16; m; m >>= 1) v += kvm_shfl_xor(v, m);
  return v;
}
__device__ __forceinline__ uint64_t kvm_warp_min(uint64_t v) {
#pragma unroll
  for (int m = 16; m; m >>= 1) {
    uint64_t o = kvm_shfl_xor(v, m);
    v = o < v ? o : v;
  }
  return v;
}
__device__ __forceinline__ uint64_t kvm_warp_max(uint64_t v) {
#pragma unroll
  for (int m = 16; m; m >>= 1) {
    uint64_t o = kvm_shfl_xor(v, m);
    v = o > v ? o : v;
  }
  return v;
}

// ------------------------------------------------------------------ double-double
// Error-free transformations; compiled with -fmad=false so nothing here is contracted.
struct kvm_dd {
  double hi, lo;
};
__host__ __device__ __forceinline__ kvm_dd kvm_two_sum(double a, double b) {
  double s = a + b;
  double bb = s - a;
  double e = (a - (s - bb)) + (b - bb);
  return {s, e};
}
__host__ __device__ __forceinline__ kvm_dd kvm_dd_add(kvm_dd a, double b) {
  kvm_dd s = kvm_two_sum(a.hi, b);
  double lo = s.lo + a.lo;
  double hi = s.hi + lo;
  lo = lo - (hi - s.hi);
  return {hi, lo};
}
__host__ __device__ __forceinline__ kvm_dd kvm_dd_add(kvm_dd a, kvm_dd b) {
  kvm_dd s = kvm_two_sum(a.hi, b.hi);
  kvm_dd t = kvm_two_sum(a.lo, b.lo);
  double c = s.lo + t.hi;
  double vh = s.hi + c;
  double vl = c - (vh - s.hi);
  c = t.lo + vl;
  double hi = vh + c;
  double lo = c - (hi - vh);
  return {hi, lo};
}
__host__ __device__ __forceinline__ kvm_dd kvm_dd_sub(kvm_dd a, kvm_dd b) {
  return kvm_dd_add(a, kvm_dd{-b.hi, -b.lo});
}
__device__ __forceinline__ kvm_dd kvm_warp_sum(kvm_dd v) {
#pragma unroll
  for (int m = 16; m; m >>= 1) {
    kvm_dd o{kvm_shfl_xor(v.hi, m), kvm_shfl_xor(v.lo, m)};
    // both lanes of a pair must add in the same operand order to stay bit-identical
    kvm_dd a = (threadIdx.x & m) ? o : v;
    kvm_dd b = (threadIdx.x & m) ? v : o;
    v = kvm_dd_add(a, b);
  }
  return v;
}

// ------------------------------------------------------------------ grid sizing
// Persistent-style grids: ctas_per_sm * SM count, capped by the amount of work.
static inline int kvm_grid_for(int64_t items, int items_per_cta, int ctas_per_sm) {
  int64_t want = (items + items_per_cta - 1) / items_per_cta;
  int64_t cap = (int64_t)kvm_num_sms() * ctas_per_sm;
  if (want < 1) want = 1;
  return (int)(want < cap ? want : cap);
}
// Grid of a streaming kernel that fills the machine with two CTAs per SM, minus `kvm_tail_slots()`
// CTAs: that many SMs then hold ONE of its CTAs and keep half their registers and shared memory
// free, which is where the few-CTA kernels of ANOTHER analysis in flight (plan, resolve, finish: 512
// threads, 64 registers) run at the same time instead of waiting for the whole grid to drain.
int kvm_tail_slots();
static inline int kvm_stream_grid(int64_t items, int items_per_cta) {
  const int full = 2 * kvm_num_sms();
  int g = kvm_grid_for(items, items_per_cta, 2);
  if (g == full && full - kvm_tail_slots() >= full / 2) g = full - kvm_tail_slots();
  return g;
}
