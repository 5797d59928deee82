// bw_probe.cu - which load form / work partition streams two 8-byte columns fastest on this GPU?
// (measurement tool; results are summarised under profiles/)
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
struct u4 { unsigned long long x, y, z, w; };
__device__ __forceinline__ u4 ld256_ef(const void* p) { u4 r; asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v4.b64 {%0,%1,%2,%3}, [%4];" : "=l"(r.x), "=l"(r.y), "=l"(r.z), "=l"(r.w) : "l"(p)); return r; }
__device__ __forceinline__ u4 ld256(const void* p) { u4 r; asm volatile("ld.global.nc.L1::no_allocate.v4.b64 {%0,%1,%2,%3}, [%4];" : "=l"(r.x), "=l"(r.y), "=l"(r.z), "=l"(r.w) : "l"(p)); return r; }
__device__ __forceinline__ void ld128(const void* p, unsigned long long& a, unsigned long long& b) { asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0,%1}, [%2];" : "=l"(a), "=l"(b) : "l"(p)); }
__device__ __forceinline__ void st_keep(void* p, uint4 v, unsigned long long pol) { asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(pol) : "memory"); }
// LOAD: 0 = 2x128-bit per column (4 rec/iter), 1 = 2x256 plain (8 rec), 2 = 2x256 evict_first (8 rec)
// PART: 0 = grid-stride, 1 = contiguous range per CTA.  STORE: 0 none, 1 plain, 2 keep policy
template <int LOAD, int PART, int STORE>
__global__ void __launch_bounds__(512, 2) k(const double* __restrict__ v, const long long* __restrict__ s, long long n, unsigned short* __restrict__ rec, unsigned long long* out) {
  constexpr int R = LOAD == 0 ? 4 : 8;
  const long long units = n / R;
  long long u, uend, step;
  if (PART == 0) { u = (long long)blockIdx.x * 512 + threadIdx.x; uend = units; step = (long long)gridDim.x * 512; }
  else { const long long chunk = (units + gridDim.x - 1) / gridDim.x; u = blockIdx.x * chunk + threadIdx.x; uend = min(units, (blockIdx.x + 1) * chunk); step = 512; }
  unsigned long long acc = 0, pol = 0;
  if (STORE == 2) asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  for (; u < uend; u += step) {
    const long long i = u * R;
    unsigned long long a[8], b[8];
    if (LOAD == 0) { ld128(v + i, a[0], a[1]); ld128(v + i + 2, a[2], a[3]); ld128(s + i, b[0], b[1]); ld128(s + i + 2, b[2], b[3]); }
    else {
      u4 x0 = LOAD == 2 ? ld256_ef(v + i) : ld256(v + i), x1 = LOAD == 2 ? ld256_ef(v + i + 4) : ld256(v + i + 4);
      u4 y0 = LOAD == 2 ? ld256_ef(s + i) : ld256(s + i), y1 = LOAD == 2 ? ld256_ef(s + i + 4) : ld256(s + i + 4);
      a[0]=x0.x;a[1]=x0.y;a[2]=x0.z;a[3]=x0.w;a[4]=x1.x;a[5]=x1.y;a[6]=x1.z;a[7]=x1.w;
      b[0]=y0.x;b[1]=y0.y;b[2]=y0.z;b[3]=y0.w;b[4]=y1.x;b[5]=y1.y;b[6]=y1.z;b[7]=y1.w;
    }
    unsigned d[8];
#pragma unroll
    for (int j = 0; j < R; ++j) { d[j] = b[j] == 200ull ? (unsigned)(a[j] >> 40) & 0x1fffu : 0xffffu; acc += d[j]; }
    if (STORE) {
      if (R == 4) { uint2 o; o.x = d[0] | d[1] << 16; o.y = d[2] | d[3] << 16; *reinterpret_cast<uint2*>(rec + i) = o; }
      else { uint4 o; o.x = d[0] | d[1] << 16; o.y = d[2] | d[3] << 16; o.z = d[4] | d[5] << 16; o.w = d[6] | d[7] << 16;
             if (STORE == 2) st_keep(rec + i, o, pol); else *reinterpret_cast<uint4*>(rec + i) = o; }
    }
  }
  if (acc == 0x123456789ull) out[0] = acc;
}
template <int L, int P, int S> void run(const char* name, const double* v, const long long* s, long long n, unsigned short* rec, unsigned long long* out, int grid) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 3; ++i) k<L, P, S><<<grid, 512>>>(v, s, n, rec, out);
  cudaEventRecord(e0);
  const int reps = 20;
  for (int i = 0; i < reps; ++i) k<L, P, S><<<grid, 512>>>(v, s, n, rec, out);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= reps;
  printf("{\"variant\": \"%s\", \"n\": %lld, \"grid\": %d, \"us\": %.2f, \"alg_gbs\": %.1f, \"err\": \"%s\"}\n", name, n, grid, ms * 1e3, 16.0 * n / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
}
int main(int argc, char** argv) {
  for (long long n : {10000000ll, 100000000ll}) {
    double* v; long long* s; unsigned short* rec; unsigned long long* out;
    cudaMalloc(&v, n * 8); cudaMalloc(&s, n * 8); cudaMalloc(&rec, n * 2 + 64); cudaMalloc(&out, 8);
    cudaMemset(v, 0x3f, n * 8); cudaMemset(s, 0, n * 8);
    for (int grid : {296, 592}) {
      run<0, 0, 0>("ld128 stride nostore", v, s, n, rec, out, grid);
      run<0, 0, 1>("ld128 stride store", v, s, n, rec, out, grid);
      run<1, 0, 1>("ld256 stride store", v, s, n, rec, out, grid);
      run<2, 0, 1>("ld256ef stride store", v, s, n, rec, out, grid);
      run<2, 0, 2>("ld256ef stride keepstore", v, s, n, rec, out, grid);
      run<0, 1, 1>("ld128 contig store", v, s, n, rec, out, grid);
      run<2, 1, 1>("ld256ef contig store", v, s, n, rec, out, grid);
      run<2, 1, 2>("ld256ef contig keepstore", v, s, n, rec, out, grid);
      run<1, 1, 1>("ld256 contig store", v, s, n, rec, out, grid);
    }
    cudaFree(v); cudaFree(s); cudaFree(rec); cudaFree(out);
  }
  return 0;
}
