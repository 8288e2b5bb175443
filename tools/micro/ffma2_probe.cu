// FFMA vs FFMA2 (fma.rn.f32x2) issue/throughput probe on sm_100a, alone and interleaved with broadcast LDS.128.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2_probe ffma2_probe.cu && ./ffma2_probe
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ float fma1(float a, float b, float c) { float r; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }

template <int MODE>   // 0: 16 FFMA chains; 1: 8 FFMA2 chains (=16 lanes); 2/3: same + one LDS.128 per 8 FMA-lanes
__global__ void __launch_bounds__(512) k(float* out, int iters) {
  __shared__ __align__(16) float sh[1024];
  for (int i = threadIdx.x; i < 1024; i += 512) sh[i] = 1.0f + i * 1e-6f;
  __syncthreads();
  float a[16];
  u64 p[8];
  for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 1e-3f + i;
  for (int i = 0; i < 8; ++i) p[i] = pk(a[2 * i], a[2 * i + 1]);
  float w = 0.999f; u64 w2 = pk(w, w);
  for (int it = 0; it < iters; ++it) {
    if (MODE >= 2) {
      const float4 v = *reinterpret_cast<const float4*>(sh + ((it * 4) & 1020));
      const float4 v2 = *reinterpret_cast<const float4*>(sh + ((it * 4 + 512) & 1020));
      w = v.x + v2.y; w2 = pk(v.z, v2.w);
    }
    if (MODE == 0 || MODE == 2) {
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int i = 0; i < 16; ++i) a[i] = fma1(a[i], w, 0.5f);
    } else {
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int i = 0; i < 8; ++i) p[i] = fma2(p[i], w2, w2);
    }
  }
  float s = 0;
  for (int i = 0; i < 16; ++i) s += a[i];
  for (int i = 0; i < 8; ++i) { float lo, hi; asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p[i])); s += lo + hi; }
  out[blockIdx.x * 512 + threadIdx.x] = s;
}
template <int MODE> void run(const char* name, float* out) {
  const int iters = 20000, grid = 148 * 2;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<grid, 512>>>(out, 100);
  cudaEventRecord(e0); k<MODE><<<grid, 512>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double fma = (double)grid * 512 * iters * 64;
  printf("%-28s %8.3f ms  %7.2f TFMA/s  (%.1f FMA/clk/SM at 1.9 GHz)\n", name, ms, fma / ms * 1e-9, fma / (ms * 1e-3) / 148 / 1.9e9);
}
int main() {
  float* out; cudaMalloc(&out, 148 * 2 * 512 * 4);
  run<0>("FFMA", out); run<1>("FFMA2", out); run<2>("FFMA + 2 LDS.128 / 64", out); run<3>("FFMA2 + 2 LDS.128 / 64", out);
  return 0;
}
