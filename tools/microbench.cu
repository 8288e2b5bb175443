// microbench.cu -- B200 pipe-rate probes that size the fused ELBO kernels:
//   mma.sync.m16n8k8 TF32 issue rate per SM (legacy tensor path), FFMA rate, and the mix used by
//   the 3xTF32 split (MMA + LOP/IADD/FADD split ops).  Build: nvcc -gencode arch=compute_100a,code=sm_100a
//   -O3 -o gpurun_out/microbench tools/microbench.cu ; run on the GPU box.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

// NACC independent accumulators per warp, ITER rounds
template <int NACC>
__global__ void k_mma(float* out, int iters, long long* cyc) {
  float acc[NACC][4];
  uint32_t a[4], b[2];
  for (int i = 0; i < 4; ++i) a[i] = __float_as_uint(1.0f + threadIdx.x * 1e-3f + i);
  for (int i = 0; i < 2; ++i) b[i] = __float_as_uint(0.5f + threadIdx.x * 1e-3f + i);
  for (int n = 0; n < NACC; ++n)
    for (int i = 0; i < 4; ++i) acc[n][i] = 0.f;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int n = 0; n < NACC; ++n) mma_tf32(acc[n], a, b);
  }
  long long t1 = clock64();
  float s = 0.f;
  for (int n = 0; n < NACC; ++n)
    for (int i = 0; i < 4; ++i) s += acc[n][i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

// 3xTF32 pattern: per B fragment 2 values split (IADD+LOP+FADD each), then 3 MMAs per m-tile
template <int MT>
__global__ void k_mma3(float* out, const float* in, int iters, long long* cyc) {
  float accM[MT][4][4], accC[MT][4][4];
  uint32_t ahi[MT][4], alo[MT][4];
  for (int m = 0; m < MT; ++m)
    for (int i = 0; i < 4; ++i) {
      float v = 1.0f + threadIdx.x * 1e-3f + i + m;
      uint32_t h = (__float_as_uint(v) + 0x1000u) & 0xffffe000u;
      ahi[m][i] = h;
      alo[m][i] = __float_as_uint(v - __uint_as_float(h));
    }
  for (int m = 0; m < MT; ++m)
    for (int n = 0; n < 4; ++n)
      for (int i = 0; i < 4; ++i) accM[m][n][i] = accC[m][n][i] = 0.f;
  __shared__ float sm[8 * 132];
  for (int i = threadIdx.x; i < 8 * 132; i += blockDim.x) sm[i] = in[i % 64] + i;
  __syncthreads();
  const int g = (threadIdx.x & 31) >> 2, t4 = threadIdx.x & 3;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int n = 0; n < 4; ++n) {
      float b0 = sm[(2 * t4) * 132 + n * 8 + g + (it & 3)];
      float b1 = sm[(2 * t4 + 1) * 132 + n * 8 + g + (it & 3)];
      uint32_t bh[2], bl[2];
      bh[0] = (__float_as_uint(b0) + 0x1000u) & 0xffffe000u;
      bh[1] = (__float_as_uint(b1) + 0x1000u) & 0xffffe000u;
      bl[0] = __float_as_uint(b0 - __uint_as_float(bh[0]));
      bl[1] = __float_as_uint(b1 - __uint_as_float(bh[1]));
#pragma unroll
      for (int m = 0; m < MT; ++m) {
        mma_tf32(accC[m][n], alo[m], bh);
        mma_tf32(accC[m][n], ahi[m], bl);
        mma_tf32(accM[m][n], ahi[m], bh);
      }
    }
  }
  long long t1 = clock64();
  float s = 0.f;
  for (int m = 0; m < MT; ++m)
    for (int n = 0; n < 4; ++n)
      for (int i = 0; i < 4; ++i) s += accM[m][n][i] + accC[m][n][i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int NACC>
__global__ void k_ffma(float* out, int iters, long long* cyc) {
  float acc[NACC];
  float a = 1.0f + threadIdx.x * 1e-6f, b = 0.999f;
  for (int n = 0; n < NACC; ++n) acc[n] = n;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int n = 0; n < NACC; ++n) acc[n] = fmaf(acc[n], a, b);
  }
  long long t1 = clock64();
  float s = 0.f;
  for (int n = 0; n < NACC; ++n) s += acc[n];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

template <typename F>
static int run(const char* name, F launch, int threads, int blocks_per_sm, double macs_per_warp_iter, int iters) {
  int nsm = 148;
  float* out; long long* cyc;
  CK(cudaMalloc(&out, (size_t)nsm * blocks_per_sm * threads * 4));
  CK(cudaMalloc(&cyc, (size_t)nsm * blocks_per_sm * 8));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  launch(nsm * blocks_per_sm, threads, out, iters, cyc);   // warm-up
  CK(cudaDeviceSynchronize());
  cudaEventRecord(e0);
  launch(nsm * blocks_per_sm, threads, out, iters, cyc);
  cudaEventRecord(e1);
  CK(cudaDeviceSynchronize());
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long h[4];
  CK(cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost));
  double warps = threads / 32.0 * blocks_per_sm;
  double mac_per_clk_sm = macs_per_warp_iter * iters * warps / (double)h[0];
  double total = macs_per_warp_iter * iters * warps * nsm;
  printf("%-34s thr=%4d cta/sm=%d  cycles=%lld  %.1f MAC/clk/SM  %.1f TMAC/s (%.3f ms, eff clk %.0f MHz)\n", name, threads,
         blocks_per_sm, h[0], mac_per_clk_sm, total / (ms * 1e-3) / 1e12, ms, h[0] / (ms * 1e-3) / 1e6);
  cudaFree(out); cudaFree(cyc);
  return 0;
}

int main() {
  const int iters = 20000;
  float* in; cudaMalloc(&in, 64 * 4); cudaMemset(in, 0, 64 * 4);
  for (int thr : {128, 256, 512, 1024}) {
    run("mma.sync tf32 m16n8k8 x8 acc", [&](int g, int t, float* o, int it, long long* c) { k_mma<8><<<g, t>>>(o, it, c); }, thr, 1, 8 * 1024.0, iters);
  }
  run("mma.sync tf32 m16n8k8 x16 acc", [&](int g, int t, float* o, int it, long long* c) { k_mma<16><<<g, t>>>(o, it, c); }, 256, 1, 16 * 1024.0, iters);
  run("mma.sync tf32 m16n8k8 x2 acc", [&](int g, int t, float* o, int it, long long* c) { k_mma<2><<<g, t>>>(o, it, c); }, 256, 1, 2 * 1024.0, iters);
  run("mma.sync tf32 x8 acc, 2 cta/sm", [&](int g, int t, float* o, int it, long long* c) { k_mma<8><<<g, t>>>(o, it, c); }, 256, 2, 8 * 1024.0, iters);
  for (int thr : {256, 512}) {
    run("3xTF32 pattern MT=2 NT=4 (LDS+split)", [&](int g, int t, float* o, int it, long long* c) { k_mma3<2><<<g, t>>>(o, in, it, c); }, thr, 1, 2 * 4 * 3 * 1024.0, iters / 4);
    run("3xTF32 pattern MT=1 NT=4 (LDS+split)", [&](int g, int t, float* o, int it, long long* c) { k_mma3<1><<<g, t>>>(o, in, it, c); }, thr, 1, 1 * 4 * 3 * 1024.0, iters / 4);
    run("3xTF32 pattern MT=4 NT=4 (LDS+split)", [&](int g, int t, float* o, int it, long long* c) { k_mma3<4><<<g, t>>>(o, in, it, c); }, thr > 256 ? 256 : thr, 1, 4 * 4 * 3 * 1024.0, iters / 4);
  }
  for (int thr : {256, 512, 1024}) {
    run("FFMA x16 acc", [&](int g, int t, float* o, int it, long long* c) { k_ffma<16><<<g, t>>>(o, it, c); }, thr, 1, 16 * 32.0, iters);
  }
  return 0;
}
