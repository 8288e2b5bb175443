// tc5_probe.cu -- standalone probe of the tcgen05 building blocks the fused decoder kernel uses:
//   * no-swizzle ("interleave") shared-memory operand descriptors, K-major and MN-major, kind::tf32
//   * a conv tap as a +16-byte shift of the descriptor start address (act[c/4][t][4] layout)
//   * TMEM allocation, tcgen05.commit -> mbarrier, tcgen05.ld epilogue (lane = row, column = n)
//   * rounding of the fp32 accumulate inside the tensor core
//   * issue rate of M=128 N=64 K=8 MMAs fed from shared memory
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/bin/tc5_probe tools/tc5_probe.cu
// Every test checks itself against a CPU evaluation of the same products (inputs are small
// integers, exact in TF32) and prints PASS/FAIL.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(2); } } while (0)

struct Job {
  uint32_t a_off, b_off;          // byte offsets of the operand starts in the smem image
  uint32_t a_lbo, a_sbo, b_lbo, b_sbo;
  uint32_t a_kstep, b_kstep;      // byte advance per K=8 MMA
  uint32_t idesc;
  int nk, reps, N;
  uint32_t image_bytes;
  uint32_t a_layout, b_layout;    // descriptor layout_type (0 none, 2 = 128B swizzle); base_offset is derived from the address
  int nacc;                       // MMAs rotate over this many accumulators (0/1: one)
  int kind;                       // 0: kind::tf32, 1: kind::f16 (bf16 operands)
};

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;   // descriptor version 1 (Blackwell); layout_type 0 = no swizzle
  return d;
}
__device__ __forceinline__ uint64_t make_desc_sw(uint32_t saddr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
  uint64_t d = make_desc(saddr, lbo, sbo);
  if (layout) { d |= (uint64_t)layout << 61; d |= (uint64_t)((saddr >> 7) & 7) << 49; }
  return d;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mma_tf32(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n"
      :: "r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}

__device__ __forceinline__ void mma_f16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
      :: "r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}

__device__ __forceinline__ bool mbar_wait(uint32_t bar, uint32_t parity) {
  for (int it = 0; it < (1 << 22); ++it) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    if (ok) return true;
  }
  return false;
}

// one CTA of 128 threads: thread 0 issues, all 4 warps read back their 32 TMEM lanes
__global__ void __launch_bounds__(128, 1) probe_kernel(Job job, const uint8_t* __restrict__ image, float* __restrict__ D,
                                                       long long* __restrict__ cycles, int* __restrict__ status) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base_sh;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (uint32_t i = tid * 16; i < job.image_bytes; i += 128 * 16)
    *reinterpret_cast<uint4*>(smem + i) = *reinterpret_cast<const uint4*>(image + i);
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(smem_u32(&tmem_base_sh)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // generic-proxy writes of the operands must be visible to the async (tensor core) proxy
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_base_sh;
  long long t0 = 0, t1 = 0;
  if (warp == 0) {
    // the whole warp runs the (uniform) descriptor arithmetic so that it stays on the uniform datapath;
    // one elected lane issues
    const uint32_t sb = smem_u32(smem);
    const uint64_t da0 = make_desc_sw(sb + job.a_off, job.a_lbo, job.a_sbo, job.a_layout), db0 = make_desc_sw(sb + job.b_off, job.b_lbo, job.b_sbo, job.b_layout);
    const uint32_t ak = job.a_kstep >> 4, bk = job.b_kstep >> 4, idesc = job.idesc;
    const int nk = job.nk, reps = job.reps, nacc = job.nacc > 1 ? job.nacc : 1;
    const uint32_t N = job.N;
    const bool f16 = job.kind != 0;
    uint32_t elected;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(elected));
    t0 = clock64();
    uint32_t first = 0;
    for (int r = 0; r < reps; ++r) {
      uint32_t acc_i = 0;
#pragma unroll 8
      for (int k = 0; k < nk; ++k) {
        const uint64_t da = da0 + (uint64_t)(k * ak), db = db0 + (uint64_t)(k * bk);
        const uint32_t td = tmem + acc_i * N;
        const uint32_t acc = (r > 0 || k >= nacc) ? 1u : 0u;
        if (elected) { if (f16) mma_f16(td, da, db, idesc, acc); else mma_tf32(td, da, db, idesc, acc); }
        acc_i = (acc_i + 1 == (uint32_t)nacc) ? 0 : acc_i + 1;
      }
    }
    (void)first;
    if (elected) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(&bar)) : "memory");
    __syncwarp();
  }
  // only the issuing thread polls the mbarrier (127 spinning threads would steal shared-memory
  // bandwidth from the tensor core's operand reads); everybody else blocks in bar.sync
  __shared__ int ok_sh;
  if (tid == 0) {
    const bool okw = mbar_wait(smem_u32(&bar), 0);
    t1 = clock64(); cycles[0] = t1 - t0;
    ok_sh = okw ? 1 : 0;
    if (!okw) atomicExch(status, 1);
  }
  __syncthreads();
  const bool ok = ok_sh != 0;
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (ok) {
    for (int c0 = 0; c0 < job.N; c0 += 8) {
      uint32_t v[8];
      const uint32_t addr = tmem + ((uint32_t)(warp * 32) << 16) + c0;
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                   : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                   : "r"(addr));
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      for (int j = 0; j < 8; ++j) D[(size_t)tid * job.N + c0 + j] = __uint_as_float(v[j]);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tmem) : "memory");
}

// ---- host ---------------------------------------------------------------------------------------
static uint32_t idesc_tf32(int M, int N, int a_mn, int b_mn, int a_neg = 0) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_neg << 13) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// element (ch, row) of a [ch/4][rows][4] fp32 operand buffer
static void put(std::vector<uint8_t>& im, uint32_t base, uint32_t plane, int ch, int row, float v) {
  const size_t o = (size_t)base + (size_t)(ch / 4) * plane + (size_t)row * 16 + (size_t)(ch % 4) * 4;
  if (o + 4 > im.size()) { printf("image overflow\n"); exit(3); }
  memcpy(&im[o], &v, 4);
}
static uint8_t* d_image; static float* d_D; static long long* d_cyc; static int* d_status;
static int run(Job job, const std::vector<uint8_t>& im, std::vector<float>& D, long long* cyc) {
  job.image_bytes = (uint32_t)im.size();
  CK(cudaMemcpy(d_image, im.data(), im.size(), cudaMemcpyHostToDevice));
  CK(cudaMemset(d_status, 0, 4));
  CK(cudaMemset(d_D, 0, 128 * 256 * 4));
  probe_kernel<<<1, 128, im.size()>>>(job, d_image, d_D, d_cyc, d_status);
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());
  int st = 0;
  CK(cudaMemcpy(&st, d_status, 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(cyc, d_cyc, 8, cudaMemcpyDeviceToHost));
  D.resize(128 * job.N);
  CK(cudaMemcpy(D.data(), d_D, D.size() * 4, cudaMemcpyDeviceToHost));
  return st;
}
static float ival(int a, int b, int salt) { return (float)(((a * 7 + b * 13 + salt * 5 + (a * b) % 11) % 9) - 4); }

static const int ROWS = 136;                 // rows of an activation plane
static const uint32_t PLANE = ROWS * 16;     // bytes of one 4-channel plane
static const uint32_t WPLANE = 64 * 16;      // weight plane: 64 rows x 4 k

static double check(const std::vector<float>& D, const std::vector<double>& ref, int N) {
  double e = 0;
  for (int m = 0; m < 128; ++m)
    for (int n = 0; n < N; ++n) e = fmax(e, fabs((double)D[m * N + n] - ref[m * N + n]));
  return e;
}

int main(int argc, char** argv) {
  const int only = argc > 1 ? atoi(argv[1]) : 0;      // run one test group (0 = all)
  const int nswap = argc > 2 ? atoi(argv[2]) + 1 : 1;  // 2: also try LBO/SBO swapped (may fault)
  CK(cudaSetDevice(0));
  CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  CK(cudaMalloc(&d_image, 200 * 1024)); CK(cudaMalloc(&d_D, 128 * 256 * 4)); CK(cudaMalloc(&d_cyc, 8)); CK(cudaMalloc(&d_status, 4));
  std::vector<float> D; long long cyc = 0; int nfail = 0;

  if (only == 0 || only == 1)
  // ---- test 1: forward conv tap.  D[t][o] = sum_c act[c][t+shift] * W[o][c]; A K-major (rows = t), B K-major
  for (int swap = 0; swap < nswap; ++swap)
    for (int shift = 0; shift < 3; ++shift) {
      std::vector<uint8_t> im(16 * PLANE + 16 * WPLANE, 0);
      const uint32_t wb = 16 * PLANE;
      for (int c = 0; c < 64; ++c) for (int r = 0; r < ROWS; ++r) put(im, 0, PLANE, c, r, ival(c, r, 1));
      for (int c = 0; c < 64; ++c) for (int o = 0; o < 64; ++o) put(im, wb, WPLANE, c, o, ival(o, c, 2));
      Job j{}; j.a_off = shift * 16; j.b_off = wb; j.a_kstep = 2 * PLANE; j.b_kstep = 2 * WPLANE; j.nk = 8; j.reps = 1; j.N = 64;
      j.a_lbo = swap ? 128 : PLANE; j.a_sbo = swap ? PLANE : 128; j.b_lbo = swap ? 128 : WPLANE; j.b_sbo = swap ? WPLANE : 128;
      j.idesc = idesc_tf32(128, 64, 0, 0);
      const int st = run(j, im, D, &cyc);
      std::vector<double> ref(128 * 64, 0.0);
      for (int t = 0; t < 128; ++t) for (int o = 0; o < 64; ++o) { double s = 0; for (int c = 0; c < 64; ++c) s += (double)ival(c, t + shift, 1) * ival(o, c, 2); ref[t * 64 + o] = s; }
      const double e = check(D, ref, 64);
      printf("test1 K-major fwd tap shift=%d lbo/sbo-swapped=%d: status=%d maxerr=%g cycles=%lld %s\n", shift, swap, st, e, cyc, (st == 0 && e == 0) ? "PASS" : "FAIL");
      if (!swap && !(st == 0 && e == 0)) ++nfail;
    }

  if (only == 0 || only == 2)
  // ---- test 2: weight gradient.  D[o][c] = sum_t dg[o][t] * g1[c][t+shift]; A and B MN-major, K = time
  for (int swap = 0; swap < nswap; ++swap)
    for (int shift = 0; shift < 3; ++shift) {
      std::vector<uint8_t> im(48 * PLANE, 0);
      const uint32_t bb = 32 * PLANE;
      for (int o = 0; o < 128; ++o) for (int r = 0; r < ROWS; ++r) put(im, 0, PLANE, o, r, ival(o, r, 3));
      for (int c = 0; c < 64; ++c) for (int r = 0; r < ROWS; ++r) put(im, bb, PLANE, c, r, ival(c, r, 4));
      Job j{}; j.a_off = 0; j.b_off = bb + shift * 16; j.a_kstep = 128; j.b_kstep = 128; j.nk = 16; j.reps = 1; j.N = 64;
      j.a_lbo = swap ? PLANE : 128; j.a_sbo = swap ? 128 : PLANE; j.b_lbo = j.a_lbo; j.b_sbo = j.a_sbo;
      j.idesc = idesc_tf32(128, 64, 1, 1);
      const int st = run(j, im, D, &cyc);
      std::vector<double> ref(128 * 64, 0.0);
      for (int o = 0; o < 128; ++o) for (int c = 0; c < 64; ++c) { double s = 0; for (int t = 0; t < 128; ++t) s += (double)ival(o, t, 3) * ival(c, t + shift, 4); ref[o * 64 + c] = s; }
      const double e = check(D, ref, 64);
      printf("test2 MN-major wgrad shift=%d lbo/sbo-swapped=%d: status=%d maxerr=%g cycles=%lld %s\n", shift, swap, st, e, cyc, (st == 0 && e == 0) ? "PASS" : "FAIL");
      if (!swap && !(st == 0 && e == 0)) ++nfail;
    }

  if (only == 0 || only == 3)
  // ---- test 3: input gradient.  D[t][c] = sum_o dg2[o][t+shift] * W[o][c]; A K-major, B = the forward weight buffer read MN-major
  for (int swap = 0; swap < nswap; ++swap)
    for (int shift = 0; shift < 3; ++shift) {
      std::vector<uint8_t> im(16 * PLANE + 16 * WPLANE, 0);
      const uint32_t wb = 16 * PLANE;
      for (int o = 0; o < 64; ++o) for (int r = 0; r < ROWS; ++r) put(im, 0, PLANE, o, r, ival(o, r, 5));
      for (int c = 0; c < 64; ++c) for (int o = 0; o < 64; ++o) put(im, wb, WPLANE, c, o, ival(o, c, 6));
      Job j{}; j.a_off = shift * 16; j.b_off = wb; j.a_kstep = 2 * PLANE; j.b_kstep = 128; j.nk = 8; j.reps = 1; j.N = 64;
      j.a_lbo = PLANE; j.a_sbo = 128; j.b_lbo = swap ? WPLANE : 128; j.b_sbo = swap ? 128 : WPLANE;
      j.idesc = idesc_tf32(128, 64, 0, 1);
      const int st = run(j, im, D, &cyc);
      std::vector<double> ref(128 * 64, 0.0);
      for (int t = 0; t < 128; ++t) for (int c = 0; c < 64; ++c) { double s = 0; for (int o = 0; o < 64; ++o) s += (double)ival(o, t + shift, 5) * ival(o, c, 6); ref[t * 64 + c] = s; }
      const double e = check(D, ref, 64);
      printf("test3 dgrad (B MN-major) shift=%d swapped=%d: status=%d maxerr=%g cycles=%lld %s\n", shift, swap, st, e, cyc, (st == 0 && e == 0) ? "PASS" : "FAIL");
      if (!swap && !(st == 0 && e == 0)) ++nfail;
    }

  if (only == 0 || only == 4)
  // ---- test 4: rounding of the accumulate.  D = 1 + sum of 16 addends s*f*2^-23 (one per MMA, or 8 equal parts per MMA)
  for (int variant = 0; variant < 2; ++variant) {
    const int NKR = 17;
    std::vector<uint8_t> im(2 * NKR * PLANE + 2 * NKR * WPLANE, 0);
    const uint32_t wb = 2 * NKR * PLANE;
    for (int r = 0; r < ROWS; ++r) {
      put(im, 0, PLANE, 0, r, 1.f);
      const float delta = ldexpf((float)(r % 16 + 1) / 8.f, -23) / (variant ? 8.f : 1.f);
      for (int i = 1; i < NKR; ++i) for (int jj = 0; jj < (variant ? 8 : 1); ++jj) put(im, 0, PLANE, 8 * i + jj, r, delta);
    }
    for (int o = 0; o < 64; ++o) {
      put(im, wb, WPLANE, 0, o, 1.f);
      for (int i = 1; i < NKR; ++i) for (int jj = 0; jj < (variant ? 8 : 1); ++jj) put(im, wb, WPLANE, 8 * i + jj, o, (o & 1) ? -1.f : 1.f);
    }
    Job j{}; j.a_off = 0; j.b_off = wb; j.a_kstep = 2 * PLANE; j.b_kstep = 2 * WPLANE; j.nk = NKR; j.reps = 1; j.N = 64;
    j.a_lbo = PLANE; j.a_sbo = 128; j.b_lbo = WPLANE; j.b_sbo = 128; j.idesc = idesc_tf32(128, 64, 0, 0);
    const int st = run(j, im, D, &cyc);
    printf("test4 rounding variant=%d (%s) status=%d: f, exact 16f, measured (D-1)*2^23 for +, for -\n", variant, variant ? "8 products per MMA" : "1 product per MMA", st);
    for (int t = 0; t < 16; ++t)
      printf("   f=%.3f exact=%+.2f  plus=%+.3f  minus=%+.3f\n", (t + 1) / 8.0, 16.0 * (t + 1) / 8.0,
             ((double)D[t * 64 + 0] - 1.0) * 8388608.0, ((double)D[t * 64 + 1] - 1.0) * 8388608.0);
  }

  if (only == 0 || only == 5)
  // ---- test 5: issue rate (MMAs are chained on one accumulator; operands re-read from shared memory)
  {
    std::vector<uint8_t> im(48 * PLANE, 0);
    Job j{}; j.N = 64; j.nk = 8; j.reps = 512;
    j.a_off = 0; j.b_off = 16 * PLANE; j.a_kstep = 2 * PLANE; j.b_kstep = 2 * WPLANE; j.a_lbo = PLANE; j.a_sbo = 128; j.b_lbo = WPLANE; j.b_sbo = 128;
    j.idesc = idesc_tf32(128, 64, 0, 0);
    int st = run(j, im, D, &cyc);
    printf("test5 rate K-major x K-major   M128 N64 K8: status=%d %d MMAs %lld cycles -> %.1f cyc/MMA, %.0f MAC/clk\n", st, j.nk * j.reps, cyc, (double)cyc / (j.nk * j.reps), 128.0 * 64 * 8 * j.nk * j.reps / cyc);
    j.a_kstep = 128; j.b_kstep = 128; j.nk = 16; j.reps = 256; j.a_lbo = 128; j.a_sbo = PLANE; j.b_lbo = 128; j.b_sbo = PLANE; j.b_off = 32 * PLANE; j.idesc = idesc_tf32(128, 64, 1, 1);
    st = run(j, im, D, &cyc);
    printf("test5 rate MN-major x MN-major M128 N64 K8: status=%d %d MMAs %lld cycles -> %.1f cyc/MMA, %.0f MAC/clk\n", st, j.nk * j.reps, cyc, (double)cyc / (j.nk * j.reps), 128.0 * 64 * 8 * j.nk * j.reps / cyc);
    j.a_kstep = 2 * PLANE; j.b_kstep = 128; j.nk = 8; j.reps = 512; j.a_lbo = PLANE; j.a_sbo = 128; j.b_lbo = 128; j.b_sbo = WPLANE; j.b_off = 16 * PLANE; j.idesc = idesc_tf32(128, 64, 0, 1);
    st = run(j, im, D, &cyc);
    printf("test5 rate K-major x MN-major  M128 N64 K8: status=%d %d MMAs %lld cycles -> %.1f cyc/MMA, %.0f MAC/clk\n", st, j.nk * j.reps, cyc, (double)cyc / (j.nk * j.reps), 128.0 * 64 * 8 * j.nk * j.reps / cyc);
  }

  if (only == 6) {
    // ---- test 6: which shared-memory word does an MN-major descriptor read for element (mn, k)?
    // the probed operand region holds its own word index; the other operand is an identity (K-major, verified by test 1)
    const uint32_t lbo = argc > 3 ? (uint32_t)atoi(argv[3]) : 128, sbo = argc > 4 ? (uint32_t)atoi(argv[4]) : PLANE;
    const int which = argc > 5 ? atoi(argv[5]) : 0;   // 0: probe A, 1: probe B
    const uint32_t RB = 180 * 1024;                  // probed region [0, RB) holds its word index, identity after it
    std::vector<float> lo, hi;
    for (int pass = 0; pass < 2; ++pass) {
      std::vector<uint8_t> im(RB + 2 * 2048, 0);
      for (uint32_t w = 0; w < RB / 4; ++w) { const float v = pass ? (float)(w >> 11) : (float)(w & 2047); memcpy(&im[4 * (size_t)w], &v, 4); }
      for (int k = 0; k < 8; ++k) put(im, RB, 2048, k, k, 1.f);   // identity: rows 0..7, [k/4][128 rows][4]
      Job j{}; j.nk = 1; j.reps = 1; j.N = 64; j.a_kstep = j.b_kstep = 0;
      if (which == 0 || which == 2) { j.a_off = 0; j.a_lbo = lbo; j.a_sbo = sbo; j.b_off = RB; j.b_lbo = 2048; j.b_sbo = 128; j.idesc = idesc_tf32(128, 64, which == 0, 0); }
      else            { j.b_off = 0; j.b_lbo = lbo; j.b_sbo = sbo; j.a_off = RB; j.a_lbo = 2048; j.a_sbo = 128; j.idesc = idesc_tf32(128, 64, 0, 1); }
      const int st = run(j, im, D, &cyc);
      printf("test6 pass %d status=%d cycles=%lld D[0..3]=%g %g %g %g D[64..67]=%g %g %g %g\n", pass, st, cyc, D[0], D[1], D[2], D[3], D[64], D[65], D[66], D[67]);
      (pass ? hi : lo) = D;
    }
    printf("test6 probe %s MN-major lbo=%u sbo=%u: byte address read for (mn, k=0..7)\n", which ? "B" : "A", lbo, sbo);
    const int mns[] = {0, 1, 2, 3, 4, 5, 7, 8, 9, 12, 16, 32, 63, 64, 127};
    for (int mi = 0; mi < 15; ++mi) {
      const int mn = mns[mi];
      if (which == 1 && mn >= 64) continue;
      printf("   mn=%3d:", mn);
      for (int k = 0; k < 8; ++k) {
        const float l = which != 1 ? lo[mn * 64 + k] : lo[k * 64 + mn], h = which != 1 ? hi[mn * 64 + k] : hi[k * 64 + mn];
        printf(" %6d", 4 * ((int)l + 2048 * (int)h));
      }
      printf("\n");
    }
  }

  if (only == 7) {
    // ---- test 7: same address discovery for bf16 operands (kind::f16, K = 16 per MMA); 16-byte rows hold 8 elements
    const uint32_t lbo = argc > 3 ? (uint32_t)atoi(argv[3]) : 128, sbo = argc > 4 ? (uint32_t)atoi(argv[4]) : PLANE;
    const int which = argc > 5 ? atoi(argv[5]) : 0;   // 0: A MN-major, 1: B MN-major, 2: A K-major
    const uint32_t RB = 180 * 1024;
    auto bf16 = [](float f) { uint32_t u; memcpy(&u, &f, 4); return (uint16_t)(u >> 16); };
    std::vector<float> res[3];
    for (int pass = 0; pass < 3; ++pass) {
      std::vector<uint8_t> im(RB + 2 * 2048, 0);
      for (uint32_t e = 0; e < RB / 2; ++e) { const uint16_t v = bf16((float)((e >> (8 * pass)) & 255)); memcpy(&im[2 * (size_t)e], &v, 2); }
      for (int k = 0; k < 16; ++k) { const uint16_t one = bf16(1.f); memcpy(&im[RB + (k / 8) * 2048 + k * 16 + (k % 8) * 2], &one, 2); }   // identity [k/8][128 rows][8]
      Job j{}; j.kind = 1; j.nk = 1; j.reps = 1; j.N = 64; j.a_kstep = j.b_kstep = 0;
      const uint32_t fmt = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      if (which != 1) { j.a_off = 0; j.a_lbo = lbo; j.a_sbo = sbo; j.b_off = RB; j.b_lbo = 2048; j.b_sbo = 128; j.idesc = fmt | ((which == 0 ? 1u : 0u) << 15); }
      else            { j.b_off = 0; j.b_lbo = lbo; j.b_sbo = sbo; j.a_off = RB; j.a_lbo = 2048; j.a_sbo = 128; j.idesc = fmt | (1u << 16); }
      const int st = run(j, im, D, &cyc);
      printf("test7 pass %d status=%d D[0..3]=%g %g %g %g\n", pass, st, D[0], D[1], D[2], D[3]);
      res[pass] = D;
    }
    printf("test7 bf16 probe %s lbo=%u sbo=%u: byte address read for (mn, k=0..15)\n", which == 0 ? "A MN-major" : which == 1 ? "B MN-major" : "A K-major", lbo, sbo);
    const int mns[] = {0, 1, 2, 3, 4, 7, 8, 9, 15, 16, 17, 32, 63, 64, 127};
    for (int mi = 0; mi < 15; ++mi) {
      const int mn = mns[mi];
      if (which == 1 && mn >= 64) continue;
      printf("   mn=%3d:", mn);
      for (int k = 0; k < 16; ++k) {
        int e = 0;
        for (int pass = 0; pass < 3; ++pass) e += (int)(which != 1 ? res[pass][mn * 64 + k] : res[pass][k * 64 + mn]) << (8 * pass);
        printf(" %6d", 2 * e);
      }
      printf("\n");
    }
  }

  if (only == 8) {
    // ---- test 8: bf16 (kind::f16, K=16) issue rates for several N, chained on one accumulator or rotating over several
    std::vector<uint8_t> im(190 * 1024, 0);
    const int Ns[] = {64, 128, 192, 256, 64, 64, 128};
    const int naccs[] = {1, 1, 1, 1, 2, 4, 2};
    for (int i = 0; i < 7; ++i) {
      Job j{}; j.kind = 1; j.N = Ns[i]; j.nacc = naccs[i]; j.nk = 8; j.reps = 256;
      j.a_off = 0; j.a_lbo = PLANE; j.a_sbo = 128; j.a_kstep = 2 * PLANE;          // A: 128 rows, 16 planes of 8 channels
      j.b_off = 40 * 1024; j.b_lbo = 4096; j.b_sbo = 128; j.b_kstep = 2 * 4096;      // B: up to 256 rows per plane
      j.idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(j.N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      const int st = run(j, im, D, &cyc);
      const double n = (double)j.nk * j.reps;
      printf("test8 bf16 M128 N%-3d K16 nacc=%d: status=%d %.0f MMAs %lld cycles -> %.1f cyc/MMA, %.0f MAC/clk\n", j.N, j.nacc, st, n, cyc, cyc / n, 128.0 * j.N * 16 * n / cyc);
    }
    // MN-major A and B (weight-gradient shape), N = 64 and 128
    for (int N = 64; N <= 128; N += 64) {
      Job j{}; j.kind = 1; j.N = N; j.nacc = 1; j.nk = 8; j.reps = 256;
      j.a_off = 0; j.a_lbo = 128; j.a_sbo = PLANE; j.a_kstep = 256; j.b_off = 40 * 1024; j.b_lbo = 128; j.b_sbo = PLANE; j.b_kstep = 256;
      j.idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
      const int st = run(j, im, D, &cyc);
      const double n = (double)j.nk * j.reps;
      printf("test8 bf16 MNxMN M128 N%-3d K16: status=%d %.0f MMAs %lld cycles -> %.1f cyc/MMA, %.0f MAC/clk\n", N, st, n, cyc, cyc / n, 128.0 * N * 16 * n / cyc);
    }
  }
  if (only == 9) {
    // ---- test 9: sign dependence of the accumulate rounding (bf16 operands): acc = +-1, then 16 MMAs adding +-f*2^-23 each
    auto bf16 = [](float f) { uint32_t u; memcpy(&u, &f, 4); return (uint16_t)(u >> 16); };
    const int NKR = 17;
    std::vector<uint8_t> im(130 * 1024, 0);
    const uint32_t wb = 90 * 1024;
    auto putb = [&](uint32_t base, uint32_t plane, int ch, int row, float v) { const uint16_t h = bf16(v); memcpy(&im[base + (ch / 8) * plane + row * 16 + (ch % 8) * 2], &h, 2); };
    for (int r = 0; r < 128; ++r) {
      putb(0, PLANE, 0, r, 1.f);
      const float delta = ldexpf((float)(r % 16 + 1) / 8.f, -23);
      for (int i = 1; i < NKR; ++i) putb(0, PLANE, 16 * i, r, delta);
    }
    for (int o = 0; o < 64; ++o) {   // column o: bit0 = sign of the addends, bit1 = sign of the accumulator start
      putb(wb, 1024, 0, o, (o & 2) ? -1.f : 1.f);
      for (int i = 1; i < NKR; ++i) putb(wb, 1024, 16 * i, o, (o & 1) ? -1.f : 1.f);
    }
    Job j{}; j.kind = 1; j.N = 64; j.nacc = 1; j.nk = NKR; j.reps = 1;
    j.a_off = 0; j.a_lbo = PLANE; j.a_sbo = 128; j.a_kstep = 2 * PLANE; j.b_off = wb; j.b_lbo = 1024; j.b_sbo = 128; j.b_kstep = 2048;
    j.idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const int st = run(j, im, D, &cyc);
    printf("test9 rounding status=%d: f | exact 16f | (D-s)*2^23 for start +1 (add +, add -) and start -1 (add +, add -)\n", st);
    for (int t = 0; t < 16; ++t)
      printf("   f=%.3f exact=%5.2f | %+8.3f %+8.3f | %+8.3f %+8.3f\n", (t + 1) / 8.0, 16.0 * (t + 1) / 8.0,
             ((double)D[t * 64 + 0] - 1.0) * 8388608.0, ((double)D[t * 64 + 1] - 1.0) * 8388608.0,
             ((double)D[t * 64 + 2] + 1.0) * 8388608.0, ((double)D[t * 64 + 3] + 1.0) * 8388608.0);
  }

  if (only == 10) {
    // ---- test 10: 128-byte-swizzled bf16 operands.  Buffers are [row][64 ch] with 128-byte rows; the 16-byte chunk index
    // of a row is XORed with (absolute byte address >> 7) & 7.  The same buffer is K-major (rows = M/N, K = channel) and
    // MN-major (MN = channel, K = row).  (a) forward tap (b) weight gradient (c) rates
    auto bf16 = [](float f) { uint32_t u; memcpy(&u, &f, 4); return (uint16_t)(u >> 16); };
    const uint32_t smem_base = argc > 3 ? (uint32_t)atoi(argv[3]) : 1024;   // dynamic smem base address assumed for the absolute swizzle (printed by the kernel?)
    (void)smem_base;
    std::vector<uint8_t> im(64 * 1024, 0);
    auto putsw = [&](uint32_t base, int row, int ch, float v) {   // base must be 1024-aligned relative to a 1024-aligned smem base
      const uint32_t off = base + row * 128;
      const uint32_t chunk = (uint32_t)(ch / 8) ^ ((off >> 7) & 7);
      const uint16_t h = bf16(v); memcpy(&im[off + chunk * 16 + (ch % 8) * 2], &h, 2);
    };
    const uint32_t AB = 0, WB = 24 * 1024, GB = 40 * 1024;
    for (int r = 0; r < 136; ++r) for (int c = 0; c < 64; ++c) putsw(AB, r, c, ival(c, r, 1));
    for (int o = 0; o < 64; ++o) for (int c = 0; c < 64; ++c) putsw(WB, o, c, ival(o, c, 2));
    for (int r = 0; r < 136; ++r) for (int c = 0; c < 64; ++c) putsw(GB, r, c, ival(c, r, 4));
    const uint32_t fmt = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    for (int shift = 0; shift < 3; ++shift) {   // (a) D[t][o] = sum_c act[c][t+shift] W[o][c]
      Job j{}; j.kind = 1; j.N = 64; j.nacc = 1; j.nk = 4; j.reps = 1; j.a_layout = j.b_layout = 2;
      j.a_off = AB + shift * 128; j.a_lbo = 16; j.a_sbo = 1024; j.a_kstep = 32; j.b_off = WB; j.b_lbo = 16; j.b_sbo = 1024; j.b_kstep = 32; j.idesc = fmt;
      const int st = run(j, im, D, &cyc);
      std::vector<double> ref(128 * 64, 0.0);
      for (int t = 0; t < 128; ++t) for (int o = 0; o < 64; ++o) { double s2 = 0; for (int c = 0; c < 64; ++c) s2 += (double)ival(c, t + shift, 1) * ival(o, c, 2); ref[t * 64 + o] = s2; }
      const double e = check(D, ref, 64);
      printf("test10a SW128 K-major fwd tap shift=%d: status=%d maxerr=%g %s\n", shift, st, e, (st == 0 && e == 0) ? "PASS" : "FAIL");
    }
    for (int shift = 0; shift < 3; ++shift) {   // (b) D[o][c] = sum_{t<128} act[o][t] g[c][t+shift]  (M = 64 channels of act twice: rows 64..127 read the next 64-channel block = W region junk, ignored)
      Job j{}; j.kind = 1; j.N = 64; j.nacc = 1; j.nk = 8; j.reps = 1; j.a_layout = j.b_layout = 2;
      j.a_off = AB; j.a_lbo = 24 * 1024; j.a_sbo = 2048; j.a_kstep = 2048;     // MN-major: LBO = stride between 64-element MN blocks, SBO = stride between 8-row K groups... K=16 -> 2048 B per step
      j.b_off = GB + shift * 128; j.b_lbo = 1024; j.b_sbo = 2048; j.b_kstep = 2048;
      j.a_sbo = 1024; j.b_sbo = 1024;
      j.idesc = fmt | (1u << 15) | (1u << 16);
      const int st = run(j, im, D, &cyc);
      std::vector<double> ref(128 * 64, 0.0);
      for (int o = 0; o < 64; ++o) for (int c = 0; c < 64; ++c) { double s2 = 0; for (int t = 0; t < 128; ++t) s2 += (double)ival(o, t, 1) * ival(c, t + shift, 4); ref[o * 64 + c] = s2; }
      double e = 0; for (int o = 0; o < 64; ++o) for (int c = 0; c < 64; ++c) e = fmax(e, fabs((double)D[o * 64 + c] - ref[o * 64 + c]));
      printf("test10b SW128 MN-major wgrad shift=%d: status=%d maxerr(rows 0..63)=%g %s\n", shift, st, e, (st == 0 && e == 0) ? "PASS" : "FAIL");
    }
    {   // (c) rates
      std::vector<uint8_t> z(160 * 1024, 0);
      const int Ns[] = {64, 128, 192, 256};
      for (int i = 0; i < 4; ++i) {
        Job j{}; j.kind = 1; j.N = Ns[i]; j.nacc = 1; j.nk = 4; j.reps = 512; j.a_layout = j.b_layout = 2;
        j.a_off = 0; j.a_lbo = 16; j.a_sbo = 1024; j.a_kstep = 32; j.b_off = 32 * 1024; j.b_lbo = 16; j.b_sbo = 1024; j.b_kstep = 32;
        j.idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(j.N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        const int st = run(j, z, D, &cyc);
        const double n = (double)j.nk * j.reps;
        printf("test10c SW128 KxK bf16 M128 N%-3d K16: status=%d %.0f MMAs %lld cycles -> %.1f cyc/MMA, %.0f MAC/clk\n", j.N, st, n, cyc, cyc / n, 128.0 * j.N * 16 * n / cyc);
      }
    }
  }
  printf("tc5_probe: %d failing test group(s)\n", nfail);
  return nfail ? 1 : 0;
}
