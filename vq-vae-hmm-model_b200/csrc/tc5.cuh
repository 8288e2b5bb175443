// tc5.cuh -- thin inline-PTX layer over the 5th-generation tensor core (tcgen05) as used by the
// fused decoder kernel: shared-memory operand descriptors (no-swizzle "interleave" layout),
// single-thread MMA issue (kind::f16), commit -> mbarrier, TMEM allocation and TMEM <-> register
// moves.  Every fact used here was checked on a B200 with tools/tc5_probe.cu (profiles/r1_tc5_probe.md):
//   * canonical no-swizzle layouts for 16-bit elements.  A 16-byte row holds 8 elements.
//       K-major  (rows = M or N, 8 k per 16 B): addr(r,k)  = (r%8)*16 + (r/8)*SBO + (k%8)*2 + (k/8)*LBO
//       MN-major (8 mn per 16 B, rows = k)     : addr(mn,k) = (mn%8)*2 + (mn/8)*SBO + (k%8)*16 + (k/8)*LBO
//     so ONE buffer act[ch/8][row][8] is a K-major operand (rows = time, K = channel: LBO = plane,
//     SBO = 128) and an MN-major operand (MN = channel, K = time: SBO = plane, LBO = 128), and a
//     conv tap is a +16-byte shift of the start address.
//   * MN-major kind::tf32 operands read zeros in this layout -> the kernel uses 16-bit operands.
//   * the fp32 accumulate inside the tensor core rounds toward zero (sign-symmetric truncation).
//   * one M=128 MMA costs >= ~115 cycles whatever N <= 128 is (operand fetch), so N is stacked.
#pragma once
#include <stdint.h>
#include <cuda_fp16.h>

namespace vhmm {
namespace tc5 {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// no-swizzle shared-memory matrix descriptor (descriptor version 1 = Blackwell)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) | ((uint64_t)1 << 46);
}

// instruction descriptor, kind::f16 with fp16 operands, fp32 accumulate, M = 128
__host__ __device__ constexpr uint32_t idesc_f16(int N, int a_mn_major, int b_mn_major) {
  return (1u << 4) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) | ((uint32_t)(N >> 3) << 17) |
         ((uint32_t)(128 >> 4) << 24);
}

// D[tmem] (+)= A[smem] * B[smem]; issued by ONE thread
__device__ __forceinline__ void mma_f16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
      :: "r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}

// all MMAs issued so far by this thread arrive on the mbarrier when they have completed
__device__ __forceinline__ void commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_init_fence() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }

// lane 0 of every warp polls (bounded: a lost arrival traps instead of hanging the GPU), the
// other lanes wait at the warp barrier
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  if ((threadIdx.x & 31) == 0) {
    const uint32_t a = smem_u32(bar);
    uint32_t ok = 0;
    for (int it = 0; it < (1 << 24) && !ok; ++it)
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                   : "=r"(ok) : "r"(a), "r"(parity) : "memory");
    if (!ok) __trap();
  }
  __syncwarp();
}

__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// generic-proxy shared-memory writes -> visible to the tensor core's (async proxy) operand reads
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ bool elect_one() {
  uint32_t e;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(e));
  return e != 0;
}

// one warp allocates all 512 TMEM columns of the SM (1 CTA per SM) / frees them
__device__ __forceinline__ void tmem_alloc512(uint32_t* dst_smem) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(smem_u32(dst_smem)) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_free512(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(taddr) : "memory");
}

// 16 consecutive fp32 columns of this thread's TMEM lane (lane = 32*(warp%4) + laneid)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
// 8 consecutive fp32 columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float (&v)[8]) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const float (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
      :: "r"(taddr), "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])),
         "r"(__float_as_uint(v[3])), "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])),
         "r"(__float_as_uint(v[7])), "r"(__float_as_uint(v[8])), "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])),
         "r"(__float_as_uint(v[11])), "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])),
         "r"(__float_as_uint(v[15]))
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// x*s = h + m (+ 2^-22 relative): two fp16 parts of a scaled fp32 value
__device__ __forceinline__ void split_f16(float xs, __half& h, __half& m) {
  h = __float2half_rn(xs);
  m = __float2half_rn(xs - __half2float(h));
}

// packed fp32 pairs (Blackwell add / mul .f32x2): one issue slot for two lanes of work (the conversion loops are issue-bound)
typedef unsigned long long f2_t;
__device__ __forceinline__ f2_t f2_pack(float a, float b) { f2_t r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void f2_unpack(f2_t v, float& a, float& b) { asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ f2_t f2_add(f2_t a, f2_t b) { f2_t r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f2_t f2_sub(f2_t a, f2_t b) { f2_t r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f2_t f2_mul(f2_t a, f2_t b) { f2_t r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
// eight fp32 values (scaled by s2 = {s, s}) -> their fp16 h and m parts as two 16-byte rows (same split as tc5::split_f16)
__device__ __forceinline__ void split8(const float* v, f2_t s2, uint4& h, uint4& m) {
  uint32_t hh[4], mm[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const f2_t xs = f2_mul(f2_pack(v[2 * e], v[2 * e + 1]), s2);
    float x0, x1;
    f2_unpack(xs, x0, x1);
    const __half2 h2 = __floats2half2_rn(x0, x1);
    const float2 hf = __half22float2(h2);
    float r0, r1;
    f2_unpack(f2_sub(xs, f2_pack(hf.x, hf.y)), r0, r1);
    const __half2 m2 = __floats2half2_rn(r0, r1);
    hh[e] = *reinterpret_cast<const uint32_t*>(&h2);
    mm[e] = *reinterpret_cast<const uint32_t*>(&m2);
  }
  h = make_uint4(hh[0], hh[1], hh[2], hh[3]);
  m = make_uint4(mm[0], mm[1], mm[2], mm[3]);
}

// power-of-two scale that maps |x| <= bound into [2^13, 2^14) (fp16 max = 65504); 1 for bound == 0
__device__ __forceinline__ float pow2_scale(float bound) {
  // integer form (this runs on every thread after each per-tile maximum): exponent field -> 2^(13 - e)
  const uint32_t bits = __float_as_uint(bound) & 0x7fffffffu;
  if (bits == 0u || bits >= 0x7f800000u) return 1.f;        // 0, inf, nan
  int e = (int)(bits >> 23) - 127;                            // bound in [2^e, 2^(e+1)); subnormals read as -127
  int k = 13 - e;
  k = k > 40 ? 40 : (k < -40 ? -40 : k);   // products of two scales stay far inside fp32's range
  return __uint_as_float((uint32_t)(k + 127) << 23);
}

}  // namespace tc5
}  // namespace vhmm
