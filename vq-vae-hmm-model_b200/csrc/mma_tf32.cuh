// mma_tf32.cuh -- warp-level 3xTF32 GEMM building blocks of the fused ELBO kernels.
//
// Every channel-mixing op of the VAE_HMM path (VQ_VAE_HMM_fixed.py:34-36, :53-57, :77-79, :83) is
// one of two GEMM shapes over a time tile held in shared memory as act[channel][PT]:
//   gemm_cols : out[m][j]        = sum_tap sum_k A_tap[m][k] * In[k][j + tap - tapc]   (fwd, dgrad)
//   gemm_time : acc[tap][m][n]  += sum_j     Y[m][j]   * X[n][j + tap - tapc]         (wgrad)
// Both run on mma.sync.m16n8k8 TF32 with the fp32 operands split on the fly into hi + lo TF32
// parts (x = hi + lo, hi = RN_tf32(x)).  The tensor core's accumulate truncates (floor) instead of
// rounding to nearest, so each hi*hi product block is formed with C = 0 and added to the running
// sum with FADD; only the two small cross terms lo*hi, hi*lo are chained inside the tensor core
// (their truncation error is 2^-11 further down).  lo*lo ~ 2^-22 is dropped: fp32-level results.
//
// Bank-conflict plan (PT == 4 mod 32, weight pitch PW == 8 mod 32):
//   gemm_cols uses the k-slot permutation slot s<4 -> k = 2s, slot s+4 -> k = 2s+1, so a B fragment
//   load touches banks 8*t4 + g (+4): conflict-free; A (weights) are 8-byte loads.
//   gemm_time uses the natural order: banks 4*g + t4.
#pragma once
#include <stdint.h>

namespace vhmm {
namespace fz {

__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

// d = a*b (C = 0): used for the hi*hi products so that the tensor core's truncating accumulate
// (measured on B200: d = floor-to-fp32(c + a*b), a systematic -0.5 ulp per chained MMA, see
// tools/mma_round.cu) never touches a running sum; the running sum is kept with FADD (RN).
__device__ __forceinline__ void mma_tf32_z(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};\n"
      : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]), "f"(0.f));
}

// x = hi + lo exactly; hi has a 10-bit mantissa (round to nearest, ties away), lo is left as a
// full fp32 (the tensor core reads its top 19 bits).
__device__ __forceinline__ void split_tf32(float v, uint32_t& hi, uint32_t& lo) {
  hi = (__float_as_uint(v) + 0x1000u) & 0xffffe000u;
  lo = __float_as_uint(v - __uint_as_float(hi));
}

// ---------------------------------------------------------------------------------------
// gemm_cols: one warp computes rows [m0, m0+16*MT) x columns [j0, j0+8*NT) of
//   out[m][j] = sum_{tap<TAPS} sum_{k<8*KSTEPS} A_tap(m,k) * In[k*PT + j + tap - (TAPS==3)]
// TRANS_A == false : A_tap(m,k) = W[tap*tapStride + m*PW + k]                (forward conv)
// TRANS_A == true  : A_tap(m,k) = W[(TAPS-1-tap)*tapStride + k*PW + m]       (input gradient)
// epi(row, col, v0, v1) receives out[row][col], out[row][col+1].
// ---------------------------------------------------------------------------------------
template <int MT, int NT, int KSTEPS, int TAPS, bool TRANS_A, int PT, typename Epi>
__device__ __forceinline__ void gemm_cols(const float* __restrict__ W, int tapStride, int PW,
                                          const float* __restrict__ In, int m0, int j0, Epi&& epi) {
  const int lane = threadIdx.x & 31, g = lane >> 2, t4 = lane & 3;
  float accM[MT][NT][4], accC[MT][NT][4];
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int i = 0; i < 4; ++i) accM[mt][nt][i] = accC[mt][nt][i] = 0.f;

#pragma unroll 1
  for (int tap = 0; tap < TAPS; ++tap) {
    const float* Wt = W + (TRANS_A ? (TAPS - 1 - tap) : tap) * tapStride;
    const float* Bt = In + j0 + g + tap - (TAPS == 3 ? 1 : 0);
#pragma unroll 2
    for (int ks = 0; ks < KSTEPS; ++ks) {
      const int k = ks * 8 + 2 * t4;
      uint32_t ahi[MT][4], alo[MT][4];
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) {
        const int m = m0 + mt * 16 + g;
        float a0, a1, a2, a3;
        if (!TRANS_A) {
          const float2 v0 = *reinterpret_cast<const float2*>(Wt + m * PW + k);
          const float2 v1 = *reinterpret_cast<const float2*>(Wt + (m + 8) * PW + k);
          a0 = v0.x; a2 = v0.y; a1 = v1.x; a3 = v1.y;
        } else {
          a0 = Wt[k * PW + m]; a1 = Wt[k * PW + m + 8];
          a2 = Wt[(k + 1) * PW + m]; a3 = Wt[(k + 1) * PW + m + 8];
        }
        split_tf32(a0, ahi[mt][0], alo[mt][0]);
        split_tf32(a1, ahi[mt][1], alo[mt][1]);
        split_tf32(a2, ahi[mt][2], alo[mt][2]);
        split_tf32(a3, ahi[mt][3], alo[mt][3]);
      }
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        uint32_t bhi[2], blo[2];
        split_tf32(Bt[k * PT + nt * 8], bhi[0], blo[0]);
        split_tf32(Bt[(k + 1) * PT + nt * 8], bhi[1], blo[1]);
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
          mma_tf32(accC[mt][nt], alo[mt], bhi);
          mma_tf32(accC[mt][nt], ahi[mt], blo);
          float d[4];
          mma_tf32_z(d, ahi[mt], bhi);
#pragma unroll
          for (int i = 0; i < 4; ++i) accM[mt][nt][i] += d[i];
        }
      }
    }
  }
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int r = m0 + mt * 16 + g, c = j0 + nt * 8 + 2 * t4;
      epi(r, c, accM[mt][nt][0] + accC[mt][nt][0], accM[mt][nt][1] + accC[mt][nt][1]);
      epi(r + 8, c, accM[mt][nt][2] + accC[mt][nt][2], accM[mt][nt][3] + accC[mt][nt][3]);
    }
}

// ---------------------------------------------------------------------------------------
// gemm_time: one warp accumulates, for rows [0,16*MT) of Y and rows [0,8) of X (both pointers
// already offset to their first row),
//   acc[tap*accTapStride + m*PA + n] += sum_{j in [J0, J0+8*KSTEPS)} Y[m*PT + j] * X[n*PT + j + tap - (TAPS==3)]
// The per-tile sum is formed in registers from zero and then added (RN) into the fp32 accumulators
// in shared memory, which persist across the tiles of a CTA.
// ---------------------------------------------------------------------------------------
template <int MT, int TAPS, int KSTEPS, int J0, int PT>
__device__ __forceinline__ void gemm_time(const float* __restrict__ Y, const float* __restrict__ X,
                                          float* __restrict__ acc, int accTapStride, int PA) {
  const int lane = threadIdx.x & 31, g = lane >> 2, t4 = lane & 3;
  float tM[MT][TAPS][4], tC[MT][TAPS][4];
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int tp = 0; tp < TAPS; ++tp)
#pragma unroll
      for (int i = 0; i < 4; ++i) tM[mt][tp][i] = tC[mt][tp][i] = 0.f;
  const float* Yl = Y + g * PT + J0 + t4;
  const float* Xl = X + g * PT + J0 + t4 - (TAPS == 3 ? 1 : 0);
#pragma unroll 1
  for (int ks = 0; ks < KSTEPS; ++ks) {
    uint32_t ahi[MT][4], alo[MT][4];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
      const float* p = Yl + mt * 16 * PT + ks * 8;
      split_tf32(p[0], ahi[mt][0], alo[mt][0]);
      split_tf32(p[8 * PT], ahi[mt][1], alo[mt][1]);
      split_tf32(p[4], ahi[mt][2], alo[mt][2]);
      split_tf32(p[8 * PT + 4], ahi[mt][3], alo[mt][3]);
    }
#pragma unroll
    for (int tp = 0; tp < TAPS; ++tp) {
      uint32_t bhi[2], blo[2];
      split_tf32(Xl[ks * 8 + tp], bhi[0], blo[0]);
      split_tf32(Xl[ks * 8 + tp + 4], bhi[1], blo[1]);
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) {
        mma_tf32(tC[mt][tp], alo[mt], bhi);
        mma_tf32(tC[mt][tp], ahi[mt], blo);
        float d[4];
        mma_tf32_z(d, ahi[mt], bhi);
#pragma unroll
        for (int i = 0; i < 4; ++i) tM[mt][tp][i] += d[i];
      }
    }
  }
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int tp = 0; tp < TAPS; ++tp) {
      float2* p0 = reinterpret_cast<float2*>(acc + tp * accTapStride + (mt * 16 + g) * PA + 2 * t4);
      float2* p1 = reinterpret_cast<float2*>(acc + tp * accTapStride + (mt * 16 + g + 8) * PA + 2 * t4);
      float2 v0 = *p0, v1 = *p1;
      v0.x += tM[mt][tp][0] + tC[mt][tp][0]; v0.y += tM[mt][tp][1] + tC[mt][tp][1];
      v1.x += tM[mt][tp][2] + tC[mt][tp][2]; v1.y += tM[mt][tp][3] + tC[mt][tp][3];
      *p0 = v0; *p1 = v1;
    }
}

// acc[r] += sum_{j in [J0, J0+4*NV)} buf[r*PT + j] for rows r = warp, warp+NW, ... < nrows
template <int J0, int NV, int PT, int NW>
__device__ __forceinline__ void row_sums(const float* __restrict__ buf, int nrows, float* __restrict__ acc) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int r = w; r < nrows; r += NW) {
    float s = 0.f;
    if (lane < NV) {
      const float4 v = *reinterpret_cast<const float4*>(buf + r * PT + J0 + 4 * lane);
      s = (v.x + v.y) + (v.z + v.w);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) acc[r] += s;
  }
}

}  // namespace fz
}  // namespace vhmm
