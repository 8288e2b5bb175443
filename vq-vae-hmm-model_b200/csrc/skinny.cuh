// skinny.cuh -- FP32-FFMA version of the one "skinny" weight gradient where it beats the padded
// tensor-core helper of mma_tf32.cuh: dWt0 (128 x 4 outputs, K = the 120 owned columns), one thread
// per output with 16-byte operand loads.  (Measured on B200: the same treatment of the 128 -> 9
// layer and its gradients is shared-memory-bandwidth bound -- ~1.7 FMA per LDS wavefront -- and
// SLOWER than the mma.sync path, so those stay on the tensor core.)  Buffers are act[row][PT] fp32
// with PT == 4 mod 32: a 16-byte piece of 8 consecutive rows is bank-conflict free.
// Included inside namespace vhmm::fz.
#pragma once

// acc[h*8 + c] += sum_{j owned} act[h][j] * in[c][j] for c < 4, h < 128; bsum[h] += sum_{j owned} act[h][j]
__device__ __forceinline__ void sk_wgrad_in4(const float* __restrict__ act, const float* __restrict__ in,
                                             float* __restrict__ acc, float* __restrict__ bsum) {
  const int h = threadIdx.x & 127, c = threadIdx.x / 128;
  float s = 0.f, sb = 0.f;
#pragma unroll 2
  for (int q = 0; q < TT / 4; ++q) {
    const float4 a = *reinterpret_cast<const float4*>(act + h * PT + J0 + 4 * q);
    const float4 v = *reinterpret_cast<const float4*>(in + c * PT + J0 + 4 * q);
    s = fmaf(v.w, a.w, fmaf(v.z, a.z, fmaf(v.y, a.y, fmaf(v.x, a.x, s))));
    sb += (a.x + a.y) + (a.z + a.w);
  }
  acc[h * 8 + c] += s;
  if (c == 0) bsum[h] += sb;
}
