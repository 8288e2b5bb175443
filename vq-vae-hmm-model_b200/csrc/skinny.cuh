// skinny.cuh -- register-blocked FP32-FFMA forms of the "skinny" layers of the fused ELBO kernels: layers with
// 3..10 rows on one side (encoder conv1 5 -> 64, folded decoder conv1 3 -> 64, to_logits 32 -> 3, to_params
// 64 -> 10, the transition MLP 4 -> 128 -> 9) and their gradients.  On the padded mma.sync helpers of
// mma_tf32.cuh these are latency- or issue-bound (rows padded to 8/16, operands split on the fly, one dependent
// MMA chain per warp).  Common scheme: thread = (the long-side row, a time slice), the few short-side values in
// registers, 16-byte operand loads of which all but one are warp broadcasts, slice partials in a scratch region
// that is dead at that point, one barrier, an ordered (deterministic) reduction.  Buffers are act[row][PT] fp32
// with PT == 4 mod 32: a 16-byte piece of 8 consecutive rows is bank-conflict free.
// What stays on mma.sync after measurement (B200): the 128 -> 9 transition logits forward (kernel A) and the
// to_params forward / input gradient (decoder) -- long contraction with BOTH operands in shared memory, ~1.7 FMA
// per LDS wavefront as FFMA.  Included inside namespace vhmm::fz.
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

// ---- "skinny" k=3 conv gradients with K_Q <= 5 input rows (folded decoder conv1: 3 -> 64, encoder
// conv1: 5 -> 64).  On the padded mma.sync helpers these are latency-bound (rows padded 3 -> 8/16, a
// 15-step dependent MMA chain per warp); as register-blocked FFMA they are ~4x shorter.
//
// sk_wgrad_taps: thread (o = tid % HH, s = tid / HH) sums its slice of the owned columns
//   part[s][o][tau*KQ + k] = sum_{j in slice s} dY[o][j] * X[k][j + tau - 1],  part[s][o][3 KQ] = sum_j dY[o][j]
// into scr[(s*HH + o)*PS + .] (PS odd: conflict-free); after a __syncthreads sk_wgrad_taps_reduce adds
// the NTH/HH slices in a fixed order into accW[(tau*HH + o)*8 + k] and accB[o] (deterministic).
template <int KQ, int HH, int NS_ = NTH / HH>
struct SkTaps {
  static constexpr int NS = NS_;                 // time slices (threads with tid / HH >= NS idle)
  static constexpr int NO = 3 * KQ + 1;          // outputs per (o, slice)
  static constexpr int PS = NO | 1;              // odd pitch
  static constexpr int FLOATS = NS * HH * PS;
};

template <int KQ, int HH, int NS = NTH / HH>
__device__ __forceinline__ void sk_wgrad_taps(const float* __restrict__ dY, const float* __restrict__ X,
                                              float* __restrict__ scr) {
  using SK = SkTaps<KQ, HH, NS>;
  const int o = threadIdx.x % HH, s = threadIdx.x / HH;
  if (s >= NS) return;
  float a0[KQ], a1[KQ], a2[KQ], sb = 0.f;
#pragma unroll
  for (int k = 0; k < KQ; ++k) a0[k] = a1[k] = a2[k] = 0.f;
  for (int q = s; q < TT / 4; q += SK::NS) {
    const int j = J0 + 4 * q;
    const float4 d = *reinterpret_cast<const float4*>(dY + o * PT + j);
    sb += (d.x + d.y) + (d.z + d.w);
#pragma unroll
    for (int k = 0; k < KQ; ++k) {
      const float* xr = X + k * PT + j;
      const float xm = xr[-1], xp = xr[4];
      const float4 xv = *reinterpret_cast<const float4*>(xr);
      a0[k] = fmaf(d.w, xv.z, fmaf(d.z, xv.y, fmaf(d.y, xv.x, fmaf(d.x, xm, a0[k]))));
      a1[k] = fmaf(d.w, xv.w, fmaf(d.z, xv.z, fmaf(d.y, xv.y, fmaf(d.x, xv.x, a1[k]))));
      a2[k] = fmaf(d.w, xp, fmaf(d.z, xv.w, fmaf(d.y, xv.z, fmaf(d.x, xv.y, a2[k]))));
    }
  }
  float* p = scr + (s * HH + o) * SK::PS;
#pragma unroll
  for (int k = 0; k < KQ; ++k) { p[k] = a0[k]; p[KQ + k] = a1[k]; p[2 * KQ + k] = a2[k]; }
  p[3 * KQ] = sb;
}

template <int KQ, int HH, int NS = NTH / HH>
__device__ __forceinline__ void sk_wgrad_taps_reduce(const float* __restrict__ scr, float* __restrict__ accW,
                                                     float* __restrict__ accB) {
  using SK = SkTaps<KQ, HH, NS>;
  for (int i = threadIdx.x; i < HH * SK::NO; i += NTH) {
    const int o = i / SK::NO, c = i - o * SK::NO;
    float sum = 0.f;
#pragma unroll
    for (int s = 0; s < SK::NS; ++s) sum += scr[(s * HH + o) * SK::PS + c];
    if (c < 3 * KQ) accW[((c / KQ) * HH + o) * 8 + (c % KQ)] += sum;
    else accB[o] += sum;
  }
}

// sk_dgrad_taps: input gradient of the same conv, out[k][j] = sum_tau sum_o W[tau][o][k] * dY[o][j + 1 - tau]
// (W as [3][HH][8] floats, k < KQ <= 4).  Thread (p = tid % 64 < TT/2, g = tid / 64 < NG) handles the
// column pair j = J0 + 2p, j+1 over the HH/NG rows of group g -> scr[(g*KQ + k)*NC + j]; threads with
// g >= NG skip.  sk_dgrad_taps_store (after a __syncthreads) adds the groups and writes the owned
// columns with t < T to dq[k*T + t] (dq already offset to the sequence).
template <int KQ, int HH, int NG>
__device__ __forceinline__ void sk_dgrad_taps(const float* __restrict__ Wf, const float* __restrict__ dY,
                                              float* __restrict__ scr) {
  static_assert(KQ <= 4, "weights are read as one float4 per (tau, o)");
  const int p = threadIdx.x & 63, g = threadIdx.x >> 6;
  if (g >= NG || p >= TT / 2) return;
  const int j = J0 + 2 * p;
  float o0[4] = {0.f, 0.f, 0.f, 0.f}, o1[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
  for (int o = g * (HH / NG); o < (g + 1) * (HH / NG); ++o) {
    const float* r = dY + o * PT + j;
    const float dm = r[-1], dp = r[2];
    const float2 d = *reinterpret_cast<const float2*>(r);
    const float4 w0 = *reinterpret_cast<const float4*>(Wf + o * 8);
    const float4 w1 = *reinterpret_cast<const float4*>(Wf + (HH + o) * 8);
    const float4 w2 = *reinterpret_cast<const float4*>(Wf + (2 * HH + o) * 8);
    const float a0[4] = {w0.x, w0.y, w0.z, w0.w}, a1[4] = {w1.x, w1.y, w1.z, w1.w}, a2[4] = {w2.x, w2.y, w2.z, w2.w};
#pragma unroll
    for (int k = 0; k < KQ; ++k) {
      o0[k] = fmaf(a2[k], dm, fmaf(a1[k], d.x, fmaf(a0[k], d.y, o0[k])));
      o1[k] = fmaf(a2[k], d.x, fmaf(a1[k], d.y, fmaf(a0[k], dp, o1[k])));
    }
  }
#pragma unroll
  for (int k = 0; k < KQ; ++k) *reinterpret_cast<float2*>(scr + (g * KQ + k) * NC + j) = make_float2(o0[k], o1[k]);
}

template <int KQ, int NG>
__device__ __forceinline__ void sk_dgrad_taps_store(const float* __restrict__ scr, float* __restrict__ dq, int t0, int T) {
  for (int i = threadIdx.x; i < KQ * TT; i += NTH) {
    const int k = i / TT, c = J0 + (i - k * TT);
    float sum = 0.f;
#pragma unroll
    for (int g = 0; g < NG; ++g) sum += scr[(g * KQ + k) * NC + c];
    const int t = t0 - HL + c;
    if (t < T) dq[(int64_t)k * T + t] = sum;
  }
}

// ---- prior MLP backward (transition_net: U -> TH -> K2 logits), all three gradients in one pass.
// Given dM[r][j] (d loss / d transition logits, zero outside the owned columns), a[h][j] (hidden
// activations) and u[c][j]:
//   da[h][j]    = relu'(a[h][j]) * sum_r W2[r][h] dM[r][j]          (never stored: u is data)
//   dW2[r][h]  += sum_j dM[r][j] a[h][j]
//   dW0[h][c]  += sum_j da[h][j] u[c][j],   db0[h] += sum_j da[h][j]
// Thread (h = tid % THH, s = tid / THH < NS) walks the 4-column groups q = s, s+NS, ... of the owned
// columns: per group 1 + K2 + U 16-byte loads (all but the first are warp broadcasts) feed
// 4 (2 K2 + U) FMAs.  Slice partials -> scr[(s*THH + h)*PS + .]; sk_prior_bwd_reduce adds the slices in
// a fixed order.  Replaces three padded 3xTF32 mma.sync phases (37 % of kernel A).
template <int KK2, int UU, int THH, int NS>
struct SkPrior {
  static constexpr int NO = KK2 + UU + 1;
  static constexpr int PS = NO | 1;
  static constexpr int FLOATS = NS * THH * PS;
};

template <int KK2, int UU, int THH, int NS, int PW2>
__device__ __forceinline__ void sk_prior_bwd(const float* __restrict__ a, const float* __restrict__ dM,
                                             const float* __restrict__ W2, const float* __restrict__ u,
                                             float* __restrict__ scr) {
  using SK = SkPrior<KK2, UU, THH, NS>;
  // 2 NS time slices are computed; the two slices of a warp (lanes 0..15 / 16..31, 16 hidden rows each)
  // are combined with one shuffle before the store, so only NS partials per row reach the scratch.
  static_assert(THH * NS * 2 == NTH && THH % 16 == 0, "thread mapping: NTH = 2 NS THH");
  const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
  const int h = (wp % (THH / 16)) * 16 + (lane & 15);
  const int sg = wp / (THH / 16);                 // slice pair kept in the scratch
  const int s = 2 * sg + (lane >> 4);             // time slice of this thread, < 2 NS
  float w2[KK2], g2[KK2], g0[UU], sb = 0.f;
#pragma unroll
  for (int r = 0; r < KK2; ++r) { w2[r] = W2[r * PW2 + h]; g2[r] = 0.f; }
#pragma unroll
  for (int c = 0; c < UU; ++c) g0[c] = 0.f;
#pragma unroll 1
  for (int q = s; q < TT / 4; q += 2 * NS) {
    const int j = J0 + 4 * q;
    const float4 av = *reinterpret_cast<const float4*>(a + h * PT + j);
    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int r = 0; r < KK2; ++r) {
      const float4 d = *reinterpret_cast<const float4*>(dM + r * PT + j);
      t.x = fmaf(w2[r], d.x, t.x); t.y = fmaf(w2[r], d.y, t.y); t.z = fmaf(w2[r], d.z, t.z); t.w = fmaf(w2[r], d.w, t.w);
      g2[r] = fmaf(d.w, av.w, fmaf(d.z, av.z, fmaf(d.y, av.y, fmaf(d.x, av.x, g2[r]))));
    }
    t.x = av.x > 0.f ? t.x : 0.f; t.y = av.y > 0.f ? t.y : 0.f; t.z = av.z > 0.f ? t.z : 0.f; t.w = av.w > 0.f ? t.w : 0.f;
    sb += (t.x + t.y) + (t.z + t.w);
#pragma unroll
    for (int c = 0; c < UU; ++c) {
      const float4 v = *reinterpret_cast<const float4*>(u + c * PT + j);
      g0[c] = fmaf(t.w, v.w, fmaf(t.z, v.z, fmaf(t.y, v.y, fmaf(t.x, v.x, g0[c]))));
    }
  }
#pragma unroll
  for (int r = 0; r < KK2; ++r) g2[r] += __shfl_xor_sync(0xffffffffu, g2[r], 16);
#pragma unroll
  for (int c = 0; c < UU; ++c) g0[c] += __shfl_xor_sync(0xffffffffu, g0[c], 16);
  sb += __shfl_xor_sync(0xffffffffu, sb, 16);
  if (lane < 16) {
    float* p = scr + (sg * THH + h) * SK::PS;
#pragma unroll
    for (int r = 0; r < KK2; ++r) p[r] = g2[r];
#pragma unroll
    for (int c = 0; c < UU; ++c) p[KK2 + c] = g0[c];
    p[KK2 + UU] = sb;
  }
}

template <int KK2, int UU, int THH, int NS, int PW2>
__device__ __forceinline__ void sk_prior_bwd_reduce(const float* __restrict__ scr, float* __restrict__ accW2,
                                                    float* __restrict__ accW0, float* __restrict__ accB0) {
  using SK = SkPrior<KK2, UU, THH, NS>;
  for (int i = threadIdx.x; i < THH * SK::NO; i += NTH) {
    const int h = i / SK::NO, c = i - h * SK::NO;
    float sum = 0.f;
#pragma unroll
    for (int s = 0; s < NS; ++s) sum += scr[(s * THH + h) * SK::PS + c];
    if (c < KK2) accW2[c * PW2 + h] += sum;
    else if (c < KK2 + UU) accW0[h * 8 + (c - KK2)] += sum;
    else accB0[h] += sum;
  }
}

// ---- weight gradient of a 1x1 layer with few output rows (decoder.to_params: 2C = 10 rows):
//   acc[p*PA + o] += sum_{j owned} dP[p][j] * Y[o][j],  p < NP, o < HH
// thread (o = tid % HH, s = tid / HH): 1 + NP 16-byte loads (the dP ones are warp broadcasts) per 4 NP FMAs.
template <int NP, int HH, int NS = NTH / HH>
struct SkRows {
  static constexpr int PS = NP | 1;
  static constexpr int FLOATS = NS * HH * PS;
};

template <int NP, int HH, int NS = NTH / HH>
__device__ __forceinline__ void sk_wgrad_rows(const float* __restrict__ dP, const float* __restrict__ Y,
                                              float* __restrict__ scr) {
  using SK = SkRows<NP, HH, NS>;
  const int o = threadIdx.x % HH, s = threadIdx.x / HH;
  if (s >= NS) return;
  float g[NP];
#pragma unroll
  for (int p = 0; p < NP; ++p) g[p] = 0.f;
  for (int q = s; q < TT / 4; q += NS) {
    const int j = J0 + 4 * q;
    const float4 y = *reinterpret_cast<const float4*>(Y + o * PT + j);
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      const float4 d = *reinterpret_cast<const float4*>(dP + p * PT + j);
      g[p] = fmaf(d.w, y.w, fmaf(d.z, y.z, fmaf(d.y, y.y, fmaf(d.x, y.x, g[p]))));
    }
  }
  float* dst = scr + (s * HH + o) * SK::PS;
#pragma unroll
  for (int p = 0; p < NP; ++p) dst[p] = g[p];
}

template <int NP, int HH, int NS = NTH / HH>
__device__ __forceinline__ void sk_wgrad_rows_reduce(const float* __restrict__ scr, float* __restrict__ accW, int PA) {
  using SK = SkRows<NP, HH, NS>;
  for (int i = threadIdx.x; i < HH * NP; i += NTH) {
    const int o = i / NP, p = i - o * NP;
    float sum = 0.f;
#pragma unroll
    for (int s = 0; s < NS; ++s) sum += scr[(s * HH + o) * SK::PS + p];
    accW[p * PA + o] += sum;
  }
}

// (Measured and dropped: the input gradients of the same 1x1 layers, dg2 = relu'(g2) * Wp^T dp and
// dh2 = relu'(h2) * W3^T dl, as FFMA over all 128 columns were not faster than the mma.sync helper --
// 35.99 vs 35.57 ms for the decoder kernel -- because there every output column is needed, so the
// tensor-core tile is not padded in the long dimension.)

// ---- forward "skinny" layers (few input rows, or few output rows) as FP32 FFMA -----------------
// On the padded mma.sync helper these are one long dependent MMA chain per warp (3xTF32, rows padded to
// 8/16): 3.2 k cycles per tile for encoder conv1 (5 -> 64, k = 3) and 1.3 k for the 32 -> K logits layer.
//
// sk_conv_taps_fwd: out[r][c] = relu(bias[r] + sum_{tau,k} W[(tau*KR + k)*HH + r] * X[k][c + tau - 1]) for
// columns [1, NC-1) inside the sequence, 0 elsewhere.  Lane <-> output row (its 3 KQ weights live in
// registers), warp <-> 16-column slice, so every X load is a warp broadcast and the float4 stores of rows
// r..r+7 tile the banks (PT == 4 mod 32).  Returns this thread's max output (for the fp16 scale).
template <int HH>
struct SkConv {
  static constexpr int RW = HH / 32, NSL = NW / RW, CS = NC / NSL;   // row warps, column slices, columns per slice
};
// core: the masked outputs of this thread's (row r, columns c0..c0+CS) stay in registers (o[]); returns their max
template <int KQ, int HH, int KR = 8>
__device__ __forceinline__ float sk_conv_taps_regs(const float* __restrict__ W, const float* __restrict__ bias,
                                                   const float* __restrict__ X, int t0, int T,
                                                   float (&o)[SkConv<HH>::CS], int& r_out, int& c0_out) {
  constexpr int RW = HH / 32, NSL = NW / RW, CS = NC / NSL;
  static_assert(HH % 32 == 0 && NW % RW == 0 && NC % NSL == 0 && CS % 4 == 0, "row / slice split");
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = 32 * (w % RW) + lane, c0 = CS * (w / RW);
  float wt[3][KQ];
#pragma unroll
  for (int tau = 0; tau < 3; ++tau)
#pragma unroll
    for (int k = 0; k < KQ; ++k) wt[tau][k] = W[(tau * KR + k) * HH + r];
  float acc[CS];
  const float bb = bias[r];
#pragma unroll
  for (int j = 0; j < CS; ++j) acc[j] = bb;
#pragma unroll
  for (int k = 0; k < KQ; ++k) {
    const float* xr = X + k * PT + c0;
    float xv[CS + 2];
    xv[0] = xr[-1];                     // column -1 of slice 0 only feeds the (zeroed) edge column
    xv[CS + 1] = xr[CS];                // column NC of the last slice likewise (row padding)
#pragma unroll
    for (int q = 0; q < CS / 4; ++q) {
      const float4 v = *reinterpret_cast<const float4*>(xr + 4 * q);
      xv[4 * q + 1] = v.x; xv[4 * q + 2] = v.y; xv[4 * q + 3] = v.z; xv[4 * q + 4] = v.w;
    }
#pragma unroll
    for (int tau = 0; tau < 3; ++tau)
#pragma unroll
      for (int j = 0; j < CS; ++j) acc[j] = fmaf(wt[tau][k], xv[j + tau], acc[j]);
  }
  // live columns: [max(1, HL - t0), min(NC - 1, T - t0 + HL)) -- one unsigned compare per element
  const int lo = max(1, HL - t0), hi = min(NC - 1, T - t0 + HL);
  const unsigned span = (unsigned)max(hi - lo, 0);
  const int d0 = c0 - lo;
  float lm = 0.f;
#pragma unroll
  for (int j = 0; j < CS; ++j) {
    o[j] = ((unsigned)(d0 + j) < span) ? fmaxf(acc[j], 0.f) : 0.f;
    lm = fmaxf(lm, o[j]);
  }
  r_out = r; c0_out = c0;
  return lm;
}
// the same, stored as fp32 out[r][c]
template <int KQ, int HH, int KR = 8>
__device__ __forceinline__ float sk_conv_taps_fwd(const float* __restrict__ W, const float* __restrict__ bias,
                                                  const float* __restrict__ X, float* __restrict__ out, int t0, int T) {
  constexpr int CS = SkConv<HH>::CS;
  float o[CS];
  int r, c0;
  const float lm = sk_conv_taps_regs<KQ, HH, KR>(W, bias, X, t0, T, o, r, c0);
#pragma unroll
  for (int q = 0; q < CS / 4; ++q)
    *reinterpret_cast<float4*>(out + r * PT + c0 + 4 * q) = make_float4(o[4 * q], o[4 * q + 1], o[4 * q + 2], o[4 * q + 3]);
  return lm;
}
// registers -> the two fp16 parts of the tcgen05 operand layout act[part][r/8][col+1][8]: 2-byte stores; the lanes of
// a warp (32 rows = 4 planes x 8 channels) hit 16 distinct words in 16 distinct banks (plane pitch == 8 words mod 32)
template <int CS>
__device__ __forceinline__ void sk_store_parts(const float (&o)[CS], float scale, uint8_t* __restrict__ dst, int plane_bytes,
                                               int part_bytes, int r, int c0) {
  uint8_t* p = dst + (r >> 3) * plane_bytes + (c0 + 1) * 16 + (r & 7) * 2;
#pragma unroll
  for (int j = 0; j < CS; ++j) {
    __half h, m;
    tc5::split_f16(o[j] * scale, h, m);
    *reinterpret_cast<__half*>(p + j * 16) = h;
    *reinterpret_cast<__half*>(p + part_bytes + j * 16) = m;
  }
}

// sk_rows_fwd: out[k][c] = bias[k] + sum_{o < KK} W[k*PW + o] * In[o][c] for k < NR (NR * NC <= NTH threads,
// thread = (column, output row); weights are warp broadcasts, In reads run along the lanes).
template <int NR, int KK>
__device__ __forceinline__ void sk_rows_fwd(const float* __restrict__ W, int PW, const float* __restrict__ bias,
                                            const float* __restrict__ In, float* __restrict__ out) {
  static_assert(NR * NC <= NTH && KK % 4 == 0, "one thread per output");
  const int c = threadIdx.x % NC, k = threadIdx.x / NC;
  if (k >= NR) return;
  float a0 = bias[k], a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
  for (int q = 0; q < KK / 4; ++q) {
    const float4 wv = *reinterpret_cast<const float4*>(W + k * PW + 4 * q);
    a0 = fmaf(wv.x, In[(4 * q + 0) * PT + c], a0);
    a1 = fmaf(wv.y, In[(4 * q + 1) * PT + c], a1);
    a2 = fmaf(wv.z, In[(4 * q + 2) * PT + c], a2);
    a3 = fmaf(wv.w, In[(4 * q + 3) * PT + c], a3);
  }
  out[k * PT + c] = (a0 + a1) + (a2 + a3);
}

// sk_rowsT_mask: input gradient of a layer with few output rows, fused with the ReLU mask of the activation
// it flows into:  buf[o][c] = buf[o][c] > 0 ? sum_{p < NP} W[p*PW + o] * D[p][c] : 0   (o < HH, every column);
// returns the thread's max |buf|.
// Lane <-> o (its NP weights in registers, read along the lanes), warp <-> column slice, D loads are broadcasts.
template <int NP, int HH>
__device__ __forceinline__ float sk_rowsT_mask(const float* __restrict__ W, int PW, const float* __restrict__ D,
                                               float* __restrict__ buf) {
  constexpr int RW = HH / 32, NSL = NW / RW, CS = NC / NSL;
  static_assert(HH % 32 == 0 && NW % RW == 0 && NC % NSL == 0 && CS % 4 == 0, "row / slice split");
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int o = 32 * (w % RW) + lane, c0 = CS * (w / RW);
  float wt[NP];
#pragma unroll
  for (int p = 0; p < NP; ++p) wt[p] = W[p * PW + o];
  float acc[CS];
#pragma unroll
  for (int j = 0; j < CS; ++j) acc[j] = 0.f;
#pragma unroll
  for (int p = 0; p < NP; ++p)
#pragma unroll
    for (int q = 0; q < CS / 4; ++q) {
      const float4 d = *reinterpret_cast<const float4*>(D + p * PT + c0 + 4 * q);
      acc[4 * q + 0] = fmaf(wt[p], d.x, acc[4 * q + 0]);
      acc[4 * q + 1] = fmaf(wt[p], d.y, acc[4 * q + 1]);
      acc[4 * q + 2] = fmaf(wt[p], d.z, acc[4 * q + 2]);
      acc[4 * q + 3] = fmaf(wt[p], d.w, acc[4 * q + 3]);
    }
  float lm = 0.f;
#pragma unroll
  for (int q = 0; q < CS / 4; ++q) {
    float4* pb = reinterpret_cast<float4*>(buf + o * PT + c0 + 4 * q);
    const float4 a = *pb;
    const float4 o = make_float4(a.x > 0.f ? acc[4 * q] : 0.f, a.y > 0.f ? acc[4 * q + 1] : 0.f, a.z > 0.f ? acc[4 * q + 2] : 0.f,
                                 a.w > 0.f ? acc[4 * q + 3] : 0.f);
    *pb = o;
    lm = fmaxf(fmaxf(lm, fmaxf(fabsf(o.x), fabsf(o.y))), fmaxf(fabsf(o.z), fabsf(o.w)));
  }
  return lm;   // this thread's max |out| (scale of the fp16 parts)
}

// sk_rows_in_fwd: a layer with few INPUT rows and no taps:  out[h][c] = relu(bias[h] + sum_{k < KQ} W[h*8 + k] * X[k][c])
// inside the sequence, 0 outside (every column).  Lane <-> h, warp <-> column slice, processed 16 columns at a time.
template <int KQ, int HH>
__device__ __forceinline__ void sk_rows_in_fwd(const float* __restrict__ W, const float* __restrict__ bias,
                                               const float* __restrict__ X, float* __restrict__ out, int t0, int T) {
  constexpr int RW = HH / 32, NSL = NW / RW, CS = NC / NSL;
  static_assert(HH % 32 == 0 && NW % RW == 0 && NC % NSL == 0 && CS % 16 == 0 && KQ <= 4, "row / slice split");
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int h = 32 * (w % RW) + lane, c0 = CS * (w / RW);
  const float4 w4 = *reinterpret_cast<const float4*>(W + h * 8);
  const float wt[4] = {w4.x, w4.y, w4.z, w4.w};
  const float bb = bias[h];
  const int lo = max(0, HL - t0), hi = min(NC, T - t0 + HL);
  const unsigned span = (unsigned)max(hi - lo, 0);
#pragma unroll
  for (int cb = 0; cb < CS; cb += 16) {
    float acc[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) acc[j] = bb;
#pragma unroll
    for (int k = 0; k < KQ; ++k)
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 v = *reinterpret_cast<const float4*>(X + k * PT + c0 + cb + 4 * q);
        acc[4 * q + 0] = fmaf(wt[k], v.x, acc[4 * q + 0]);
        acc[4 * q + 1] = fmaf(wt[k], v.y, acc[4 * q + 1]);
        acc[4 * q + 2] = fmaf(wt[k], v.z, acc[4 * q + 2]);
        acc[4 * q + 3] = fmaf(wt[k], v.w, acc[4 * q + 3]);
      }
    const int d0 = c0 + cb - lo;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float o[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) o[e] = ((unsigned)(d0 + 4 * q + e) < span) ? fmaxf(acc[4 * q + e], 0.f) : 0.f;
      *reinterpret_cast<float4*>(out + h * PT + c0 + cb + 4 * q) = make_float4(o[0], o[1], o[2], o[3]);
    }
  }
}
