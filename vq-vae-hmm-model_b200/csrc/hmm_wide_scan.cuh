// hmm_wide_scan.cuh -- log-likelihood and posteriors of FEW LONG sequences for 4 < K <= 32, parallel over time (SURVEY.md
// Appendix B.5, B.3).
//
// hx::fwd_kernel (hmm_wide.cuh) walks a sequence with one warp, one step after the other, each step waiting for its own
// K x K tile of log_A from HBM: 1.4 us per step at K = 32 whatever the batch -- 22 ms for T = 16384.  A batch of thousands of
// sequences hides that; four sequences do not.  Here the sum-product recursion is re-associated like the small-K scan
// (hmm_scan.cuh), in scaled linear space:
//   seg_product_kernel  one warp per (sequence, 128-step segment): P_seg = prod_t M_t, M_t[k][j] = exp(log_A_t[k][j] +
//                       log_b_t[j] - max_j log_b_t).  Lane i owns ROW i of the running product (K registers); M_t is built by
//                       the whole warp (coalesced loads one step ahead, K*K/32 ex2 per lane) into a double-buffered
//                       shared-memory tile and read back as broadcast 16-byte words: K*K FMAs per lane and step.  Every row is
//                       renormalised by ITS OWN power of two after every step (exact; exponent per row), so rows of very
//                       different weight -- a state that the segment makes 1e-60 times less likely than another -- keep their
//                       full mantissa: the incoming vector may sit on any of them.
//   seg_carry_kernel    one warp per sequence: alpha_0 from log_pi and log_b_0, then v <- v P_seg over the segments (rows
//                       weighted by 2^(exponent - max), next segment's rows prefetched), log-normalisers in double.
//   seg_carry_back_kernel  the same backwards: beta_hat at the last step of every segment.  With alpha_hat before and beta_hat
//                       after every segment known, hx::fwd_kernel / hx::bwd_kernel run on the 128-step ROWS (hx::RowView) as if
//                       they were sequences of their own: gamma and xi of few long sequences on B*T/128 warps.
// K^3 instead of K^2 work per step, but over B*T/128 warps instead of B.  A carry that comes out zero or not finite (only
// underflow inside a row: transitions below e^-87) flags the sequence, and hx::fwd_kernel redoes the flagged ones.
#pragma once
#include "hmm_wide.cuh"

namespace vhmm {
namespace hx {

constexpr int WS_SEG = 128;   // steps per segment row
constexpr int WS_WPB = 4;     // warps (rows) per CTA

// x * 2^-e and e, with x in [0.5, 1) for normal x > 0 (e clamped so that 2^-e is a normal float; 0 stays 0)
__device__ __forceinline__ float pow2_inv_of(float mx, int& e) {
  const int be = (__float_as_int(mx) >> 23) & 0xff;      // biased exponent; 0 for zero / subnormal
  e = (be == 0) ? -126 : be - 126;                       // mx = f * 2^e, f in [0.5, 1)
  if (mx == 0.f) e = 0;
  return __int_as_float((127 - e) << 23);                // 2^-e: 127 - e in [1, 253]
}

template <int KP>
__global__ void __launch_bounds__(WS_WPB * 32)
seg_product_kernel(const float* __restrict__ A, const float* __restrict__ Bm, const int64_t* __restrict__ lengths,
                   int64_t rows, int nseg, int T, int K, float* __restrict__ Pmat, int* __restrict__ Pex,
                   float* __restrict__ mbsum) {
  constexpr int NE = KP * KP / 32;                       // elements of M_t a lane builds
  __shared__ __align__(16) float Msm[WS_WPB][2][KP * KP];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int64_t r = (int64_t)blockIdx.x * WS_WPB + wib;
  if (r >= rows) return;
  const int64_t b = r / nseg;
  const int s = (int)(r - b * nseg);
  const int64_t L64 = lengths[b];
  const int L = (int)(L64 < 0 ? 0 : (L64 > T ? T : L64));
  const int t_lo = max(1, s * WS_SEG), t_hi = min((s + 1) * WS_SEG, L);     // step 0 belongs to the carry kernel
  const int KK = K * K;
  const float* Ab = A + b * (int64_t)T * KK;
  const float* Bb = Bm + b * (int64_t)T * K;
  float* M0 = Msm[wib][0];
  for (int e = lane; e < 2 * KP * KP; e += 32) M0[e] = 0.f;                 // columns / rows beyond K stay zero
  __syncwarp();
  float P[KP];
#pragma unroll
  for (int j = 0; j < KP; ++j) P[j] = (j == lane) ? 1.f : 0.f;
  int ex = 0;
  float mbs = 0.f;
  // element e = lane + 32 m of the K x K tile: row k = e / K, column j = e % K
  int ek[NE], ej[NE];
#pragma unroll
  for (int m = 0; m < NE; ++m) {
    const int e = lane + 32 * m;
    ek[m] = e / K; ej[m] = e - ek[m] * K;
  }
  float an[NE], bn = -INFINITY;
  if (t_lo < t_hi) {
#pragma unroll
    for (int m = 0; m < NE; ++m) an[m] = (lane + 32 * m < KK) ? __ldcs(Ab + (int64_t)t_lo * KK + lane + 32 * m) : -INFINITY;
    bn = (lane < K) ? __ldcs(Bb + (int64_t)t_lo * K + lane) : -INFINITY;
  }
  for (int t = t_lo; t < t_hi; ++t) {
    float ac[NE];
#pragma unroll
    for (int m = 0; m < NE; ++m) ac[m] = an[m];
    const float bc = bn;
    if (t + 1 < t_hi) {
#pragma unroll
      for (int m = 0; m < NE; ++m) an[m] = (lane + 32 * m < KK) ? __ldcs(Ab + (int64_t)(t + 1) * KK + lane + 32 * m) : -INFINITY;
      bn = (lane < K) ? __ldcs(Bb + (int64_t)(t + 1) * K + lane) : -INFINITY;
    }
    const float mb = wmax(bc);
    mbs += mb;
    const float cj = (lane < K) ? (bc - mb) * LOG2E : -INFINITY;
    float* Mt = Msm[wib][t & 1];
#pragma unroll
    for (int m = 0; m < NE; ++m) {
      const float c = __shfl_sync(0xffffffffu, cj, ej[m]);
      if (lane + 32 * m < KK) Mt[ek[m] * KP + ej[m]] = ex2(fmaf(ac[m], LOG2E, c));
    }
    __syncwarp();
    float N[KP];
#pragma unroll
    for (int j = 0; j < KP; ++j) N[j] = 0.f;
#pragma unroll
    for (int k = 0; k < KP; ++k) {
      if (k < K) {
        const float4* row = reinterpret_cast<const float4*>(Mt + k * KP);
#pragma unroll
        for (int q = 0; q < KP / 4; ++q) {
          const float4 v = row[q];
          N[4 * q] = fmaf(P[k], v.x, N[4 * q]);
          N[4 * q + 1] = fmaf(P[k], v.y, N[4 * q + 1]);
          N[4 * q + 2] = fmaf(P[k], v.z, N[4 * q + 2]);
          N[4 * q + 3] = fmaf(P[k], v.w, N[4 * q + 3]);
        }
      }
    }
    float mx = N[0];
#pragma unroll
    for (int j = 1; j < KP; ++j) mx = fmaxf(mx, N[j]);
    int e;
    const float sc = pow2_inv_of(mx, e);
    ex += e;
#pragma unroll
    for (int j = 0; j < KP; ++j) P[j] = N[j] * sc;
    // (the tile of step t+1 goes to the other buffer; the one after that waits at the __syncwarp above of step t+1, which
    //  every lane reaches only after its reads of this step)
  }
  if (lane < K) {
    float* dst = Pmat + r * (int64_t)KK + lane * K;
#pragma unroll
    for (int j = 0; j < KP; ++j)
      if (j < K) dst[j] = P[j];
    Pex[r * K + lane] = ex;
  }
  if (lane == 0) mbsum[r] = mbs;
}

// one warp per sequence; lane = state
__global__ void __launch_bounds__(WPB * 32)
seg_carry_kernel(const float* __restrict__ log_pi, const float* __restrict__ Bm, const int64_t* __restrict__ lengths,
                 int64_t nB, int nseg, int T, int K, const float* __restrict__ Pmat, const int* __restrict__ Pex,
                 const float* __restrict__ mbsum, float* __restrict__ loglik, int* __restrict__ suspect,
                 float* __restrict__ vin, double* __restrict__ ll_in, float* __restrict__ carried) {
  const int lane = threadIdx.x & 31;
  const int64_t b = (int64_t)blockIdx.x * WPB + (threadIdx.x >> 5);
  if (b >= nB) return;
  const int64_t L64 = lengths[b];
  const int L = (int)(L64 < 0 ? 0 : (L64 > T ? T : L64));
  const bool on = lane < K;
  if (lane == 0) suspect[b] = 0;
  if (L == 0) {
    if (lane == 0 && loglik) loglik[b] = 0.f;
    if (carried && on) carried[b * K + lane] = logf(0.f);      // what hx::fwd_kernel carries for an empty sequence
    return;
  }
  const int KK = K * K;
  const float x = on ? log_pi[lane] + Bm[b * (int64_t)T * K + lane] : -INFINITY;
  const float m = wmax(x);
  float v = on ? ex2((x - m) * LOG2E) : 0.f;
  const float s0 = wsum(v);
  v *= 1.f / s0;
  double ll = (double)m + (double)logf(s0);
  const int ns = (L + WS_SEG - 1) / WS_SEG;             // segments with at least one valid step
  const float* Pb = Pmat + b * (int64_t)nseg * KK;
  float pn[32];
  int en = 0;
#pragma unroll
  for (int i = 0; i < 32; ++i) pn[i] = (on && i < K) ? Pb[i * K + lane] : 0.f;
  en = on ? Pex[(b * nseg) * (int64_t)K + lane] : 0;
  bool bad = false;
  for (int s = 0; s < ns; ++s) {
    float pc[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) pc[i] = pn[i];
    const int ec = en;
    if (s + 1 < ns) {
      const float* Pn = Pb + (int64_t)(s + 1) * KK;
#pragma unroll
      for (int i = 0; i < 32; ++i) pn[i] = (on && i < K) ? Pn[i * K + lane] : 0.f;
      en = on ? Pex[(b * nseg + s + 1) * (int64_t)K + lane] : 0;
    }
    if (vin && on) vin[(b * nseg + s) * (int64_t)K + lane] = v;      // alpha_hat at the step before segment s
    if (ll_in && lane == 0) ll_in[b * nseg + s] = ll;                // ... and its log-normaliser
    // weight of row `lane`: v * 2^ex, brought to the largest exponent among the rows that carry mass
    const int ve = (v > 0.f) ? (((__float_as_int(v) >> 23) & 0xff) - 126 + ec) : INT_MIN / 2;
    int E = ve;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) E = max(E, __shfl_xor_sync(0xffffffffu, E, o));
    const float u = (v > 0.f) ? ldexpf(v, ec - E) : 0.f;
    float w0 = 0.f, w1 = 0.f, w2 = 0.f, w3 = 0.f;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      const float ui = __shfl_sync(0xffffffffu, u, i);
      if ((i & 3) == 0) w0 = fmaf(ui, pc[i], w0);
      else if ((i & 3) == 1) w1 = fmaf(ui, pc[i], w1);
      else if ((i & 3) == 2) w2 = fmaf(ui, pc[i], w2);
      else w3 = fmaf(ui, pc[i], w3);
    }
    const float w = on ? (w0 + w1) + (w2 + w3) : 0.f;
    const float sm = wsum(w);
    if (!(sm > 0.f) || !(sm < INFINITY)) { bad = true; break; }
    v = w * (1.f / sm);
    ll += (double)logf(sm) + (double)E * 0.6931471805599453 + (double)mbsum[b * nseg + s];
  }
  if (lane == 0) {
    if (loglik) loglik[b] = (float)ll;
    if (bad) suspect[b] = 1;
  }
  if (carried && on) carried[b * K + lane] = logf(v) + (float)ll;
}

// beta_hat at the last valid step of every segment row, backwards: bin[last valid row] = 1, bin[s-1] = normalise(P_s bin[s]).
// One warp per sequence, lane = state i: (P b)[i] = 2^ex[i] * sum_j Phat[i][j] b[j].
__global__ void __launch_bounds__(WPB * 32)
seg_carry_back_kernel(const int64_t* __restrict__ lengths, int64_t nB, int nseg, int T, int K, const float* __restrict__ Pmat,
                      const int* __restrict__ Pex, float* __restrict__ bin, int* __restrict__ suspect) {
  const int lane = threadIdx.x & 31;
  const int64_t b = (int64_t)blockIdx.x * WPB + (threadIdx.x >> 5);
  if (b >= nB) return;
  const int64_t L64 = lengths[b];
  const int L = (int)(L64 < 0 ? 0 : (L64 > T ? T : L64));
  if (L == 0) return;
  const bool on = lane < K;
  const int KK = K * K;
  const int ns = (L + WS_SEG - 1) / WS_SEG;
  float be = on ? 1.f : 0.f;
  // row `lane` of a segment's product, fetched one segment ahead of the dependent chain
  float pn[32];
  int en = 0;
  auto fetch = [&](int s) {
    const float* Ps = Pmat + (b * nseg + s) * (int64_t)KK + lane * K;
    if ((K & 3) == 0) {
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (on && 4 * q < K) v = *reinterpret_cast<const float4*>(Ps + 4 * q);
        pn[4 * q] = v.x; pn[4 * q + 1] = v.y; pn[4 * q + 2] = v.z; pn[4 * q + 3] = v.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) pn[j] = (on && j < K) ? Ps[j] : 0.f;
    }
    en = on ? Pex[(b * nseg + s) * (int64_t)K + lane] : 0;
  };
  if (ns - 1 >= 1) fetch(ns - 1);
  for (int s = ns - 1; s >= 1; --s) {
    if (on) bin[(b * nseg + s) * (int64_t)K + lane] = be;
    float pc[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) pc[j] = pn[j];
    const int ec = en;
    if (s - 1 >= 1) fetch(s - 1);
    float d0 = 0.f, d1 = 0.f, d2 = 0.f, d3 = 0.f;
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      d0 = fmaf(pc[j], __shfl_sync(0xffffffffu, be, j), d0);
      d1 = fmaf(pc[j + 1], __shfl_sync(0xffffffffu, be, j + 1), d1);
      d2 = fmaf(pc[j + 2], __shfl_sync(0xffffffffu, be, j + 2), d2);
      d3 = fmaf(pc[j + 3], __shfl_sync(0xffffffffu, be, j + 3), d3);
    }
    const float d = on ? (d0 + d1) + (d2 + d3) : 0.f;
    const int ve = (d > 0.f) ? (((__float_as_int(d) >> 23) & 0xff) - 126 + ec) : INT_MIN / 2;
    int E = ve;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) E = max(E, __shfl_xor_sync(0xffffffffu, E, o));
    if (E == INT_MIN / 2) {                              // every row vanished: flag, keep something finite
      if (lane == 0) suspect[b] = 1;
      be = on ? 1.f : 0.f;
      continue;
    }
    be = (d > 0.f) ? ldexpf(d, ec - E) : 0.f;
  }
  if (on) bin[(b * nseg) * (int64_t)K + lane] = be;
}

}  // namespace hx
}  // namespace vhmm
