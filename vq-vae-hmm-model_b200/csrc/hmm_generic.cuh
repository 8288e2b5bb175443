// hmm_generic.cuh -- exact-HMM recursions for any K <= 32: one WARP per sequence, lane = state.
// Log-space, straight from global memory (rows of log_A are contiguous in the lane index, so
// every load is coalesced).  This is the correctness path for state counts without a tuned
// kernel in hmm_small.cuh (e.g. the scaled config's K = 32); maths: SURVEY.md Appendix B.
#pragma once
#include "common.cuh"

namespace vhmm {
namespace hg {

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// logsumexp over lanes < K of v (lanes >= K must pass -inf)
__device__ __forceinline__ float warp_lse(float v) {
  const float m = warp_max(v);
  const float mm = (m == -INFINITY) ? 0.f : m;
  const float s = warp_sum(expf(v - mm));
  return mm + logf(s);
}

// forward recursion; optional alpha (B,T,K) log-space (also the stash for the backward pass).
// VIT: max-plus recursion with back-pointers psi (B,T,K) uint8 and back-tracking by lane 0.
template <bool VIT>
__global__ void __launch_bounds__(128)
fwd_kernel(const float* __restrict__ log_pi, const float* __restrict__ A, const float* __restrict__ Bm,
           const int64_t* __restrict__ lengths, int64_t nB, int T, int K, float* __restrict__ loglik,
           float* __restrict__ alpha, int stash, uint8_t* __restrict__ psi, uint8_t* __restrict__ path,
           float* __restrict__ score) {
  const int lane = threadIdx.x & 31;
  const int64_t b = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (b >= nB) return;
  int64_t L64 = lengths[b];
  const int L = (int)(L64 < 0 ? 0 : (L64 > T ? T : L64));
  const bool on = lane < K;
  const float* Ab = A + b * (int64_t)T * K * K;
  const float* Bb = Bm + b * (int64_t)T * K;
  // sum-product: `al` holds the NORMALISED log forward variable (logsumexp_k al = 0) and `ll`
  // the running log normaliser in double, so magnitudes stay O(10) and fp32 keeps ~1e-7.
  // `stash` != 0 stores the normalised values (for bwd_kernel), else alpha = al + ll (B.2).
  float al = -INFINITY;
  double ll = 0.0;
  if (L > 0 && on) al = log_pi[lane] + Bb[lane];
  if (!VIT && L > 0) {
    const float z = warp_lse(on ? al : -INFINITY);
    al -= z;
    ll = (double)z;
    if (alpha && on) alpha[(b * T) * K + lane] = stash ? al : al + z;
  }
  for (int t = 1; t < L; ++t) {
    const float* At = Ab + (int64_t)t * K * K;
    float nv;
    if (!VIT) {
      float m = -INFINITY;
      for (int i = 0; i < K; ++i) {
        const float ai = __shfl_sync(0xffffffffu, al, i);
        if (on) m = fmaxf(m, ai + At[i * K + lane]);
      }
      float s = 0.f;
      const float mm = (m == -INFINITY) ? 0.f : m;
      for (int i = 0; i < K; ++i) {
        const float ai = __shfl_sync(0xffffffffu, al, i);
        if (on) s += expf(ai + At[i * K + lane] - mm);
      }
      nv = on ? Bb[(int64_t)t * K + lane] + mm + logf(s) : -INFINITY;
      const float z = warp_lse(nv);
      nv -= z;
      ll += (double)z;
      if (alpha && on) alpha[(b * T + t) * K + lane] = stash ? nv : nv + (float)ll;
    } else {
      float best = -INFINITY;
      int arg = 0;
      for (int i = 0; i < K; ++i) {
        const float di = __shfl_sync(0xffffffffu, al, i);
        if (on) {
          const float c = __fadd_rn(di, At[i * K + lane]);
          if (i == 0 || c > best) { best = c; arg = i; }
        }
      }
      nv = on ? __fadd_rn(Bb[(int64_t)t * K + lane], best) : -INFINITY;
      if (on) psi[(b * T + t) * K + lane] = (uint8_t)arg;
    }
    al = nv;
  }
  if (!VIT) {
    if (lane == 0 && loglik) loglik[b] = (L > 0) ? (float)ll : 0.f;
    // carried values for t >= L (oracle.forward keeps alpha unchanged)
    if (alpha && on && !stash)
      for (int t = (L > 0 ? L : 0); t < T; ++t) alpha[(b * T + t) * K + lane] = al + (float)ll;
  } else {
    // final state: lowest index attaining the max
    const float v = on ? al : -INFINITY;
    const float m = warp_max(v);
    const unsigned ball = __ballot_sync(0xffffffffu, on && v == m);
    int state = __ffs(ball) - 1;
    if (lane == 0) {
      if (score) score[b] = (L > 0) ? m : 0.f;
      for (int t = T - 1; t >= L; --t) path[b * T + t] = VHMM_PAD_STATE;
    }
    __syncwarp();
    if (lane == 0) {
      for (int t = L - 1; t >= 0; --t) {
        path[b * T + t] = (uint8_t)state;
        if (t >= 1) state = psi[(b * T + t) * K + state];
      }
    }
  }
}

// backward recursion with gamma / xi from the log-space alpha stash
__global__ void __launch_bounds__(128)
bwd_kernel(const float* __restrict__ A, const float* __restrict__ Bm, const int64_t* __restrict__ lengths,
           int64_t nB, int T, int K, const float* __restrict__ alpha /* normalised stash */,
           float* __restrict__ gamma, float* __restrict__ xi) {
  const int lane = threadIdx.x & 31;
  const int64_t b = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (b >= nB) return;
  int64_t L64 = lengths[b];
  const int L = (int)(L64 < 0 ? 0 : (L64 > T ? T : L64));
  const bool on = lane < K;
  const float* Ab = A + b * (int64_t)T * K * K;
  const float* Bb = Bm + b * (int64_t)T * K;
  for (int t = T - 1; t >= L; --t) {
    if (on) gamma[(b * T + t) * K + lane] = 0.f;
    if (xi) for (int e = lane; e < K * K; e += 32) xi[(b * T + t) * (int64_t)K * K + e] = 0.f;
  }
  float be = on ? 0.f : -INFINITY;   // beta_{L-1} = 0
  for (int t = L - 1; t >= 0; --t) {
    const float at = on ? alpha[(b * T + t) * K + lane] : -INFINITY;
    {
      const float gl = on ? at + be : -INFINITY;
      const float zg = warp_lse(gl);
      if (on) gamma[(b * T + t) * K + lane] = expf(gl - zg);
    }
    if (t == 0) {
      if (xi) for (int e = lane; e < K * K; e += 32) xi[(b * T) * (int64_t)K * K + e] = 0.f;
      break;
    }
    const float* At = Ab + (int64_t)t * K * K;
    const float w = on ? Bb[(int64_t)t * K + lane] + be : -INFINITY;  // b_t[j] + beta_t[j], lane = j
    const float ap = on ? alpha[(b * T + t - 1) * K + lane] : -INFINITY;  // alpha_{t-1}[i], lane = i
    // xi_t[i,j]: lane = j, loop i
    if (xi) {
      float mx = -INFINITY;
      for (int i = 0; i < K; ++i) {
        const float ai = __shfl_sync(0xffffffffu, ap, i);
        if (on) mx = fmaxf(mx, ai + At[i * K + lane]);
      }
      const float mmx = (mx == -INFINITY) ? 0.f : mx;
      float sx = 0.f;
      for (int i = 0; i < K; ++i) {
        const float ai = __shfl_sync(0xffffffffu, ap, i);
        if (on) sx += expf(ai + At[i * K + lane] - mmx);
      }
      const float zx = warp_lse(on ? mmx + logf(sx) + w : -INFINITY);   // log sum_ij of xi terms
      for (int i = 0; i < K; ++i) {
        const float ai = __shfl_sync(0xffffffffu, ap, i);
        if (on) xi[((b * T + t) * (int64_t)K + i) * K + lane] = expf(ai + At[i * K + lane] + w - zx);
      }
    }
    // beta_{t-1}[i] = lse_j(A[i,j] + w[j]): lane = i reads row i
    float m = -INFINITY;
    for (int j = 0; j < K; ++j) {
      const float wj = __shfl_sync(0xffffffffu, w, j);
      if (on) m = fmaxf(m, At[lane * K + j] + wj);
    }
    const float mm = (m == -INFINITY) ? 0.f : m;
    float s = 0.f;
    for (int j = 0; j < K; ++j) {
      const float wj = __shfl_sync(0xffffffffu, w, j);
      if (on) s += expf(At[lane * K + j] + wj - mm);
    }
    be = on ? mm + logf(s) : -INFINITY;
    be -= warp_lse(be);   // keep beta normalised too (gamma / xi are renormalised per step)
  }
}

}  // namespace hg
}  // namespace vhmm
