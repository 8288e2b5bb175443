// hmm_inputs.cu -- the front end of the exact-HMM stages, fused: from raw (x, u) to the materialised
// log_A (B,T,K,K) and log_b (B,T,K) the recursion kernels consume.
//
//   log_A[b,t] = log_softmax_rows( W2 relu(W0 u[b,:,t] + b0) + b2 )      Prior.forward, VQ_VAE_HMM_fixed.py:59-71
//   log_b[b,t,k] = -0.5 sum_c [ log(2 pi v) + (x - mu)^2 / v ],  v = max(exp(logvar), 1e-8)   (:118-119)
//
// The layer-wise path needs three launches that stage the 128-wide hidden layer through HBM (95 ms at
// B = 2^20, T = 256) plus an emission kernel that re-evaluates exp/log of the (B-independent) table for
// every sequence (9.8 ms).  Here one thread owns 4 consecutive steps of one sequence: u and x come in with
// 16-byte loads, the hidden units are produced one at a time in registers (never stored) and immediately
// consumed by the K*K logits, so per hidden unit 2 + PK2/4 broadcast LDS.128 feed 4 (U + 1 + K*K) FP32
// operations -- FFMA bound (1,664 MAC per step), no activation traffic at all.  The emission term reads a
// pre-reduced table (mu, 1/v, log(2 pi v)) built once per call by a tiny kernel.
#include <atomic>
#include "common.cuh"
#include "hmm_warp.cuh"

namespace vhmm {

template <int K>
struct HiCfg {
  static constexpr int K2 = K * K;
  static constexpr int PK2 = (K2 + 3) & ~3;
};

// tab3 = [3][K][C][T]: mu, 1/v, log(2 pi v)
__global__ void __launch_bounds__(256)
emission_tab3_kernel(const float* __restrict__ mu, const float* __restrict__ lv, int64_t n, float* __restrict__ tab3) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  const float v = fmaxf(expf(lv[i]), 1e-8f);
  tab3[i] = mu[i];
  tab3[n + i] = 1.f / v;
  tab3[2 * n + i] = logf(6.283185307179586f * v);
}

// shared-memory image of the transition MLP: sW0[h][0..U-1] weights, [h][U] bias; sW2[h][r]; sb2[r]
template <int K, int U, int TH>
__device__ __forceinline__ void hi_stage_weights(const float* __restrict__ W0, const float* __restrict__ b0,
                                                 const float* __restrict__ W2, const float* __restrict__ b2,
                                                 float* sW0, float* sW2, float* sb2) {
  constexpr int K2 = HiCfg<K>::K2, PK2 = HiCfg<K>::PK2;
  for (int i = threadIdx.x; i < TH * 8; i += blockDim.x) {
    const int h = i >> 3, c = i & 7;
    sW0[i] = (c < U) ? W0[h * U + c] : (c == U ? b0[h] : 0.f);
  }
  for (int i = threadIdx.x; i < TH * PK2; i += blockDim.x) {
    const int h = i / PK2, r = i - h * PK2;
    sW2[i] = (r < K2) ? W2[r * TH + h] : 0.f;
  }
  if (threadIdx.x < PK2) sb2[threadIdx.x] = (threadIdx.x < K2) ? b2[threadIdx.x] : 0.f;
}

// log_A of 4 consecutive steps: out[s*K2 + i*K + j] = log_softmax_j( W2 relu(W0 u_s + b0) + b2 )[i,j].
// (A packed fma.rn.f32x2 form of this loop -- steps paired in the halves of 64-bit registers, weights
// duplicated in shared memory -- was measured and dropped: FFMA2 halves the issue slots but not the FMA-pipe
// cycles (tools/micro/ffma2_probe.cu: 131.7 vs 128.4 FMA/clk/SM) and its two-chain first layer is latency
// bound; 24.3 ms against 22.7 ms for this scalar form at B = 2^20, T = 256.)
template <int K, int U, int TH>
__device__ __forceinline__ void hi_log_A_quad(const float* sW0, const float* sW2, const float* sb2,
                                              const float* __restrict__ urow /* u + (b*U)*T + t */, int T, float* out) {
  constexpr int K2 = HiCfg<K>::K2, PK2 = HiCfg<K>::PK2;
  float uv[U][4];
#pragma unroll
  for (int c = 0; c < U; ++c) {
    const float4 v = __ldcs(reinterpret_cast<const float4*>(urow + c * (int64_t)T));
    uv[c][0] = v.x; uv[c][1] = v.y; uv[c][2] = v.z; uv[c][3] = v.w;
  }
  float M[4][PK2];
#pragma unroll
  for (int s = 0; s < 4; ++s)
#pragma unroll
    for (int r = 0; r < PK2; ++r) M[s][r] = sb2[r];
#pragma unroll 2
  for (int h = 0; h < TH; ++h) {
    const float4 wa = *reinterpret_cast<const float4*>(sW0 + h * 8);
    const float4 wb = *reinterpret_cast<const float4*>(sW0 + h * 8 + 4);
    const float w0[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
    float a[4];
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      float p = w0[U];
#pragma unroll
      for (int c = 0; c < U; ++c) p = fmaf(w0[c], uv[c][s], p);
      a[s] = fmaxf(p, 0.f);
    }
    float w2[PK2];
#pragma unroll
    for (int g = 0; g < PK2 / 4; ++g) {
      const float4 v = *reinterpret_cast<const float4*>(sW2 + h * PK2 + 4 * g);
      w2[4 * g] = v.x; w2[4 * g + 1] = v.y; w2[4 * g + 2] = v.z; w2[4 * g + 3] = v.w;
    }
#pragma unroll
    for (int s = 0; s < 4; ++s)
#pragma unroll
      for (int r = 0; r < K2; ++r) M[s][r] = fmaf(w2[r], a[s], M[s][r]);
  }
#pragma unroll
  for (int s = 0; s < 4; ++s)
#pragma unroll
    for (int i = 0; i < K; ++i) {
      float m = M[s][i * K];
#pragma unroll
      for (int j = 1; j < K; ++j) m = fmaxf(m, M[s][i * K + j]);
      float sum = 0.f;
#pragma unroll
      for (int j = 0; j < K; ++j) sum += expf(M[s][i * K + j] - m);
      const float lse = m + logf(sum);
#pragma unroll
      for (int j = 0; j < K; ++j) out[s * K2 + i * K + j] = M[s][i * K + j] - lse;
    }
}

// log_b of 4 consecutive steps: ob[s*K + k]
template <int K, int C>
__device__ __forceinline__ void hi_log_b_quad(const float* __restrict__ xrow /* x + (b*C)*T + t */, int T,
                                              const float* __restrict__ tab3, int64_t nkct, int t, float* ob) {
  float lb[4][K];
#pragma unroll
  for (int s = 0; s < 4; ++s)
#pragma unroll
    for (int k = 0; k < K; ++k) lb[s][k] = 0.f;
#pragma unroll
  for (int c = 0; c < C; ++c) {
    const float4 xv4 = __ldcs(reinterpret_cast<const float4*>(xrow + c * (int64_t)T));
    const float xv[4] = {xv4.x, xv4.y, xv4.z, xv4.w};
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const int64_t o = ((int64_t)k * C + c) * T + t;
      const float4 mu = __ldg(reinterpret_cast<const float4*>(tab3 + o));
      const float4 iv = __ldg(reinterpret_cast<const float4*>(tab3 + nkct + o));
      const float4 lc = __ldg(reinterpret_cast<const float4*>(tab3 + 2 * nkct + o));
      const float m4[4] = {mu.x, mu.y, mu.z, mu.w}, i4[4] = {iv.x, iv.y, iv.z, iv.w}, l4[4] = {lc.x, lc.y, lc.z, lc.w};
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        const float d = xv[s] - m4[s];
        lb[s][k] += fmaf(d * d, i4[s], l4[s]);
      }
    }
  }
#pragma unroll
  for (int s = 0; s < 4; ++s)
#pragma unroll
    for (int k = 0; k < K; ++k) ob[s * K + k] = -0.5f * lb[s][k];
}

template <int K, int C, int U, int TH>
__global__ void __launch_bounds__(256)
hmm_inputs_kernel(const float* __restrict__ W0, const float* __restrict__ b0, const float* __restrict__ W2,
                  const float* __restrict__ b2, const float* __restrict__ x, const float* __restrict__ u,
                  int64_t nquads, int T, const float* __restrict__ tab3, float* __restrict__ log_A,
                  float* __restrict__ log_b) {
  constexpr int K2 = HiCfg<K>::K2;
  static_assert(U <= 7, "W0 row + bias are packed into 8 floats");
  __shared__ __align__(16) float sW0[TH * 8];
  __shared__ __align__(16) float sW2[TH * HiCfg<K>::PK2];
  __shared__ float sb2[HiCfg<K>::PK2];
  hi_stage_weights<K, U, TH>(W0, b0, W2, b2, sW0, sW2, sb2);
  __syncthreads();
  const int T4 = T >> 2;
  const int64_t nkct = (int64_t)K * C * T;
  for (int64_t q = (int64_t)blockIdx.x * 256 + threadIdx.x; q < nquads; q += (int64_t)gridDim.x * 256) {
    const int64_t b = q / T4;
    const int t = (int)(q - b * T4) * 4;
    float out[4 * K2], ob[4 * K];
    hi_log_A_quad<K, U, TH>(sW0, sW2, sb2, u + (b * U) * (int64_t)T + t, T, out);
    float4* dA = reinterpret_cast<float4*>(log_A + (b * T + t) * (int64_t)K2);   // 16-byte aligned: 4 K2 floats per quad
#pragma unroll
    for (int i = 0; i < K2; ++i) __stcs(dA + i, make_float4(out[4 * i], out[4 * i + 1], out[4 * i + 2], out[4 * i + 3]));
    hi_log_b_quad<K, C>(x + (b * C) * (int64_t)T + t, T, tab3, nkct, t, ob);
    float4* dB = reinterpret_cast<float4*>(log_b + (b * T + t) * (int64_t)K);
#pragma unroll
    for (int i = 0; i < K; ++i) __stcs(dB + i, make_float4(ob[4 * i], ob[4 * i + 1], ob[4 * i + 2], ob[4 * i + 3]));
  }
}

// log_A (or A = exp(log_A): PROBS) alone, from u: Prior.forward (VQ_VAE_HMM_fixed.py:59-71) and the `transition_probs` accessor
// without the layer-wise path's hidden layer in HBM (B*TH*T floats written and read back: 1.6 ms at B = 64, T = 65536)
template <int K, int U, int TH, bool PROBS>
__global__ void __launch_bounds__(256)
transitions_kernel(const float* __restrict__ W0, const float* __restrict__ b0, const float* __restrict__ W2,
                   const float* __restrict__ b2, const float* __restrict__ u, int64_t nquads, int T, float* __restrict__ out_A) {
  constexpr int K2 = HiCfg<K>::K2;
  static_assert(U <= 7, "W0 row + bias are packed into 8 floats");
  __shared__ __align__(16) float sW0[TH * 8];
  __shared__ __align__(16) float sW2[TH * HiCfg<K>::PK2];
  __shared__ float sb2[HiCfg<K>::PK2];
  hi_stage_weights<K, U, TH>(W0, b0, W2, b2, sW0, sW2, sb2);
  __syncthreads();
  const int T4 = T >> 2;
  for (int64_t q = (int64_t)blockIdx.x * 256 + threadIdx.x; q < nquads; q += (int64_t)gridDim.x * 256) {
    const int64_t b = q / T4;
    const int t = (int)(q - b * T4) * 4;
    float out[4 * K2];
    hi_log_A_quad<K, U, TH>(sW0, sW2, sb2, u + (b * U) * (int64_t)T + t, T, out);
    if (PROBS) {
#pragma unroll
      for (int i = 0; i < 4 * K2; ++i) out[i] = expf(out[i]);
    }
    float4* dA = reinterpret_cast<float4*>(out_A + (b * T + t) * (int64_t)K2);
#pragma unroll
    for (int i = 0; i < K2; ++i) __stcs(dA + i, make_float4(out[4 * i], out[4 * i + 1], out[4 * i + 2], out[4 * i + 3]));
  }
}

// ---- fused transition-softmax + emission-logprob + forward recursion -----------------------------------
// log-likelihood of every sequence straight from (x, u): one WARP per sequence, lane l owns the steps
// [l*S, l*S+S) (S = 4 ceil(T/128)); per 4-step quad it forms log_A and log_b in registers (above), turns them
// into the scaled transfer matrices P_t = exp(log_A_t + log_b_t - max log_b_t), multiplies them into its chunk
// product, and the warp-shuffle prefix scan of hmm_warp.cuh yields log p(x | u).  Neither log_A nor log_b nor
// alpha ever leaves the registers; HBM traffic is the 36 B/step of x and u.
template <int K, int C, int U, int TH>
__global__ void __launch_bounds__(128, 4)
hmm_loglik_fused_kernel(const float* __restrict__ W0, const float* __restrict__ b0, const float* __restrict__ W2,
                        const float* __restrict__ b2, const float* __restrict__ log_pi, const float* __restrict__ x,
                        const float* __restrict__ u, const int64_t* __restrict__ lengths, int64_t B, int T,
                        const float* __restrict__ tab3, float* __restrict__ loglik) {
  constexpr int K2 = HiCfg<K>::K2;
  constexpr float LOG2E = 1.4426950408889634f, LN2 = 0.6931471805599453f;
  __shared__ __align__(16) float sW0[TH * 8];
  __shared__ __align__(16) float sW2[TH * HiCfg<K>::PK2];
  __shared__ float sb2[HiCfg<K>::PK2];
  hi_stage_weights<K, U, TH>(W0, b0, W2, b2, sW0, sW2, sb2);
  __syncthreads();
  const int lane = threadIdx.x & 31, wpb = blockDim.x >> 5, wib = threadIdx.x >> 5;
  const int S = 4 * ((T + 127) / 128);
  const int t0 = lane * S;
  const int64_t nkct = (int64_t)K * C * T;
  float pi2[K];
#pragma unroll
  for (int j = 0; j < K; ++j) pi2[j] = log_pi[j] * LOG2E;
  for (int64_t b = (int64_t)blockIdx.x * wpb + wib; b < B; b += (int64_t)gridDim.x * wpb) {
    int64_t Ll = lengths ? lengths[b] : T;
    const int L = (int)(Ll < 0 ? 0 : (Ll > T ? T : Ll));
    const int nv = max(0, min(S, L - t0));
    float Mc[K2], v0[K];
#pragma unroll
    for (int i = 0; i < K2; ++i) Mc[i] = (i / K == i % K) ? 1.f : 0.f;
#pragma unroll
    for (int j = 0; j < K; ++j) v0[j] = 0.f;
    int ex = 0;
    float mbsum = 0.f;
    for (int g = 0; g * 4 < nv; ++g) {
      const int t = t0 + 4 * g;
      float la[4 * K2], lb[4 * K];
      hi_log_b_quad<K, C>(x + (b * C) * (int64_t)T + t, T, tab3, nkct, t, lb);
      hi_log_A_quad<K, U, TH>(sW0, sW2, sb2, u + (b * U) * (int64_t)T + t, T, la);
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        if (4 * g + s < nv) {
          const float mb = hw::max_of<K>(lb + s * K);
          mbsum += mb;
          float cj[K];
#pragma unroll
          for (int j = 0; j < K; ++j) cj[j] = (lb[s * K + j] - mb) * LOG2E;
          if (t + s == 0) {
#pragma unroll
            for (int j = 0; j < K; ++j) v0[j] = hw::ex2f(pi2[j] + cj[j]);
          } else {
            float P[K2], R[K2];
#pragma unroll
            for (int i = 0; i < K; ++i)
#pragma unroll
              for (int j = 0; j < K; ++j) P[i * K + j] = hw::ex2f(fmaf(la[s * K2 + i * K + j], LOG2E, cj[j]));
            hw::mat_mul<K>(R, Mc, P);
#pragma unroll
            for (int i = 0; i < K2; ++i) Mc[i] = R[i];
          }
        }
        if (s & 1) {
          const float sc = hw::pow2_rescale(hw::max_of<K2>(Mc), &ex);
#pragma unroll
          for (int i = 0; i < K2; ++i) Mc[i] *= sc;
        }
      }
    }
    float pre[K];
    int pex;
    hw::warp_prefix_scan<K>(Mc, ex, v0, lane, pre, pex);
    const double ms = warp_sum((double)mbsum);
    if (lane == 31) {
      float sm = pre[0];
#pragma unroll
      for (int j = 1; j < K; ++j) sm += pre[j];
      loglik[b] = (L > 0) ? (float)((double)logf(sm) + (double)pex * (double)LN2 + ms) : 0.f;
    }
  }
}

bool hmm_inputs_supported(const vhmm_dims& d, int64_t T) {
  return d.K == 3 && d.C == 5 && d.U == 4 && d.TH == 128 && (T % 4) == 0;
}

bool hmm_transitions_supported(const vhmm_dims& d, int64_t T, const float* u, const float* out) {
  return d.K == 3 && d.U == 4 && d.TH == 128 && (T % 4) == 0 && ((((uintptr_t)u | (uintptr_t)out) & 15) == 0);
}

int hmm_transitions(const vhmm_dims& d, const vhmm_params& P, const float* u, int64_t B, int64_t T, float* out, int as_probs,
                    cudaStream_t st) {
  VHMM_REQUIRE(hmm_transitions_supported(d, T, u, out), VHMM_E_DIMS, "hmm_transitions: needs K=3,U=4,TH=128, T %% 4 == 0, 16-byte aligned u / out");
  const int64_t nquads = B * (T / 4);
  int64_t grid = cdiv(nquads, 256);
  if (grid > 148 * 8) grid = 148 * 8;
  if (as_probs)
    transitions_kernel<3, 4, 128, true><<<(unsigned)grid, 256, 0, st>>>(P.p[P_T0W], P.p[P_T0B], P.p[P_T2W], P.p[P_T2B], u, nquads, (int)T, out);
  else
    transitions_kernel<3, 4, 128, false><<<(unsigned)grid, 256, 0, st>>>(P.p[P_T0W], P.p[P_T0B], P.p[P_T2W], P.p[P_T2B], u, nquads, (int)T, out);
  VHMM_LAUNCH_CHECK();
  return 0;
}

size_t hmm_inputs_workspace_bytes(const vhmm_dims& d, int64_t T) {
  return align_up((size_t)3 * d.K * d.C * T * sizeof(float), 256) + 256;
}

int hmm_inputs(const vhmm_dims& d, const vhmm_params& P, const float* x, const float* u, int64_t B, int64_t T,
               const float* mu_tab, const float* logvar_tab, float* log_A, float* log_b, void* ws, size_t ws_bytes,
               cudaStream_t st) {
  VHMM_REQUIRE(hmm_inputs_supported(d, T), VHMM_E_DIMS, "hmm_inputs: fused front end needs C=5,K=3,U=4,TH=128 and T %% 4 == 0");
  VHMM_REQUIRE((((uintptr_t)x | (uintptr_t)u | (uintptr_t)log_A | (uintptr_t)log_b) & 15) == 0, VHMM_E_ALIGN,
               "hmm_inputs: x, u, log_A, log_b must be 16-byte aligned");
  Arena ar(ws, ws_bytes);
  const int64_t n = (int64_t)d.K * d.C * T;
  float* tab3 = ar.take<float>((size_t)3 * n);
  VHMM_REQUIRE(ar.ok, VHMM_E_WORKSPACE, "hmm_inputs: workspace %zu < %zu", ws_bytes, ar.used);
  emission_tab3_kernel<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(mu_tab, logvar_tab, n, tab3);
  VHMM_LAUNCH_CHECK();
  const int64_t nquads = B * (T / 4);
  int64_t grid = cdiv(nquads, 256);
  if (grid > 148 * 8) grid = 148 * 8;
  hmm_inputs_kernel<3, 5, 4, 128><<<(unsigned)grid, 256, 0, st>>>(P.p[P_T0W], P.p[P_T0B], P.p[P_T2W], P.p[P_T2B], x, u,
                                                                  nquads, (int)T, tab3, log_A, log_b);
  VHMM_LAUNCH_CHECK();
  return 0;
}

int hmm_loglik_fused(const vhmm_dims& d, const vhmm_params& P, const float* log_pi, const float* x, const float* u,
                     const int64_t* lengths, int64_t B, int64_t T, const float* mu_tab, const float* logvar_tab,
                     float* loglik, void* ws, size_t ws_bytes, cudaStream_t st) {
  VHMM_REQUIRE(hmm_inputs_supported(d, T), VHMM_E_DIMS, "hmm_loglik_fused: needs C=5,K=3,U=4,TH=128 and T %% 4 == 0");
  VHMM_REQUIRE((((uintptr_t)x | (uintptr_t)u) & 15) == 0, VHMM_E_ALIGN, "hmm_loglik_fused: x, u must be 16-byte aligned");
  Arena ar(ws, ws_bytes);
  const int64_t n = (int64_t)d.K * d.C * T;
  float* tab3 = ar.take<float>((size_t)3 * n);
  VHMM_REQUIRE(ar.ok, VHMM_E_WORKSPACE, "hmm_loglik_fused: workspace %zu < %zu", ws_bytes, ar.used);
  emission_tab3_kernel<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(mu_tab, logvar_tab, n, tab3);
  VHMM_LAUNCH_CHECK();
  // persistent grid = exactly the resident CTAs (a partial second wave of equal-work CTAs would double the time)
  static std::atomic<int> per_sm_dev[VHMM_MAX_DEVICES];   // resident CTAs per SM, per device
  const int sms = device_sm_count();
  int per_sm = per_sm_dev[current_device()].load(std::memory_order_relaxed);
  if (!per_sm) {
    int n_cta = 0;
    VHMM_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n_cta, hmm_loglik_fused_kernel<3, 5, 4, 128>, 128, 0));
    per_sm = n_cta < 1 ? 1 : n_cta;
    per_sm_dev[current_device()].store(per_sm, std::memory_order_relaxed);
  }
  int64_t grid = cdiv(B, 4);
  if (grid > (int64_t)sms * per_sm) grid = (int64_t)sms * per_sm;
  hmm_loglik_fused_kernel<3, 5, 4, 128><<<(unsigned)grid, 128, 0, st>>>(P.p[P_T0W], P.p[P_T0B], P.p[P_T2W], P.p[P_T2B], log_pi, x, u,
                                                                        lengths, B, (int)T, tab3, loglik);
  VHMM_LAUNCH_CHECK();
  return 0;
}

}  // namespace vhmm
