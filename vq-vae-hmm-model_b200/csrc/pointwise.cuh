// pointwise.cuh -- per-(sequence, step) kernels of the layer-wise path: softmax, transition
// row log-softmax, ELBO terms + gradient seeds, d logits, small deterministic reductions.
// Formulas: SURVEY.md Appendix A.2 (forward) and VQ_VAE_HMM_fixed.py:111-137 (loss).
#pragma once
#include "common.cuh"

namespace vhmm {

constexpr int KMAX = 32;  // max states handled by the layer-wise pointwise kernels

// q = softmax_k(logits), logq = log_softmax_k(logits); (B,K,T) layout; one thread per (b,t).
// VQ_VAE_HMM_fixed.py:114 and :134.  q / logq may be null.
__global__ void softmax_kernel(const float* __restrict__ logits, float* __restrict__ q,
                               float* __restrict__ logq, int K, int T) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  const int64_t base = (int64_t)blockIdx.y * K * T + t;
  float m = -INFINITY;
  for (int k = 0; k < K; ++k) m = fmaxf(m, logits[base + (int64_t)k * T]);
  float s = 0.f;
  for (int k = 0; k < K; ++k) s += expf(logits[base + (int64_t)k * T] - m);
  const float ls = logf(s), inv = 1.f / s;
  for (int k = 0; k < K; ++k) {
    const float z = logits[base + (int64_t)k * T] - m;
    if (q) q[base + (int64_t)k * T] = expf(z) * inv;
    if (logq) logq[base + (int64_t)k * T] = z - ls;
  }
}

// in-place row log-softmax of transition logits stored as (B, K(i), K(j), T):
// M[b,i,:,t] -= logsumexp_j.  One thread per (b,i,t).  VQ_VAE_HMM_fixed.py:69.
__global__ void row_logsoftmax_kernel(float* __restrict__ M, int K, int T) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  const int64_t base = ((int64_t)blockIdx.z * K + blockIdx.y) * K * (int64_t)T + t;
  float m = -INFINITY;
  for (int j = 0; j < K; ++j) m = fmaxf(m, M[base + (int64_t)j * T]);
  float s = 0.f;
  for (int j = 0; j < K; ++j) s += expf(M[base + (int64_t)j * T] - m);
  const float lse = m + logf(s);
  for (int j = 0; j < K; ++j) M[base + (int64_t)j * T] -= lse;
}

// (B,K*K,T) raw transition logits -> log_A (B,T,K,K) with row log-softmax (public prior()).
// grid (ceil(T/128), K(i), B); block 128.
__global__ void logA_out_kernel(const float* __restrict__ M, float* __restrict__ log_A, int K, int T, int as_probs) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  const int i = blockIdx.y;
  const int64_t b = blockIdx.z;
  const int64_t base = (b * K + i) * K * (int64_t)T + t;
  float m = -INFINITY;
  for (int j = 0; j < K; ++j) m = fmaxf(m, M[base + (int64_t)j * T]);
  float s = 0.f;
  for (int j = 0; j < K; ++j) s += expf(M[base + (int64_t)j * T] - m);
  const float lse = m + logf(s);
  float* o = log_A + ((b * T + t) * K + i) * (int64_t)K;
  for (int j = 0; j < K; ++j) {
    const float v = M[base + (int64_t)j * T] - lse;
    o[j] = as_probs ? expf(v) : v;      // as_probs: A = exp(log_A), the (B,T,K,K) tensor of regime_utilities.py:42-53
  }
}

// log_pi = log_softmax(log_prior)   (K <= KMAX; single warp)
__global__ void log_pi_kernel(const float* __restrict__ log_prior, float* __restrict__ log_pi, int K) {
  const int k = threadIdx.x;
  float v = (k < K) ? log_prior[k] : -INFINITY;
  float m = v;
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  float e = (k < K) ? expf(v - m) : 0.f;
  e = warp_sum(e);
  if (k < K) log_pi[k] = v - m - logf(e);
}

// ELBO terms and gradient seeds; one thread per (b,t); grid (ceil(T/256), B), block 256.
//   p    (B,2C,T) decoder output (mu rows then logvar rows)      q,(B,K,T)  logq (B,K,T)
//   logA (B,K,K,T) row-normalised                                lengths (B) int64
// out: partial[blk*4 + {0,1,2,3}] = block sums of {nll*mask, trans term, entropy*mask, valid}
//      dp (B,2C,T), dq_prior (B,K,T), dM (B,K,K,T)   (only when dp != null)
__global__ void __launch_bounds__(256)
loss_seed_kernel(const float* __restrict__ x, const float* __restrict__ p,
                 const float* __restrict__ q, const float* __restrict__ logq,
                 const float* __restrict__ logA, const float* __restrict__ log_pi,
                 const int64_t* __restrict__ lengths, const double* __restrict__ norms, float beta,
                 int C, int K, int T, double* __restrict__ partial, float* __restrict__ dp,
                 float* __restrict__ dq_prior, float* __restrict__ dM) {
  __shared__ double red[32];
  const int t = blockIdx.x * 256 + threadIdx.x;
  const int64_t b = blockIdx.y;
  const int64_t L = lengths[b];
  const bool in = t < T;
  const bool m = in && (int64_t)t < L;
  const float aR = (float)(1.0 / norms[0]);
  const float aP = (float)(-(double)beta / norms[1]);
  float nll_sum = 0.f, ent = 0.f, trans = 0.f;
  if (in) {
    // reconstruction, VQ_VAE_HMM_fixed.py:118-120
    for (int c = 0; c < C; ++c) {
      const float mu = p[(b * 2 * C + c) * (int64_t)T + t];
      const float lv = p[(b * 2 * C + C + c) * (int64_t)T + t];
      const float xv = x[(b * C + c) * (int64_t)T + t];
      const float ev = expf(lv);
      const float v = fmaxf(ev, 1e-8f);
      const float d = mu - xv;
      const float r = d * d / v;
      if (m) nll_sum += 0.5f * (logf(6.283185307179586f * v) + r);
      if (dp) {
        dp[(b * 2 * C + c) * (int64_t)T + t] = m ? aR * d / v : 0.f;
        dp[(b * 2 * C + C + c) * (int64_t)T + t] = (m && ev >= 1e-8f) ? aR * 0.5f * (1.f - r) : 0.f;
      }
    }
    float qc[KMAX], qp[KMAX], qn[KMAX];
    const bool tm = m && t >= 1;                        // :128 (mask is a prefix)
    const bool tm_next = (t + 1 < T) && (int64_t)(t + 1) < L;
    for (int k = 0; k < K; ++k) {
      const int64_t o = (b * K + k) * (int64_t)T + t;
      qc[k] = q[o];
      qp[k] = tm ? q[o - 1] : 0.f;
      qn[k] = tm_next ? q[o + 1] : 0.f;
      if (m) ent -= qc[k] * logq[o];                    // :134
    }
    float dq[KMAX];
    for (int k = 0; k < K; ++k) dq[k] = (t == 0) ? log_pi[k] : 0.f;   // :123
    for (int i = 0; i < K; ++i) {
      for (int j = 0; j < K; ++j) {
        const int64_t o = ((b * K + i) * K + j) * (int64_t)T + t;
        if (tm) {
          const float la = logA[o];
          trans += qp[i] * qc[j] * la;                  // :127
          dq[j] += qp[i] * la;
          if (dM) dM[o] = aP * qp[i] * (qc[j] - expf(la));
        } else if (dM) {
          dM[o] = 0.f;
        }
        if (tm_next) dq[i] += qn[j] * logA[o + 1];
      }
    }
    if (dq_prior)
      for (int k = 0; k < K; ++k) dq_prior[(b * K + k) * (int64_t)T + t] = aP * dq[k];
  }
  const int64_t blk = (int64_t)blockIdx.y * gridDim.x + blockIdx.x;
  double s0 = block_sum<double>((double)nll_sum, red);
  double s1 = block_sum<double>((double)trans, red);
  double s2 = block_sum<double>((double)ent, red);
  double s3 = block_sum<double>(m ? 1.0 : 0.0, red);
  if (threadIdx.x == 0) {
    partial[blk * 4 + 0] = s0; partial[blk * 4 + 1] = s1;
    partial[blk * 4 + 2] = s2; partial[blk * 4 + 3] = s3;
  }
}

// sums[j] += sum_blk partial[blk*4+j]  (one block of 1024 threads, fixed order)
__global__ void __launch_bounds__(1024)
sum_partials_kernel(const double* __restrict__ partial, int64_t nblk, double* __restrict__ sums) {
  __shared__ double red[32];
  for (int j = 0; j < 4; ++j) {
    double s = 0.0;
    for (int64_t i = threadIdx.x; i < nblk; i += 1024) s += partial[i * 4 + j];
    s = block_sum<double>(s, red);
    if (threadIdx.x == 0) sums[j] += s;
  }
}

// qsum[k] += sum_b q[b,k,0]   (one block of 1024 threads)
__global__ void __launch_bounds__(1024)
q0_sum_kernel(const float* __restrict__ q, int64_t B, int K, int T, double* __restrict__ qsum) {
  __shared__ double red[32];
  for (int k = 0; k < K; ++k) {
    double s = 0.0;
    for (int64_t b = threadIdx.x; b < B; b += 1024) s += (double)q[(b * K + k) * (int64_t)T];
    s = block_sum<double>(s, red);
    if (threadIdx.x == 0) qsum[k] += s;
  }
}

// after all chunks: sums[1] += sum_k log_pi[k]*qsum[k]  (init term, :123) and the gradient of
// log_prior through log_softmax (:71): g_k = aP*qsum[k]; d = g - softmax(log_prior)*sum(g).
__global__ void finalize_prior_kernel(const float* __restrict__ log_pi, const double* __restrict__ qsum,
                                      const double* __restrict__ norms, float beta, int K,
                                      double* __restrict__ sums, float* __restrict__ g_logprior) {
  if (threadIdx.x != 0) return;
  double init = 0.0, gs = 0.0;
  const double aP = -(double)beta / norms[1];
  for (int k = 0; k < K; ++k) {
    init += (double)log_pi[k] * qsum[k];
    gs += aP * qsum[k];
  }
  sums[1] += init;
  if (g_logprior)
    for (int k = 0; k < K; ++k)
      g_logprior[k] = (float)(aP * qsum[k] - exp((double)log_pi[k]) * gs);
}

// d loss / d logits from dq (decoder + prior parts) and the entropy term; thread per (b,t).
__global__ void __launch_bounds__(256)
dlogits_kernel(const float* __restrict__ q, const float* __restrict__ logq,
               const float* __restrict__ dq_dec, const float* __restrict__ dq_prior,
               const int64_t* __restrict__ lengths, const double* __restrict__ norms, float beta,
               int K, int T, float* __restrict__ dl) {
  const int t = blockIdx.x * 256 + threadIdx.x;
  if (t >= T) return;
  const int64_t b = blockIdx.y;
  const bool m = (int64_t)t < lengths[b];
  const float aE = (float)((double)beta / norms[1]);   // d(-beta*E/B)/dl = +(beta/B) q (logq + H)
  float qv[KMAX], lq[KMAX], dq[KMAX];
  float s = 0.f, H = 0.f;
  for (int k = 0; k < K; ++k) {
    const int64_t o = (b * K + k) * (int64_t)T + t;
    qv[k] = q[o]; lq[k] = logq[o];
    dq[k] = dq_dec[o] + dq_prior[o];
    s += qv[k] * dq[k];
    H -= qv[k] * lq[k];
  }
  for (int k = 0; k < K; ++k) {
    float g = qv[k] * (dq[k] - s);
    if (m) g += aE * qv[k] * (lq[k] + H);
    dl[(b * K + k) * (int64_t)T + t] = g;
  }
}

// regimes[b,t] = argmax_k q[b,k,t], first index on ties (backtesting.py:155).
__global__ void argmax_kernel(const float* __restrict__ q, int K, int T, int64_t* __restrict__ out) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  const int64_t b = blockIdx.y;
  float best = q[(b * K) * (int64_t)T + t];
  int arg = 0;
  for (int k = 1; k < K; ++k) {
    const float v = q[(b * K + k) * (int64_t)T + t];
    if (v > best) { best = v; arg = k; }
  }
  out[b * T + t] = arg;
}

}  // namespace vhmm
