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
// Wide-K build of loss_seed_kernel (K > 8, e.g. the scaled config's K = 32): one thread per (b, t, state s), warps along t.
// Thread (t, s) owns ROW s of the K x K transition tile (pair term, d/dM, and the part of d/dq[s] that comes from the next
// step's transition) and COLUMN s (the part of d/dq[s] from the transition into t: a second, L1/L2-resident read of
// logA): no K-sized private array at all, every global access coalesced over t, fixed summation order (deterministic).
// block = 32 * K threads (t_local = tid & 31, s = tid >> 5); grid (ceil(T/32), B); partial: 4 doubles per block.
__global__ void __launch_bounds__(1024)
loss_seed_rows_kernel(const float* __restrict__ x, const float* __restrict__ p, const float* __restrict__ q,
                      const float* __restrict__ logq, const float* __restrict__ logA, const float* __restrict__ log_pi,
                      const int64_t* __restrict__ lengths, const double* __restrict__ norms, float beta, int C, int K, int T,
                      double* __restrict__ partial, float* __restrict__ dp, float* __restrict__ dq_prior,
                      float* __restrict__ dM) {
  __shared__ double red[32];
  const int t = blockIdx.x * 32 + (threadIdx.x & 31), s = threadIdx.x >> 5;
  const int64_t b = blockIdx.y;
  const int64_t L = lengths[b];
  const bool in = t < T;
  const bool m = in && (int64_t)t < L;
  const float aR = (float)(1.0 / norms[0]);
  const float aP = (float)(-(double)beta / norms[1]);
  float nll_sum = 0.f, ent = 0.f, trans = 0.f;
  if (in) {
    for (int c = s; c < C; c += K) {     // reconstruction (:118-120): channel c on the threads of state s = c mod K
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
    const bool tm = m && t >= 1;                                   // :128
    const bool tm_next = (t + 1 < T) && (int64_t)(t + 1) < L;
    const float* qb = q + b * K * (int64_t)T + t;                  // q[b, k, t] = qb[k * T]
    const int64_t os = (b * K + s) * (int64_t)T + t;
    if (m) ent = -q[os] * logq[os];                                // :134
    const float qps = tm ? q[os - 1] : 0.f;
    float dqs = (t == 0) ? log_pi[s] : 0.f;                        // :123
    const int64_t orow = ((b * K + s) * K) * (int64_t)T + t;       // logA[b, s, j, t] = logA[orow + j * T]
    for (int j = 0; j < K; ++j) {
      const int64_t o = orow + (int64_t)j * T;
      if (tm) {
        const float la = logA[o];
        const float qcj = qb[(int64_t)j * T];
        trans += qps * qcj * la;                                   // :127
        if (dM) dM[o] = aP * qps * (qcj - expf(la));
      } else if (dM) {
        dM[o] = 0.f;
      }
      if (tm_next) dqs += qb[(int64_t)j * T + 1] * logA[o + 1];    // row s of the transition into t + 1
    }
    if (tm) {                                                      // column s of the transition into t
      const int64_t ocol = ((b * K) * K + s) * (int64_t)T + t;     // logA[b, i, s, t] = logA[ocol + i * K * T]
      for (int i = 0; i < K; ++i) dqs += qb[(int64_t)i * T - 1] * logA[ocol + (int64_t)i * K * T];
    }
    if (dq_prior) dq_prior[os] = aP * dqs;
  }
  const int64_t blk = (int64_t)blockIdx.y * gridDim.x + blockIdx.x;
  double s0 = block_sum<double>((double)nll_sum, red);
  double s1 = block_sum<double>((double)trans, red);
  double s2 = block_sum<double>((double)ent, red);
  double s3 = block_sum<double>((m && s == 0) ? 1.0 : 0.0, red);
  if (threadIdx.x == 0) {
    partial[blk * 4 + 0] = s0; partial[blk * 4 + 1] = s1;
    partial[blk * 4 + 2] = s2; partial[blk * 4 + 3] = s3;
  }
}

// The same kernel with FOUR consecutive steps per thread (128-bit loads / stores; T % 4 == 0, 16-byte aligned tensors): the
// scalar build above spends ~15 instructions per (row, column, step) on 64-bit address arithmetic and 4-byte accesses and
// runs at IPC ~0.5 (15 ms per 2048 sequences of cfg 5 against 1.3 ms of HBM time).
// block = 8 * K threads (quad = tid & 7 -> steps t0..t0+3, s = tid >> 3); grid (ceil(T/32), B); partial: 4 doubles per block.
__global__ void __launch_bounds__(256)
loss_seed_rows4_kernel(const float* __restrict__ x, const float* __restrict__ p, const float* __restrict__ q,
                       const float* __restrict__ logq, const float* __restrict__ logA, const float* __restrict__ log_pi,
                       const int64_t* __restrict__ lengths, const double* __restrict__ norms, float beta, int C, int K, int T,
                       double* __restrict__ partial, float* __restrict__ dp, float* __restrict__ dq_prior,
                       float* __restrict__ dM) {
  __shared__ double red[32];
  const int t0 = blockIdx.x * 32 + (threadIdx.x & 7) * 4, s = threadIdx.x >> 3;
  const int64_t b = blockIdx.y;
  const int L = (int)min((int64_t)T, max((int64_t)0, lengths[b]));
  const bool in = t0 < T;                             // T % 4 == 0: a quad is inside or outside as a whole
  const float aR = (float)(1.0 / norms[0]);
  const float aP = (float)(-(double)beta / norms[1]);
  float nll_sum = 0.f, ent = 0.f, trans = 0.f, nvalid = 0.f;
  if (in) {
    bool m[4], tm[4], tmn[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int t = t0 + e;
      m[e] = t < L; tm[e] = m[e] && t >= 1; tmn[e] = (t + 1 < T) && (t + 1 < L);
      if (s == 0 && m[e]) nvalid += 1.f;
    }
    for (int c = s; c < C; c += K) {                  // reconstruction (:118-120)
      const int64_t om = (b * 2 * C + c) * (int64_t)T + t0, ol = (b * 2 * C + C + c) * (int64_t)T + t0;
      const float4 mu4 = *reinterpret_cast<const float4*>(p + om), lv4 = *reinterpret_cast<const float4*>(p + ol);
      const float4 x4 = *reinterpret_cast<const float4*>(x + (b * C + c) * (int64_t)T + t0);
      const float mu[4] = {mu4.x, mu4.y, mu4.z, mu4.w}, lv[4] = {lv4.x, lv4.y, lv4.z, lv4.w}, xv[4] = {x4.x, x4.y, x4.z, x4.w};
      float dmu[4], dlv[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float ev = expf(lv[e]), v = fmaxf(ev, 1e-8f), d = mu[e] - xv[e], r = d * d / v;
        if (m[e]) nll_sum += 0.5f * (logf(6.283185307179586f * v) + r);
        dmu[e] = m[e] ? aR * d / v : 0.f;
        dlv[e] = (m[e] && ev >= 1e-8f) ? aR * 0.5f * (1.f - r) : 0.f;
      }
      if (dp) {
        *reinterpret_cast<float4*>(dp + om) = make_float4(dmu[0], dmu[1], dmu[2], dmu[3]);
        *reinterpret_cast<float4*>(dp + ol) = make_float4(dlv[0], dlv[1], dlv[2], dlv[3]);
      }
    }
    const float* qb = q + b * K * (int64_t)T + t0;       // q[b, k, t0 + e] = qb[k * T + e]
    const bool more = t0 + 4 < T;                        // column t0 + 4 exists
    const int64_t os = (b * K + s) * (int64_t)T + t0;
    float qps[4], dqs[4];
    {
      const float4 qs4 = *reinterpret_cast<const float4*>(q + os), lq4 = *reinterpret_cast<const float4*>(logq + os);
      const float qv[4] = {qs4.x, qs4.y, qs4.z, qs4.w}, lq[4] = {lq4.x, lq4.y, lq4.z, lq4.w};
      const float qprev = (t0 >= 1) ? q[os - 1] : 0.f;
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        if (m[e]) ent -= qv[e] * lq[e];                  // :134
        qps[e] = tm[e] ? (e ? qv[e - 1] : qprev) : 0.f;
        dqs[e] = (t0 + e == 0) ? log_pi[s] : 0.f;        // :123
      }
    }
    // (logA == nullptr: the pair term and its gradients come from the fused prior kernel, prior_tc5.cuh)
    const float* rowp = logA + ((b * K + s) * K) * (int64_t)T + t0;     // logA[b, s, j, t0 + e] = rowp[j * T + e]
    float* dmp = dM ? dM + ((b * K + s) * K) * (int64_t)T + t0 : nullptr;
    for (int j = 0; logA && j < K; ++j) {
      const float4 la4 = *reinterpret_cast<const float4*>(rowp + (int64_t)j * T);
      const float4 qc4 = *reinterpret_cast<const float4*>(qb + (int64_t)j * T);
      const float la5 = more ? rowp[(int64_t)j * T + 4] : 0.f, qc5 = more ? qb[(int64_t)j * T + 4] : 0.f;
      const float la[5] = {la4.x, la4.y, la4.z, la4.w, la5}, qc[5] = {qc4.x, qc4.y, qc4.z, qc4.w, qc5};
      float dm[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        dm[e] = 0.f;
        if (tm[e]) {
          trans += qps[e] * qc[e] * la[e];               // :127
          dm[e] = aP * qps[e] * (qc[e] - expf(la[e]));
        }
        if (tmn[e]) dqs[e] += qc[e + 1] * la[e + 1];     // row s of the transition into t + 1
      }
      if (dmp) *reinterpret_cast<float4*>(dmp + (int64_t)j * T) = make_float4(dm[0], dm[1], dm[2], dm[3]);
    }
    if (logA) {   // column s of the transition into t: sum_i q[i, t-1] logA[i, s, t]
      const float* colp = logA + ((b * K) * K + s) * (int64_t)T + t0;   // logA[b, i, s, t0 + e] = colp[i * K * T + e]
      for (int i = 0; i < K; ++i) {
        const float4 la4 = *reinterpret_cast<const float4*>(colp + (int64_t)i * K * T);
        const float4 qi4 = *reinterpret_cast<const float4*>(qb + (int64_t)i * T);
        const float qm1 = (t0 >= 1) ? qb[(int64_t)i * T - 1] : 0.f;
        const float la[4] = {la4.x, la4.y, la4.z, la4.w}, qp[4] = {qm1, qi4.x, qi4.y, qi4.z};
#pragma unroll
        for (int e = 0; e < 4; ++e) if (tm[e]) dqs[e] += qp[e] * la[e];
      }
    }
    if (dq_prior) *reinterpret_cast<float4*>(dq_prior + os) = make_float4(aP * dqs[0], aP * dqs[1], aP * dqs[2], aP * dqs[3]);
  }
  const int64_t blk = (int64_t)blockIdx.y * gridDim.x + blockIdx.x;
  double s0 = block_sum<double>((double)nll_sum, red);
  double s1 = block_sum<double>((double)trans, red);
  double s2 = block_sum<double>((double)ent, red);
  double s3 = block_sum<double>((double)nvalid, red);
  if (threadIdx.x == 0) {
    partial[blk * 4 + 0] = s0; partial[blk * 4 + 1] = s1;
    partial[blk * 4 + 2] = s2; partial[blk * 4 + 3] = s3;
  }
}

// KT: compile-time bound of K (4 / 8 / 16 / 32): the K-indexed vectors that the j loops touch stay in REGISTERS (the
// first build indexed four KMAX-sized arrays with run-time K: local memory, 6x off the HBM time at K = 32)
template <int KT>
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
    float qc[KT], qn[KT], dqj[KT];
    const bool tm = m && t >= 1;                        // :128 (mask is a prefix)
    const bool tm_next = (t + 1 < T) && (int64_t)(t + 1) < L;
#pragma unroll
    for (int k = 0; k < KT; ++k) {
      qc[k] = 0.f; qn[k] = 0.f; dqj[k] = 0.f;
      if (k < K) {
        const int64_t o = (b * K + k) * (int64_t)T + t;
        qc[k] = q[o];
        qn[k] = tm_next ? q[o + 1] : 0.f;
        if (m) ent -= qc[k] * logq[o];                  // :134
        dqj[k] = (t == 0) ? log_pi[k] : 0.f;            // :123
      }
    }
    for (int i = 0; i < K; ++i) {
      const float qpi = tm ? q[(b * K + i) * (int64_t)T + t - 1] : 0.f;
      float dqi = 0.f;                                  // the part of d/dq[i] that comes from the NEXT step's transition
      const int64_t oi = ((b * K + i) * K) * (int64_t)T + t;
#pragma unroll
      for (int j = 0; j < KT; ++j) {
        if (j < K) {
          const int64_t o = oi + (int64_t)j * T;
          if (tm) {
            const float la = logA[o];
            trans += qpi * qc[j] * la;                  // :127
            dqj[j] += qpi * la;
            if (dM) dM[o] = aP * qpi * (qc[j] - expf(la));
          } else if (dM) {
            dM[o] = 0.f;
          }
          if (tm_next) dqi += qn[j] * logA[o + 1];
        }
      }
      if (dq_prior) dq_prior[(b * K + i) * (int64_t)T + t] = dqi;   // completed below
    }
    if (dq_prior) {
#pragma unroll
      for (int k = 0; k < KT; ++k)
        if (k < K) {
          float* d = dq_prior + (b * K + k) * (int64_t)T + t;
          *d = aP * (*d + dqj[k]);
        }
    }
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
               int K, int T, float* __restrict__ dl, const float* __restrict__ dq_row = nullptr,
               const float* __restrict__ dq_col = nullptr) {
  const int t = blockIdx.x * 256 + threadIdx.x;
  if (t >= T) return;
  const int64_t b = blockIdx.y;
  const bool m = (int64_t)t < lengths[b];
  const float aE = (float)((double)beta / norms[1]);   // d(-beta*E/B)/dl = +(beta/B) q (logq + H)
  const float aP = -aE;                                // scale of the un-scaled pair-term parts of the fused prior kernel
  float qv[KMAX], lq[KMAX], dq[KMAX];
  float s = 0.f, H = 0.f;
  for (int k = 0; k < K; ++k) {
    const int64_t o = (b * K + k) * (int64_t)T + t;
    qv[k] = q[o]; lq[k] = logq[o];
    dq[k] = dq_dec[o] + dq_prior[o];
    if (dq_row) dq[k] += aP * (dq_row[o] + dq_col[o]);
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
