// hmm_scan.cuh -- parallel scan over T for long sequences with a small batch.
//
// The scan element of step t is the K x K transfer matrix M_t[i,j] = A_t[i,j] * b_t[j]; the
// log-semiring product (P (x) Q)[i,j] = logsumexp_k(P[i,k] + Q[k,j]) is carried in the scaled
// representation (P / s, log s): an ordinary matrix product of the normalised factors plus a
// sum of log-scales -- the same semiring element, but 27 FMA per combine instead of
// 27 exp + 9 log.  Three phases:
//   1. scan_local_kernel : one thread per chunk of CH steps -> chunk product (matrix, log-scale)
//   2. scan_carry_kernel : one warp per sequence: exclusive scan of the chunk products applied
//                          to alpha_0 -> incoming vector + log-scale of every chunk, loglik
//   3. hs::fwd_kernel    : (only if all-prefix alpha is requested) re-run each chunk from its
//                          incoming vector
#pragma once
#include "hmm_small.cuh"

namespace vhmm {
namespace hs {

struct ScanArgs {
  float* Pmat;       // (nrows, K*K) normalised chunk products
  double* Pscale;    // (nrows)
  float* vin;        // (nrows, K)
  double* ll_in;     // (nrows)
  float* loglik;     // (B)
  int* suspect;      // (B) set to 1 by the carry scan when a product it needed had vanished (see scan_fixup_kernel)
};

// A product of transfer matrices is carried with ONE scale for the whole K x K matrix.  With structural zeros
// (log_A = -inf: left-to-right chains) or near-absorbing states under sharply discriminative emissions the ROWS of
// a long product differ by more than fp32's range, and the rows that matter for a given start vector can underflow
// while the largest row stays O(1).  Everything the scan does is exact up to powers of two until a value drops below
// 2^-126 of the largest entry, so the condition is detectable: the row-weighted sum `cs` the carry scan divides by
// falls below 2^-100.  Such sequences are marked and recomputed by scan_fixup_kernel with the per-step-normalised
// sequential recursion (which only ever carries a vector).
constexpr float SCAN_VANISH = 7.8886090522101181e-31f;   // 2^-100

// P <- normalise(P * M), *scale += log(sum)
template <int K>
__device__ __forceinline__ void matmul_norm(float* P, const float* M, double* scale) {
  float R[K * K], s = 0.f;
#pragma unroll
  for (int i = 0; i < K; ++i)
#pragma unroll
    for (int j = 0; j < K; ++j) {
      float acc = 0.f;
#pragma unroll
      for (int k = 0; k < K; ++k) acc = fmaf(P[i * K + k], M[k * K + j], acc);
      R[i * K + j] = acc;
      s += acc;
    }
  s = fmaxf(s, 1.17549435e-38f);   // a vanished product (structural zeros / underflow) must not turn into NaN
  const float inv = 1.f / s;
#pragma unroll
  for (int i = 0; i < K * K; ++i) P[i] = R[i] * inv;
  *scale += (double)logf(s);
}

template <int K, int TC, int NT>
__global__ void __launch_bounds__(NT) scan_local_kernel(Args a, ScanArgs sa) {
  using TL = Tile<K, TC, NT>;
  extern __shared__ __align__(16) float smem[];
  __shared__ int64_t row_step0[NT];
  __shared__ int row_n[NT];
  __shared__ int max_n;
  const int tid = threadIdx.x;
  const int64_t r = (int64_t)blockIdx.x * NT + tid;
  int64_t b = 0;
  int t_start = 0, n = 0;
  if (tid == 0) max_n = 0;
  if (r < a.nrows) {
    b = r / a.rows_per_seq;
    t_start = (int)(r - b * a.rows_per_seq) * a.CH;
    const int cap = (int)min((int64_t)a.CH, a.T - t_start);
    int64_t L = a.lengths[b];
    L = L < 0 ? 0 : (L > a.T ? a.T : L);
    n = (int)max((int64_t)0, min((int64_t)cap, L - t_start));
  }
  row_step0[tid] = b * a.T + t_start;
  row_n[tid] = n;
  __syncthreads();
  atomicMax(&max_n, n);
  __syncthreads();
  const int nch = (max_n + TC - 1) / TC;
  float P[K * K];
#pragma unroll
  for (int i = 0; i < K * K; ++i) P[i] = (i / K == i % K) ? 1.f : 0.f;  // the identity, exactly (chunks beyond the length keep it)
  double scale = 0.0;
  if (nch > 0) {
    stage_slab<K, TC, NT>(smem, a, row_step0, row_n, 0);
    cp_commit();
  }
  for (int c = 0; c < nch; ++c) {
    float* cur = smem + (c & 1) * TL::STAGE;
    if (c + 1 < nch) stage_slab<K, TC, NT>(smem + ((c + 1) & 1) * TL::STAGE, a, row_step0, row_n, c + 1);
    cp_commit();
    cp_wait<1>();
    __syncthreads();
    if (r < a.nrows) {
#pragma unroll
      for (int g = 0; g < TC / 4; ++g) {
        const int tl0 = c * TC + g * 4;
        if (tl0 >= n) break;
        float Av[4 * K * K], Bv[4 * K];
        load4<K>(cur + tid * TL::ROW, cur + tid * TL::ROW + TL::PA, g, Av, Bv);
#pragma unroll
        for (int s = 0; s < 4; ++s) {
          const int tl = tl0 + s;
          if (tl >= n || (t_start == 0 && tl == 0)) continue;   // step 0 is alpha_0, not a transfer
          const float* la = Av + s * K * K;
          const float* lb = Bv + s * K;
          float mb = lb[0];
#pragma unroll
          for (int j = 1; j < K; ++j) mb = fmaxf(mb, lb[j]);
          float M[K * K];
#pragma unroll
          for (int j = 0; j < K; ++j) {
            const float bj = __expf(lb[j] - mb);
#pragma unroll
            for (int i = 0; i < K; ++i) M[i * K + j] = __expf(la[i * K + j]) * bj;
          }
          matmul_norm<K>(P, M, &scale);
          scale += (double)mb;
        }
      }
    }
    __syncthreads();
  }
  if (r < a.nrows) {
#pragma unroll
    for (int i = 0; i < K * K; ++i) sa.Pmat[r * K * K + i] = P[i];
    sa.Pscale[r] = scale;
  }
}

// P <- (P * M) * 2^-e with e = floor(log2 max entry), *scale += e ln 2.  Power-of-two normalisation is EXACT: identities (the
// chunks beyond a sequence's length) combine without any rounding, however many threads take part in the scan -- with the
// sum-normalised form above every combine of two identities leaves ~6e-8 in the log-scale, 2.5e-5 over 256 threads.
template <int K>
__device__ __forceinline__ void matmul_pow2(float* P, const float* M, double* scale) {
  float R[K * K], m = 0.f;
#pragma unroll
  for (int i = 0; i < K; ++i)
#pragma unroll
    for (int j = 0; j < K; ++j) {
      float acc = 0.f;
#pragma unroll
      for (int k = 0; k < K; ++k) acc = fmaf(P[i * K + k], M[k * K + j], acc);
      R[i * K + j] = acc;
      m = fmaxf(m, acc);
    }
  const unsigned bits = __float_as_uint(m) & 0x7f800000u;     // (a vanished product: bits == 0 -> zeros stay zeros, no NaN)
  const float sc = __uint_as_float(0x7f000000u - bits);         // 2^-e
#pragma unroll
  for (int i = 0; i < K * K; ++i) P[i] = R[i] * sc;
  *scale += (double)((int)(bits >> 23) - 127) * 0.6931471805599453;
}

// One CTA of CARRY_W warps per sequence; thread i owns chunks [i*per, (i+1)*per).  (The first build used ONE warp per
// sequence: at cfg 4 -- 64 sequences, 256 chunk products each -- every lane walked 8 chunks through dependent global loads,
// and the kernel took as long as the pass that reads the 201 MB of inputs.)
constexpr int CARRY_W = 8;
template <int K>
__global__ void __launch_bounds__(32 * CARRY_W) scan_carry_kernel(Args a, ScanArgs sa) {
  constexpr int KK = K * K;
  constexpr int NTH = 32 * CARRY_W;
  __shared__ float wQ[CARRY_W][KK];
  __shared__ double wS[CARRY_W];
  const int64_t b = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int64_t L = a.lengths[b];
  if (L <= 0) { if (tid == 0) { sa.loglik[b] = 0.f; if (sa.suspect) sa.suspect[b] = 0; } return; }   // (uniform per CTA)
  const int nck = a.rows_per_seq;
  const int per = (nck + NTH - 1) / NTH;
  const int c_lo = min(tid * per, nck), c_hi = min(c_lo + per, nck);
  const int64_t row0 = b * nck;
  // (1) thread-local product of this thread's chunk matrices
  float Q[KK];
#pragma unroll
  for (int i = 0; i < KK; ++i) Q[i] = (i / K == i % K) ? 1.f : 0.f;
  double qs = 0.0;
  for (int c = c_lo; c < c_hi; ++c) {
    float M[KK];
#pragma unroll
    for (int i = 0; i < KK; ++i) M[i] = sa.Pmat[(row0 + c) * KK + i];
    matmul_pow2<K>(Q, M, &qs);
    qs += sa.Pscale[row0 + c];
  }
  // (2) inclusive Kogge-Stone scan of (Q, qs) across the lanes of a warp: X_l = Q_0 * ... * Q_l
  for (int d = 1; d < 32; d <<= 1) {
    float Pm[KK];
#pragma unroll
    for (int i = 0; i < KK; ++i) Pm[i] = __shfl_up_sync(0xffffffffu, Q[i], d);
    const double ps = __shfl_up_sync(0xffffffffu, qs, d);
    if (lane >= d) {
      double sc = ps + qs;
      matmul_pow2<K>(Pm, Q, &sc);   // Pm <- Pm * Q  (earlier factors on the left)
#pragma unroll
      for (int i = 0; i < KK; ++i) Q[i] = Pm[i];
      qs = sc;
    }
  }
  // exclusive prefix inside the warp = inclusive of lane-1
  float X[KK];
#pragma unroll
  for (int i = 0; i < KK; ++i) X[i] = __shfl_up_sync(0xffffffffu, Q[i], 1);
  double xs = __shfl_up_sync(0xffffffffu, qs, 1);
  // (2b) totals of the warps -> the product of all EARLIER warps, times the lane's exclusive prefix
  if (lane == 31) {
#pragma unroll
    for (int i = 0; i < KK; ++i) wQ[w][i] = Q[i];
    wS[w] = qs;
  }
  __syncthreads();
  if (w > 0) {
    float Wp[KK];
#pragma unroll
    for (int i = 0; i < KK; ++i) Wp[i] = wQ[0][i];
    double ws = wS[0];
    for (int k = 1; k < w; ++k) {
      float M[KK];
#pragma unroll
      for (int i = 0; i < KK; ++i) M[i] = wQ[k][i];
      ws += wS[k];
      matmul_pow2<K>(Wp, M, &ws);
    }
    if (lane == 0) {
#pragma unroll
      for (int i = 0; i < KK; ++i) X[i] = Wp[i];
      xs = ws;
    } else {
      ws += xs;
      matmul_pow2<K>(Wp, X, &ws);
#pragma unroll
      for (int i = 0; i < KK; ++i) X[i] = Wp[i];
      xs = ws;
    }
  }
  // alpha_0 (normalised) and its log-scale
  float v[K];
  double ll;
  {
    float m = -INFINITY;
#pragma unroll
    for (int j = 0; j < K; ++j) { v[j] = a.log_pi[j] + a.Bm[b * a.T * K + j]; m = fmaxf(m, v[j]); }
    float cs = 0.f;
#pragma unroll
    for (int j = 0; j < K; ++j) { v[j] = __expf(v[j] - m); cs += v[j]; }
#pragma unroll
    for (int j = 0; j < K; ++j) v[j] /= cs;
    ll = (double)(logf(cs) + m);
  }
  bool vanished = false;
  if (tid > 0) {
    float wv[K], cs = 0.f;
#pragma unroll
    for (int j = 0; j < K; ++j) {
      float acc = 0.f;
#pragma unroll
      for (int i = 0; i < K; ++i) acc = fmaf(v[i], X[i * K + j], acc);
      wv[j] = acc; cs += acc;
    }
    vanished |= (c_lo < c_hi) && !(cs >= SCAN_VANISH);
    cs = fmaxf(cs, 1.17549435e-38f);
#pragma unroll
    for (int j = 0; j < K; ++j) v[j] = wv[j] / cs;
    ll += xs + (double)logf(cs);
  }
  // (3) walk this thread's chunks: record the incoming vector, then apply the chunk product
  for (int c = c_lo; c < c_hi; ++c) {
#pragma unroll
    for (int j = 0; j < K; ++j) sa.vin[(row0 + c) * K + j] = v[j];
    sa.ll_in[row0 + c] = ll;
    float wv[K], cs = 0.f;
#pragma unroll
    for (int j = 0; j < K; ++j) {
      float acc = 0.f;
#pragma unroll
      for (int i = 0; i < K; ++i) acc = fmaf(v[i], sa.Pmat[(row0 + c) * KK + i * K + j], acc);
      wv[j] = acc; cs += acc;
    }
    // (only chunks that hold a valid transition count: beyond the length the products are identities)
    vanished |= ((int64_t)c * a.CH < L) && !(cs >= SCAN_VANISH);
    cs = fmaxf(cs, 1.17549435e-38f);
#pragma unroll
    for (int j = 0; j < K; ++j) v[j] = wv[j] / cs;
    ll += sa.Pscale[row0 + c] + (double)logf(cs);
  }
  const int any_vanished = __syncthreads_or(vanished ? 1 : 0);
  if (tid == 0 && sa.suspect) sa.suspect[b] = any_vanished ? 1 : 0;
  // the thread that owns the last chunk holds the total
  const int last_tid = (nck - 1) / per;
  if (tid == last_tid) sa.loglik[b] = (float)ll;
}


// Backward carry over the chunk products of a sequence: bin[c] = P_{c+1} P_{c+2} ... P_{last} 1 (normalised to max 1: only
// the direction matters, gamma and xi are normalised per step) -- beta after the last step of chunk c.  One warp per sequence;
// the lanes fetch 32 products at a time, then the vector walks through them in reverse order by shuffles.
template <int K>
__global__ void __launch_bounds__(32) scan_carry_back_kernel(Args a, ScanArgs sa, float* __restrict__ bin) {
  constexpr int KK = K * K;
  const int64_t b = blockIdx.x;
  const int lane = threadIdx.x;
  const int nck = a.rows_per_seq;
  const int64_t row0 = b * nck;
  float r[K];
#pragma unroll
  for (int j = 0; j < K; ++j) r[j] = 1.f;
  for (int base = ((nck - 1) / 32) * 32; base >= 0; base -= 32) {
    const int c = base + lane;
    float M[KK];
#pragma unroll
    for (int i = 0; i < KK; ++i) M[i] = (c < nck) ? sa.Pmat[(row0 + c) * KK + i] : ((i / K == i % K) ? 1.f : 0.f);
    for (int l = 31; l >= 0; --l) {
      if (base + l >= nck) continue;                     // (uniform)
      if (lane == l) {
#pragma unroll
        for (int j = 0; j < K; ++j) bin[(row0 + c) * K + j] = r[j];
      }
      float n[K], m = 0.f;
#pragma unroll
      for (int i = 0; i < K; ++i) {
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < K; ++j) acc = fmaf(__shfl_sync(0xffffffffu, M[i * K + j], l), r[j], acc);
        n[i] = acc;
        m = fmaxf(m, acc);
      }
      const float inv = 1.f / fmaxf(m, 1.17549435e-38f);
#pragma unroll
      for (int i = 0; i < K; ++i) r[i] = n[i] * inv;
    }
  }
}

// Sequences marked by the carry scan: per-step-normalised forward recursion (B.2), one warp per sequence.  The lanes
// form the transfer matrices of 32 consecutive steps (coalesced loads, exps in parallel), lane 0 applies them in
// order.  Writes loglik and, if requested, the log-space alpha rows (carried unchanged beyond the length).
template <int K>
__global__ void __launch_bounds__(32) scan_fixup_kernel(Args a, ScanArgs sa, float* __restrict__ alpha) {
  const int64_t b = blockIdx.x;
  if (!sa.suspect[b]) return;
  __shared__ float Pm[32][K * K + 1];
  __shared__ float mb[32];
  const int lane = threadIdx.x;
  const int T = (int)a.T;
  int64_t L64 = a.lengths[b];
  const int L = (int)(L64 < 0 ? 0 : (L64 > T ? T : L64));
  const float* Ab = a.A + b * (int64_t)T * K * K;
  const float* Bb = a.Bm + b * (int64_t)T * K;
  float v[K];
  double ll = 0.0;
  if (lane == 0) {
    float m = -INFINITY, cs = 0.f;
#pragma unroll
    for (int j = 0; j < K; ++j) { v[j] = a.log_pi[j] + Bb[j]; m = fmaxf(m, v[j]); }
#pragma unroll
    for (int j = 0; j < K; ++j) { v[j] = expf(v[j] - m); cs += v[j]; }
#pragma unroll
    for (int j = 0; j < K; ++j) v[j] /= cs;
    ll = (double)m + (double)logf(cs);
    if (alpha)
#pragma unroll
      for (int j = 0; j < K; ++j) alpha[(b * T) * K + j] = logf(v[j]) + (float)ll;
  }
  for (int tb = 1; tb < L; tb += 32) {
    const int t = tb + lane;
    if (t < L) {
      float lb[K], m = -INFINITY;
#pragma unroll
      for (int j = 0; j < K; ++j) { lb[j] = Bb[(int64_t)t * K + j]; m = fmaxf(m, lb[j]); }
#pragma unroll
      for (int i = 0; i < K; ++i)
#pragma unroll
        for (int j = 0; j < K; ++j) Pm[lane][i * K + j] = expf(Ab[(int64_t)t * K * K + i * K + j] + lb[j] - m);
      mb[lane] = m;
    }
    __syncwarp();
    if (lane == 0) {
      const int n = min(32, L - tb);
      for (int s = 0; s < n; ++s) {
        float w[K], cs = 0.f;
#pragma unroll
        for (int j = 0; j < K; ++j) {
          float acc = 0.f;
#pragma unroll
          for (int i = 0; i < K; ++i) acc = fmaf(v[i], Pm[s][i * K + j], acc);
          w[j] = acc; cs += acc;
        }
        const float inv = 1.f / cs;          // cs == 0: the data has probability zero -> -inf, as the oracle gives
#pragma unroll
        for (int j = 0; j < K; ++j) v[j] = w[j] * inv;
        ll += (double)mb[s] + (double)logf(cs);
        if (alpha)
#pragma unroll
          for (int j = 0; j < K; ++j) alpha[(b * T + tb + s) * K + j] = logf(v[j]) + (float)ll;
      }
    }
    __syncwarp();
  }
  if (lane == 0) {
    sa.loglik[b] = (L > 0) ? (float)ll : 0.f;
    if (alpha)
      for (int t = (L > 0 ? L : 1); t < T; ++t)
#pragma unroll
        for (int j = 0; j < K; ++j) alpha[(b * T + t) * K + j] = logf(v[j]) + (float)ll;
  }
}

}  // namespace hs
}  // namespace vhmm
