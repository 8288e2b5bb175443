// hmm_small.cuh -- exact-HMM recursions for small state counts (K <= 8), one THREAD per row.
//
// A "row" is a run of consecutive time steps of one sequence: the whole sequence for the
// plain kernels, a chunk of CH steps for the scan phases.  A CTA of NT threads owns NT rows and
// walks them TC steps at a time; every TC-step slab of log_A / log_b (NT x TC x (K*K+K) floats,
// each row's part contiguous in HBM) is staged into shared memory with cp.async (16-byte LDGSTS,
// double buffered) and then read back conflict-free by its owning thread with LDS.128 (row
// pitches are odd multiples of 16 bytes).
//
// Sum-product recursions run in scaled linear space (alpha_hat normalised every step, the log
// normalisers accumulated in double): per step 12 ex2 + 1 log instead of the 27 exp + 3 log of
// a log-space logsumexp, and the exps are off the dependency chain.  Viterbi runs in log space
// with exactly the operation order of oracle/hmm.py:viterbi (adds and strict compares only), so
// paths are bit-exact.  Maths: SURVEY.md Appendix B.2-B.5.
#pragma once
#include "common.cuh"

namespace vhmm {
namespace hs {

constexpr int pitch4(int n) {
  int p = (n + 3) & ~3;
  return (((p >> 2) & 1) == 0) ? p + 4 : p;
}

template <int K, int TC, int NT>
struct Tile {
  static constexpr int NA = TC * K * K, NB = TC * K;
  static constexpr int PA = (NA + 3) & ~3;                  // offset of the log_b part in a row
  static constexpr int ROW = pitch4(PA + ((NB + 3) & ~3));  // row pitch: odd multiple of 16 B
  static constexpr int STAGE = NT * ROW;
  static constexpr size_t SMEM = 2 * (size_t)STAGE * sizeof(float);
};

__device__ __forceinline__ void cp_async16(void* s, const void* g) {
  unsigned a = (unsigned)__cvta_generic_to_shared(s);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(a), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_async4(void* s, const void* g) {
  unsigned a = (unsigned)__cvta_generic_to_shared(s);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(a), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

struct Args {
  const float* log_pi;
  const float* A;          // (B,T,K,K)
  const float* Bm;         // (B,T,K)
  const int64_t* lengths;  // (B)
  int64_t nrows;           // B * rows_per_seq
  int64_t T;
  int rows_per_seq;        // 1: a row is a sequence
  int CH;                  // steps per row (== T when rows_per_seq == 1)
  int vec_in;              // 16-byte staging allowed (alignment checked by the host)
  int vec_out;             // 16-byte stores allowed for (B,T,K) outputs / stash
  int vec_xi;              // 16-byte stores allowed for xi
  int full;                // visit every step (outputs are defined for t >= L too)
  int vec_path;            // 4-byte stores into `path` allowed (T % 4 == 0 and the caller's pointer is 4-byte aligned)
  // forward outputs
  float* loglik;           // (B)            plain rows only
  float* alpha_log;        // (B,T,K) or null
  float* ahat;             // (B,T,K) or null: normalised forward variables (stash)
  // scan phase 3 inputs (rows = chunks): incoming normalised vector and log-scale per row
  const float* vin;        // (nrows,K) or null
  const double* ll_in;     // (nrows) or null
  // viterbi
  uint8_t* psi;            // (B,T,PSIB)
  uint8_t* path;           // (B,T)
  float* score;            // (B) or null
  // backward outputs
  float* gamma;            // (B,T,K)
  float* xi;               // (B,T,K,K) or null
};

// cooperative staging of slab `c` (local steps [c*TC, c*TC+TC)) for the CTA's NT rows
template <int K, int TC, int NT>
__device__ __forceinline__ void stage_slab(float* stage, const Args& a, const int64_t* row_step0,
                                           const int* row_n, int c) {
  using TL = Tile<K, TC, NT>;
  constexpr int NA4 = TL::NA / 4, NB4 = TL::NB / 4;
  static_assert(TL::NA % 4 == 0 && TL::NB % 4 == 0, "TC must be a multiple of 4");
  for (int idx = threadIdx.x; idx < NT * NA4; idx += NT) {
    const int r = idx / NA4, part = idx - r * NA4;
    int n = row_n[r] - c * TC;
    if (n <= 0) continue;
    n = n < TC ? n : TC;
    const float* src = a.A + (row_step0[r] + (int64_t)c * TC) * (K * K);
    float* dst = stage + r * TL::ROW;
    if (a.vec_in && n == TC) {
      cp_async16(dst + part * 4, src + part * 4);
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e)
        if (part * 4 + e < n * K * K) cp_async4(dst + part * 4 + e, src + part * 4 + e);
    }
  }
  for (int idx = threadIdx.x; idx < NT * NB4; idx += NT) {
    const int r = idx / NB4, part = idx - r * NB4;
    int n = row_n[r] - c * TC;
    if (n <= 0) continue;
    n = n < TC ? n : TC;
    const float* src = a.Bm + (row_step0[r] + (int64_t)c * TC) * K;
    float* dst = stage + r * TL::ROW + TL::PA;
    if (a.vec_in && n == TC) {
      cp_async16(dst + part * 4, src + part * 4);
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e)
        if (part * 4 + e < n * K) cp_async4(dst + part * 4 + e, src + part * 4 + e);
    }
  }
}

// 4 steps of A (4*K*K floats) and b (4*K floats) of this thread's row -> registers (LDS.128)
template <int K>
__device__ __forceinline__ void load4(const float* rowA, const float* rowB, int g, float* Av, float* Bv) {
#pragma unroll
  for (int i = 0; i < K * K; ++i) {
    float4 v = *reinterpret_cast<const float4*>(rowA + g * 4 * K * K + i * 4);
    Av[i * 4 + 0] = v.x; Av[i * 4 + 1] = v.y; Av[i * 4 + 2] = v.z; Av[i * 4 + 3] = v.w;
  }
#pragma unroll
  for (int i = 0; i < K; ++i) {
    float4 v = *reinterpret_cast<const float4*>(rowB + g * 4 * K + i * 4);
    Bv[i * 4 + 0] = v.x; Bv[i * 4 + 1] = v.y; Bv[i * 4 + 2] = v.z; Bv[i * 4 + 3] = v.w;
  }
}

// store n (<=4) steps x W floats starting at dst (row-contiguous); 16-byte path when allowed
template <int W>
__device__ __forceinline__ void store_steps(float* dst, const float* v, int n, bool vec) {
  if (vec && n == 4) {
#pragma unroll
    for (int i = 0; i < W; ++i)
      *reinterpret_cast<float4*>(dst + i * 4) = make_float4(v[i * 4], v[i * 4 + 1], v[i * 4 + 2], v[i * 4 + 3]);
  } else {
#pragma unroll
    for (int i = 0; i < 4 * W; ++i)
      if (i < n * W) dst[i] = v[i];
  }
}

template <int K>
struct Psi {
  static constexpr int BYTES = (K <= 4) ? 1 : K;
};

// ---------------------------------------------------------------------------------------
// forward sweep: sum-product (SP) and/or Viterbi (VIT)
// ---------------------------------------------------------------------------------------
template <int K, int TC, int NT, bool SP, bool VIT>
__global__ void __launch_bounds__(NT) fwd_kernel(Args a) {
  using TL = Tile<K, TC, NT>;
  extern __shared__ __align__(16) float smem[];
  __shared__ int64_t row_step0[NT];
  __shared__ int row_n[NT];
  __shared__ int max_n;
  const int tid = threadIdx.x;
  const int64_t r = (int64_t)blockIdx.x * NT + tid;
  int64_t b = 0;
  int t_start = 0, n = 0, cap = 0;
  if (tid == 0) max_n = 0;
  if (r < a.nrows) {
    b = r / a.rows_per_seq;
    t_start = (int)(r - b * a.rows_per_seq) * a.CH;
    cap = (int)min((int64_t)a.CH, a.T - t_start);           // steps that exist in this row
    int64_t L = a.lengths[b];
    L = L < 0 ? 0 : (L > a.T ? a.T : L);
    n = (int)max((int64_t)0, min((int64_t)cap, L - t_start)); // valid steps in this row
  }
  row_step0[tid] = b * a.T + t_start;
  row_n[tid] = n;
  __syncthreads();
  atomicMax(&max_n, a.full ? cap : n);
  __syncthreads();
  const int nch = (max_n + TC - 1) / TC;
  if (nch == 0) {
    if (r < a.nrows && a.rows_per_seq == 1) {
      if (SP && a.loglik) a.loglik[b] = 0.f;
      if (VIT && a.score) a.score[b] = 0.f;
    }
    return;
  }
  const float* rowA = smem + tid * TL::ROW;
  const float* rowB = rowA + TL::PA;

  float ah[K];
  float dl[K];
  double ll = 0.0;
  const bool first_row = (t_start == 0);
  if (SP && !first_row && r < a.nrows) {
#pragma unroll
    for (int j = 0; j < K; ++j) ah[j] = a.vin[r * K + j];
    ll = a.ll_in[r];
  } else {
#pragma unroll
    for (int j = 0; j < K; ++j) ah[j] = 1.f / K;
  }
#pragma unroll
  for (int j = 0; j < K; ++j) dl[j] = 0.f;

  stage_slab<K, TC, NT>(smem, a, row_step0, row_n, 0);
  cp_commit();
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
        if (tl0 >= (a.full ? cap : n)) break;
        float Av[4 * K * K], Bv[4 * K];
        load4<K>(cur + tid * TL::ROW, cur + tid * TL::ROW + TL::PA, g, Av, Bv);
        float out_ah[4 * K], out_al[4 * K];
        uint8_t psib[4 * Psi<K>::BYTES];
#pragma unroll
        for (int s = 0; s < 4; ++s) {
          const int tl = tl0 + s;
          const float* la = Av + s * K * K;
          const float* lb = Bv + s * K;
          if (tl < n) {
            if (first_row && tl == 0) {
              // B.2 / B.4 initial step
              if (SP) {
                float v[K], m = -INFINITY;
#pragma unroll
                for (int j = 0; j < K; ++j) { v[j] = a.log_pi[j] + lb[j]; m = fmaxf(m, v[j]); }
                float csum = 0.f;
#pragma unroll
                for (int j = 0; j < K; ++j) { v[j] = __expf(v[j] - m); csum += v[j]; }
                const float inv = 1.f / csum;
#pragma unroll
                for (int j = 0; j < K; ++j) ah[j] = v[j] * inv;
                ll = (double)(logf(csum) + m);
              }
              if (VIT) {
#pragma unroll
                for (int j = 0; j < K; ++j) dl[j] = a.log_pi[j] + lb[j];
#pragma unroll
                for (int j = 0; j < Psi<K>::BYTES; ++j) psib[s * Psi<K>::BYTES + j] = 0;
              }
            } else {
              if (SP) {
                float mb = lb[0];
#pragma unroll
                for (int j = 1; j < K; ++j) mb = fmaxf(mb, lb[j]);
                float a2[K], csum = 0.f;
#pragma unroll
                for (int j = 0; j < K; ++j) {
                  float sj = 0.f;
#pragma unroll
                  for (int i = 0; i < K; ++i) sj = fmaf(ah[i], __expf(la[i * K + j]), sj);
                  a2[j] = sj * __expf(lb[j] - mb);
                  csum += a2[j];
                }
                const float inv = 1.f / csum;
#pragma unroll
                for (int j = 0; j < K; ++j) ah[j] = a2[j] * inv;
                ll += (double)(logf(csum) + mb);
              }
              if (VIT) {
                float nd[K];
                unsigned packed = 0;
#pragma unroll
                for (int j = 0; j < K; ++j) {
                  float best = __fadd_rn(dl[0], la[j]);
                  int arg = 0;
#pragma unroll
                  for (int i = 1; i < K; ++i) {
                    const float cnd = __fadd_rn(dl[i], la[i * K + j]);
                    if (cnd > best) { best = cnd; arg = i; }
                  }
                  nd[j] = __fadd_rn(lb[j], best);
                  if constexpr (K <= 4) packed |= (unsigned)arg << (2 * j);
                  else psib[s * Psi<K>::BYTES + j] = (uint8_t)arg;
                }
                if constexpr (K <= 4) psib[s * Psi<K>::BYTES] = (uint8_t)packed;
#pragma unroll
                for (int j = 0; j < K; ++j) dl[j] = nd[j];
              }
            }
          } else if (VIT) {
#pragma unroll
            for (int j = 0; j < Psi<K>::BYTES; ++j) psib[s * Psi<K>::BYTES + j] = 0;
          }
          if (SP) {
#pragma unroll
            for (int j = 0; j < K; ++j) {
              out_ah[s * K + j] = ah[j];
              out_al[s * K + j] = logf(ah[j]) + (float)ll;
            }
          }
        }
        const int nst = min(4, (a.full ? cap : n) - tl0);
        const int64_t step = row_step0[tid] + tl0;
        if (SP && a.ahat) store_steps<K>(a.ahat + step * K, out_ah, nst, a.vec_out);
        if (SP && a.alpha_log) store_steps<K>(a.alpha_log + step * K, out_al, nst, a.vec_out);
        if (VIT) {
          uint8_t* pp = a.psi + step * Psi<K>::BYTES;
          if (Psi<K>::BYTES == 1 && nst == 4 && ((a.T & 3) == 0)) {
            *reinterpret_cast<uint32_t*>(pp) = (uint32_t)psib[0] | ((uint32_t)psib[1] << 8) |
                                               ((uint32_t)psib[2] << 16) | ((uint32_t)psib[3] << 24);
          } else {
#pragma unroll
            for (int i = 0; i < 4 * Psi<K>::BYTES; ++i)
              if (i < nst * Psi<K>::BYTES) pp[i] = psib[i];
          }
        }
      }
    }
    __syncthreads();
  }
  if (r >= a.nrows) return;
  if (SP && a.loglik && a.rows_per_seq == 1) a.loglik[b] = (float)ll;
  if (VIT) {
    // final state: lowest j attaining the max (B.4); back-track through psi
    float best = dl[0];
    int state = 0;
#pragma unroll
    for (int j = 1; j < K; ++j)
      if (dl[j] > best) { best = dl[j]; state = j; }
    if (a.score) a.score[b] = (n > 0) ? best : 0.f;
    uint8_t* prow = a.path + b * a.T;
    const uint8_t* psirow = a.psi + b * a.T * Psi<K>::BYTES;
    const bool v4 = ((a.T & 3) == 0);          // psi rows (library workspace, 256-byte aligned)
    const bool p4 = v4 && a.vec_path;          // path rows (caller's pointer)
    const int T = (int)a.T;
    // pad region
    for (int t = T - 1; t >= n; --t) {
      if (p4 && (t & 3) == 3 && t - 3 >= n) { *reinterpret_cast<uint32_t*>(prow + t - 3) = 0xFFFFFFFFu; t -= 3; }
      else prow[t] = VHMM_PAD_STATE;
    }
    int t = n - 1;
    while (t >= 0) {
      if (v4 && (t & 3) == 3 && Psi<K>::BYTES == 1) {
        const uint32_t w = *reinterpret_cast<const uint32_t*>(psirow + t - 3);
        uint32_t out = 0;
#pragma unroll
        for (int s = 3; s >= 0; --s) {
          out |= (uint32_t)state << (8 * s);
          if (t - 3 + s >= 1) state = (int)((w >> (8 * s + 2 * state)) & 3u);
        }
        if (p4) *reinterpret_cast<uint32_t*>(prow + t - 3) = out;
        else { prow[t - 3] = (uint8_t)out; prow[t - 2] = (uint8_t)(out >> 8); prow[t - 1] = (uint8_t)(out >> 16); prow[t] = (uint8_t)(out >> 24); }
        t -= 4;
      } else {
        prow[t] = (uint8_t)state;
        if (t >= 1) {
          if (Psi<K>::BYTES == 1) state = (psirow[t] >> (2 * state)) & 3;
          else state = psirow[(int64_t)t * Psi<K>::BYTES + state];
        }
        --t;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------
// backward sweep: beta_hat recursion, gamma and xi (B.3); needs the alpha_hat stash
// ---------------------------------------------------------------------------------------
template <int K, int TC, int NT>
__global__ void __launch_bounds__(NT) bwd_kernel(Args a) {
  using TL = Tile<K, TC, NT>;
  extern __shared__ __align__(16) float smem[];
  __shared__ int64_t row_step0[NT];
  __shared__ int row_n[NT];
  const int tid = threadIdx.x;
  const int64_t r = (int64_t)blockIdx.x * NT + tid;
  const int64_t b = r;
  int n = 0;
  const int T = (int)a.T;
  if (r < a.nrows) {
    int64_t L = a.lengths[b];
    n = (int)(L < 0 ? 0 : (L > a.T ? a.T : L));
  }
  row_step0[tid] = b * a.T;
  row_n[tid] = n;
  __syncthreads();
  const int nch = (T + TC - 1) / TC;  // outputs are defined (zero) for t >= L too
  float bh[K];
#pragma unroll
  for (int j = 0; j < K; ++j) bh[j] = 1.f;

  stage_slab<K, TC, NT>(smem, a, row_step0, row_n, nch - 1);
  cp_commit();
  for (int it = 0; it < nch; ++it) {
    const int c = nch - 1 - it;
    float* cur = smem + (it & 1) * TL::STAGE;
    if (it + 1 < nch) stage_slab<K, TC, NT>(smem + ((it + 1) & 1) * TL::STAGE, a, row_step0, row_n, c - 1);
    cp_commit();
    cp_wait<1>();
    __syncthreads();
    if (r < a.nrows) {
#pragma unroll
      for (int g = TC / 4 - 1; g >= 0; --g) {
        const int tl0 = c * TC + g * 4;
        if (tl0 >= T) continue;
        const int nst = min(4, T - tl0);
        float gam[4 * K], xiv[4 * K * K];
        if (tl0 >= n) {
          // fully padded group: zeros
#pragma unroll
          for (int i = 0; i < 4 * K; ++i) gam[i] = 0.f;
#pragma unroll
          for (int i = 0; i < 4 * K * K; ++i) xiv[i] = 0.f;
        } else {
          float Av[4 * K * K], Bv[4 * K];
          load4<K>(cur + tid * TL::ROW, cur + tid * TL::ROW + TL::PA, g, Av, Bv);
          // alpha_hat for steps tl0-1 .. tl0+3 (this thread's own row, straight from HBM/L2)
          float ahv[5 * K];
          const float* ap = a.ahat + (row_step0[tid] + tl0) * K;
          if (a.vec_out && tl0 + 4 <= T) {
#pragma unroll
            for (int i = 0; i < K; ++i) {
              float4 v = *reinterpret_cast<const float4*>(ap + i * 4);
              ahv[K + i * 4 + 0] = v.x; ahv[K + i * 4 + 1] = v.y; ahv[K + i * 4 + 2] = v.z; ahv[K + i * 4 + 3] = v.w;
            }
          } else {
#pragma unroll
            for (int i = 0; i < 4 * K; ++i) ahv[K + i] = (tl0 + i / K < n) ? ap[i] : 0.f;
          }
#pragma unroll
          for (int j = 0; j < K; ++j) ahv[j] = (tl0 >= 1) ? ap[j - K] : 0.f;
#pragma unroll
          for (int s = 3; s >= 0; --s) {
            const int tl = tl0 + s;
            const float* la = Av + s * K * K;
            const float* lb = Bv + s * K;
            float* gm = gam + s * K;
            float* xv = xiv + s * K * K;
            if (tl >= n) {
#pragma unroll
              for (int j = 0; j < K; ++j) gm[j] = 0.f;
#pragma unroll
              for (int i = 0; i < K * K; ++i) xv[i] = 0.f;
              continue;
            }
            // gamma_t ~ alpha_hat_t * beta_hat_t
            float gs = 0.f;
#pragma unroll
            for (int j = 0; j < K; ++j) { gm[j] = ahv[(s + 1) * K + j] * bh[j]; gs += gm[j]; }
            const float ginv = 1.f / gs;
#pragma unroll
            for (int j = 0; j < K; ++j) gm[j] *= ginv;
            if (tl >= 1) {
              float mb = lb[0];
#pragma unroll
              for (int j = 1; j < K; ++j) mb = fmaxf(mb, lb[j]);
              float w[K];                                   // b_t[j] * beta_hat_t[j]
#pragma unroll
              for (int j = 0; j < K; ++j) w[j] = __expf(lb[j] - mb) * bh[j];
              float nb[K], xs = 0.f, bs = 0.f;
#pragma unroll
              for (int i = 0; i < K; ++i) {
                float acc = 0.f;
#pragma unroll
                for (int j = 0; j < K; ++j) {
                  const float m = __expf(la[i * K + j]) * w[j];
                  acc += m;
                  xv[i * K + j] = ahv[s * K + i] * m;
                  xs += xv[i * K + j];
                }
                nb[i] = acc;
                bs += acc;
              }
              const float xinv = 1.f / xs, binv = 1.f / bs;
#pragma unroll
              for (int i = 0; i < K * K; ++i) xv[i] *= xinv;
#pragma unroll
              for (int i = 0; i < K; ++i) bh[i] = nb[i] * binv;
            } else {
#pragma unroll
              for (int i = 0; i < K * K; ++i) xv[i] = 0.f;
            }
          }
        }
        const int64_t step = row_step0[tid] + tl0;
        store_steps<K>(a.gamma + step * K, gam, nst, a.vec_out);
        if (a.xi) store_steps<K * K>(a.xi + step * K * K, xiv, nst, a.vec_xi);
      }
    }
    __syncthreads();
  }
}

}  // namespace hs
}  // namespace vhmm
