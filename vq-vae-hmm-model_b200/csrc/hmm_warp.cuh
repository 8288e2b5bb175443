// hmm_warp.cuh -- one-pass posteriors (gamma, xi, log-likelihood) for small K: one WARP per
// sequence, the whole sequence resident in shared memory, time parallelised over the lanes.
//
// Why: the two-sweep kernels of hmm_small.cuh move 132 B per step for 60 algorithmic bytes
// (inputs read twice, alpha_hat stash written and read).  Here every input byte crosses HBM
// once and every output byte once (gamma 60 B/step, gamma+xi 96 B/step).
//
// Lane l owns the S consecutive steps [l*S, l*S+S), S = 4*ceil(T/128).  Per sequence:
//   load   cp.async (16 B, coalesced over the warp) of log_A / log_b into per-lane chunks whose
//          pitches are odd multiples of 16 B, so that each lane reads its own chunk back with
//          conflict-free LDS.128.
//   pass 1 P_t[i,j] = exp(log_A_t[i,j] + log_b_t[j] - max_j log_b_t) written back in place
//          (the exps are paid once); chunk product M_l = prod_t P_t, rescaled by powers of two.
//          Step 0 has no transition: P_0 = I and v0 = pi * b~_0 enters as the start vector.
//   scan   Kogge-Stone over the lanes: prefix products (forward start vector of every chunk,
//          log-likelihood from the tracked exponents) and suffix products (backward start vector).
//   pass 2 forward sweep inside the chunk: alpha_hat_t stored over the log_b slots.
//   pass 3 backward sweep: gamma_t (over the alpha_hat slot), xi_t (over the P_t slot), beta_hat.
//   store  coalesced 16-byte stores of the gamma / xi rows (zeros for t >= L).
// All arithmetic is on non-negative numbers (no cancellation); rescaling is by exact powers of
// two.  Maths: SURVEY.md Appendix B.2/B.3; conventions of oracle/hmm.py:posteriors.
#pragma once
#include "hmm_small.cuh"
#pragma nv_diag_suppress 128   // LL_ONLY instantiations leave the sweeps unreachable on purpose

namespace vhmm {
namespace hw {

struct Args {
  const float* log_pi;
  const float* A;          // (B,T,K,K)
  const float* Bm;         // (B,T,K)
  const int64_t* lengths;  // (B)
  int64_t B, T;
  int S;                   // steps per lane, multiple of 4, 32*S >= T
  int SA, SB;              // per-lane chunk pitches (floats) of the A / b regions
  unsigned magicA, magicB; // ceil(2^32 / granules per chunk)
  float* loglik;           // (B) or null
  float* gamma;            // (B,T,K)
  float* xi;               // (B,T,K,K) or null (XI kernels only)
  // ALPHA kernels (all-prefix log alpha; rows = segments of T steps as in SEG; `gamma` receives the alpha rows):
  const float* vin;        // (rows, K) incoming vector of every row (alpha before its first step, any scale), or null: one
  const double* ll_in;     // (rows) ... and its natural-log scale            row per sequence, alpha_0 formed here
  // gamma / xi kernels on segment rows (nseg > 1: long sequences cut into T-step rows; `vin` as above):
  const float* bin_in;     // (rows, K) beta after the row's last step (any scale), from the backward carry scan
  int64_t seqT;            // segment rows: steps per SEQUENCE in memory (0: nseg * T); the last row of a sequence may be shorter
  uint8_t* path;           // (B,T) Viterbi path (VIT kernels only)
  int path_vec16;          // T % 16 == 0 and path 16-byte aligned: the path rows leave as 16-byte stores
  // SEG kernels (long sequences cut into segments of T steps; B = sequences * nseg rows, `lengths` per SEQUENCE):
  int nseg;                // segments per sequence
  float* Pmat;             // (rows, K*K) segment transfer products, scaled
  double* Pscale;          // (rows) natural-log scale: true product = Pmat * exp(Pscale)
};

__host__ __device__ constexpr int pitch_odd16(int n) { return (((n >> 2) & 1) == 0) ? n + 4 : n; }

// 2^-e with e = floor(log2(m)) (m > 0, finite, < 2^127); *ex += e
__device__ __forceinline__ float pow2_rescale(float m, int* ex) {
  const unsigned b = __float_as_uint(m) & 0x7f800000u;
  *ex += (int)(b >> 23) - 127;
  return __uint_as_float(0x7f000000u - b);
}
__device__ __forceinline__ float pow2_rescale(float m) {
  const unsigned b = __float_as_uint(m) & 0x7f800000u;
  return __uint_as_float(0x7f000000u - b);
}
template <int N>
__device__ __forceinline__ float max_of(const float* v) {
  float m = v[0];
#pragma unroll
  for (int i = 1; i < N; ++i) m = fmaxf(m, v[i]);
  return m;
}
__device__ __forceinline__ float ex2f(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// R = X * Y (K x K)
template <int K>
__device__ __forceinline__ void mat_mul(float* R, const float* X, const float* Y) {
#pragma unroll
  for (int i = 0; i < K; ++i)
#pragma unroll
    for (int j = 0; j < K; ++j) {
      float acc = X[i * K] * Y[j];
#pragma unroll
      for (int k = 1; k < K; ++k) acc = fmaf(X[i * K + k], Y[k * K + j], acc);
      R[i * K + j] = acc;
    }
}

template <int N>
__device__ __forceinline__ void lds4(const float* p, float* v) {
#pragma unroll
  for (int i = 0; i < N; ++i) {
    const float4 q = *reinterpret_cast<const float4*>(p + i * 4);
    v[i * 4 + 0] = q.x; v[i * 4 + 1] = q.y; v[i * 4 + 2] = q.z; v[i * 4 + 3] = q.w;
  }
}
template <int N>
__device__ __forceinline__ void sts4(float* p, const float* v) {
#pragma unroll
  for (int i = 0; i < N; ++i)
    *reinterpret_cast<float4*>(p + i * 4) = make_float4(v[i * 4], v[i * 4 + 1], v[i * 4 + 2], v[i * 4 + 3]);
}

// Prefix scan over the lanes of the chunk products M_l (scaled by 2^-ex_l) applied to the start vector v0 of
// lane 0: on return lane l holds pre = v0 M_0 ... M_l scaled by 2^-pex.  Lane 0 contributes the rank-one matrix
// 1 * (v0 M_0), so that every prefix has identical rows; only that row is carried once a segment reaches lane 0.
template <int K>
__device__ __forceinline__ void warp_prefix_scan(const float* M, int ex, const float* v0, int lane, float* pre, int& pex) {
  constexpr int KK = K * K;
  pex = ex;
  {
    float Pm[KK];
    if (lane == 0) {
#pragma unroll
      for (int j = 0; j < K; ++j) {
        float acc = v0[0] * M[j];
#pragma unroll
        for (int k = 1; k < K; ++k) acc = fmaf(v0[k], M[k * K + j], acc);
        pre[j] = acc;
      }
    }
#pragma unroll
    for (int i = 0; i < KK; ++i) Pm[i] = M[i];
    // vector-times-matrix scan: lane l holds (pre_l, Pm_l); combining the segment ending at
    // l-d with the segment ending at l: pre <- pre_other * Pm_l, Pm <- Pm_other * Pm_l.
    // Lanes whose segment already starts at lane 0 (lane < d) are final.
    // Three Kogge-Stone rounds inside groups of 8 lanes (matrix products), then the carry vector of each
    // group is handed to the next one serially (3 cheap vector x matrix steps instead of 2 matrix rounds).
    bool has = (lane == 0);   // 'pre' valid (segment reaches lane 0)
    const int li = lane & 7;
#pragma unroll
    for (int d = 1; d < 8; d <<= 1) {
      float opre[K], oPm[KK];
#pragma unroll
      for (int j = 0; j < K; ++j) opre[j] = __shfl_up_sync(0xffffffffu, pre[j], d);
#pragma unroll
      for (int i = 0; i < KK; ++i) oPm[i] = __shfl_up_sync(0xffffffffu, Pm[i], d);
      const int oex = __shfl_up_sync(0xffffffffu, pex, d);
      const bool ohas = __shfl_up_sync(0xffffffffu, (int)has, d) != 0;
      if (li >= d && !has) {
        pex += oex;
        if (ohas) {
#pragma unroll
          for (int j = 0; j < K; ++j) {
            float acc = opre[0] * Pm[j];
#pragma unroll
            for (int k = 1; k < K; ++k) acc = fmaf(opre[k], Pm[k * K + j], acc);
            pre[j] = acc;
          }
          has = true;
          const float sc = pow2_rescale(fmaxf(max_of<K>(pre), 1e-37f), &pex);
#pragma unroll
          for (int j = 0; j < K; ++j) pre[j] *= sc;
        } else {
          float R[KK];
          mat_mul<K>(R, oPm, Pm);
          const float sc = pow2_rescale(fmaxf(max_of<KK>(R), 1e-37f), &pex);
#pragma unroll
          for (int i = 0; i < KK; ++i) Pm[i] = R[i] * sc;
        }
      }
    }
#pragma unroll
    for (int g = 1; g < 4; ++g) {
      float vpre[K];
#pragma unroll
      for (int j = 0; j < K; ++j) vpre[j] = __shfl_sync(0xffffffffu, pre[j], 8 * g - 1);
      const int vex = __shfl_sync(0xffffffffu, pex, 8 * g - 1);
      if ((lane >> 3) == g) {
        float n[K];
#pragma unroll
        for (int j = 0; j < K; ++j) {
          float acc = vpre[0] * Pm[j];
#pragma unroll
          for (int k = 1; k < K; ++k) acc = fmaf(vpre[k], Pm[k * K + j], acc);
          n[j] = acc;
        }
        pex += vex;
        const float sc = pow2_rescale(fmaxf(max_of<K>(n), 1e-37f), &pex);
#pragma unroll
        for (int j = 0; j < K; ++j) pre[j] = n[j] * sc;
      }
    }
  }
}

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(nthreads) : "memory");
}

// VIT (cfg-3 decode: Viterbi path + gamma from ONE read of the inputs).  The Viterbi recursion is bit-exact max-plus in log
// space (hmm_small.cuh: adds and strict compares in the oracle's order), so it cannot be re-associated into a scan: it is a
// serial chain of T steps per sequence (~16 cycles each).  Run by a lane of the one-pass kernel's own warp it would use 1/32
// of an issue slot per instruction (5.8 ms by instruction count); as a kernel of its own it reads the 48 B/step a second time
// (3.4 ms).  Here the CTA has `wpb` posterior warps (one sequence each, as before) plus ONE helper warp whose lane s walks
// sequence s of the CTA on the raw log_A / log_b that the posterior warp has just landed in shared memory, BEFORE pass 1
// overwrites them with P_t (named barriers 1: inputs landed, 2: helper done reading).  The back-track and the path stores
// run after barrier 2, under the posterior passes.  The regions of consecutive warps are skewed by 16 bytes so that the
// helper's 16-byte loads (one region per lane) are bank-conflict free.
// MEASURED (B = 2^20, T = 256, K = 3): 14.5 ms against 7.0 ms for the two kernels -- NOT the default.  The barriers alone
// cost little (4.4 ms with the helper's work removed, the posterior kernel alone takes 4.0), but the helper's forward sweep
// takes 19 us per trip instead of the ~4 us of its dependent chain: the posterior passes of the SM's other CTA are issue-bound
// and the helper gets a third of its scheduler's slots (230 instructions per 4 steps at IPC 0.4); the back-track adds 3.7 us.
// The Viterbi recursion needs ~2 k warp-instructions per sequence at 8 active lanes, half of the 4.5 k issue slots the SM has
// per sequence at the posterior kernel's pace, so no arrangement of this kind gets under ~6.5 ms.
// helper_viterbi_forward: returns the final state; psi words (one byte per step, 2 bits per target state) -> psw.
template <int K>
__device__ __forceinline__ int helper_viterbi_forward(const float* __restrict__ RA, const float* __restrict__ RB, int S, int SA, int SB,
                                                       const float* __restrict__ log_pi, int n, uint32_t* __restrict__ psw) {
  constexpr int KK = K * K;
  float dl[K];
#pragma unroll
  for (int j = 0; j < K; ++j) dl[j] = 0.f;
  for (int t = 0; t < n; t += 4) {
    const int c = t / S, r = t - c * S;              // S % 4 == 0: the four steps lie in one chunk
    float Av[4 * KK], Bv[4 * K];
    lds4<KK>(RA + c * SA + r * KK, Av);
    lds4<K>(RB + c * SB + r * K, Bv);
    uint32_t word = 0;
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      const float* la = Av + s * KK;
      const float* lb = Bv + s * K;
      if (t + s < n) {
        if (t + s == 0) {
#pragma unroll
          for (int j = 0; j < K; ++j) dl[j] = __fadd_rn(log_pi[j], lb[j]);
        } else {
          float nd[K];
          uint32_t packed = 0;
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
            packed |= (uint32_t)arg << (2 * j);
          }
#pragma unroll
          for (int j = 0; j < K; ++j) dl[j] = nd[j];
          word |= packed << (8 * s);
        }
      }
    }
    psw[t >> 2] = word;
  }
  float best = dl[0];
  int state = 0;
#pragma unroll
  for (int j = 1; j < K; ++j)
    if (dl[j] > best) { best = dl[j]; state = j; }
  return state;
}
// back-track IN PLACE: the psi bytes of a row become its path bytes (0xFF beyond the length)
__device__ __forceinline__ void helper_backtrack(uint32_t* __restrict__ psw, int T, int n, int state) {
  const int nw = (T + 3) >> 2;
  for (int g = nw - 1; g >= 0; --g) {
    const uint32_t w = (g * 4 < n) ? psw[g] : 0u;
    uint32_t out = 0;
#pragma unroll
    for (int s = 3; s >= 0; --s) {
      const int t = g * 4 + s;
      if (t >= n) { out |= 0xFFu << (8 * s); continue; }
      out |= (uint32_t)state << (8 * s);
      if (t >= 1) state = (int)((w >> (8 * s + 2 * state)) & 3u);
    }
    psw[g] = out;
  }
}

// LL_ONLY: forward log-likelihood only (hmm_forward without all-prefix alpha): load, pass 1 without the
// write-back of P_t, prefix scan, log-likelihood -- no suffix scan, no sweeps, no stores.
// SEG (with LL_ONLY): the row is one T-step segment of a longer sequence; instead of a log-likelihood the kernel
// emits the segment's transfer-matrix product (conventions of hmm_scan.cuh: step 0 of a SEQUENCE is alpha_0, not a
// transfer) for the per-sequence carry scan -- the time-parallel scan of hmm_scan at one-pass bandwidth.
// ALPHA: all-prefix log alpha instead of gamma (SURVEY B.2 with the optional alpha output; the fill pass of vhmm_hmm_scan):
// rows are T-step segments of a sequence as in SEG, each with an incoming vector (vin, ll_in) from the carry scan -- or whole
// sequences (nseg = 1, vin == null: alpha_0 is formed here).  Load, pass 1 and the prefix scan as for gamma; pass 2 tracks
// the exact scale of alpha_hat (powers of two from the rescales + the subtracted emission maxima + ll_in) and writes
// log alpha_t = log(alpha_hat_t) + scale_t; rows beyond the length repeat the last valid row (the convention of hs::fwd_kernel).
template <int K, bool XI, bool LL_ONLY = false, bool SEG = false, bool VIT = false, bool ALPHA = false>
__global__ void __launch_bounds__(VIT ? 288 : 128) posterior_kernel(Args a) {
  static_assert(!SEG || LL_ONLY, "segment products come from the log-likelihood-only pass");
  static_assert(!VIT || (!XI && !LL_ONLY && !SEG && K <= 4), "decode: gamma + path, 2-bit back-pointers");
  static_assert(!ALPHA || (!XI && !LL_ONLY && !SEG && !VIT), "alpha rows replace gamma");
  constexpr bool ROWS = SEG || ALPHA;                      // rows are segments of sequences
  constexpr int KK = K * K;
  constexpr float LOG2E = 1.4426950408889634f, LN2 = 0.6931471805599453f;
  extern __shared__ __align__(16) float smem[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int wpb = (blockDim.x >> 5) - (VIT ? 1 : 0);      // posterior warps (VIT: the last warp is the Viterbi helper)
  const int S = a.S, SA = a.SA, SB = a.SB;
  const int REG = 32 * (SA + SB) + (VIT ? 4 : 0);         // floats per warp region (VIT: skewed by 16 bytes)
  float* regA = smem + (size_t)wib * REG;
  float* regB = regA + 32 * SA;
  float* myA = regA + lane * SA;
  float* myB = regB + lane * SB;
  const int T = (int)a.T;
  const int t0 = lane * S;
  const int GA = S * KK / 4, GB = S * K / 4;   // 16-byte granules per chunk
  float pi[K];
#pragma unroll
  for (int j = 0; j < K; ++j) pi[j] = a.log_pi[j] * LOG2E;

  const int64_t nwarps = (int64_t)gridDim.x * wpb;
  const int TP = (int)((a.T + 15) & ~(int64_t)15);        // VIT: bytes of one psi / path row in shared memory
  uint8_t* psi_sm = reinterpret_cast<uint8_t*>(smem + (size_t)wpb * REG);
  // (VIT: every warp of the CTA makes the same number of trips -- the condition is on the CTA's first sequence)
  for (int64_t b = (int64_t)blockIdx.x * wpb + wib; VIT ? (b - wib < a.B) : (b < a.B); b += nwarps) {
    if constexpr (VIT) {
      const int NTH = (wpb + 1) * 32;
      if (wib == wpb) {                                   // ---- the Viterbi helper warp: lane s <-> sequence base + s
        const int64_t base = b - wib;
        const int64_t bs = base + lane;
        const bool act = lane < wpb && bs < a.B;
        int n = 0;
        if (act) {
          const int64_t Lh = a.lengths[bs];
          n = (int)(Lh < 0 ? 0 : (Lh > a.T ? a.T : Lh));
        }
        uint32_t* psw = reinterpret_cast<uint32_t*>(psi_sm + (size_t)(act ? lane : 0) * TP);
        const float* RA = smem + (size_t)(act ? lane : 0) * REG;
        named_bar_sync(1, NTH);                           // the raw inputs of this trip are in shared memory
        int state = 0;
        if (act) state = helper_viterbi_forward<K>(RA, RA + 32 * SA, S, SA, SB, a.log_pi, n, psw);
        named_bar_sync(2, NTH);                           // ... and may be overwritten (pass 1)
        if (act) helper_backtrack(psw, (int)a.T, n, state);
        __syncwarp();
        const int nact = (int)((a.B - base < wpb) ? a.B - base : wpb);
        if (a.path_vec16) {
          const int gpr = (int)(a.T >> 4);                // 16-byte granules per row (T % 16 == 0: TP == T)
          for (int i = lane; i < nact * gpr; i += 32) {
            const int sq = i / gpr, g = i - sq * gpr;
            const uint4 v = *reinterpret_cast<const uint4*>(psi_sm + (size_t)sq * TP + g * 16);
            __stcs(reinterpret_cast<uint4*>(a.path + (base + sq) * a.T) + g, v);
          }
        } else {
          for (int sq = 0; sq < nact; ++sq)
            for (int i = lane; i < (int)a.T; i += 32) a.path[(base + sq) * a.T + i] = psi_sm[(size_t)sq * TP + i];
        }
        __syncwarp();
        continue;
      }
      if (b >= a.B) {                                     // a posterior warp without a sequence on the last trip
        named_bar_sync(1, NTH);
        named_bar_sync(2, NTH);
        continue;
      }
    }
    int64_t Ll;
    bool first = true;                               // row starts at step 0 of its sequence
    int64_t Lseq = 0;                                // ALPHA: length of the whole sequence
    int64_t step0 = b * (int64_t)T;                  // first step of the row in the (.., T_total, ..) tensors
    int capT = T;                                    // steps of the row that exist in memory
    if (ROWS || (!LL_ONLY && a.nseg > 1)) {          // (gamma / xi kernels take segment rows at run time)
      const int64_t seq = b / a.nseg;
      const int seg = (int)(b - seq * a.nseg);
      first = (seg == 0);
      Lseq = a.lengths[seq];
      Ll = Lseq - (int64_t)seg * a.T;
      const int64_t ST = a.seqT ? a.seqT : (int64_t)a.nseg * a.T;
      step0 = seq * ST + (int64_t)seg * a.T;
      const int64_t left = ST - (int64_t)seg * a.T;
      capT = (int)(left < a.T ? left : a.T);
    } else {
      Ll = a.lengths[b];
    }
    const int L = (int)(Ll < 0 ? 0 : (Ll > capT ? capT : Ll));
    const int rowGA = capT * KK / 4, rowGB = capT * K / 4;   // 16-byte granules of the row's output (capT * K % 4 == 0)
    const int nv = max(0, min(S, L - t0));          // valid steps of my chunk
    // ---- load: granules that hold a valid step --------------------------------------
    {
      const float* gA = a.A + step0 * KK;
      const float* gB = a.Bm + step0 * K;
      const int nA = (L * KK + 3) >> 2, nB = (L * K + 3) >> 2;
      for (int g = lane; g < nA; g += 32) {
        const int c = (int)__umulhi((unsigned)g, a.magicA);
        hs::cp_async16(regA + c * SA + (g - c * GA) * 4, gA + (int64_t)g * 4);
      }
      for (int g = lane; g < nB; g += 32) {
        const int c = (int)__umulhi((unsigned)g, a.magicB);
        hs::cp_async16(regB + c * SB + (g - c * GB) * 4, gB + (int64_t)g * 4);
      }
      hs::cp_commit();
      hs::cp_wait<0>();
      __syncwarp();
    }
    if constexpr (VIT) {
      const int NTH = (wpb + 1) * 32;
      named_bar_sync(1, NTH);            // inputs landed: the helper reads them
      named_bar_sync(2, NTH);            // the helper is through: pass 1 may overwrite them
    }
    // ---- pass 1: P_t in place, chunk product ----------------------------------------
    float M[KK];
#pragma unroll
    for (int i = 0; i < KK; ++i) M[i] = (i / K == i % K) ? 1.f : 0.f;
    int ex = 0;
    float mbsum = 0.f;
    float mb0 = 0.f;                                 // ALPHA without vin: the emission maximum of step 0 (scale of alpha_0)
    float v0[K];
#pragma unroll
    for (int j = 0; j < K; ++j) v0[j] = 0.f;
    for (int g = 0; g * 4 < nv; ++g) {
      float Av[4 * KK], Bv[4 * K];
      lds4<KK>(myA + g * 4 * KK, Av);
      lds4<K>(myB + g * 4 * K, Bv);
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        float* la = Av + s * KK;
        const float* lb = Bv + s * K;
        if (g * 4 + s < nv) {
          const float mb = max_of<K>(lb);
          const bool is0 = (t0 + g * 4 + s == 0) && first;
          if (!(ROWS && is0)) mbsum += mb;           // (SEG / ALPHA: the emission of step 0 belongs to alpha_0)
          float cj[K];
#pragma unroll
          for (int j = 0; j < K; ++j) cj[j] = (lb[j] - mb) * LOG2E;
          if (is0) {
            if (ALPHA) mb0 = mb;
#pragma unroll
            for (int j = 0; j < K; ++j) v0[j] = ex2f(pi[j] + cj[j]);
#pragma unroll
            for (int i = 0; i < KK; ++i) la[i] = (i / K == i % K) ? 1.f : 0.f;
          } else {
#pragma unroll
            for (int i = 0; i < K; ++i)
#pragma unroll
              for (int j = 0; j < K; ++j) la[i * K + j] = ex2f(fmaf(la[i * K + j], LOG2E, cj[j]));
            float R[KK];
            mat_mul<K>(R, M, la);
#pragma unroll
            for (int i = 0; i < KK; ++i) M[i] = R[i];
          }
        }
        if (s & 1) {
          const float sc = pow2_rescale(max_of<KK>(M), &ex);
#pragma unroll
          for (int i = 0; i < KK; ++i) M[i] *= sc;
        }
      }
      if (!LL_ONLY) sts4<KK>(myA + g * 4 * KK, Av);
    }
    if (SEG) {   // ordered product of the 32 chunk products (tree over the lanes), then out
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        float O[KK];
#pragma unroll
        for (int i = 0; i < KK; ++i) O[i] = __shfl_down_sync(0xffffffffu, M[i], d);
        const int oex = __shfl_down_sync(0xffffffffu, ex, d);
        if ((lane & (2 * d - 1)) == 0) {
          float R[KK];
          mat_mul<K>(R, M, O);
          ex += oex;
          const float sc = pow2_rescale(fmaxf(max_of<KK>(R), 1e-37f), &ex);
#pragma unroll
          for (int i = 0; i < KK; ++i) M[i] = R[i] * sc;
        }
      }
      const double ms = warp_sum((double)mbsum);
      if (lane == 0) {
#pragma unroll
        for (int i = 0; i < KK; ++i) a.Pmat[b * KK + i] = M[i];
        a.Pscale[b] = (double)ex * (double)LN2 + ms;
      }
      __syncwarp();
      continue;
    }
    // ---- scan over the lanes ---------------------------------------------------------
    if (!LL_ONLY && a.vin) {                         // the row's incoming vector replaces alpha_0
#pragma unroll
      for (int j = 0; j < K; ++j) v0[j] = a.vin[b * K + j];
    }
    float pre[K];
    int pex;
    warp_prefix_scan<K>(M, ex, v0, lane, pre, pex);
    // suffix: only the row sums (suffix product applied to the all-ones vector) are needed:
    // w_l = M_l M_{l+1} ... M_31 1.  Carry (w, Sm) with Sm the segment product.
    float suf[K];
    if (!LL_ONLY && !ALPHA) {
      float Sm[KK];
#pragma unroll
      for (int i = 0; i < KK; ++i) Sm[i] = M[i];
      bool has = (lane == 31);
      float rin[K];                                  // beta after the row's last step: ones, or the backward carry of later rows
#pragma unroll
      for (int j = 0; j < K; ++j) rin[j] = a.bin_in ? a.bin_in[b * K + j] : 1.f;
#pragma unroll
      for (int i = 0; i < K; ++i) {
        float acc = M[i * K] * rin[0];
#pragma unroll
        for (int j = 1; j < K; ++j) acc = fmaf(M[i * K + j], rin[j], acc);
        suf[i] = acc;
      }
      const int li = lane & 7;
#pragma unroll
      for (int d = 1; d < 8; d <<= 1) {
        float osuf[K], oSm[KK];
#pragma unroll
        for (int j = 0; j < K; ++j) osuf[j] = __shfl_down_sync(0xffffffffu, suf[j], d);
#pragma unroll
        for (int i = 0; i < KK; ++i) oSm[i] = __shfl_down_sync(0xffffffffu, Sm[i], d);
        const bool ohas = __shfl_down_sync(0xffffffffu, (int)has, d) != 0;
        if (li + d < 8 && !has) {
          if (ohas) {
#pragma unroll
            for (int i = 0; i < K; ++i) {
              float acc = Sm[i * K] * osuf[0];
#pragma unroll
              for (int k = 1; k < K; ++k) acc = fmaf(Sm[i * K + k], osuf[k], acc);
              suf[i] = acc;
            }
            has = true;
            const float sc = pow2_rescale(fmaxf(max_of<K>(suf), 1e-37f));
#pragma unroll
            for (int j = 0; j < K; ++j) suf[j] *= sc;
          } else {
            float R[KK];
            mat_mul<K>(R, Sm, oSm);
            const float sc = pow2_rescale(fmaxf(max_of<KK>(R), 1e-37f));
#pragma unroll
            for (int i = 0; i < KK; ++i) Sm[i] = R[i] * sc;
          }
        }
      }
#pragma unroll
      for (int g = 2; g >= 0; --g) {   // hand the carry of group g+1 (its first lane) to group g
        float vs[K];
#pragma unroll
        for (int j = 0; j < K; ++j) vs[j] = __shfl_sync(0xffffffffu, suf[j], 8 * g + 8);
        if ((lane >> 3) == g) {
          float n[K];
#pragma unroll
          for (int i = 0; i < K; ++i) {
            float acc = Sm[i * K] * vs[0];
#pragma unroll
            for (int k = 1; k < K; ++k) acc = fmaf(Sm[i * K + k], vs[k], acc);
            n[i] = acc;
          }
          const float sc = pow2_rescale(fmaxf(max_of<K>(n), 1e-37f));
#pragma unroll
          for (int j = 0; j < K; ++j) suf[j] = n[j] * sc;
        }
      }
    }
    // log-likelihood: lane 31 holds v0 M_0 ... M_31 (scaled by 2^-pex)
    if (a.loglik) {
      double ms = warp_sum((double)mbsum + (double)mb0);   // (mb0: ALPHA keeps the step-0 maximum apart; zero otherwise)
      if (lane == 31) {
        float s = pre[0];
#pragma unroll
        for (int j = 1; j < K; ++j) s += pre[j];
        a.loglik[b] = (L > 0) ? (float)((double)logf(s) + (double)pex * (double)LN2 + ms) : 0.f;
      }
    }
    if (LL_ONLY) { __syncwarp(); continue; }
    if constexpr (ALPHA) {
      // ---- all-prefix log alpha ------------------------------------------------------------
      // start vector of my chunk = prefix of the previous lane, true value fin * 2^E * exp(MB + LLin)
      float fin[K];
      int E;
      {
        const int pe = __shfl_up_sync(0xffffffffu, pex, 1);
#pragma unroll
        for (int j = 0; j < K; ++j) {
          const float f = __shfl_up_sync(0xffffffffu, pre[j], 1);
          fin[j] = (lane == 0) ? v0[j] : f;
        }
        E = (lane == 0) ? 0 : pe;
        const float sf = pow2_rescale(fmaxf(max_of<K>(fin), 1e-37f), &E);
#pragma unroll
        for (int j = 0; j < K; ++j) fin[j] *= sf;
      }
      // emission maxima of the steps before my chunk (exclusive scan over the lanes) + the row's incoming log-scale
      double MB = (double)mbsum;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const double o = __shfl_up_sync(0xffffffffu, MB, d);
        if (lane >= d) MB += o;
      }
      MB -= (double)mbsum;
      const double LLin = a.vin ? a.ll_in[b] : (double)__shfl_sync(0xffffffffu, mb0, 0);
      MB += LLin;
      float last[K];                                   // my last valid row (carried over the padding steps)
#pragma unroll
      for (int j = 0; j < K; ++j) last[j] = 0.f;
      {
        float al[K];
#pragma unroll
        for (int j = 0; j < K; ++j) al[j] = fin[j];
        double mbr = 0.0;
        for (int g = 0; g * 4 < nv; ++g) {
          float Av[4 * KK], Bv[4 * K], Ov[4 * K];
          lds4<KK>(myA + g * 4 * KK, Av);
          lds4<K>(myB + g * 4 * K, Bv);
#pragma unroll
          for (int s = 0; s < 4; ++s) {
            const float* P = Av + s * KK;
            if (g * 4 + s < nv) {
              const bool is0 = (t0 + g * 4 + s == 0) && first;
              if (!is0) mbr += (double)max_of<K>(Bv + s * K);
              float n[K];
#pragma unroll
              for (int j = 0; j < K; ++j) {
                float acc = al[0] * P[j];
#pragma unroll
                for (int k = 1; k < K; ++k) acc = fmaf(al[k], P[k * K + j], acc);
                n[j] = acc;
              }
              const float sc = pow2_rescale(fmaxf(max_of<K>(n), 1e-37f), &E);
              const float off = (float)((double)E * (double)LN2 + MB + mbr);
#pragma unroll
              for (int j = 0; j < K; ++j) { al[j] = n[j] * sc; last[j] = logf(al[j]) + off; }
            }
#pragma unroll
            for (int j = 0; j < K; ++j) Ov[s * K + j] = last[j];
          }
          sts4<K>(myB + g * 4 * K, Ov);
        }
      }
      // padding steps (t >= L inside the row): the last valid row of the SEQUENCE -- held by the lane of step L-1, or, for a
      // row that lies entirely beyond the length, the incoming vector itself; a sequence of length 0 reads log(1/K)
      {
        const int lv = (L > 0) ? (L - 1) / S : 0;
        float fill[K];
#pragma unroll
        for (int j = 0; j < K; ++j) fill[j] = __shfl_sync(0xffffffffu, last[j], lv);
        if (L == 0) {
#pragma unroll
          for (int j = 0; j < K; ++j)
            fill[j] = (a.vin && Lseq > 0) ? logf(fmaxf(a.vin[b * K + j], 1e-37f)) + (float)a.ll_in[b] : -logf((float)K);
        }
        const int cap = max(0, min(S, capT - t0));     // steps of my chunk that exist
        for (int sidx = nv; sidx < cap; ++sidx)
#pragma unroll
          for (int j = 0; j < K; ++j) myB[sidx * K + j] = fill[j];
      }
      __syncwarp();
      {
        float* gG = a.gamma + step0 * K;
        for (int g = lane; g < rowGB; g += 32) {
          const int c = (int)__umulhi((unsigned)g, a.magicB);
          const float4 v = *reinterpret_cast<const float4*>(regB + c * SB + (g - c * GB) * 4);
          __stcs(reinterpret_cast<float4*>(gG) + g, v);
        }
      }
      __syncwarp();
      continue;
    }
    // start vectors of my chunk
    float fin[K], bin[K];
#pragma unroll
    for (int j = 0; j < K; ++j) {
      const float f = __shfl_up_sync(0xffffffffu, pre[j], 1);
      const float w = __shfl_down_sync(0xffffffffu, suf[j], 1);
      fin[j] = (lane == 0) ? v0[j] : f;
      bin[j] = (lane == 31) ? (a.bin_in ? a.bin_in[b * K + j] : 1.f) : w;
    }
    {
      const float sf = pow2_rescale(fmaxf(max_of<K>(fin), 1e-37f));
      const float sb = pow2_rescale(fmaxf(max_of<K>(bin), 1e-37f));
#pragma unroll
      for (int j = 0; j < K; ++j) { fin[j] *= sf; bin[j] *= sb; }
    }
    // ---- pass 2: forward sweep, alpha_hat over the log_b slots -------------------------
    {
      float al[K];
#pragma unroll
      for (int j = 0; j < K; ++j) al[j] = fin[j];
      for (int g = 0; g * 4 < nv; ++g) {
        float Av[4 * KK], Ov[4 * K];
        lds4<KK>(myA + g * 4 * KK, Av);
#pragma unroll
        for (int s = 0; s < 4; ++s) {
          const float* P = Av + s * KK;
          if (g * 4 + s < nv) {
            float n[K];
#pragma unroll
            for (int j = 0; j < K; ++j) {
              float acc = al[0] * P[j];
#pragma unroll
              for (int k = 1; k < K; ++k) acc = fmaf(al[k], P[k * K + j], acc);
              n[j] = acc;
            }
            const float sc = pow2_rescale(fmaxf(max_of<K>(n), 1e-37f));
#pragma unroll
            for (int j = 0; j < K; ++j) al[j] = n[j] * sc;
          }
#pragma unroll
          for (int j = 0; j < K; ++j) Ov[s * K + j] = al[j];
        }
        sts4<K>(myB + g * 4 * K, Ov);
      }
    }
    // ---- pass 3: backward sweep: gamma, xi, beta_hat ------------------------------------
    {
      float bh[K];
#pragma unroll
      for (int j = 0; j < K; ++j) bh[j] = bin[j];
      const int cap = max(0, min(S, capT - t0));     // steps of my chunk that exist
      for (int g = (cap + 3) / 4 - 1; g >= 0; --g) {
        float Gv[4 * K];
        if (g * 4 >= nv) {
          // group entirely beyond the sequence length: zeros
#pragma unroll
          for (int i = 0; i < 4 * K; ++i) Gv[i] = 0.f;
          sts4<K>(myB + g * 4 * K, Gv);
          if (XI) {
            float Z[4 * KK];
#pragma unroll
            for (int i = 0; i < 4 * KK; ++i) Z[i] = 0.f;
            sts4<KK>(myA + g * 4 * KK, Z);
          }
          continue;
        }
        float Av[4 * KK], Hv[4 * K], prev[K];
        lds4<KK>(myA + g * 4 * KK, Av);
        lds4<K>(myB + g * 4 * K, Hv);
        if (XI) {
          // alpha_hat of the step before this group
          if (g == 0) {
#pragma unroll
            for (int j = 0; j < K; ++j) prev[j] = fin[j];
          } else {
#pragma unroll
            for (int j = 0; j < K; ++j) prev[j] = myB[g * 4 * K - K + j];
          }
        }
#pragma unroll
        for (int s = 3; s >= 0; --s) {
          float* P = Av + s * KK;
          float* gm = Gv + s * K;
          if (g * 4 + s >= nv) {
#pragma unroll
            for (int j = 0; j < K; ++j) gm[j] = 0.f;
            if (XI) {
#pragma unroll
              for (int i = 0; i < KK; ++i) P[i] = 0.f;
            }
            continue;
          }
          const float* ah = Hv + s * K;
          float gs = 0.f;
#pragma unroll
          for (int j = 0; j < K; ++j) { gm[j] = ah[j] * bh[j]; gs += gm[j]; }
          const float ginv = __frcp_rn(gs);
#pragma unroll
          for (int j = 0; j < K; ++j) gm[j] *= ginv;
          float nb[K];
#pragma unroll
          for (int i = 0; i < K; ++i) {
            float acc = P[i * K] * bh[0];
#pragma unroll
            for (int j = 1; j < K; ++j) acc = fmaf(P[i * K + j], bh[j], acc);
            nb[i] = acc;
          }
          if (XI) {
            if (t0 + g * 4 + s == 0 && first) {      // (step 0 of the SEQUENCE has no incoming transition)
#pragma unroll
              for (int i = 0; i < KK; ++i) P[i] = 0.f;
            } else {
              const float* ap = (s == 0) ? prev : (Hv + (s - 1) * K);
              float xs = 0.f;
#pragma unroll
              for (int i = 0; i < K; ++i)
#pragma unroll
                for (int j = 0; j < K; ++j) {
                  P[i * K + j] = ap[i] * P[i * K + j] * bh[j];
                  xs += P[i * K + j];
                }
              const float xinv = __frcp_rn(xs);
#pragma unroll
              for (int i = 0; i < KK; ++i) P[i] *= xinv;
            }
          }
          const float sc = pow2_rescale(fmaxf(max_of<K>(nb), 1e-37f));
#pragma unroll
          for (int j = 0; j < K; ++j) bh[j] = nb[j] * sc;
        }
        sts4<K>(myB + g * 4 * K, Gv);
        if (XI) sts4<KK>(myA + g * 4 * KK, Av);
      }
    }
    __syncwarp();
    // ---- store: coalesced rows ------------------------------------------------------
    {
      float* gG = a.gamma + step0 * K;
      for (int g = lane; g < rowGB; g += 32) {
        const int c = (int)__umulhi((unsigned)g, a.magicB);
        const float4 v = *reinterpret_cast<const float4*>(regB + c * SB + (g - c * GB) * 4);
        __stcs(reinterpret_cast<float4*>(gG) + g, v);
      }
      if (XI) {
        float* gX = a.xi + step0 * KK;
        for (int g = lane; g < rowGA; g += 32) {
          const int c = (int)__umulhi((unsigned)g, a.magicA);
          const float4 v = *reinterpret_cast<const float4*>(regA + c * SA + (g - c * GA) * 4);
          __stcs(reinterpret_cast<float4*>(gX) + g, v);
        }
      }
    }
    __syncwarp();
  }
}

}  // namespace hw
}  // namespace vhmm
