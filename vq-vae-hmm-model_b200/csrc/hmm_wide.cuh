// hmm_wide.cuh -- exact-HMM recursions for 4 < K <= 32 (the scaled config's K = 32): one WARP per sequence,
// lane = state, HBM-bound by design (4(K^2+K) = 4,224 B per step at K = 32).
//
// Replaces the first-round correctness kernels of hmm_generic.cuh, which walked log-space recursions straight out
// of global memory (two dependent sweeps over the K rows per step, no prefetch: 13 us per step and warp at K = 32,
// 4..10 % of the HBM roofline).  Here:
//   * lane j reads COLUMN j of log_A_t (row i = 32 consecutive floats over the lanes: every load is one coalesced
//     128-byte line) into registers, one step AHEAD of the recursion (register double buffer), so the recursion's
//     dependent chain never waits for HBM;
//   * sum-product runs in scaled linear space exactly like the small-K kernels: P_t[i,j] = exp(log_A_t[i,j] +
//     log_b_t[j] - max_j log_b_t) (the K exps per lane are independent of the recursion), alpha_hat_t =
//     normalise(alpha_hat_{t-1} P_t) with K shuffle-broadcasts + FMAs on four accumulators, log-normalisers summed
//     in double.  No logsumexp chain, no cancellation (all terms non-negative);
//   * backward: beta_hat_{t-1}[i] = sum_j P_t[i,j] beta_hat_t[j] needs ROW sums of what the lanes hold as columns:
//     each lane scales its column by beta_hat_t[j], the warp writes the K x K tile to shared memory (pitch K+1) and
//     lane i sums row i -- conflict-free both ways; xi_t[i,j] = alpha_hat_{t-1}[i] P_t[i,j] beta_hat_t[j] / Z is
//     formed in the column layout, i.e. stored with coalesced 128-byte rows;
//   * Viterbi keeps the oracle's operation order (one rounded add per candidate, strict '>' scanning i upwards, one
//     rounded add of log_b): bit-exact paths; back-pointers one byte per state, back-tracking in blocks of 32 steps
//     staged through shared memory (the dependent chain runs on LDS latency, not on HBM latency).
// Maths: SURVEY.md Appendix B.2-B.4; conventions of oracle/hmm.py.
#pragma once
#include "common.cuh"

namespace vhmm {
namespace hx {

constexpr float LOG2E = 1.4426950408889634f;
constexpr int WPB = 4;   // warps per CTA

__device__ __forceinline__ float wmax(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float wsum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// column `lane` of log_A_t (K x K row-major at At) -> a[i], i < K; -inf for i >= K or lanes >= K
template <int KP>
__device__ __forceinline__ void load_col(const float* __restrict__ At, int K, int lane, bool on, float (&a)[KP]) {
#pragma unroll
  for (int i = 0; i < KP; ++i) a[i] = (on && i < K) ? __ldcs(At + i * K + lane) : -INFINITY;
}

// ---------------------------------------------------------------------------------------------------------
// forward: loglik (B); optional log-space alpha (B,T,K) (carried unchanged for t >= L, B.2) and/or the
// normalised linear alpha_hat (B,T,K) (stash for bwd_kernel)
// ---------------------------------------------------------------------------------------------------------
// A warp's unit of work is a whole sequence (nseg <= 1) or ROW s of a sequence cut into `seg`-step rows (hmm_wide_scan.cuh:
// few long sequences): the row starts from the normalised alpha_hat of the step before it (vin) and, in the backward kernel,
// from the beta_hat of its last step (bin); both come from the carry kernels over the segment products.
struct RowView {
  int nseg, seg;
  const float* vin;      // (rows, K) alpha_hat at the step before the row (row 0: unused)
  const float* bin;      // (rows, K) beta_hat at the row's last valid step (last valid row: unused, ones)
  const int* only;       // per sequence: 0 = skip (redo pass for flagged sequences), or null
  const double* ll_in;   // (rows) log-normaliser accumulated before the row   } all-prefix log alpha by rows
  const float* carried;  // (B, K) log alpha of the sequence's last valid step  }
};

template <int KP, bool ROWS = false>
__global__ void __launch_bounds__(WPB * 32)      // (a 128-register cap for 16 instead of 12 warps per SM measured slower: 72.6 vs 77.6 %)
fwd_kernel(const float* __restrict__ log_pi, const float* __restrict__ A, const float* __restrict__ Bm,
           const int64_t* __restrict__ lengths, int64_t nW, int T, int K, float* __restrict__ loglik,
           float* __restrict__ alpha, float* __restrict__ ahat, RowView rv) {
  const int lane = threadIdx.x & 31;
  const int64_t w = (int64_t)blockIdx.x * WPB + (threadIdx.x >> 5);
  if (w >= nW) return;
  int64_t b = w;
  int t0 = 0;
  if (ROWS) { b = w / rv.nseg; t0 = (int)(w - b * rv.nseg) * rv.seg; }
  if (rv.only && !rv.only[b]) return;
  const int64_t L64 = lengths[b];
  const int Ls = (int)(L64 < 0 ? 0 : (L64 > T ? T : L64));
  const int L = ROWS ? max(0, min(min(rv.seg, T - t0), Ls - t0)) : Ls;    // valid steps of this row
  const bool first = ROWS ? (t0 == 0) : true;
  const bool on = lane < K;
  const int KK = K * K;
  const int64_t s0 = b * (int64_t)T + t0;                                            // first step of the row
  const float* Ab = A + s0 * KK;
  const float* Bb = Bm + s0 * K;
  float v = 0.f;
  double ll = 0.0;
  if (L > 0 && first) {
    const float x = on ? log_pi[lane] + Bb[lane] : -INFINITY;
    const float m = wmax(x);
    v = on ? ex2((x - m) * LOG2E) : 0.f;
    const float s = wsum(v);
    v *= 1.f / s;
    ll = (double)m + (double)logf(s);
    if (on) {
      if (alpha) alpha[s0 * K + lane] = logf(v) + (float)ll;
      if (ahat) ahat[s0 * K + lane] = v;
    }
  } else if (ROWS && L > 0) {
    v = on ? rv.vin[w * K + lane] : 0.f;          // a later row: its step 0 is a recursion step like any other
    if (alpha) ll = rv.ll_in[w];
  }
  const int tb = first ? 1 : 0;
  float an[KP], bn = -INFINITY;
  if (L > tb) {
    load_col<KP>(Ab + (int64_t)tb * KK, K, lane, on, an);
    bn = on ? __ldcs(Bb + (int64_t)tb * K + lane) : -INFINITY;
  }
  for (int t = tb; t < L; ++t) {
    float ac[KP];
#pragma unroll
    for (int i = 0; i < KP; ++i) ac[i] = an[i];
    const float bc = bn;
    if (t + 1 < L) {   // next step's inputs are in flight while this step computes
      load_col<KP>(Ab + (int64_t)(t + 1) * KK, K, lane, on, an);
      bn = on ? __ldcs(Bb + (int64_t)(t + 1) * K + lane) : -INFINITY;
    }
    const float mb = wmax(bc);
    const float cj = on ? (bc - mb) * LOG2E : -INFINITY;
    float w0 = 0.f, w1 = 0.f, w2 = 0.f, w3 = 0.f;
#pragma unroll
    for (int i = 0; i < KP; ++i) {
      const float p = ex2(fmaf(ac[i], LOG2E, cj));          // 0 for i >= K, lanes >= K and structural zeros
      const float vi = __shfl_sync(0xffffffffu, v, i);
      if ((i & 3) == 0) w0 = fmaf(vi, p, w0);
      else if ((i & 3) == 1) w1 = fmaf(vi, p, w1);
      else if ((i & 3) == 2) w2 = fmaf(vi, p, w2);
      else w3 = fmaf(vi, p, w3);
    }
    const float wv = on ? (w0 + w1) + (w2 + w3) : 0.f;
    const float s = wsum(wv);
    v = wv * (1.f / s);
    ll += (double)mb + (double)logf(s);
    if (on) {
      if (alpha) alpha[(s0 + t) * K + lane] = logf(v) + (float)ll;
      if (ahat) ahat[(s0 + t) * K + lane] = v;
    }
  }
  if (ROWS) {                                      // rows: steps beyond the length repeat the sequence's last valid row (B.2)
    if (alpha && on) {
      const float carried = rv.carried[b * K + lane];
      const int cap = min(rv.seg, T - t0);
      for (int t = L; t < cap; ++t) alpha[(s0 + t) * K + lane] = carried;
    }
    return;
  }
  if (lane == 0 && loglik) loglik[b] = (L > 0) ? (float)ll : 0.f;
  if (alpha && on) {
    const float carried = logf(v) + (float)ll;
    for (int t = (L > 0 ? L : 0); t < T; ++t) alpha[(b * T + t) * K + lane] = carried;
  }
}

// ---------------------------------------------------------------------------------------------------------
// backward: gamma (B,T,K) and, if XI, xi (B,T,K,K) from the alpha_hat stash (zeros for t >= L; xi[:,0] = 0)
// ---------------------------------------------------------------------------------------------------------
template <int KP, bool XI, bool ROWS = false>
__global__ void __launch_bounds__(WPB * 32)
bwd_kernel(const float* __restrict__ A, const float* __restrict__ Bm, const int64_t* __restrict__ lengths,
           int64_t nW, int T, int K, const float* __restrict__ ahat, float* __restrict__ gamma,
           float* __restrict__ xi, RowView rv) {
  __shared__ float tr[WPB][KP][KP + 1];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int64_t w = (int64_t)blockIdx.x * WPB + wib;
  if (w >= nW) return;
  int64_t b = w;
  int t0 = 0;
  if (ROWS) { b = w / rv.nseg; t0 = (int)(w - b * rv.nseg) * rv.seg; }
  if (rv.only && !rv.only[b]) return;
  const int64_t L64 = lengths[b];
  const int Ls = (int)(L64 < 0 ? 0 : (L64 > T ? T : L64));
  const int cap = ROWS ? min(rv.seg, T - t0) : T;          // steps that exist in this row
  const int L = ROWS ? max(0, min(cap, Ls - t0)) : Ls;      // valid steps of this row
  const bool first = ROWS ? (t0 == 0) : true;
  const bool on = lane < K;
  const int KK = K * K;
  const int64_t s0 = b * (int64_t)T + t0;
  const float* Ab = A + s0 * KK;
  const float* Bb = Bm + s0 * K;
  const float* ah = ahat + s0 * K;                // (a later row reads ah[-1 * K ..]: the previous row's last alpha_hat)
  float* gm = gamma + s0 * K;
  float* xb = XI ? xi + s0 * KK : nullptr;
  // steps beyond the length (and xi of step 0): zeros
  for (int64_t e = (int64_t)L * K + lane; e < (int64_t)cap * K; e += 32) gm[e] = 0.f;
  if (XI) {
    for (int64_t e = (int64_t)L * KK + lane; e < (int64_t)cap * KK; e += 32) xb[e] = 0.f;
    if (L > 0 && first) for (int e = lane; e < KK; e += 32) xb[e] = 0.f;
  }
  if (L <= 0) return;
  float (*mytr)[KP + 1] = tr[wib];
  // beta_hat at the row's last valid step (any positive scale: gamma / xi are normalised per step)
  float be = on ? 1.f : 0.f;
  if (ROWS && t0 + L < Ls) be = on ? rv.bin[w * K + lane] : 0.f;
  const int tb = first ? 1 : 0;                   // steps tb .. L-1 have a transition INTO them inside this row's reach
  float an[KP], bn = -INFINITY, apn = 0.f;
  float at = on ? ah[(int64_t)(L - 1) * K + lane] : 0.f;
  if (L - 1 >= tb) {
    load_col<KP>(Ab + (int64_t)(L - 1) * KK, K, lane, on, an);
    bn = on ? __ldcs(Bb + (int64_t)(L - 1) * K + lane) : -INFINITY;
    apn = on ? ah[(int64_t)(L - 2) * K + lane] : 0.f;
  }
  for (int t = L - 1; t >= 0; --t) {
    {
      const float g = at * be;
      const float sg = wsum(g);
      if (on) gm[(int64_t)t * K + lane] = g * (1.f / sg);
    }
    if (t < tb) break;
    float ac[KP];
#pragma unroll
    for (int i = 0; i < KP; ++i) ac[i] = an[i];
    const float bc = bn, ap = apn;
    if (t - 1 >= tb) {
      load_col<KP>(Ab + (int64_t)(t - 1) * KK, K, lane, on, an);
      bn = on ? __ldcs(Bb + (int64_t)(t - 1) * K + lane) : -INFINITY;
      apn = on ? ah[(int64_t)(t - 2) * K + lane] : 0.f;
    }
    const float mb = wmax(bc);
    const float cj = on ? (bc - mb) * LOG2E : -INFINITY;
    float q[KP];                                   // q[i] = P_t[i, lane] * beta_hat_t[lane]
#pragma unroll
    for (int i = 0; i < KP; ++i) q[i] = ex2(fmaf(ac[i], LOG2E, cj)) * be;
    if (XI) {
      float zs = 0.f;
#pragma unroll
      for (int i = 0; i < KP; ++i) zs = fmaf(__shfl_sync(0xffffffffu, ap, i), q[i], zs);
      const float iz = 1.f / wsum(zs);
      float* xr = xb + (int64_t)t * KK;
#pragma unroll
      for (int i = 0; i < KP; ++i) {
        const float ai = __shfl_sync(0xffffffffu, ap, i);
        if (on && i < K) xr[i * K + lane] = ai * q[i] * iz;
      }
    }
    if (ROWS && t == 0) break;                     // first step of a later row: its xi only (beta of the step before is the previous row's bin)
    // row sums through shared memory: tile[i][j] = q_j[i]; lane i sums row i
    __syncwarp();
    if (lane < KP) {
#pragma unroll
      for (int i = 0; i < KP; ++i) mytr[i][lane] = q[i];
    }
    __syncwarp();
    float s0a = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    if (lane < KP) {
#pragma unroll
      for (int j = 0; j < KP; j += 4) {
        s0a += mytr[lane][j]; s1 += mytr[lane][j + 1]; s2 += mytr[lane][j + 2]; s3 += mytr[lane][j + 3];
      }
    }
    const float bs = on ? (s0a + s1) + (s2 + s3) : 0.f;
    be = bs * (1.f / wmax(bs));                    // keep beta_hat in range (power-free rescale; any positive factor)
    at = ap;
  }
}

// ---------------------------------------------------------------------------------------------------------
// Viterbi (B.4): psi (B,T,K) uint8 workspace, path (B,T) uint8 (255 for t >= L), score (B) or null
// ---------------------------------------------------------------------------------------------------------
template <int KP>
__global__ void __launch_bounds__(WPB * 32, 5)   // <= 102 registers: 20 instead of 16 warps per SM, 58 -> 67 % of the HBM peak at K = 32
vit_kernel(const float* __restrict__ log_pi, const float* __restrict__ A, const float* __restrict__ Bm,
           const int64_t* __restrict__ lengths, int64_t nB, int T, int K, uint8_t* __restrict__ psi,
           uint8_t* __restrict__ path, float* __restrict__ score) {
  __shared__ uint8_t pb[WPB][32 * KP];
  __shared__ uint8_t ob[WPB][32];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int64_t b = (int64_t)blockIdx.x * WPB + wib;
  if (b >= nB) return;
  const int64_t L64 = lengths[b];
  const int L = (int)(L64 < 0 ? 0 : (L64 > T ? T : L64));
  const bool on = lane < K;
  const int KK = K * K;
  const float* Ab = A + b * (int64_t)T * KK;
  const float* Bb = Bm + b * (int64_t)T * K;
  uint8_t* ps = psi + b * (int64_t)T * K;
  uint8_t* pr = path + b * (int64_t)T;
  for (int t = L + lane; t < T; t += 32) pr[t] = VHMM_PAD_STATE;
  float dl = (L > 0 && on) ? __fadd_rn(log_pi[lane], Bb[lane]) : -INFINITY;
  float an[KP], bn = -INFINITY;
  if (L > 1) {
    load_col<KP>(Ab + KK, K, lane, on, an);
    bn = on ? __ldcs(Bb + K + lane) : -INFINITY;
  }
  for (int t = 1; t < L; ++t) {
    float ac[KP];
#pragma unroll
    for (int i = 0; i < KP; ++i) ac[i] = an[i];
    const float bc = bn;
    if (t + 1 < L) {
      load_col<KP>(Ab + (int64_t)(t + 1) * KK, K, lane, on, an);
      bn = on ? __ldcs(Bb + (int64_t)(t + 1) * K + lane) : -INFINITY;
    }
    float best = __fadd_rn(__shfl_sync(0xffffffffu, dl, 0), ac[0]);
    int arg = 0;
#pragma unroll
    for (int i = 1; i < KP; ++i) {
      const float c = __fadd_rn(__shfl_sync(0xffffffffu, dl, i), ac[i]);
      if (i < K && c > best) { best = c; arg = i; }      // strict: the lowest i wins ties
    }
    dl = on ? __fadd_rn(bc, best) : -INFINITY;
    if (on) ps[(int64_t)t * K + lane] = (uint8_t)arg;
  }
  const float m = wmax(dl);
  const unsigned ball = __ballot_sync(0xffffffffu, on && dl == m);
  int state = ball ? __ffs(ball) - 1 : 0;             // lowest index attaining the max
  if (lane == 0 && score) score[b] = (L > 0) ? m : 0.f;
  __syncwarp();                                        // psi written by the lanes is visible to the warp
  for (int hi = L - 1; hi >= 0; hi -= 32) {
    const int lo = hi - 31 > 0 ? hi - 31 : 0;
    const int n = hi - lo + 1;
    for (int e = lane; e < n * K; e += 32) pb[wib][e] = ps[(int64_t)lo * K + e];
    __syncwarp();
    if (lane == 0) {
      for (int t = hi; t >= lo; --t) {
        ob[wib][t - lo] = (uint8_t)state;
        if (t >= 1) state = pb[wib][(t - lo) * K + state];
      }
    }
    __syncwarp();
    state = __shfl_sync(0xffffffffu, state, 0);
    if (lane < n) pr[lo + lane] = ob[wib][lane];
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------------------------------
// Viterbi of FEW sequences: vit_kernel above keeps ONE step of log_A in flight per warp, which a batch of thousands hides and
// four sequences do not (1.1 us per step at K = 32: the HBM latency of the step's own tile).  Same recursion, same order of
// operations (bit-identical paths and scores), but one warp per CTA and the inputs streamed by lane 0 into a shared-memory
// ring with cp.async.bulk + mbarrier, 4 .. 32 steps per stage and VR_NST - 1 stages ahead of the dependent chain.  Needs K % 4 == 0
// (every step's tile and row 16-byte aligned) and 16-byte aligned log_A / log_b.
// ---------------------------------------------------------------------------------------------------------
constexpr int VR_NST = 8;                                               // stages of the ring (7 in flight)
template <int KP> struct VitRing { static constexpr int S = KP >= 32 ? 4 : (KP >= 16 ? 16 : 32); };   // steps per stage: 9 .. 17 KB
template <int KP>
constexpr size_t vit_ring_smem() { return (size_t)VR_NST * VitRing<KP>::S * (KP * KP + KP) * sizeof(float); }

template <int KP, bool EXACT>        // EXACT: K == KP (compile-time strides)
__global__ void __launch_bounds__(32)
vit_ring_kernel(const float* __restrict__ log_pi, const float* __restrict__ A, const float* __restrict__ Bm,
                const int64_t* __restrict__ lengths, int64_t nB, int T, int K, uint8_t* __restrict__ psi,
                uint8_t* __restrict__ path, float* __restrict__ score) {
  constexpr int VR_S = VitRing<KP>::S;
  extern __shared__ __align__(16) float ring[];
  __shared__ __align__(8) uint64_t full[VR_NST];
  __shared__ __align__(16) uint8_t pb[32 * KP];
  __shared__ uint8_t ob[32];
  __shared__ __align__(16) float dls[32];
  const int lane = threadIdx.x;
  const int64_t b = blockIdx.x;
  if (b >= nB) return;
  const int64_t L64 = lengths[b];
  const int L = (int)(L64 < 0 ? 0 : (L64 > T ? T : L64));
  const bool on = lane < K;
  const int KK = K * K;
  const float* Ab = A + b * (int64_t)T * KK;
  const float* Bb = Bm + b * (int64_t)T * K;
  uint8_t* ps = psi + b * (int64_t)T * K;
  uint8_t* pr = path + b * (int64_t)T;
  for (int t = L + lane; t < T; t += 32) pr[t] = VHMM_PAD_STATE;
  float dl = (L > 0 && on) ? __fadd_rn(log_pi[lane], Bb[lane]) : -INFINITY;
  const int stage_floats = VR_S * (KK + K);            // A tiles of the stage's steps, then their b rows
  auto s32 = [](const void* p) { return (uint32_t)__cvta_generic_to_shared(p); };
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < VR_NST; ++i)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(s32(&full[i])), "r"(1) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  const int nst = (L + VR_S - 1) / VR_S;               // stage c holds steps [c*VR_S, min(L, (c+1)*VR_S))
  auto issue = [&](int c) {
    float* st = ring + (c % VR_NST) * stage_floats;
    const int n = min(VR_S, L - c * VR_S);
    const uint32_t ba = (uint32_t)n * KK * 4u, bb = (uint32_t)n * K * 4u;
    uint64_t* bar = &full[c % VR_NST];
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(s32(bar)), "r"(ba + bb) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(s32(st)), "l"(Ab + (int64_t)c * VR_S * KK), "r"(ba), "r"(s32(bar)) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(s32(st + VR_S * KK)), "l"(Bb + (int64_t)c * VR_S * K), "r"(bb), "r"(s32(bar)) : "memory");
  };
  if (lane == 0)
    for (int c = 0; c < VR_NST - 1 && c < nst; ++c) issue(c);
  for (int c = 0; c < nst; ++c) {
    __syncwarp();                                      // every lane is done with the stage that is refilled now
    if (lane == 0 && c + VR_NST - 1 < nst) issue(c + VR_NST - 1);
    {
      const uint32_t a = s32(&full[c % VR_NST]), parity = (uint32_t)(c / VR_NST) & 1u;
      uint32_t ok = 0;
      for (int it = 0; it < (1 << 24) && !ok; ++it)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                     : "=r"(ok) : "r"(a), "r"(parity) : "memory");
      if (!ok) __trap();
    }
    const float* sA = ring + (c % VR_NST) * stage_floats;
    const float* sB = sA + VR_S * KK;
    const int n = min(VR_S, L - c * VR_S);
    for (int s = 0; s < n; ++s) {
      const int t = c * VR_S + s;
      if (t == 0) continue;                            // step 0 is the initial vector
      // every lane needs the whole vector dl: through shared memory (one store, KP/4 broadcast 16-byte loads) instead of
      // KP shuffles
      dls[lane] = dl;
      __syncwarp();
      float dv[KP];
#pragma unroll
      for (int q = 0; q < KP / 4; ++q) {
        const float4 v = reinterpret_cast<const float4*>(dls)[q];
        dv[4 * q] = v.x; dv[4 * q + 1] = v.y; dv[4 * q + 2] = v.z; dv[4 * q + 3] = v.w;
      }
      __syncwarp();
      // column `lane` of the step's tile (lanes >= K read neighbouring words of the stage: unused)
      const int Kc = EXACT ? KP : K;
      const float* col = sA + s * (Kc * Kc) + lane;
      const float bc = sB[s * Kc + (on ? lane : 0)];
      // max_i (dl[i] + a[i]) and the LOWEST i attaining it.  vit_kernel scans i upwards with a strict '>' -- a dependent chain
      // of K compare/select pairs; here a tree of (value, index) pairs whose nodes take the right operand only when it is
      // strictly greater: the same result (lowest index on ties), depth log2 K.
      float cv[KP];
      int ci[KP];
#pragma unroll
      for (int i = 0; i < KP; ++i) {
        cv[i] = (EXACT || i < K) ? __fadd_rn(dv[i], col[i * Kc]) : -INFINITY;
        ci[i] = i;
      }
#pragma unroll
      for (int w = 1; w < KP; w <<= 1) {
#pragma unroll
        for (int i = 0; i + w < KP; i += 2 * w) {
          const bool r = cv[i + w] > cv[i];
          cv[i] = r ? cv[i + w] : cv[i];
          ci[i] = r ? ci[i + w] : ci[i];
        }
      }
      dl = on ? __fadd_rn(bc, cv[0]) : -INFINITY;
      if (on) ps[(int64_t)t * K + lane] = (uint8_t)ci[0];
    }
  }
  const float m = wmax(dl);
  const unsigned ball = __ballot_sync(0xffffffffu, on && dl == m);
  int state = ball ? __ffs(ball) - 1 : 0;             // lowest index attaining the max
  if (lane == 0 && score) score[b] = (L > 0) ? m : 0.f;
  __syncwarp();                                        // psi written by the lanes is visible to the warp
  // back-track in 32-step blocks ALIGNED to multiples of 32 steps: a block's back-pointers are 32*K contiguous bytes that the
  // warp fetches with 16-byte loads one block AHEAD of the walk, lane 0 walks the block in shared memory
  const bool wide_ok = (((int64_t)T * K) & 15) == 0;  // (K % 4 == 0: a block is 2 K 16-byte words, at most 64)
  const int nv = 2 * K;
  uint4 pre[2];
  auto fetch = [&](int lo, int n) {                   // registers <- back-pointers of steps [lo, lo+n)
    if (wide_ok && n == 32) {
      const uint4* src = reinterpret_cast<const uint4*>(ps + (int64_t)lo * K);
      if (lane < nv) pre[0] = src[lane];
      if (32 + lane < nv) pre[1] = src[32 + lane];
    }
  };
  int hi = L - 1;
  int lo = hi >= 0 ? (hi & ~31) : 0;
  if (hi >= 0) fetch(lo, hi - lo + 1);
  while (hi >= 0) {
    const int n = hi - lo + 1;
    if (wide_ok && n == 32) {
      if (lane < nv) reinterpret_cast<uint4*>(pb)[lane] = pre[0];
      if (32 + lane < nv) reinterpret_cast<uint4*>(pb)[32 + lane] = pre[1];
    } else {
      for (int e = lane; e < n * K; e += 32) pb[e] = ps[(int64_t)lo * K + e];
    }
    if (lo >= 32) fetch(lo - 32, 32);                 // the next block, in flight during the walk
    __syncwarp();
    if (lane == 0) {
      for (int t = hi; t >= lo; --t) {
        ob[t - lo] = (uint8_t)state;
        if (t >= 1) state = pb[(t - lo) * K + state];
      }
    }
    __syncwarp();
    state = __shfl_sync(0xffffffffu, state, 0);
    if (lane < n) pr[lo + lane] = ob[lane];
    __syncwarp();
    hi = lo - 1;
    lo -= 32;
  }
}

}  // namespace hx
}  // namespace vhmm
