// elbo_fused.cu -- fused, recompute-in-tile implementation of VAE_HMM.compute_loss forward +
// backward (VQ_VAE_HMM_fixed.py:106-137 and loss.backward() at :156) for model dimensions that
// fit the tensor-core tiling (H multiple of 32, H2 of 16, TH of 64; C, U, K <= 8; K*K, 2C <= 16).
//
// No layer activation ever goes to HBM.  Three persistent kernels walk (sequence, 120-step tile)
// work items; a tile stages a 128-column window (4-step halo each side) of its inputs in shared
// memory as act[channel][132] and evaluates every layer on it (maths: SURVEY.md Appendix A.2):
//   A  enc_prior : x -> h1 -> h2 -> logits -> q (HBM, 12 B/step); entropy and init terms;
//                  u -> a -> M -> log_A; pairwise transition term; d loss/d M; d loss/d q of the
//                  prior (HBM); gradients of transition_net.{0,2}.
//   B  decoder   : q -> g1 -> g2 -> (mu, logvar); Gaussian NLL; backward through the decoder to
//                  d loss/d q (HBM); gradients of decoder.{conv1 (folded with the embedding),
//                  conv2, to_params}.
//   C  enc_bwd   : recompute h1, h2, q; d loss/d logits (decoder + prior + entropy parts);
//                  backward through the encoder; gradients of encoder.{conv1, conv2, to_logits}.
// The embedding is folded into decoder.conv1 (W'[o,k,tau] = sum_d W[o,d,tau] E[k,d], exact up to
// fp32 re-association) and unfolded again for the gradients of E and W.
// Weight gradients are accumulated per CTA in shared memory (fp32) over its tiles, written as
// per-CTA partials and reduced in double in a fixed order: results are run-to-run deterministic.
// All GEMM-shaped work runs on mma.sync TF32 with a 3-term hi/lo split (mma_tf32.cuh).
//
// Padding semantics (SURVEY.md finding 4): every layer is evaluated for all t in [0,T) of the
// padded batch; activations are zero only outside [0,T) (conv zero padding); `lengths` masks the
// loss terms only.  Tiles beyond the dependency cone of the valid steps are skipped (their
// gradients are exactly zero) and the consumers substitute zeros for what was not written.
#include <atomic>
#include <stdlib.h>
#include <string.h>
#include "common.cuh"
#include "mma_tf32.cuh"
#include "tc5.cuh"
#include "tma.cuh"

namespace vhmm {
namespace fz {

constexpr int TT = 120;    // owned steps per tile
constexpr int HL = 4;      // halo columns on each side
constexpr int NC = 128;    // columns per tile: column j <-> t = t0 - HL + j
constexpr int PT = 132;    // activation row pitch (floats), == 4 mod 32
constexpr int NTH = 512;   // threads per CTA (16 warps, <= 128 registers each)
constexpr int NW = NTH / 32;
constexpr int J0 = HL;     // first owned column
constexpr int KST = TT / 8;  // k-steps of a time GEMM over the owned columns
constexpr int LP = 16;     // doubles of loss partials per CTA

// optional per-phase cycle accounting (tools: -DVHMM_FUSED_PROFILE): thread 0 of CTA 0 prints the
// cycles between consecutive phase barriers at kernel exit
#ifdef VHMM_FUSED_PROFILE
#define PROF_DECL long long prof_t[20] = {0}; long long prof_prev = clock64();
#define PROF(id) do { if (threadIdx.x == 0) { long long n_ = clock64(); prof_t[id] += n_ - prof_prev; prof_prev = n_; } } while (0)
#define PROF_DUMP(name, n) do { if (threadIdx.x == 0 && blockIdx.x == 0) { long long tot_ = 0; for (int i_ = 0; i_ < n; ++i_) tot_ += prof_t[i_]; \
  printf("%s phases (%% of %lld cyc):", name, tot_); for (int i_ = 0; i_ < n; ++i_) printf(" %d:%.1f", i_, 100.0 * prof_t[i_] / tot_); printf("\n"); } } while (0)
#define PROF_PARAMS , long long* prof_t, long long& prof_prev
#define PROF_ARGS , prof_t, prof_prev
#else
#define PROF_DECL
#define PROF(id)
#define PROF_DUMP(name, n)
#define PROF_PARAMS
#define PROF_ARGS
#endif

template <int C_, int H_, int H2_, int K_, int U_, int TH_>
struct Cfg {
  static constexpr int C = C_, H = H_, H2 = H2_, K = K_, U = U_, TH = TH_, K2 = K_ * K_;
  static_assert(H % 32 == 0 && H2 % 16 == 0 && TH % 32 == 0, "channel counts must tile the MMA shapes");
  static_assert(C <= 8 && U <= 8 && K <= 8 && K2 <= 16 && 2 * C <= 16, "small dims exceed the padded tiles");
  static constexpr int PW_H = H + 8, PW_H2 = H2 + 8, PW_TH = TH + 8;   // weight pitches
  static_assert(PW_H % 32 == 8 && PW_H2 % 32 == 8 && PW_TH % 32 == 8, "weight pitch must be 8 mod 32");
  // per-CTA gradient accumulator layouts (floats); identical to the staged weight layouts
  // kernel A: [Wt2 16 x PW_TH][Wt0 TH x 8][bt2 16][bt0 TH]
  static constexpr int A_WT2 = 0, A_WT0 = A_WT2 + 16 * PW_TH, A_BT2 = A_WT0 + TH * 8, A_BT0 = A_BT2 + 16,
                       ACC_A = A_BT0 + TH;
  // kernel B: [W2d 3 x H x PW_H][W' 3 x H x 8][Wp 16 x PW_H][bd2 H][bd1 H][bp 16]
  static constexpr int B_W2 = 0, B_WF = B_W2 + 3 * H * PW_H, B_WP = B_WF + 3 * H * 8, B_BD2 = B_WP + 16 * PW_H,
                       B_BD1 = B_BD2 + H, B_BP = B_BD1 + H, ACC_B = B_BP + 16;
  // kernel C: [W2e 3 x H2 x PW_H][W1 3 x H x 8][W3 16 x PW_H2][b2 H2][b1 H][b3 16]
  static constexpr int C_W2 = 0, C_W1 = C_W2 + 3 * H2 * PW_H, C_W3 = C_W1 + 3 * H * 8, C_B2 = C_W3 + 16 * PW_H2,
                       C_B1 = C_B2 + H2, C_B3 = C_B1 + H, ACC_C = C_B3 + 16;
  static constexpr int ACC_ALL = ACC_A + ACC_B + ACC_C;
  // shared memory (floats)
  static constexpr int R1 = (H + H2 > TH) ? H + H2 : TH;   // h1|h2 region, reused for the prior's hidden layer
  static constexpr int SM_A = 3 * H * 8 + 3 * H2 * PW_H + 16 * PW_H2 + TH * 8 + 16 * PW_TH + (H + H2 + 16 + TH + 16 + 8) +
                              (16 + R1 + 16 + 16 + 16) * PT + ACC_A + 16 * PT + H2 * PT;   // + second input buffer (prefetch) + h2 (tcgen05 build)
  static constexpr int SM_B = 3 * H * 8 + 3 * H * PW_H + 16 * PW_H + (H + H + 16) + (16 + H + H + 16) * PT + ACC_B;
  static constexpr int SM_C = 3 * H * 8 + 3 * H2 * PW_H + 16 * PW_H2 + (H + H2 + 16) + (16 + H + H2 + 16 + 16) * PT + ACC_C + 32 * PT;   // + second input buffers (prefetch)
};

// ---- helpers --------------------------------------------------------------------------------
// conv weight (O, Cin, KW) -> dst[tau][o][PW] (rows [O, Opad) and columns [Cin, PW) stay zero)
__device__ __forceinline__ void stage_conv_w(float* dst, const float* __restrict__ src, int O, int Cin, int KW,
                                             int Opad, int PW) {
  for (int i = threadIdx.x; i < O * Cin * KW; i += NTH) {
    const int tau = i % KW, c = (i / KW) % Cin, o = i / (KW * Cin);
    dst[(tau * Opad + o) * PW + c] = src[i];
  }
}
// conv weight (O, Cin, 3) -> dst[tau][c][O] (8 input rows per tap, rows [Cin, 8) stay zero): the layout
// sk_conv_taps_fwd reads with lane <-> o, bank-conflict free
__device__ __forceinline__ void stage_conv_w_T(float* dst, const float* __restrict__ src, int O, int Cin) {
  for (int i = threadIdx.x; i < O * Cin * 3; i += NTH) {
    const int tau = i % 3, c = (i / 3) % Cin, o = i / (3 * Cin);
    dst[(tau * 8 + c) * O + o] = src[i];
  }
}
__device__ __forceinline__ void stage_vec(float* dst, const float* __restrict__ src, int n) {
  for (int i = threadIdx.x; i < n; i += NTH) dst[i] = src[i];
}

struct Common {
  const float* x;          // (B,C,T)
  const float* u;          // (B,U,T)
  const int64_t* lengths;  // (B)
  int64_t B;
  int T;
  int tiles_per_seq;
  int64_t ntiles;
  const double* norms;     // {max(N_valid*C,1), B}
  float beta;
  const double* beta_dev;  // when non-null the KL weight is read from device memory (graph-replayed training step)
  int use_tma;             // input windows by tensor-map TMA (T % 4 == 0, aligned tensors), else cp.async
  const float* log_pi;     // (K) log_softmax(log_prior)
  const float* wfold;      // (3,H,8) folded decoder.conv1 weights, staged layout
  float* q;                // (B,K,T)  written by A
  float* dq_prior;         // (B,K,T)  written by A
  float* lg;               // (B,K,T)  encoder logits, written by A (tcgen05 build) so that C need not recompute them
  float* dq_dec;           // (B,K,T)  written by B
  double* loss_part;       // [grid][LP]: 0 nll (B); 1 init+trans, 2 entropy, 3 valid steps, 4.. sum_b q[b,k,0] (A)
  double* acc_part;        // [grid][ACC_ALL], zeroed by the host; each CTA adds its smem accumulators every FLUSH tiles
};

__device__ __forceinline__ double beta_of(const Common& cm) { return cm.beta_dev ? *cm.beta_dev : (double)cm.beta; }

// tensor maps of the input tensors of the tile kernels (tma.cuh): boxes of PT columns x all rows x one sequence
struct TmaMaps {
  CUtensorMap x, u, q;
};

struct TileInfo {
  int64_t b;
  int t0, L, Leff;
};

__device__ __forceinline__ TileInfo tile_info(const Common& cm, int64_t tile) {
  TileInfo ti;
  ti.b = tile / cm.tiles_per_seq;
  ti.t0 = (int)(tile - ti.b * cm.tiles_per_seq) * TT;
  const int64_t L = cm.lengths[ti.b];
  ti.L = (int)(L < 0 ? 0 : (L > cm.T ? cm.T : L));
  ti.Leff = ti.L < 1 ? 1 : ti.L;
  return ti;
}

// rows [0,nrows) of a (B,nrows,T) tensor -> dst[r][j]; zero for t outside [0, tmax)
__device__ __forceinline__ void load_rows(float* dst, const float* __restrict__ src, int nrows, int64_t b, int T,
                                          int t0, int tmax) {
  for (int i = threadIdx.x; i < nrows * NC; i += NTH) {
    const int r = i / NC, j = i - r * NC, t = t0 - HL + j;
    dst[r * PT + j] = (t >= 0 && t < tmax) ? src[(b * nrows + r) * (int64_t)T + t] : 0.f;
  }
}

// asynchronous variant (cp.async, 4-byte granules so that any T / alignment works; columns outside
// [0, tmax) are zero-filled): used to fetch the NEXT tile's inputs while the current tile computes
__device__ __forceinline__ void prefetch_rows(float* dst, const float* __restrict__ src, int nrows, int64_t b, int T,
                                              int t0, int tmax) {
  if (((T & 3) == 0) && ((reinterpret_cast<uintptr_t>(src) & 15) == 0)) {
    // 16-byte granules: t0 - HL is a multiple of 4 (TT and HL are), rows start 16-byte aligned when T % 4 == 0;
    // a granule that straddles tmax is zero-filled beyond it through the src-size operand
    for (int i = threadIdx.x; i < nrows * (NC / 4); i += NTH) {
      const int r = i / (NC / 4), j = (i - r * (NC / 4)) * 4, t = t0 - HL + j;
      int nval = (t < 0) ? 0 : (tmax - t);
      nval = nval < 0 ? 0 : (nval > 4 ? 4 : nval);
      const float* g = src + (b * nrows + r) * (int64_t)T + (nval ? t : 0);
      const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst + r * PT + j);
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(d), "l"(g), "r"(4 * nval) : "memory");
    }
    return;
  }
  for (int i = threadIdx.x; i < nrows * NC; i += NTH) {
    const int r = i / NC, j = i - r * NC, t = t0 - HL + j;
    const bool ok = (t >= 0 && t < tmax);
    const float* g = src + (b * nrows + r) * (int64_t)T + (ok ? t : 0);
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst + r * PT + j);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" :: "r"(d), "l"(g), "r"(ok ? 4 : 0) : "memory");
  }
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_prev() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }

// first tile >= `tile` of this CTA's round-robin sequence that is inside the dependency cone of
// its sequence's valid steps (t0 < Leff + SKIP); -1 when there is none
template <int SKIP>
__device__ __forceinline__ int64_t next_tile(const Common& cm, int64_t tile, TileInfo& ti) {
  for (; tile < cm.ntiles; tile += gridDim.x) {
    ti = tile_info(cm, tile);
    if (ti.t0 < ti.Leff + SKIP) return tile;
  }
  return -1;
}

// Software-pipelined tile selection for the persistent tile kernels: the length of the candidate tile
// AFTER the next one is loaded at the top of an iteration (tile_pre) and only looked at once the
// current tile is done (tile_resolve), so the dependent global load -- and the index division -- are
// off the critical path of every tile (they were 3-6 % of the three kernels).
struct TilePre {
  int64_t cand, b, L;
  int t0;
};
__device__ __forceinline__ TilePre tile_pre(const Common& cm, int64_t cand) {
  TilePre p;
  p.cand = cand; p.b = 0; p.L = 0; p.t0 = 0;
  if (cand < cm.ntiles) {
    if (cm.ntiles <= 0x7fffffffLL) {       // 32-bit division whenever the tile count allows it
      const unsigned bb = (unsigned)cand / (unsigned)cm.tiles_per_seq;
      p.b = bb;
      p.t0 = (int)((unsigned)cand - bb * (unsigned)cm.tiles_per_seq) * TT;
    } else {
      p.b = cand / cm.tiles_per_seq;
      p.t0 = (int)(cand - p.b * cm.tiles_per_seq) * TT;
    }
    p.L = cm.lengths[p.b];
  }
  return p;
}
template <int SKIP>
__device__ __forceinline__ int64_t tile_resolve(const Common& cm, const TilePre& p, TileInfo& ti) {
  if (p.cand >= cm.ntiles) return -1;
  ti.b = p.b; ti.t0 = p.t0;
  ti.L = (int)(p.L < 0 ? 0 : (p.L > cm.T ? cm.T : p.L));
  ti.Leff = ti.L < 1 ? 1 : ti.L;
  if (ti.t0 < ti.Leff + SKIP) return p.cand;
  return next_tile<SKIP>(cm, p.cand + gridDim.x, ti);   // ragged batch: this tile is outside the dependency cone
}

__device__ __forceinline__ double block_sum_d(double v, double* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  double r = 0.0;
  if (threadIdx.x == 0)
    for (int i = 0; i < NW; ++i) r += sh[i];
  return r;
}

// part[i] += acc[i]; acc[i] = 0: the fp32 shared-memory accumulators only ever sum FLUSH tiles
// before they are folded into the CTA's double-precision partial in HBM.  The fold is a
// fire-and-forget RED.ADD.F64 (the partial is private to the CTA and every address receives
// exactly one add per flush, so the result is as deterministic as load-add-store, without the
// 64 MB of partials -- more than fits beside the inputs in L2 -- being read back through the SM:
// the load-add-store version cost 36 us per flush, 12 % of the decoder kernel)
#ifndef VHMM_FLUSH
#define VHMM_FLUSH 256   /* tools/flush_check.py: gradients at cfg 2 agree with the FLUSH = 16 build to ~1e-6 */
#endif
constexpr int FLUSH = VHMM_FLUSH;
__device__ __forceinline__ void flush_acc(double* __restrict__ part, float* acc, int n) {
  for (int i = threadIdx.x; i < n; i += NTH) {
    atomicAdd(part + i, (double)acc[i]);
    acc[i] = 0.f;
  }
  __syncthreads();
}

// out[r][c..c+1] = relu(v + bias[r]) inside the sequence, 0 outside (conv zero padding)
struct EpiReluZero {
  float* out; const float* bias; int t0, T;
  __device__ __forceinline__ void operator()(int r, int c, float v0, float v1) const {
    const int t = t0 - HL + c;
    const float b = bias[r];
    float2 o;
    o.x = (t >= 0 && t < T) ? fmaxf(v0 + b, 0.f) : 0.f;
    o.y = (t + 1 >= 0 && t + 1 < T) ? fmaxf(v1 + b, 0.f) : 0.f;
    *reinterpret_cast<float2*>(out + r * PT + c) = o;
  }
};
// out[r][c..] = v + bias[r] for r < nrows
struct EpiBiasRows {
  float* out; const float* bias; int nrows;
  __device__ __forceinline__ void operator()(int r, int c, float v0, float v1) const {
    if (r < nrows) {
      const float b = bias[r];
      *reinterpret_cast<float2*>(out + r * PT + c) = make_float2(v0 + b, v1 + b);
    }
  }
};
// buf[r][c..] = buf > 0 ? v : 0   (ReLU backward, in place over the stored activation)
struct EpiReluMask {
  float* buf;
  __device__ __forceinline__ void operator()(int r, int c, float v0, float v1) const {
    float2* p = reinterpret_cast<float2*>(buf + r * PT + c);
    const float2 a = *p;
    *p = make_float2(a.x > 0.f ? v0 : 0.f, a.y > 0.f ? v1 : 0.f);
  }
};

// EpiReluMask that also tracks the thread's max |value written| (scale of the fp16 parts that follow)
struct EpiReluMaskMax {
  float* buf; float* lm;
  __device__ __forceinline__ void operator()(int r, int c, float v0, float v1) const {
    float2* p = reinterpret_cast<float2*>(buf + r * PT + c);
    const float2 a = *p;
    const float2 o = make_float2(a.x > 0.f ? v0 : 0.f, a.y > 0.f ? v1 : 0.f);
    *p = o;
    *lm = fmaxf(*lm, fmaxf(fabsf(o.x), fabsf(o.y)));
  }
};

// out rows [0,M) x the 128 columns of the tile, split into (16*MT) x (8*NT) warp work units
template <int M, int MT, int NT, int KSTEPS, int TAPS, bool TRANS, typename Epi>
__device__ __forceinline__ void cols_op(const float* W, int tapStride, int PW, const float* In, Epi epi) {
  const int w = threadIdx.x >> 5;
  constexpr int MG = M / (16 * MT), CG = NC / (8 * NT);
  static_assert(M % (16 * MT) == 0 && NC % (8 * NT) == 0, "tile shape");
  for (int unit = w; unit < MG * CG; unit += NW)
    gemm_cols<MT, NT, KSTEPS, TAPS, TRANS, PT>(W, tapStride, PW, In, (unit % MG) * 16 * MT, (unit / MG) * 8 * NT, epi);
}

// encoder forward on the staged window (kernels A and C): XU rows [0,8) hold x.
// h1 valid on columns [1,127), h2 and the logits on [2,126).  Ends with a __syncthreads().
template <typename CF>
__device__ __forceinline__ void encoder_forward(const float* W1s, const float* b1s, const float* W2s,
                                                const float* b2s, const float* W3s, const float* b3s,
                                                const float* XU, float* H1, float* H2b, float* LQ, int t0, int T) {
  cols_op<CF::H, 2, 2, 1, 3, false>(W1s, CF::H * 8, 8, XU, EpiReluZero{H1, b1s, t0, T});
  __syncthreads();
  cols_op<CF::H2, 1, 2, CF::H / 8, 3, false>(W2s, CF::H2 * CF::PW_H, CF::PW_H, H1, EpiReluZero{H2b, b2s, t0, T});
  __syncthreads();
  cols_op<16, 1, 1, CF::H2 / 8, 1, false>(W3s, 0, CF::PW_H2, H2b, EpiBiasRows{LQ, b3s, CF::K});
  __syncthreads();
}

// =============================================================================================
// kernel A: encoder forward + prior forward/backward
// =============================================================================================
template <typename CF>
__global__ void __launch_bounds__(NTH, 1) enc_prior_kernel(Common cm, vhmm_params P, const __grid_constant__ TmaMaps /*unused: cp.async build*/) {
  constexpr int C = CF::C, H = CF::H, H2 = CF::H2, K = CF::K, U = CF::U, TH = CF::TH, K2 = CF::K2;
  extern __shared__ __align__(16) float sm[];
  __shared__ double red[NW];
  float* W1s = sm;                        // [3][H][8]
  float* W2s = W1s + 3 * H * 8;           // [3][H2][PW_H]
  float* W3s = W2s + 3 * H2 * CF::PW_H;   // [16][PW_H2]
  float* Wt0s = W3s + 16 * CF::PW_H2;     // [TH][8]
  float* Wt2s = Wt0s + TH * 8;            // [16][PW_TH]
  float* b1s = Wt2s + 16 * CF::PW_TH;
  float* b2s = b1s + H;
  float* b3s = b2s + H2;                  // [16]
  float* bt0s = b3s + 16;
  float* bt2s = bt0s + TH;                // [16]
  float* lps = bt2s + 16;                 // [8]
  float* XU0 = lps + 8;                   // rows 0..7 x, rows 8..15 u (buffer 0; buffer 1 at the end)
  float* R1 = XU0 + 16 * PT;               // h1 | h2, later the prior's hidden layer a
  float* H1 = R1;
  float* H2b = R1 + H * PT;
  float* Ab = R1;
  float* LQ = R1 + CF::R1 * PT;           // rows 0..K-1 logits -> q, rows 8.. log q
  float* Mb = LQ + 16 * PT;               // transition logits -> log_A, row i*K+j
  float* dMb = Mb + 16 * PT;
  float* acc = dMb + 16 * PT;             // ACC_A
  float* XU1 = acc + CF::ACC_A;
  for (int i = threadIdx.x; i < CF::SM_A; i += NTH) sm[i] = 0.f;
  __syncthreads();
  stage_conv_w(W1s, P.p[P_E1W], H, C, 3, H, 8);
  stage_conv_w(W2s, P.p[P_E2W], H2, H, 3, H2, CF::PW_H);
  stage_conv_w(W3s, P.p[P_E3W], K, H2, 1, 16, CF::PW_H2);
  stage_conv_w(Wt0s, P.p[P_T0W], TH, U, 1, TH, 8);
  stage_conv_w(Wt2s, P.p[P_T2W], K2, TH, 1, 16, CF::PW_TH);
  stage_vec(b1s, P.p[P_E1B], H); stage_vec(b2s, P.p[P_E2B], H2); stage_vec(b3s, P.p[P_E3B], K);
  stage_vec(bt0s, P.p[P_T0B], TH); stage_vec(bt2s, P.p[P_T2B], K2); stage_vec(lps, cm.log_pi, K);
  const float aP = (float)(-beta_of(cm) / cm.norms[1]);
  double s_trans = 0.0, s_ent = 0.0, s_valid = 0.0, s_q0[K];
#pragma unroll
  for (int k = 0; k < K; ++k) s_q0[k] = 0.0;
  const int tid = threadIdx.x, w = tid >> 5;
  PROF_DECL
  __syncthreads();

  double* part = cm.acc_part + (int64_t)blockIdx.x * CF::ACC_ALL;
  int ndone = 0;
  TileInfo ti, tn;
  int buf = 0;
  auto fetch = [&](const TileInfo& t_, float* dst) {
    prefetch_rows(dst, cm.x, C, t_.b, cm.T, t_.t0, cm.T);
    prefetch_rows(dst + 8 * PT, cm.u, U, t_.b, cm.T, t_.t0, cm.T);
  };
  int64_t tile = next_tile<6>(cm, blockIdx.x, ti);
  if (tile >= 0) fetch(ti, XU0);
  cp_async_commit();
  while (tile >= 0) {
    const int64_t nxt = next_tile<6>(cm, tile + gridDim.x, tn);
    if (nxt >= 0) fetch(tn, buf ? XU0 : XU1);
    cp_async_commit();
    cp_async_wait_prev();
    const float* XU = buf ? XU1 : XU0;
    const int t0 = ti.t0, T = cm.T, L = ti.L;
    const int64_t b = ti.b;
    __syncthreads(); PROF(0);
    encoder_forward<CF>(W1s, b1s, W2s, b2s, W3s, b3s, XU, H1, H2b, LQ, t0, T); PROF(1);
    // softmax, entropy, init term, q -> HBM
    if (tid < NC) {
      const int j = tid, t = t0 - HL + j;
      const bool inr = (j >= 2 && j < NC - 2 && t >= 0 && t < T);
      float lg[K], m = -INFINITY, s = 0.f;
#pragma unroll
      for (int k = 0; k < K; ++k) { lg[k] = LQ[k * PT + j]; m = fmaxf(m, lg[k]); }
#pragma unroll
      for (int k = 0; k < K; ++k) s += expf(lg[k] - m);
      const float ls = logf(s), inv = 1.f / s;
      const bool own = inr && j >= J0 && j < J0 + TT;
      float ent = 0.f;
#pragma unroll
      for (int k = 0; k < K; ++k) {
        const float z = lg[k] - m;
        const float qv = inr ? expf(z) * inv : 0.f, lq = inr ? z - ls : 0.f;
        LQ[k * PT + j] = qv;
        LQ[(8 + k) * PT + j] = lq;
        ent -= qv * lq;
        if (own) {
          cm.q[(b * K + k) * (int64_t)T + t] = qv;
          if (t == 0) { s_q0[k] += (double)qv; s_trans += (double)(qv * lps[k]); }   // :123 (not masked)
        }
      }
      if (own && t < L) { s_ent += (double)ent; s_valid += 1.0; }                     // :134-135
    }
    __syncthreads(); PROF(2);   // h1/h2 are dead from here: R1 becomes the prior's hidden layer
    // a = relu(W0 u + b0) on every column (VQ_VAE_HMM_fixed.py:68)
    cols_op<TH, 2, 2, 1, 1, false>(Wt0s, 0, 8, XU + 8 * PT, EpiReluZero{Ab, bt0s, t0, T});
    __syncthreads();
    PROF(3);
    cols_op<16, 1, 1, TH / 8, 1, false>(Wt2s, 0, CF::PW_TH, Ab, EpiBiasRows{Mb, bt2s, K2});
    __syncthreads(); PROF(4);
    if (tid < NC) {    // row log-softmax -> log_A (:69)
      const int j = tid;
#pragma unroll
      for (int i = 0; i < K; ++i) {
        float v[K], m = -INFINITY, s = 0.f;
#pragma unroll
        for (int jj = 0; jj < K; ++jj) { v[jj] = Mb[(i * K + jj) * PT + j]; m = fmaxf(m, v[jj]); }
#pragma unroll
        for (int jj = 0; jj < K; ++jj) s += expf(v[jj] - m);
        const float lse = m + logf(s);
#pragma unroll
        for (int jj = 0; jj < K; ++jj) Mb[(i * K + jj) * PT + j] = v[jj] - lse;
      }
    }
    __syncthreads();
    if (tid < NC) {    // pair term (:125-131), d/dM, d/dq of the prior
      const int j = tid, t = t0 - HL + j;
      const bool own = (j >= J0 && j < J0 + TT && t < T);
      const bool mk = own && t < L;
      const bool tm = mk && t >= 1;
      const bool tmn = own && (t + 1 < T) && (t + 1 < L);
      float qc[K], qp[K], qn[K], dq[K];
#pragma unroll
      for (int k = 0; k < K; ++k) {
        qc[k] = LQ[k * PT + j];
        qp[k] = tm ? LQ[k * PT + j - 1] : 0.f;
        qn[k] = tmn ? LQ[k * PT + j + 1] : 0.f;
        dq[k] = (own && t == 0) ? lps[k] : 0.f;
      }
      float trans = 0.f;
#pragma unroll
      for (int i = 0; i < K; ++i)
#pragma unroll
        for (int jj = 0; jj < K; ++jj) {
          const int r = i * K + jj;
          float dm = 0.f;
          if (tm) {
            const float la = Mb[r * PT + j];
            trans += qp[i] * qc[jj] * la;
            dq[jj] += qp[i] * la;
            dm = aP * qp[i] * (qc[jj] - expf(la));
          }
          if (tmn) dq[i] += qn[jj] * Mb[r * PT + j + 1];
          dMb[r * PT + j] = dm;
        }
      s_trans += (double)trans;
      if (own)
#pragma unroll
        for (int k = 0; k < K; ++k) cm.dq_prior[(b * K + k) * (int64_t)T + t] = aP * dq[k];
    }
    __syncthreads();
    PROF(5);
    // gradients of transition_net.2: dWt2[r][h] += sum_t dM[r][t] a[h][t]
    for (int unit = w; unit < TH / 8; unit += NW)
      gemm_time<1, 1, KST, J0, PT>(dMb, Ab + unit * 8 * PT, acc + CF::A_WT2 + unit * 8, 0, CF::PW_TH);
    row_sums<J0, TT / 4, PT, NW>(dMb, K2, acc + CF::A_BT2);
    __syncthreads(); PROF(6);
    // da = relu'(a) * Wt2^T dM, in place over a
    cols_op<TH, 2, 2, 2, 1, true>(Wt2s, 0, CF::PW_TH, dMb, EpiReluMask{Ab});
    __syncthreads(); PROF(7);
    // gradients of transition_net.0: dWt0[h][c] += sum_t da[h][t] u[c][t]
    for (int unit = w; unit < TH / 16; unit += NW)
      gemm_time<1, 1, KST, J0, PT>(Ab + unit * 16 * PT, XU + 8 * PT, acc + CF::A_WT0 + unit * 16 * 8, 0, 8);
    row_sums<J0, TT / 4, PT, NW>(Ab, TH, acc + CF::A_BT0);
    __syncthreads(); PROF(8);
    if (++ndone % FLUSH == 0) flush_acc(part, acc, CF::ACC_A);
    tile = nxt; ti = tn; buf ^= 1;
  }
  PROF_DUMP("enc_prior", 9);
  // per-CTA partials
  flush_acc(part, acc, CF::ACC_A);
  double* lp = cm.loss_part + (int64_t)blockIdx.x * LP;
  double v;
  v = block_sum_d(s_trans, red); if (tid == 0) lp[1] = v;
  v = block_sum_d(s_ent, red);   if (tid == 0) lp[2] = v;
  v = block_sum_d(s_valid, red); if (tid == 0) lp[3] = v;
#pragma unroll
  for (int k = 0; k < K; ++k) { v = block_sum_d(s_q0[k], red); if (tid == 0) lp[4 + k] = v; }
}

// =============================================================================================
// kernel B: decoder forward + backward
// =============================================================================================
template <typename CF>
__global__ void __launch_bounds__(NTH, 1) decoder_kernel(Common cm, vhmm_params P, const __grid_constant__ TmaMaps /*unused: cp.async build*/) {
  constexpr int C = CF::C, H = CF::H, K = CF::K;
  extern __shared__ __align__(16) float sm[];
  __shared__ double red[NW];
  float* WFs = sm;                        // [3][H][8]   folded conv1
  float* W2s = WFs + 3 * H * 8;           // [3][H][PW_H]
  float* WPs = W2s + 3 * H * CF::PW_H;    // [16][PW_H]
  float* bd1s = WPs + 16 * CF::PW_H;
  float* bd2s = bd1s + H;
  float* bps = bd2s + H;                  // [16]
  float* QX = bps + 16;                   // rows 0..7 q, rows 8..15 x
  float* G1 = QX + 16 * PT;
  float* G2 = G1 + H * PT;
  float* Pb = G2 + H * PT;                // rows 0..C-1 mu, C..2C-1 logvar; then their gradients
  float* acc = Pb + 16 * PT;              // ACC_B
  for (int i = threadIdx.x; i < CF::SM_B; i += NTH) sm[i] = 0.f;
  __syncthreads();
  stage_vec(WFs, cm.wfold, 3 * H * 8);
  stage_conv_w(W2s, P.p[P_D2W], H, H, 3, H, CF::PW_H);
  stage_conv_w(WPs, P.p[P_DPW], 2 * C, H, 1, 16, CF::PW_H);
  stage_vec(bd1s, P.p[P_D1B], H); stage_vec(bd2s, P.p[P_D2B], H); stage_vec(bps, P.p[P_DPB], 2 * C);
  const float aR = (float)(1.0 / cm.norms[0]);
  double s_nll = 0.0;
  PROF_DECL
  const int tid = threadIdx.x, w = tid >> 5;
  __syncthreads();

  double* part = cm.acc_part + (int64_t)blockIdx.x * CF::ACC_ALL + CF::ACC_A;
  int ndone = 0;
  for (int64_t tile = blockIdx.x; tile < cm.ntiles; tile += gridDim.x) {
    const TileInfo ti = tile_info(cm, tile);
    if (ti.t0 >= ti.Leff + 2) continue;
    const int t0 = ti.t0, T = cm.T, L = ti.L;
    const int64_t b = ti.b;
    const int qmax = (ti.Leff + 6 < T) ? ti.Leff + 6 : T;
    load_rows(QX, cm.q, K, b, T, t0, qmax);
    load_rows(QX + 8 * PT, cm.x, C, b, T, t0, T);
    __syncthreads(); PROF(0);
    // g1 = relu(conv3'(q) + b), g2 = relu(conv3(g1) + b), p = Wp g2 + bp   (:83-87)
    cols_op<H, 2, 2, 1, 3, false>(WFs, H * 8, 8, QX, EpiReluZero{G1, bd1s, t0, T});
    __syncthreads(); PROF(1);
    cols_op<H, 2, 2, H / 8, 3, false>(W2s, H * CF::PW_H, CF::PW_H, G1, EpiReluZero{G2, bd2s, t0, T});
    __syncthreads(); PROF(2);
    cols_op<16, 1, 1, H / 8, 1, false>(WPs, 0, CF::PW_H, G2, EpiBiasRows{Pb, bps, 2 * C});
    __syncthreads(); PROF(3);
    if (tid < NC) {    // Gaussian NLL (:118-120) and its gradient w.r.t. (mu, logvar)
      const int j = tid, t = t0 - HL + j;
      const bool inr = (j >= 2 && j < NC - 2 && t >= 0 && t < T);
      const bool mk = inr && t < L;
      const bool own = mk && j >= J0 && j < J0 + TT;
      float nll = 0.f;
#pragma unroll
      for (int c = 0; c < C; ++c) {
        const float mu = Pb[c * PT + j], lv = Pb[(C + c) * PT + j], xv = QX[(8 + c) * PT + j];
        const float ev = expf(lv), v = fmaxf(ev, 1e-8f), d = mu - xv, r = d * d / v;
        nll += 0.5f * (logf(6.283185307179586f * v) + r);
        Pb[c * PT + j] = mk ? aR * d / v : 0.f;
        Pb[(C + c) * PT + j] = (mk && ev >= 1e-8f) ? aR * 0.5f * (1.f - r) : 0.f;
      }
      if (own) s_nll += (double)nll;
    }
    __syncthreads(); PROF(4);
    // to_params gradients: dWp[r][c] += sum_t dp[r][t] g2[c][t]
    for (int unit = w; unit < H / 8; unit += NW)
      gemm_time<1, 1, KST, J0, PT>(Pb, G2 + unit * 8 * PT, acc + CF::B_WP + unit * 8, 0, CF::PW_H);
    row_sums<J0, TT / 4, PT, NW>(Pb, 2 * C, acc + CF::B_BP);
    __syncthreads(); PROF(5);
    // dg2 = relu'(g2) * Wp^T dp, in place
    cols_op<H, 2, 2, 2, 1, true>(WPs, 0, CF::PW_H, Pb, EpiReluMask{G2});
    __syncthreads(); PROF(6);
    // conv2 gradients: dW2[tau][o][c] += sum_t dg2[o][t] g1[c][t+tau-1]
    for (int unit = w; unit < (H / 8) * (H / 32); unit += NW) {
      const int nt = unit % (H / 8), mh = unit / (H / 8);
      gemm_time<2, 3, KST, J0, PT>(G2 + mh * 32 * PT, G1 + nt * 8 * PT, acc + CF::B_W2 + mh * 32 * CF::PW_H + nt * 8,
                                   H * CF::PW_H, CF::PW_H);
    }
    row_sums<J0, TT / 4, PT, NW>(G2, H, acc + CF::B_BD2);
    __syncthreads(); PROF(7);
    // dg1 = relu'(g1) * conv3^T(dg2), in place
    cols_op<H, 2, 2, H / 8, 3, true>(W2s, H * CF::PW_H, CF::PW_H, G2, EpiReluMask{G1});
    __syncthreads(); PROF(8);
    // folded conv1 gradients: dW'[tau][o][k] += sum_t dg1[o][t] q[k][t+tau-1]
    for (int unit = w; unit < 3 * (H / 16); unit += NW) {
      const int mt = unit % (H / 16), tap = unit / (H / 16);
      gemm_time<1, 1, KST, J0, PT>(G1 + mt * 16 * PT, QX + tap - 1, acc + CF::B_WF + tap * H * 8 + mt * 16 * 8, 0, 8);
    }
    row_sums<J0, TT / 4, PT, NW>(G1, H, acc + CF::B_BD1);
    // d loss / d q (decoder part) -> HBM for the owned columns
    {
      float* dq = cm.dq_dec;
      cols_op<16, 1, 1, H / 8, 3, true>(WFs, H * 8, 8, G1, [=](int r, int c, float v0, float v1) {
        if (r < K) {
          const int t = t0 - HL + c;
          if (c >= J0 && c < J0 + TT && t < T) dq[(b * K + r) * (int64_t)T + t] = v0;
          if (c + 1 >= J0 && c + 1 < J0 + TT && t + 1 < T) dq[(b * K + r) * (int64_t)T + t + 1] = v1;
        }
      });
    }
    __syncthreads(); PROF(9);
    if (++ndone % FLUSH == 0) flush_acc(part, acc, CF::ACC_B);
  }
  PROF_DUMP("decoder", 10);
  flush_acc(part, acc, CF::ACC_B);
  const double v = block_sum_d(s_nll, red);
  if (tid == 0) cm.loss_part[(int64_t)blockIdx.x * LP + 0] = v;
}

// kernel B on the 5th-generation tensor core (tcgen05): same interface and results as decoder_kernel
#include "decoder_tc5.cuh"

// =============================================================================================
// kernel C: encoder backward (forward recomputed in the tile)
// =============================================================================================
template <typename CF>
__global__ void __launch_bounds__(NTH, 1) enc_bwd_kernel(Common cm, vhmm_params P, const __grid_constant__ TmaMaps /*unused: cp.async build*/) {
  constexpr int C = CF::C, H = CF::H, H2 = CF::H2, K = CF::K;
  extern __shared__ __align__(16) float sm[];
  float* W1s = sm;
  float* W2s = W1s + 3 * H * 8;
  float* W3s = W2s + 3 * H2 * CF::PW_H;
  float* b1s = W3s + 16 * CF::PW_H2;
  float* b2s = b1s + H;
  float* b3s = b2s + H2;                  // [16]
  float* XU0 = b3s + 16;                  // rows 0..7 x (buffer 0)
  float* H1 = XU0 + 16 * PT;
  float* H2b = H1 + H * PT;
  float* LQ = H2b + H2 * PT;              // logits, then d loss / d logits (rows >= K stay zero)
  float* DQ0 = LQ + 16 * PT;              // d loss / d q: rows 0.. from the decoder, rows 8.. from the prior (buffer 0)
  float* acc = DQ0 + 16 * PT;             // ACC_C
  float* XU1 = acc + CF::ACC_C;
  float* DQ1 = XU1 + 16 * PT;
  for (int i = threadIdx.x; i < CF::SM_C; i += NTH) sm[i] = 0.f;
  __syncthreads();
  stage_conv_w(W1s, P.p[P_E1W], H, C, 3, H, 8);
  stage_conv_w(W2s, P.p[P_E2W], H2, H, 3, H2, CF::PW_H);
  stage_conv_w(W3s, P.p[P_E3W], K, H2, 1, 16, CF::PW_H2);
  stage_vec(b1s, P.p[P_E1B], H); stage_vec(b2s, P.p[P_E2B], H2); stage_vec(b3s, P.p[P_E3B], K);
  const float aE = (float)(beta_of(cm) / cm.norms[1]);
  const int tid = threadIdx.x, w = tid >> 5;
  PROF_DECL
  __syncthreads();

  double* part = cm.acc_part + (int64_t)blockIdx.x * CF::ACC_ALL + CF::ACC_A + CF::ACC_B;
  int ndone = 0;
  TileInfo ti, tn;
  int buf = 0;
  auto fetch = [&](const TileInfo& t_, float* xu, float* dq) {
    const int dmax = (t_.Leff + 2 < cm.T) ? t_.Leff + 2 : cm.T;
    prefetch_rows(xu, cm.x, C, t_.b, cm.T, t_.t0, cm.T);
    prefetch_rows(dq, cm.dq_dec, K, t_.b, cm.T, t_.t0, dmax);
    prefetch_rows(dq + 8 * PT, cm.dq_prior, K, t_.b, cm.T, t_.t0, dmax);
  };
  int64_t tile = next_tile<3>(cm, blockIdx.x, ti);
  if (tile >= 0) fetch(ti, XU0, DQ0);
  cp_async_commit();
  while (tile >= 0) {
    const int64_t nxt = next_tile<3>(cm, tile + gridDim.x, tn);
    if (nxt >= 0) fetch(tn, buf ? XU0 : XU1, buf ? DQ0 : DQ1);
    cp_async_commit();
    cp_async_wait_prev();
    const float* XU = buf ? XU1 : XU0;
    const float* DQ = buf ? DQ1 : DQ0;
    const int t0 = ti.t0, T = cm.T, L = ti.L;
    __syncthreads(); PROF(0);
    encoder_forward<CF>(W1s, b1s, W2s, b2s, W3s, b3s, XU, H1, H2b, LQ, t0, T); PROF(1);
    if (tid < NC) {    // d loss / d logits: softmax backward of dq plus the entropy term (:134-137)
      const int j = tid, t = t0 - HL + j;
      const bool inr = (j >= 2 && j < NC - 2 && t >= 0 && t < T);
      const bool mk = inr && t < L;
      float lg[K], m = -INFINITY, s = 0.f;
#pragma unroll
      for (int k = 0; k < K; ++k) { lg[k] = LQ[k * PT + j]; m = fmaxf(m, lg[k]); }
#pragma unroll
      for (int k = 0; k < K; ++k) s += expf(lg[k] - m);
      const float ls = logf(s), inv = 1.f / s;
      float qv[K], lq[K], dq[K], sd = 0.f, Hh = 0.f;
#pragma unroll
      for (int k = 0; k < K; ++k) {
        const float z = lg[k] - m;
        qv[k] = expf(z) * inv; lq[k] = z - ls; dq[k] = DQ[k * PT + j] + DQ[(8 + k) * PT + j];
        sd += qv[k] * dq[k];
        Hh -= qv[k] * lq[k];
      }
#pragma unroll
      for (int k = 0; k < K; ++k) {
        float gk = qv[k] * (dq[k] - sd);
        if (mk) gk += aE * qv[k] * (lq[k] + Hh);
        LQ[k * PT + j] = inr ? gk : 0.f;
      }
    }
    __syncthreads(); PROF(2);
    // to_logits gradients: dW3[k][c] += sum_t dl[k][t] h2[c][t]
    for (int unit = w; unit < H2 / 8; unit += NW)
      gemm_time<1, 1, KST, J0, PT>(LQ, H2b + unit * 8 * PT, acc + CF::C_W3 + unit * 8, 0, CF::PW_H2);
    row_sums<J0, TT / 4, PT, NW>(LQ, K, acc + CF::C_B3);
    __syncthreads(); PROF(3);
    // dh2 = relu'(h2) * W3^T dl, in place
    cols_op<H2, 1, 2, 1, 1, true>(W3s, 0, CF::PW_H2, LQ, EpiReluMask{H2b});
    __syncthreads(); PROF(4);
    // conv2 gradients: dW2[tau][o][c] += sum_t dh2[o][t] h1[c][t+tau-1]
    for (int unit = w; unit < (H / 8) * (H2 / 16); unit += NW) {
      const int nt = unit % (H / 8), mh = unit / (H / 8);
      gemm_time<1, 3, KST, J0, PT>(H2b + mh * 16 * PT, H1 + nt * 8 * PT, acc + CF::C_W2 + mh * 16 * CF::PW_H + nt * 8,
                                   H2 * CF::PW_H, CF::PW_H);
    }
    row_sums<J0, TT / 4, PT, NW>(H2b, H2, acc + CF::C_B2);
    __syncthreads(); PROF(5);
    // dh1 = relu'(h1) * conv3^T(dh2), in place
    cols_op<H, 2, 2, H2 / 8, 3, true>(W2s, H2 * CF::PW_H, CF::PW_H, H2b, EpiReluMask{H1});
    __syncthreads(); PROF(6);
    // conv1 gradients: dW1[tau][o][c] += sum_t dh1[o][t] x[c][t+tau-1]
    for (int unit = w; unit < 3 * (H / 16); unit += NW) {
      const int mt = unit % (H / 16), tap = unit / (H / 16);
      gemm_time<1, 1, KST, J0, PT>(H1 + mt * 16 * PT, XU + tap - 1, acc + CF::C_W1 + tap * H * 8 + mt * 16 * 8, 0, 8);
    }
    row_sums<J0, TT / 4, PT, NW>(H1, H, acc + CF::C_B1);
    __syncthreads(); PROF(7);
    if (++ndone % FLUSH == 0) flush_acc(part, acc, CF::ACC_C);
    tile = nxt; ti = tn; buf ^= 1;
  }
  PROF_DUMP("enc_bwd", 8);
  flush_acc(part, acc, CF::ACC_C);
}

// kernels A and C with the encoder's 64->32 convolution on tcgen05: same interfaces and results
#include "encoder_tc5.cuh"

// =============================================================================================
// small kernels around the three tile kernels
// =============================================================================================
// log_pi = log_softmax(log_prior), K <= 8 (VQ_VAE_HMM_fixed.py:71)
__global__ void fused_log_pi_kernel(const float* __restrict__ log_prior, float* __restrict__ log_pi, int K) {
  if (threadIdx.x != 0) return;
  float m = -INFINITY, s = 0.f;
  for (int k = 0; k < K; ++k) m = fmaxf(m, log_prior[k]);
  for (int k = 0; k < K; ++k) s += expf(log_prior[k] - m);
  const float lse = m + logf(s);
  for (int k = 0; k < K; ++k) log_pi[k] = log_prior[k] - lse;
}

// wfold[tau][o][k] = sum_d Wd1[o][d][tau] * E[k][d]  (columns k >= K zero)
template <typename CF>
__global__ void fold_kernel(const float* __restrict__ Wd1, const float* __restrict__ E, float* __restrict__ wfold) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 3 * CF::H * 8) return;
  const int k = i % 8, o = (i / 8) % CF::H, tau = i / (8 * CF::H);
  float s = 0.f;
  if (k < CF::K)
    for (int d = 0; d < CF::H; ++d) s = fmaf(Wd1[(o * CF::H + d) * 3 + tau], E[k * CF::H + d], s);
  wfold[i] = s;
}

// red[i] = sum over CTAs of acc_part[cta][i] (double, fixed order)
// Block = 32 consecutive elements x 8 warps; warp w sums the CTAs c = w, w + 8, ... (coalesced 256-byte rows), the eight
// partial sums meet in shared memory in a fixed order.  (A single thread per element walked 148 dependent-latency loads:
// 11-17 us of the 150 us cfg-1 training step.)
constexpr int RA_SPLIT = 8;
__global__ void __launch_bounds__(32 * RA_SPLIT) reduce_acc_kernel(const double* __restrict__ part, int ncta, int n,
                                                                   float* __restrict__ red) {
  __shared__ double sh[RA_SPLIT][32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int i = blockIdx.x * 32 + lane;
  double s = 0.0;
  if (i < n)
    for (int c = w; c < ncta; c += RA_SPLIT) s += part[(int64_t)c * n + i];
  sh[w][lane] = s;
  __syncthreads();
  if (w == 0 && i < n) {
    double t = sh[0][lane];
#pragma unroll
    for (int k = 1; k < RA_SPLIT; ++k) t += sh[k][lane];
    red[i] = (float)t;
  }
}

// loss sums, log_prior gradient and the scatter of the reduced accumulators into the flat
// gradient buffer (parameter order of include/vhmm.h)
template <typename CF>
__global__ void finish_kernel(const float* __restrict__ red, const double* __restrict__ loss_part, int ncta,
                              const float* __restrict__ log_pi, const double* __restrict__ norms, float beta,
                              const double* __restrict__ beta_dev, vhmm_params P, double* __restrict__ sums,
                              float* __restrict__ grads) {
  constexpr int C = CF::C, H = CF::H, H2 = CF::H2, K = CF::K, U = CF::U, TH = CF::TH, K2 = CF::K2;
  const int tid = blockIdx.x * blockDim.x + threadIdx.x;
  if (tid < 32) {
    // loss partials of the CTAs: lane l sums CTAs l, l + 32, ... then a fixed-order tree over the lanes (deterministic).
    // (One thread walking all CTAs through dependent loads took 40-60 us -- a quarter of the cfg-1 training step.)
    double s[4] = {0, 0, 0, 0}, q0[K];
    for (int k = 0; k < K; ++k) q0[k] = 0.0;
    for (int c = tid; c < ncta; c += 32) {
      for (int i = 0; i < 4; ++i) s[i] += loss_part[(int64_t)c * LP + i];
      for (int k = 0; k < K; ++k) q0[k] += loss_part[(int64_t)c * LP + 4 + k];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      for (int i = 0; i < 4; ++i) s[i] += __shfl_down_sync(0xffffffffu, s[i], o);
      for (int k = 0; k < K; ++k) q0[k] += __shfl_down_sync(0xffffffffu, q0[k], o);
    }
  if (tid == 0) {
    for (int i = 0; i < 4; ++i) sums[i] = s[i];
    if (grads) {   // log_prior through log_softmax (:71): g = aP*q0 - softmax * sum(aP*q0)
      const double aP = -(beta_dev ? *beta_dev : (double)beta) / norms[1];
      double gs = 0.0;
      for (int k = 0; k < K; ++k) gs += aP * q0[k];
      constexpr int off_lp = H * C * 3 + H + H2 * H * 3 + H2 + K * H2 + K;   // offset of prior.log_prior
      for (int k = 0; k < K; ++k) grads[off_lp + k] = (float)(aP * q0[k] - exp((double)log_pi[k]) * gs);
    }
  }
  }
  if (!grads) return;
  const float* rA = red;
  const float* rB = red + CF::ACC_A;
  const float* rC = rB + CF::ACC_B;
  // sizes in parameter order
  const int n[VHMM_NUM_PARAMS] = {H * C * 3, H, H2 * H * 3, H2, K * H2, K, K, TH * U, TH, K2 * TH, K2,
                                  K * H, H * H * 3, H, H * H * 3, H, 2 * C * H, 2 * C};
  int e = tid, p = 0;
  int64_t base = 0;
  while (p < VHMM_NUM_PARAMS && e >= n[p]) { e -= n[p]; base += n[p]; ++p; }
  if (p >= VHMM_NUM_PARAMS) return;
  float g = 0.f;
  switch (p) {
    case P_E1W: { const int tau = e % 3, c = (e / 3) % C, o = e / (3 * C); g = rC[CF::C_W1 + (tau * H + o) * 8 + c]; break; }
    case P_E1B: g = rC[CF::C_B1 + e]; break;
    case P_E2W: { const int tau = e % 3, c = (e / 3) % H, o = e / (3 * H); g = rC[CF::C_W2 + (tau * H2 + o) * CF::PW_H + c]; break; }
    case P_E2B: g = rC[CF::C_B2 + e]; break;
    case P_E3W: g = rC[CF::C_W3 + (e / H2) * CF::PW_H2 + (e % H2)]; break;
    case P_E3B: g = rC[CF::C_B3 + e]; break;
    case P_LOGPRIOR: return;   // written by thread 0
    case P_T0W: g = rA[CF::A_WT0 + (e / U) * 8 + (e % U)]; break;
    case P_T0B: g = rA[CF::A_BT0 + e]; break;
    case P_T2W: g = rA[CF::A_WT2 + (e / TH) * CF::PW_TH + (e % TH)]; break;
    case P_T2B: g = rA[CF::A_BT2 + e]; break;
    case P_EMB: {   // dE[k][d] = sum_{o,tau} dW'[tau][o][k] * Wd1[o][d][tau]
      const int d = e % H, k = e / H;
      double s = 0.0;
      for (int o = 0; o < H; ++o)
        for (int tau = 0; tau < 3; ++tau)
          s += (double)rB[CF::B_WF + (tau * H + o) * 8 + k] * (double)P.p[P_D1W][(o * H + d) * 3 + tau];
      g = (float)s;
      break;
    }
    case P_D1W: {   // dWd1[o][d][tau] = sum_k dW'[tau][o][k] * E[k][d]
      const int tau = e % 3, d = (e / 3) % H, o = e / (3 * H);
      double s = 0.0;
      for (int k = 0; k < K; ++k) s += (double)rB[CF::B_WF + (tau * H + o) * 8 + k] * (double)P.p[P_EMB][k * H + d];
      g = (float)s;
      break;
    }
    case P_D1B: g = rB[CF::B_BD1 + e]; break;
    case P_D2W: { const int tau = e % 3, c = (e / 3) % H, o = e / (3 * H); g = rB[CF::B_W2 + (tau * H + o) * CF::PW_H + c]; break; }
    case P_D2B: g = rB[CF::B_BD2 + e]; break;
    case P_DPW: g = rB[CF::B_WP + (e / H) * CF::PW_H + (e % H)]; break;
    case P_DPB: g = rB[CF::B_BP + e]; break;
  }
  grads[base + e] = g;
}

// ---- host side ---------------------------------------------------------------------------------
// optional per-kernel timing (vhmm_elbo_timing): CUDA events recorded on the caller's stream
// around the three tile kernels of up to TIMING_SLOTS consecutive calls
constexpr int TIMING_SLOTS = 64;
struct TimingState {
  bool enabled = false;
  int ncalls = 0;
  cudaEvent_t ev[TIMING_SLOTS][4] = {};
  bool created = false;
};
static TimingState g_timing_;
#define g_timing ::vhmm::fz::g_timing_

// input windows of the tile kernels: 0 = cp.async (default: measured 0.8 % faster at cfg 2), 1 = tensor-map TMA boxes
// (vhmm_tma_select / environment VHMM_TMA=1); both paths are covered by the tests
static std::atomic<int> g_tma_{-1};
inline bool tma_enabled() {
  int v = g_tma_.load(std::memory_order_relaxed);
  if (v < 0) {
    const char* e = getenv("VHMM_TMA");
    v = (e && e[0] == '1') ? 1 : 0;
    g_tma_.store(v, std::memory_order_relaxed);
  }
  return v != 0;
}

// which decoder kernel the fused path launches: 0 = auto (tcgen05), 1 = mma.sync 3xTF32, 2 = tcgen05 2xFP16
static std::atomic<int> g_decoder_variant_{0};
inline int decoder_variant() {
  int v = g_decoder_variant_;
  if (v == 0) {
    const char* e = getenv("VHMM_FUSED_DECODER");
    if (e && !strcmp(e, "mma")) v = 1;
    else if (e && !strcmp(e, "tc5")) v = 2;
    else v = 2;
  }
  return v;
}

template <typename CF>
struct Launcher {
  static size_t workspace_bytes(int64_t B, int64_t T, int ncta) {
    size_t s = 0;
    s += 4 * align_up((size_t)B * CF::K * T * 4, 256);             // q, dq_prior, dq_dec, logits
    s += align_up((size_t)ncta * CF::ACC_ALL * 8, 256);            // per-CTA gradient partials (double)
    s += align_up((size_t)CF::ACC_ALL * 4, 256);                   // reduced
    s += align_up((size_t)ncta * LP * 8, 256);                     // loss partials
    s += align_up((size_t)3 * CF::H * 8 * 4, 256) + 256;           // folded conv1, log_pi
    return s + 1024;
  }

  static int run(const vhmm_params& P, const float* x, const float* u, const int64_t* lengths, int64_t B, int64_t T,
                 const double* norms, double beta, double* sums, float* grads, void* ws, size_t ws_bytes, int ncta,
                 cudaStream_t st, const double* beta_dev = nullptr) {
    Arena A(ws, ws_bytes);
    Common cm{};
    cm.q = A.take<float>((size_t)B * CF::K * T);
    cm.dq_prior = A.take<float>((size_t)B * CF::K * T);
    cm.dq_dec = A.take<float>((size_t)B * CF::K * T);
    cm.lg = A.take<float>((size_t)B * CF::K * T);
    cm.acc_part = A.take<double>((size_t)ncta * CF::ACC_ALL);
    float* red = A.take<float>(CF::ACC_ALL);
    cm.loss_part = A.take<double>((size_t)ncta * LP);
    float* wfold = A.take<float>(3 * CF::H * 8);
    float* log_pi = A.take<float>(8);
    VHMM_REQUIRE(A.ok, VHMM_E_WORKSPACE, "fused elbo: workspace %zu < %zu", ws_bytes, A.used);
    cm.x = x; cm.u = u; cm.lengths = lengths; cm.B = B; cm.T = (int)T;
    cm.tiles_per_seq = (int)cdiv(T, TT);
    cm.ntiles = B * cm.tiles_per_seq;
    cm.norms = norms; cm.beta = (float)beta; cm.beta_dev = beta_dev; cm.log_pi = log_pi; cm.wfold = wfold;
    if (cm.ntiles < ncta) ncta = (int)cm.ntiles;

    TmaMaps tm;
    memset(&tm, 0, sizeof(tm));
    cm.use_tma = (tma_enabled() && ::vhmm::tma::make_map(&tm.x, x, B, CF::C, T, PT, CF::C) &&
                  ::vhmm::tma::make_map(&tm.u, u, B, CF::U, T, PT, CF::U) &&
                  ::vhmm::tma::make_map(&tm.q, cm.q, B, CF::K, T, PT, CF::K)) ? 1 : 0;
    const bool tc5e = decoder_variant() == 2;
    auto kA = tc5e ? enc_prior_tc5_kernel<CF> : enc_prior_kernel<CF>;
    const size_t smA = tc5e ? t5e::Sz<CF>::A : (size_t)CF::SM_A * 4;
    const bool tc5v = decoder_variant() == 2;
    auto kB = tc5v ? decoder_tc5_kernel<CF> : decoder_kernel<CF>;
    const size_t smB = tc5v ? t5::Lay<CF>::BYTES : (size_t)CF::SM_B * 4;
    auto kC = tc5e ? enc_bwd_tc5_kernel<CF> : enc_bwd_kernel<CF>;
    const size_t smC = tc5e ? t5e::Sz<CF>::C : (size_t)CF::SM_C * 4;
    static PerDeviceOnce attr_done;   // per device (the attribute belongs to the context); idempotent under a race
    if (attr_done.pending()) {
      VHMM_CUDA_CHECK(cudaFuncSetAttribute(enc_prior_kernel<CF>, cudaFuncAttributeMaxDynamicSharedMemorySize, CF::SM_A * 4));
      VHMM_CUDA_CHECK(cudaFuncSetAttribute(enc_prior_tc5_kernel<CF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)t5e::Sz<CF>::A));
      VHMM_CUDA_CHECK(cudaFuncSetAttribute(enc_bwd_tc5_kernel<CF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)t5e::Sz<CF>::C));
      VHMM_CUDA_CHECK(cudaFuncSetAttribute(decoder_kernel<CF>, cudaFuncAttributeMaxDynamicSharedMemorySize, CF::SM_B * 4));
      VHMM_CUDA_CHECK(cudaFuncSetAttribute(decoder_tc5_kernel<CF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)t5::Lay<CF>::BYTES));
      VHMM_CUDA_CHECK(cudaFuncSetAttribute(enc_bwd_kernel<CF>, cudaFuncAttributeMaxDynamicSharedMemorySize, CF::SM_C * 4));
      attr_done.mark();
    }
    VHMM_CUDA_CHECK(cudaMemsetAsync(cm.loss_part, 0, (size_t)ncta * LP * 8, st));
    if (grads) VHMM_CUDA_CHECK(cudaMemsetAsync(cm.acc_part, 0, (size_t)ncta * CF::ACC_ALL * 8, st));
    fused_log_pi_kernel<<<1, 32, 0, st>>>(P.p[P_LOGPRIOR], log_pi, CF::K);
    VHMM_LAUNCH_CHECK();
    fold_kernel<CF><<<(unsigned)cdiv(3 * CF::H * 8, 128), 128, 0, st>>>(P.p[P_D1W], P.p[P_EMB], wfold);
    VHMM_LAUNCH_CHECK();
    cudaEvent_t* tev = nullptr;
    if (g_timing.enabled && grads && g_timing.ncalls < TIMING_SLOTS) tev = g_timing.ev[g_timing.ncalls++];
    if (tev) VHMM_CUDA_CHECK(cudaEventRecord(tev[0], st));
    kA<<<ncta, NTH, smA, st>>>(cm, P, tm);
    VHMM_LAUNCH_CHECK();
    if (tev) VHMM_CUDA_CHECK(cudaEventRecord(tev[1], st));
    if (grads) {
      kB<<<ncta, NTH, smB, st>>>(cm, P, tm);
      VHMM_LAUNCH_CHECK();
      if (tev) VHMM_CUDA_CHECK(cudaEventRecord(tev[2], st));
      kC<<<ncta, NTH, smC, st>>>(cm, P, tm);
      VHMM_LAUNCH_CHECK();
      if (tev) VHMM_CUDA_CHECK(cudaEventRecord(tev[3], st));
      reduce_acc_kernel<<<(unsigned)cdiv(CF::ACC_ALL, 32), 32 * RA_SPLIT, 0, st>>>(cm.acc_part, ncta, CF::ACC_ALL, red);
      VHMM_LAUNCH_CHECK();
    } else {
      // loss only: the decoder kernel still evaluates the NLL (its backward half is cheap to keep)
      kB<<<ncta, NTH, smB, st>>>(cm, P, tm);
      VHMM_LAUNCH_CHECK();
    }
    vhmm_dims d{CF::C, CF::H, CF::K, CF::H2, CF::U, CF::TH};
    int64_t off[VHMM_NUM_PARAMS + 1];
    param_offsets(d, off);
    finish_kernel<CF><<<(unsigned)cdiv(off[VHMM_NUM_PARAMS], 256), 256, 0, st>>>(red, cm.loss_part, ncta, log_pi, norms,
                                                                                (float)beta, beta_dev, P, sums, grads);
    VHMM_LAUNCH_CHECK();
    return 0;
  }
};

using CfgBase = Cfg<5, 64, 32, 3, 4, 128>;   // BASELINE.json configs[0..1]

#include "forward_tc5.cuh"

// host side of the fused forward-only kernel: x != null -> encoder half, mu != null -> decoder half
template <typename CF>
struct FwdLauncher {
  static size_t workspace_bytes() { return align_up((size_t)3 * CF::H * 8 * 4, 256) + 256; }
  template <bool ENC, bool DEC>
  static int launch(const FwdArgs& fa, const vhmm_params& P, int ncta, cudaStream_t st) {
    auto k = forward_tc5_kernel<CF, ENC, DEC>;
    constexpr size_t sm = t5f::Lay<CF, ENC, DEC>::BYTES;
    static_assert(sm <= 227 * 1024, "forward tile kernel exceeds the shared memory of an SM");
    static PerDeviceOnce attr_done;
    if (attr_done.pending()) {
      VHMM_CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
      attr_done.mark();
    }
    // input windows by tensor-map TMA where the tensor qualifies (box = PT columns x all rows x 1 sequence)
    CUtensorMap tm;
    memset(&tm, 0, sizeof(tm));
    FwdArgs f2 = fa;
    f2.use_tma = (tma_enabled() && ::vhmm::tma::make_map(&tm, ENC ? fa.x : fa.q_in, fa.B, ENC ? CF::C : CF::K, fa.T, PT, ENC ? CF::C : CF::K)) ? 1 : 0;
    k<<<ncta, NTH, sm, st>>>(f2, P, tm);
    VHMM_LAUNCH_CHECK();
    return 0;
  }
  static int run(const vhmm_params& P, const float* x, const float* q_in, int64_t B, int64_t T, float* logits, float* q,
                 float* mu, float* logvar, void* ws, size_t ws_bytes, int ncta, cudaStream_t st) {
    Arena A(ws, ws_bytes);
    float* wfold = A.take<float>(3 * CF::H * 8);
    VHMM_REQUIRE(A.ok, VHMM_E_WORKSPACE, "fused forward: workspace %zu < %zu", ws_bytes, A.used);
    FwdArgs fa{};
    fa.x = x; fa.q_in = q_in; fa.B = B; fa.T = (int)T;
    fa.tiles_per_seq = (int)cdiv(T, TT);
    fa.ntiles = B * fa.tiles_per_seq;
    fa.wfold = wfold; fa.logits = logits; fa.q = q; fa.mu = mu; fa.logvar = logvar;
    if (fa.ntiles < ncta) ncta = (int)fa.ntiles;
    const bool enc = x != nullptr, dec = mu != nullptr;
    if (dec) {
      fold_kernel<CF><<<(unsigned)cdiv(3 * CF::H * 8, 128), 128, 0, st>>>(P.p[P_D1W], P.p[P_EMB], wfold);
      VHMM_LAUNCH_CHECK();
    }
    if (enc && dec) return launch<true, true>(fa, P, ncta, st);
    if (enc) return launch<true, false>(fa, P, ncta, st);
    return launch<false, true>(fa, P, ncta, st);
  }
};

}  // namespace fz

static int fused_sm_count() { return device_sm_count(); }   // of the CURRENT device

int fused_decoder_select(int variant) { fz::g_decoder_variant_ = variant; return 0; }
int fused_decoder_active() { return fz::decoder_variant(); }
int fused_tma_select(int on) { fz::g_tma_.store(on ? 1 : 0); return 0; }

int fused_timing_enable(int enable) {
  if (enable && !g_timing.created) {
    for (int i = 0; i < fz::TIMING_SLOTS; ++i)
      for (int j = 0; j < 4; ++j) VHMM_CUDA_CHECK(cudaEventCreate(&g_timing.ev[i][j]));
    g_timing.created = true;
  }
  g_timing.enabled = enable != 0;
  g_timing.ncalls = 0;
  return 0;
}

// ms_sum[0..2] = total milliseconds of kernels A, B, C over the recorded calls (waits for them)
int fused_timing_read(double* ms_sum, int* ncalls) {
  ms_sum[0] = ms_sum[1] = ms_sum[2] = 0.0;
  *ncalls = g_timing.ncalls;
  for (int i = 0; i < g_timing.ncalls; ++i) {
    VHMM_CUDA_CHECK(cudaEventSynchronize(g_timing.ev[i][3]));
    for (int j = 0; j < 3; ++j) {
      float ms = 0.f;
      VHMM_CUDA_CHECK(cudaEventElapsedTime(&ms, g_timing.ev[i][j], g_timing.ev[i][j + 1]));
      ms_sum[j] += ms;
    }
  }
  return 0;
}

bool fused_supported(const vhmm_dims& d) {
  return d.C == 5 && d.H == 64 && d.H2 == 32 && d.K == 3 && d.U == 4 && d.TH == 128;
}

size_t fused_workspace_bytes(const vhmm_dims& d, int64_t B, int64_t T) {
  if (!fused_supported(d)) return 0;
  return fz::Launcher<fz::CfgBase>::workspace_bytes(B, T, fused_sm_count());
}

size_t fused_forward_workspace_bytes(const vhmm_dims& d) {
  return fused_supported(d) ? fz::FwdLauncher<fz::CfgBase>::workspace_bytes() : 0;
}

// x (B,C,T) -> logits / q (either may be null) [-> mu, logvar]; or q_in (B,K,T) -> mu, logvar when x is null
int fused_forward(const vhmm_dims& d, const vhmm_params& P, const float* x, const float* q_in, int64_t B, int64_t T,
                  float* logits, float* q, float* mu, float* logvar, void* ws, size_t ws_bytes, cudaStream_t st) {
  VHMM_REQUIRE(fused_supported(d), VHMM_E_DIMS, "fused forward: unsupported dims");
  VHMM_REQUIRE((x != nullptr) != (q_in != nullptr), VHMM_E_ARG, "fused forward: exactly one of x, q_in");
  VHMM_REQUIRE((mu != nullptr) == (logvar != nullptr), VHMM_E_ARG, "fused forward: mu and logvar come together");
  return fz::FwdLauncher<fz::CfgBase>::run(P, x, q_in, B, T, logits, q, mu, logvar, ws, ws_bytes, fused_sm_count(), st);
}

int fused_elbo(const vhmm_dims& d, const vhmm_params& P, const float* x, const float* u, const int64_t* lengths,
               int64_t B, int64_t T, const double* norms, double beta, double* sums, float* grads, void* ws,
               size_t ws_bytes, cudaStream_t st, const double* beta_dev) {
  VHMM_REQUIRE(fused_supported(d), VHMM_E_DIMS, "fused elbo: unsupported dims");
  return fz::Launcher<fz::CfgBase>::run(P, x, u, lengths, B, T, norms, beta, sums, grads, ws, ws_bytes,
                                        fused_sm_count(), st, beta_dev);
}

}  // namespace vhmm
