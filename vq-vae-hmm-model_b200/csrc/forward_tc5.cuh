// forward_tc5.cuh -- the fused FORWARD-ONLY tile kernel of the pinned path: what
// VAE_HMM.encode / decode / forward (VQ_VAE_HMM_fixed.py:100-104, :139-143) and every regime-decoding call site of
// the reference (backtesting.py:43-44,154-155; inference_api/app.py:63-65; ...) run.  Included by elbo_fused.cu inside
// namespace vhmm::fz; it is assembled from the building blocks of the ELBO tile kernels:
//
//   ENC   x -> h1 = relu(conv3) [FFMA, skinny.cuh] -> h2 = relu(conv3) [tcgen05, 2 x fp16 split, TMEM accumulators]
//           -> logits = conv1x1 [FFMA] -> q = softmax_k                      (encoder_forward_tc5 of kernel A)
//   DEC   q -> g1 = relu(conv3 of the embedding-folded 3 -> 64 layer) [FFMA, straight to fp16 operand parts]
//           -> g2 = relu(conv3) [tcgen05] -> (mu, logvar) = conv1x1 [mma.sync 3xTF32]  (forward half of kernel B)
//
// One persistent CTA per SM walks (sequence, 120-step tile) items; a tile stages the 128-column window
// t in [t0-4, t0+124) of its input with cp.async (next tile prefetched under the current one) and evaluates the whole
// chain on chip: NO layer activation is written to HBM (the layer-wise path this replaces moved h1, h2, e, g1, g2:
// 4 x 64 + 32 channels x 4 B x 2 = 2.3 KB per step).  Algorithmic HBM bytes per step: forward() 20 (x) in, 12 (q) +
// 40 (mu, logvar) out; encode() 20 in, 12 out; decode() 12 in, 40 out.  68,580 FLOP per step for forward().
// Padding semantics as everywhere (SURVEY.md finding 4): every layer sees zeros outside [0,T) only.
#pragma once
#include "encoder_tc5.cuh"
#include "decoder_tc5.cuh"

struct FwdArgs {
  const float* x;       // (B,C,T)  ENC input
  const float* q_in;    // (B,K,T)  DEC-only input
  int64_t B;
  int T;
  int tiles_per_seq;
  int64_t ntiles;
  const float* wfold;   // (3,H,8) folded decoder.conv1 (fold_kernel), DEC
  float* logits;        // (B,K,T) or null
  float* q;             // (B,K,T) or null
  float* mu;            // (B,C,T) DEC
  float* logvar;        // (B,C,T) DEC
  int use_tma;          // the input windows arrive by tensor-map TMA (T % 4 == 0, aligned base); else cp.async
};

namespace t5f {
template <typename CF, bool ENC, bool DEC>
struct Lay {
  static constexpr int OPB = (ENC ? t5e::W_BYTES : 0) + (DEC ? t5::W_BYTES : 0) + t5::ACT_BYTES;
  static constexpr int W1 = 0;                                   // [3][8][H] transposed encoder conv1
  static constexpr int W3 = W1 + 3 * CF::H * 8;                  // [16][PW_H2]
  static constexpr int WFT = W3 + 16 * CF::PW_H2;                // [3][K][H] folded decoder conv1, transposed
  static constexpr int WFS = WFT + 3 * CF::K * CF::H;            // [3][H][8] folded conv1 (bound of g1)
  static constexpr int WP = WFS + 3 * CF::H * 8;                 // [16][PW_H]
  static constexpr int BIAS = WP + 16 * CF::PW_H;                // b1 H | b2 H2 | b3 16 | bd1 H | bd2 H | bp 16
  static constexpr int NBIAS = CF::H + CF::H2 + 16 + CF::H + CF::H + 16;
  static constexpr int PADF = BIAS + NBIAS;                      // 4 floats of slack in front of the input rows (column -1 reads)
  static constexpr int IN0 = (PADF + 4 + 31) / 32 * 32;           // [8][PT] staged input rows (x or q), buffer 0; 128-byte aligned (TMA)
  static constexpr int IN1 = IN0 + 8 * PT;                       // buffer 1
  static constexpr int LQ = IN1 + 8 * PT;                        // [16][PT] logits -> q
  static constexpr int S = LQ + 16 * PT;                         // [H][PT] h1 (fp32), later g2
  static constexpr int H2B = S + CF::H * PT;                     // [H2][PT]
  static constexpr int PB = H2B + CF::H2 * PT;                   // [16][PT] to_params output
  static constexpr int FLOATS = PB + 16 * PT;
  static constexpr size_t BYTES = (size_t)OPB + (size_t)FLOATS * 4;
};
}  // namespace t5f

template <typename CF, bool ENC, bool DEC>
__global__ void __launch_bounds__(NTH, 1) forward_tc5_kernel(FwdArgs fa, vhmm_params P, const __grid_constant__ CUtensorMap tm_in) {
  constexpr int C = CF::C, H = CF::H, H2 = CF::H2, K = CF::K;
  static_assert(H == 64 && H2 == 32, "the tcgen05 tiling is written for hidden_dim = 64, hidden_dim2 = 32");
  static_assert(ENC || DEC, "nothing to do");
  using L = t5f::Lay<CF, ENC, DEC>;
  extern __shared__ __align__(1024) uint8_t smraw[];
  __shared__ __align__(8) uint64_t bar_e, bar_d, bar_in[2];
  __shared__ uint32_t tmem_sh, amax_sh;
  uint8_t* W2e = smraw;                                        // fp16 encoder conv2 weights (ENC)
  uint8_t* W2d = W2e + (ENC ? t5e::W_BYTES : 0);               // fp16 decoder conv2 weights (DEC)
  uint8_t* RA = W2d + (DEC ? t5::W_BYTES : 0);                 // fp16 activation parts: h1, later g1
  float* fl = reinterpret_cast<float*>(smraw + L::OPB);
  float* W1s = fl + L::W1;
  float* W3s = fl + L::W3;
  float* WFT = fl + L::WFT;
  float* WFs = fl + L::WFS;
  float* WPs = fl + L::WP;
  float* b1s = fl + L::BIAS;
  float* b2s = b1s + H;
  float* b3s = b2s + H2;
  float* bd1s = b3s + 16;
  float* bd2s = bd1s + H;
  float* bps = bd2s + H;
  float* IN0 = fl + L::IN0;
  float* IN1 = fl + L::IN1;
  float* LQ = fl + L::LQ;
  float* S = fl + L::S;
  float* H2b = fl + L::H2B;
  float* Pb = fl + L::PB;
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;

  for (int i = tid; i < (int)(L::BYTES / 4); i += NTH) reinterpret_cast<uint32_t*>(smraw)[i] = 0u;
  if (tid == 0) {
    tc5::mbar_init(&bar_e, 1); tc5::mbar_init(&bar_d, 1); tc5::mbar_init(&bar_in[0], 1); tc5::mbar_init(&bar_in[1], 1);
    tc5::mbar_init_fence(); amax_sh = 0u;
  }
  if (w == 0) tc5::tmem_alloc512(&tmem_sh);
  __syncthreads();
  float sWe = 1.f, sWd = 1.f, sA = 1.f;
  if (ENC) {
    sWe = stage_w2e_f16<CF>(W2e, P.p[P_E2W], &amax_sh);        // ends with a barrier
    stage_conv_w_T(W1s, P.p[P_E1W], H, C);
    stage_conv_w(W3s, P.p[P_E3W], K, H2, 1, 16, CF::PW_H2);
    stage_vec(b1s, P.p[P_E1B], H); stage_vec(b2s, P.p[P_E2B], H2); stage_vec(b3s, P.p[P_E3B], K);
  }
  if (DEC) {
    stage_vec(WFs, fa.wfold, 3 * H * 8);
    for (int i = tid; i < 3 * K * H; i += NTH) {               // WFT[tau][k][o] = wfold[tau][o][k]
      const int o = i % H, k = (i / H) % K, tau = i / (H * K);
      WFT[i] = fa.wfold[(tau * H + o) * 8 + k];
    }
    stage_conv_w(WPs, P.p[P_DPW], 2 * C, H, 1, 16, CF::PW_H);
    stage_vec(bd1s, P.p[P_D1B], H); stage_vec(bd2s, P.p[P_D2B], H); stage_vec(bps, P.p[P_DPB], 2 * C);
    __syncthreads();
    if (tid == 0) amax_sh = 0u;
    __syncthreads();
    {   // weight scale of decoder.conv2
      float m = 0.f;
      for (int i = tid; i < H * H * 3; i += NTH) m = fmaxf(m, fabsf(P.p[P_D2W][i]));
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
      if (lane == 0) atomicMax(&amax_sh, __float_as_uint(m));
    }
    __syncthreads();
    sWd = tc5::pow2_scale(__uint_as_float(amax_sh));
    for (int i = tid; i < H * H * 3; i += NTH) {               // W2[o][c][tau] -> fp16 parts, layout of decoder_tc5.cuh
      const int tau = i % 3, c = (i / 3) % H, o = i / (3 * H);
      __half h, m;
      tc5::split_f16(P.p[P_D2W][i] * sWd, h, m);
      uint8_t* p = W2d + ((tau * 8 + (c >> 3)) * t5::WROWS + o) * 16 + (c & 7) * 2;
      *reinterpret_cast<__half*>(p) = h;
      *reinterpret_cast<__half*>(p + 64 * 16) = m;
    }
    __syncthreads();
    if (tid == 0) amax_sh = 0u;
    __syncthreads();
    if (tid < H) {   // static bound of g1 = relu(conv1'(q) + b1), q a probability vector (decoder_tc5.cuh)
      float bnd = fabsf(bd1s[tid]);
      for (int tau = 0; tau < 3; ++tau) {
        float mk = 0.f;
        for (int k = 0; k < K; ++k) mk = fmaxf(mk, fabsf(WFs[(tau * H + tid) * 8 + k]));
        bnd += mk;
      }
      atomicMax(&amax_sh, __float_as_uint(bnd));
    }
    __syncthreads();
    // decode() takes ANY q (not necessarily a probability vector): the bound is scaled by max |q| of the tile below
    sA = tc5::pow2_scale(__uint_as_float(amax_sh));
  }
  tc5::fence_async_smem();
  tc5::fence_before_sync();
  __syncthreads();
  tc5::fence_after_sync();
  const float g1_bound = DEC ? __uint_as_float(amax_sh) : 0.f;
  const uint32_t tmem = tmem_sh;
  const uint32_t sRA = tc5::smem_u32(RA), sW2e = tc5::smem_u32(W2e), sW2d = tc5::smem_u32(W2d);
  uint32_t par_e = 0, par_d = 0;
  PROF_DECL
  const uint32_t lane_base = (uint32_t)(32 * (w & 3)) << 16;
  const int jrow = 32 * (w & 3) + lane, cg = w >> 2;
  constexpr uint32_t ID_F128 = tc5::idesc_f16(128, 0, 0), ID_F64 = tc5::idesc_f16(64, 0, 0);
  __syncthreads();   // everybody has read amax_sh

  const int T = fa.T;
  constexpr int IN_ROWS = ENC ? C : K;
  uint32_t par_in[2] = {0u, 0u};
  auto fetch = [&](int64_t tile, int which) {
    float* dst = which ? IN1 : IN0;
    const int64_t b = tile / fa.tiles_per_seq;
    const int t0 = (int)(tile - b * fa.tiles_per_seq) * TT;
    if (fa.use_tma) {   // one box {132 columns from t0-4, all rows, sequence b}; zeros outside [0,T) by the TMA's own bounds fill
      if (tid == 0) {
        tc5::fence_async_smem();                     // earlier generic-proxy accesses of the buffer vs the async-proxy write
        ::vhmm::tma::expect_tx(&bar_in[which], (uint32_t)(PT * 4 * IN_ROWS));
        ::vhmm::tma::load_3d(dst, &tm_in, t0 - HL, 0, (int)b, &bar_in[which]);
      }
    } else if (ENC) prefetch_rows(dst, fa.x, C, b, T, t0, T);
    else prefetch_rows(dst, fa.q_in, K, b, T, t0, T);
  };
  int64_t tile = blockIdx.x;
  int buf = 0;
  if (tile < fa.ntiles) fetch(tile, 0);
  cp_async_commit();
  for (; tile < fa.ntiles; tile += gridDim.x) {
    const int64_t nxt = tile + gridDim.x;
    if (nxt < fa.ntiles) fetch(nxt, buf ^ 1);
    cp_async_commit();
    if (fa.use_tma) { tc5::mbar_wait(&bar_in[buf], par_in[buf]); par_in[buf] ^= 1u; }
    else cp_async_wait_prev();
    const float* IN = buf ? IN1 : IN0;
    const int64_t b = tile / fa.tiles_per_seq;
    const int t0 = (int)(tile - b * fa.tiles_per_seq) * TT;
    if (tid == 0) amax_sh = 0u;
    __syncthreads();
    const float* Q = IN;                                       // rows [0,K): q of the window (DEC)
    if (ENC) {
      float sH;
      encoder_forward_tc5<CF, true>(W1s, b1s, b2s, W3s, b3s, IN, S, H2b, LQ, RA, sRA, sW2e, tmem, &bar_e, par_e, &amax_sh, sWe,
                                    sH, t0, T, [] {} PROF_ARGS);
      if (tid < NC) {   // softmax over the K logits (:114 / :141); zero outside [0,T) (the decoder's conv padding)
        const int j = tid, t = t0 - HL + j;
        const bool inr = (j >= 2 && j < NC - 2 && t >= 0 && t < T);
        const bool own = inr && j >= J0 && j < J0 + TT;
        float lg[K], m = -INFINITY, s = 0.f;
#pragma unroll
        for (int k = 0; k < K; ++k) { lg[k] = LQ[k * PT + j]; m = fmaxf(m, lg[k]); }
#pragma unroll
        for (int k = 0; k < K; ++k) s += expf(lg[k] - m);
        const float inv = 1.f / s;
#pragma unroll
        for (int k = 0; k < K; ++k) {
          const float qv = inr ? expf(lg[k] - m) * inv : 0.f;
          LQ[k * PT + j] = qv;
          if (own) {
            if (fa.logits) fa.logits[(b * K + k) * (int64_t)T + t] = lg[k];
            if (fa.q) fa.q[(b * K + k) * (int64_t)T + t] = qv;
          }
        }
      }
      __syncthreads();
      Q = LQ;
    }
    if (DEC) {
      float sG = sA;
      if (!ENC) {   // decode(q) of an arbitrary q: the static bound of g1 scales with max |q| of the window
        float m = 0.f;
        for (int i = tid; i < K * NC; i += NTH) m = fmaxf(m, fabsf(Q[(i / NC) * PT + (i % NC)]));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (lane == 0) atomicMax(&amax_sh, __float_as_uint(m));
        __syncthreads();
        // |sum_k W'_k q_k| <= K max|q| max|W'|; pow2_scale leaves a factor 4 of fp16 headroom, of which 3 is used here so that a
        // probability-valued q (max <= 1) gets exactly the static scale of forward()
        sG = tc5::pow2_scale(g1_bound * fmaxf((float)K * __uint_as_float(amax_sh) * (1.f / 3.f), 1.f));
      }
      {   // g1 = relu(conv3'(q) + b1) as FFMA straight into the fp16 operand parts (:83-85)
        float g1v[SkConv<H>::CS];
        int r, c0;
        (void)sk_conv_taps_regs<K, H, K>(WFT, bd1s, Q, t0, T, g1v, r, c0);
        sk_store_parts(g1v, sG, RA, t5::PLANE, t5::PART, r, c0);
      }
      tc5::fence_async_smem();
      tc5::fence_before_sync();
      __syncthreads();
      if (w == 0) {   // g2 pre-activation on the tensor core (same issue pattern as decoder_tc5_kernel)
        tc5::fence_after_sync();
        const bool lead = tc5::elect_one();
#pragma unroll 1
        for (int tau = 0; tau < 3; ++tau)
#pragma unroll
          for (int ch = 0; ch < 4; ++ch) {
            const uint64_t ah = tc5::make_desc(sRA + (2 * ch) * t5::PLANE + tau * 16, t5::PLANE, 128);
            const uint64_t am = tc5::make_desc(sRA + t5::PART + (2 * ch) * t5::PLANE + tau * 16, t5::PLANE, 128);
            const uint64_t bw = tc5::make_desc(sW2d + (tau * 8 + 2 * ch) * t5::WPLANE, t5::WPLANE, 128);
            const uint32_t base = tmem + (((tau * 4 + ch) & 1) ? t5::WACC : t5::ACC_F);
            if (lead) {
              tc5::mma_f16(base, ah, bw, ID_F128, (tau * 4 + ch) >= 2 ? 1u : 0u);
              tc5::mma_f16(base + 64, am, bw, ID_F64, 1u);
            }
          }
        if (lead) tc5::commit(&bar_d);
        __syncwarp();
      }
      tc5::mbar_wait(&bar_d, par_d); par_d ^= 1u;
      tc5::fence_after_sync();
      {   // epilogue: g2 = relu(acc / (sG sW) + b2) inside the sequence, 0 outside -> S[o][j]
        float vm[16], vs[16], wm[16], ws[16];
        tc5::tmem_ld16(tmem + lane_base + t5::ACC_F + 16 * cg, vm);
        tc5::tmem_ld16(tmem + lane_base + t5::ACC_F + 64 + 16 * cg, vs);
        tc5::tmem_ld16(tmem + lane_base + t5::WACC + 16 * cg, wm);
        tc5::tmem_ld16(tmem + lane_base + t5::WACC + 64 + 16 * cg, ws);
        const float inv_f = 1.f / (sG * sWd);
        const int t = t0 - HL + jrow;
        const bool in = (t >= 0 && t < T);
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          const int o = 16 * cg + e;
          S[o * PT + jrow] = in ? fmaxf(((vm[e] + wm[e]) + (vs[e] + ws[e])) * inv_f + bd2s[o], 0.f) : 0.f;
        }
      }
      tc5::fence_before_sync();
      __syncthreads();
      // (mu | logvar) = Wp g2 + bp (:87-90)
      cols_op<16, 1, 1, H / 8, 1, false>(WPs, 0, CF::PW_H, S, EpiBiasRows{Pb, bps, 2 * C});
      __syncthreads();
      for (int i = tid; i < 2 * C * TT; i += NTH) {
        const int r = i / TT, j = J0 + (i - r * TT), t = t0 - HL + j;
        if (t < T) {
          float* dst = (r < C) ? fa.mu + (b * C + r) * (int64_t)T : fa.logvar + (b * C + (r - C)) * (int64_t)T;
          dst[t] = Pb[r * PT + j];
        }
      }
    }
    __syncthreads();
    buf ^= 1;
  }
  tc5::fence_before_sync();
  __syncthreads();
  if (w == 0) tc5::tmem_free512(tmem);
}
