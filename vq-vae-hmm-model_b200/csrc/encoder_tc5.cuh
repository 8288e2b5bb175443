// encoder_tc5.cuh -- kernels A (encoder forward + prior) and C (encoder backward) of the fused ELBO
// path with the encoder's 64->32 k=3 convolution on the 5th-generation tensor core (tcgen05.mma,
// TMEM accumulators): its forward in both kernels, its input and weight gradients in kernel C.
// Same operand scheme as decoder_tc5.cuh: fp32 operands scaled by a power of two and split into two
// fp16 parts (x*s = h + m), activations as act[part][ch/8][row][8] (K-major for the forward / input
// gradient GEMMs with a conv tap = +16 bytes, MN-major for the weight gradient), weights as
// W2h[tau][c/8][part*32+o][8] (K-major B with the parts stacked along N; read MN-major it is the B
// operand of the input gradient).  h*h terms and the small terms go to separate TMEM columns.
// Included by elbo_fused.cu inside namespace vhmm::fz.
#pragma once
#include "tc5.cuh"
#include "skinny.cuh"

namespace t5e {
constexpr int ROWS = 130;                  // row r <-> window column r-1
constexpr int PLANE = ROWS * 16;           // bytes of one 8-channel plane
constexpr int H1_PART = 8 * PLANE;         // h1: 64 channels per part
constexpr int H1_BYTES = 2 * H1_PART;
constexpr int D2_PART = 4 * PLANE;         // dh2: 32 channels per part
constexpr int D2_BYTES = 2 * D2_PART;
constexpr int WROWS = 64;                  // rows of a weight plane: part*32 + o
constexpr int WPLANE = WROWS * 16;
constexpr int W_BYTES = 3 * 8 * WPLANE;    // 24576
constexpr int SB_PLANE = 16 * 16;          // tail chunk of dh2 (non-owned columns zeroed)
constexpr int SB_BYTES = 2 * 4 * SB_PLANE;
constexpr int OPB_A = W_BYTES + H1_BYTES;                          // operand bytes of kernel A
constexpr int OPB_C = W_BYTES + H1_BYTES + D2_BYTES + SB_BYTES;    // ... of kernel C
constexpr uint32_t ACC = 0;                // [0,128): forward (two sets of [32 h*h | 32 small]) / input gradient ([64 | 64])
constexpr uint32_t WACC = 128;             // [128,256): weight-gradient tile accumulator of one tap
constexpr uint32_t TACC = 256;             // [256,448): persistent dW2 accumulators, 64 columns per tap
}  // namespace t5e

// W2e[o][c][tau] (H2 x H x 3) -> fp16 parts W2h[((tau*8 + c/8)*64 + part*32 + o)*16 + (c%8)*2]; returns the scale.
// Ends with a __syncthreads().
template <typename CF>
__device__ __forceinline__ float stage_w2e_f16(uint8_t* W2h, const float* __restrict__ W, uint32_t* amax_sh) {
  const int n = CF::H2 * CF::H * 3;
  float m = 0.f;
  for (int i = threadIdx.x; i < n; i += NTH) m = fmaxf(m, fabsf(W[i]));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) atomicMax(amax_sh, __float_as_uint(m));
  __syncthreads();
  const float sW = tc5::pow2_scale(__uint_as_float(*amax_sh));
  for (int i = threadIdx.x; i < n; i += NTH) {
    const int tau = i % 3, c = (i / 3) % CF::H, o = i / (3 * CF::H);
    __half h, mm;
    tc5::split_f16(W[i] * sW, h, mm);
    uint8_t* p = W2h + ((tau * 8 + (c >> 3)) * t5e::WROWS + o) * 16 + (c & 7) * 2;
    *reinterpret_cast<__half*>(p) = h;
    *reinterpret_cast<__half*>(p + CF::H2 * 16) = mm;
  }
  __syncthreads();
  return sW;
}

// fp32 act[ch][PT] (window columns 0..127) -> fp16 parts dst[part][ch/8][col+1][8], scaled
template <int NCH>
__device__ __forceinline__ void act_to_parts(const float* __restrict__ A, float scale, uint8_t* dst, int part_bytes) {
  for (int i = threadIdx.x; i < (NCH / 8) * NC; i += NTH) {
    const int pl = i / NC, j = i - pl * NC;
    uint4 hh[1], mm[1];
    {
      float v8[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) v8[e] = A[(pl * 8 + e) * PT + j];
      tc5::split8(v8, tc5::f2_pack(scale, scale), hh[0], mm[0]);
    }
    uint8_t* p = dst + pl * t5e::PLANE + (j + 1) * 16;
    *reinterpret_cast<uint4*>(p) = *reinterpret_cast<const uint4*>(hh);
    *reinterpret_cast<uint4*>(p + part_bytes) = *reinterpret_cast<const uint4*>(mm);
  }
}

// encoder forward with conv2 on the tensor core (kernels A and C).  XU rows [0,8) hold x.
// h1 (fp32) -> H1 and its fp16 parts -> RA (scale returned in sH); h2 -> H2b; logits -> LQ.
// *amax_sh must be 0 on entry.  Ends with a __syncthreads().
template <typename CF, bool LOGITS, typename Mid>
__device__ __forceinline__ void encoder_forward_tc5(const float* W1s, const float* b1s, const float* b2s, const float* W3s,
                                                    const float* b3s, const float* XU, float* H1, float* H2b, float* LQ,
                                                    uint8_t* RA, uint32_t sRA, uint32_t sW2, uint32_t tmem, uint64_t* bar,
                                                    uint32_t& par, uint32_t* amax_sh, float sW, float& sH, int t0, int T, Mid mid PROF_PARAMS) {
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
  {
    float lm = sk_conv_taps_fwd<CF::C, CF::H>(W1s, b1s, XU, H1, t0, T);   // conv1 + ReLU as FFMA (skinny.cuh)
    PROF(18);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) lm = fmaxf(lm, __shfl_xor_sync(0xffffffffu, lm, o));
    if (lane == 0) atomicMax(amax_sh, __float_as_uint(lm));
  }
  __syncthreads(); PROF(14);
  sH = tc5::pow2_scale(__uint_as_float(*amax_sh));
  act_to_parts<CF::H>(H1, sH, RA, t5e::H1_PART);
  tc5::fence_async_smem();
  tc5::fence_before_sync();
  __syncthreads(); PROF(15);
  if (w == 0) {   // h2 pre-activation: per accumulator set [0,32) = h1_h*W_h, [32,64) = h1_h*W_m + h1_m*W_h
    tc5::fence_after_sync();
    const bool lead = tc5::elect_one();
    constexpr uint32_t ID64 = tc5::idesc_f16(64, 0, 0), ID32 = tc5::idesc_f16(32, 0, 0);
#pragma unroll 1
    for (int tau = 0; tau < 3; ++tau)
#pragma unroll
      for (int ch = 0; ch < 4; ++ch) {
        const uint64_t ah = tc5::make_desc(sRA + (2 * ch) * t5e::PLANE + tau * 16, t5e::PLANE, 128);
        const uint64_t am = tc5::make_desc(sRA + t5e::H1_PART + (2 * ch) * t5e::PLANE + tau * 16, t5e::PLANE, 128);
        const uint64_t bw = tc5::make_desc(sW2 + (tau * 8 + 2 * ch) * t5e::WPLANE, t5e::WPLANE, 128);
        const uint32_t base = tmem + t5e::ACC + (((tau * 4 + ch) & 1) ? 64u : 0u);
        if (lead) {
          tc5::mma_f16(base, ah, bw, ID64, (tau * 4 + ch) >= 2 ? 1u : 0u);
          tc5::mma_f16(base + 32, am, bw, ID32, 1u);
        }
      }
    if (lead) tc5::commit(bar);
    __syncwarp();
  }
  mid();   // work that does not depend on h2 runs under the tensor-core latency (may use __syncthreads)
  tc5::mbar_wait(bar, par); par ^= 1u; PROF(16);
  tc5::fence_after_sync();
  {
    const uint32_t lane_base = (uint32_t)(32 * (w & 3)) << 16;
    const int jrow = 32 * (w & 3) + lane, cg = w >> 2;
    float vm[8], vs[8], wm[8], ws[8];
    tc5::tmem_ld8(tmem + lane_base + t5e::ACC + 8 * cg, vm);
    tc5::tmem_ld8(tmem + lane_base + t5e::ACC + 32 + 8 * cg, vs);
    tc5::tmem_ld8(tmem + lane_base + t5e::ACC + 64 + 8 * cg, wm);
    tc5::tmem_ld8(tmem + lane_base + t5e::ACC + 96 + 8 * cg, ws);
    const float inv = 1.f / (sH * sW);
    const int t = t0 - HL + jrow;
    const bool inside = (t >= 0 && t < T);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int o = 8 * cg + e;
      H2b[o * PT + jrow] = inside ? fmaxf(((vm[e] + wm[e]) + (vs[e] + ws[e])) * inv + b2s[o], 0.f) : 0.f;
    }
  }
  tc5::fence_before_sync();
  __syncthreads(); PROF(17);
  if (LOGITS) {
    sk_rows_fwd<CF::K, CF::H2>(W3s, CF::PW_H2, b3s, H2b, LQ);   // logits layer as FFMA
    __syncthreads();
  }
}

namespace t5e {
template <typename CF>
struct Sz {
  static constexpr size_t A = (size_t)OPB_A + 4 * (size_t)(CF::SM_A - 3 * CF::H2 * CF::PW_H) + 256;   // no fp32 copy of the conv2 weights; + 2 alignment pads (TMA)
  static constexpr size_t C = (size_t)OPB_C + 4 * (size_t)(CF::SM_C - 3 * CF::H2 * CF::PW_H);
};
}  // namespace t5e

// =============================================================================================
// kernel A (tcgen05 build): encoder forward + prior forward/backward
// =============================================================================================
template <typename CF>
__global__ void __launch_bounds__(NTH, 1) enc_prior_tc5_kernel(Common cm, vhmm_params P, const __grid_constant__ TmaMaps tm) {
  constexpr int C = CF::C, H = CF::H, H2 = CF::H2, K = CF::K, U = CF::U, TH = CF::TH, K2 = CF::K2;
  static_assert(H == 64 && H2 == 32, "the tcgen05 encoder tiling is written for hidden_dim = 64, hidden_dim2 = 32");
  static_assert(U == 4 && TH == 128 && K2 <= 12, "the FFMA transition-MLP layers are written for u_dim = 4, trans_hidden = 128");
  extern __shared__ __align__(1024) uint8_t smraw[];
  __shared__ double red[NW];
  __shared__ __align__(8) uint64_t bar_f, bar_in[2];
  __shared__ uint32_t tmem_sh, amax_sh;
  uint8_t* W2h = smraw;                   // fp16 conv2 weights
  uint8_t* RA = W2h + t5e::W_BYTES;       // h1 parts
  float* sm = reinterpret_cast<float*>(smraw + t5e::OPB_A);
  float* W1s = sm;                        // [3][H][8]
  float* W3s = W1s + 3 * H * 8;           // [16][PW_H2]
  float* Wt0s = W3s + 16 * CF::PW_H2;     // [TH][8]
  float* Wt2s = Wt0s + TH * 8;            // [16][PW_TH]
  float* b1s = Wt2s + 16 * CF::PW_TH;
  float* b2s = b1s + H;
  float* b3s = b2s + H2;                  // [16]
  float* bt0s = b3s + 16;
  float* bt2s = bt0s + TH;                // [16]
  float* lps = bt2s + 16;                 // [8]
  // the two input buffers are 128-byte aligned (TMA destinations): float offsets rounded up to multiples of 32
  float* XU0 = sm + ((int)((lps + 8) - sm) + 31) / 32 * 32;   // rows 0..7 x, rows 8..15 u (buffer 0; buffer 1 at the end)
  float* R1 = XU0 + 16 * PT;               // h1 | h2, later the prior's hidden layer a
  float* H1 = R1;
  float* Ab = R1;
  float* LQ = R1 + CF::R1 * PT;           // rows 0..K-1 logits -> q, rows 8.. log q
  float* Mb = LQ + 16 * PT;               // transition logits -> log_A, row i*K+j
  float* dMb = Mb + 16 * PT;
  float* acc = dMb + 16 * PT;             // ACC_A
  float* XU1 = sm + ((int)((acc + CF::ACC_A) - sm) + 31) / 32 * 32;
  float* H2b = XU1 + 16 * PT;             // h2 has its own rows: R1 holds the prior's hidden layer a by the time it is written
  for (int i = threadIdx.x; i < (int)(t5e::Sz<CF>::A / 4); i += NTH) reinterpret_cast<uint32_t*>(smraw)[i] = 0u;
  if (threadIdx.x == 0) { tc5::mbar_init(&bar_f, 1); tc5::mbar_init(&bar_in[0], 1); tc5::mbar_init(&bar_in[1], 1); tc5::mbar_init_fence(); amax_sh = 0u; }
  if (threadIdx.x < 32) tc5::tmem_alloc512(&tmem_sh);
  __syncthreads();
  const float sW = stage_w2e_f16<CF>(W2h, P.p[P_E2W], &amax_sh);
  stage_conv_w_T(W1s, P.p[P_E1W], H, C);   // [tau][c][H]: lane <-> output row reads are conflict-free
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
  tc5::fence_async_smem();
  tc5::fence_before_sync();
  __syncthreads();
  tc5::fence_after_sync();
  const uint32_t tmem = tmem_sh, sRA = tc5::smem_u32(RA), sW2 = tc5::smem_u32(W2h);
  uint32_t par_f = 0;

  double* part = cm.acc_part + (int64_t)blockIdx.x * CF::ACC_ALL;
  int ndone = 0;
  TileInfo ti, tn;
  int buf = 0;
  uint32_t par_in[2] = {0u, 0u};
  // the x and u windows of a tile: two TMA boxes (PT columns from t0-4, all rows, one sequence; zeros outside [0,T) by the
  // TMA's bounds fill) issued by one thread, or the cp.async path when the tensors do not qualify (T % 4, alignment)
  auto fetch = [&](const TileInfo& t_, int which) {
    float* dst = which ? XU1 : XU0;
    if (cm.use_tma) {
      if (tid == 0) {
        tc5::fence_async_smem();
        ::vhmm::tma::expect_tx(&bar_in[which], (uint32_t)(PT * 4 * (C + U)));
        ::vhmm::tma::load_3d(dst, &tm.x, t_.t0 - HL, 0, (int)t_.b, &bar_in[which]);
        ::vhmm::tma::load_3d(dst + 8 * PT, &tm.u, t_.t0 - HL, 0, (int)t_.b, &bar_in[which]);
      }
      return;
    }
    prefetch_rows(dst, cm.x, C, t_.b, cm.T, t_.t0, cm.T);
    prefetch_rows(dst + 8 * PT, cm.u, U, t_.b, cm.T, t_.t0, cm.T);
  };
  int64_t tile = next_tile<6>(cm, blockIdx.x, ti);
  if (tile >= 0) fetch(ti, 0);
  cp_async_commit();
  int64_t nxt = (tile >= 0) ? next_tile<6>(cm, tile + gridDim.x, tn) : -1;
  while (tile >= 0) {
    PROF(12);
    if (nxt >= 0) fetch(tn, buf ^ 1);
    cp_async_commit();
    PROF(9);
    const TilePre pre = tile_pre(cm, nxt >= 0 ? nxt + gridDim.x : cm.ntiles);   // consumed after the tile
    PROF(10);
    if (cm.use_tma) { tc5::mbar_wait(&bar_in[buf], par_in[buf]); par_in[buf] ^= 1u; }
    else cp_async_wait_prev();
    PROF(11);
    const float* XU = buf ? XU1 : XU0;
    const int t0 = ti.t0, T = cm.T, L = ti.L;
    const int64_t b = ti.b;
    if (tid == 0) amax_sh = 0u;
    __syncthreads(); PROF(0);
    float sH;
    (void)sizeof(sH);
    // The transition MLP depends on u only: it runs between the issue of the conv2 MMAs and the wait for them
    // (h1 is dead once its fp16 parts are staged, so its rows become the hidden layer a; h2 has rows of its own).
    auto prior_fwd = [&]() {
      PROF(13);
      sk_rows_in_fwd<U, TH>(Wt0s, bt0s, XU + 8 * PT, Ab, t0, T);   // a = relu(W0 u + b0) (:68) as FFMA
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
    };
    encoder_forward_tc5<CF, true>(W1s, b1s, b2s, W3s, b3s, XU, H1, H2b, LQ, RA, sRA, sW2, tmem, &bar_f, par_f, &amax_sh, sW, sH, t0, T, prior_fwd PROF_ARGS); PROF(1);
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
          cm.lg[(b * K + k) * (int64_t)T + t] = lg[k];   // kernel C reads the logits instead of recomputing conv3
          if (t == 0) { s_q0[k] += (double)qv; s_trans += (double)(qv * lps[k]); }   // :123 (not masked)
        }
      }
      if (own && t < L) { s_ent += (double)ent; s_valid += 1.0; }                     // :134-135
    }
    __syncthreads(); PROF(2);
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
    // gradients of transition_net.{2,0}: dWt2[r][h] += sum_t dM[r][t] a[h][t]; da = relu'(a) * Wt2^T dM (kept in
    // registers); dWt0[h][c] += sum_t da[h][t] u[c][t] and its bias -- one register-blocked FFMA pass
    // (skinny.cuh: sk_prior_bwd).  Slice partials go to LQ|Mb (q, log q and log_A are dead after the pair term).
    static_assert(SkPrior<K2, U, TH, 2>::FLOATS <= 32 * PT, "prior gradient scratch must fit LQ|Mb");
    sk_prior_bwd<K2, U, TH, 2, CF::PW_TH>(Ab, dMb, Wt2s, XU + 8 * PT, LQ);
    row_sums<J0, TT / 4, PT, NW>(dMb, K2, acc + CF::A_BT2);
    __syncthreads(); PROF(6);
    sk_prior_bwd_reduce<K2, U, TH, 2, CF::PW_TH>(LQ, acc + CF::A_WT2, acc + CF::A_WT0, acc + CF::A_BT0);
    PROF(7);
    __syncthreads(); PROF(8);
    if (++ndone % FLUSH == 0) flush_acc(part, acc, CF::ACC_A);
    tile = nxt; ti = tn; buf ^= 1;
    nxt = (tile >= 0) ? tile_resolve<6>(cm, pre, tn) : -1;
  }
  PROF_DUMP("enc_prior", 19);
  // per-CTA partials
  flush_acc(part, acc, CF::ACC_A);
  double* lp = cm.loss_part + (int64_t)blockIdx.x * LP;
  double v;
  v = block_sum_d(s_trans, red); if (tid == 0) lp[1] = v;
  v = block_sum_d(s_ent, red);   if (tid == 0) lp[2] = v;
  v = block_sum_d(s_valid, red); if (tid == 0) lp[3] = v;
#pragma unroll
  for (int k = 0; k < K; ++k) { v = block_sum_d(s_q0[k], red); if (tid == 0) lp[4 + k] = v; }
  tc5::fence_before_sync();
  __syncthreads();
  if (w == 0) tc5::tmem_free512(tmem);
}


// =============================================================================================
// kernel C (tcgen05 build): encoder backward (forward recomputed in the tile)
// =============================================================================================
template <typename CF>
__global__ void __launch_bounds__(NTH, 1) enc_bwd_tc5_kernel(Common cm, vhmm_params P, const __grid_constant__ TmaMaps tm) {
  constexpr int C = CF::C, H = CF::H, H2 = CF::H2, K = CF::K;
  static_assert(H == 64 && H2 == 32, "the tcgen05 encoder tiling is written for hidden_dim = 64, hidden_dim2 = 32");
  static_assert(K <= 4, "the stored logits ride in rows 4..7 of the dq_dec input group");
  extern __shared__ __align__(1024) uint8_t smraw[];
  __shared__ __align__(8) uint64_t bar_f, bar_d, bar_w, bar_w2;
  __shared__ uint32_t tmem_sh, amax_sh, amax2_sh;
  uint8_t* W2h = smraw;                   // fp16 conv2 weights
  uint8_t* RA = W2h + t5e::W_BYTES;       // h1 parts
  uint8_t* RB = RA + t5e::H1_BYTES;       // dh2 parts
  uint8_t* SBf = RB + t5e::D2_BYTES;      // dh2 tail chunk
  float* sm = reinterpret_cast<float*>(smraw + t5e::OPB_C);
  float* W1s = sm;
  float* W3s = W1s + 3 * H * 8;
  float* b1s = W3s + 16 * CF::PW_H2;
  float* b2s = b1s + H;
  float* b3s = b2s + H2;                  // [16]
  float* XU0 = b3s + 16;                  // rows 0..7 x (buffer 0)
  float* H1 = XU0 + 16 * PT;
  float* H2b = H1 + H * PT;
  float* LQ = H2b + H2 * PT;              // logits, then d loss / d logits (rows >= K stay zero)
  float* DQ0 = LQ + 16 * PT;              // d loss / d q: rows 0.. from the decoder, rows 8.. from the prior (buffer 0)
  float* acc = DQ0 + 16 * PT;             // ACC_C layout (the C_W2 block is unused: dW2 accumulates in TMEM)
  float* XU1 = acc + CF::ACC_C;
  float* DQ1 = XU1 + 16 * PT;
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
  for (int i = tid; i < (int)(t5e::Sz<CF>::C / 4); i += NTH) reinterpret_cast<uint32_t*>(smraw)[i] = 0u;
  if (tid == 0) {
    tc5::mbar_init(&bar_f, 1); tc5::mbar_init(&bar_d, 1); tc5::mbar_init(&bar_w, 1); tc5::mbar_init(&bar_w2, 1);
    tc5::mbar_init_fence();
    amax_sh = 0u;
  }
  if (w == 0) tc5::tmem_alloc512(&tmem_sh);
  __syncthreads();
  const float sW = stage_w2e_f16<CF>(W2h, P.p[P_E2W], &amax_sh);
  stage_conv_w_T(W1s, P.p[P_E1W], H, C);   // [tau][c][H]: lane <-> output row reads are conflict-free
  stage_conv_w(W3s, P.p[P_E3W], K, H2, 1, 16, CF::PW_H2);
  stage_vec(b1s, P.p[P_E1B], H); stage_vec(b2s, P.p[P_E2B], H2); stage_vec(b3s, P.p[P_E3B], K);
  const float aE = (float)(beta_of(cm) / cm.norms[1]);
  PROF_DECL
  tc5::fence_async_smem();
  tc5::fence_before_sync();
  __syncthreads();
  tc5::fence_after_sync();
  const uint32_t tmem = tmem_sh, sRA = tc5::smem_u32(RA), sRB = tc5::smem_u32(RB), sSB = tc5::smem_u32(SBf),
                 sW2 = tc5::smem_u32(W2h);
  uint32_t par_f = 0, par_d = 0, par_w = 0, par_w2 = 0;
  const uint32_t lane_base = (uint32_t)(32 * (w & 3)) << 16;
  const int jrow = 32 * (w & 3) + lane, cg = w >> 2;
  constexpr uint32_t ID_D64 = tc5::idesc_f16(64, 0, 1), ID_W128 = tc5::idesc_f16(128, 1, 1);

  double* part = cm.acc_part + (int64_t)blockIdx.x * CF::ACC_ALL + CF::ACC_A + CF::ACC_B;
  int ndone = 0;
  {   // persistent dW2 accumulators start at zero
    float z[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) z[e] = 0.f;
#pragma unroll
    for (int tau = 0; tau < 3; ++tau) tc5::tmem_st16(tmem + lane_base + t5e::TACC + 64 * tau + 16 * cg, z);
    tc5::tmem_st_wait();
    tc5::fence_before_sync();
  }
  __syncthreads();
  tc5::fence_after_sync();
  // TMEM lanes 0..31 hold rows o of dW2 from the h part of dh2, lanes 32..63 from its m part (64.. are unused)
  auto flush_all = [&]() {
    flush_acc(part + CF::C_W1, acc + CF::C_W1, CF::ACC_C - CF::C_W1);
    for (int pass = 0; pass < 2; ++pass) {
      if (jrow < 64 && (jrow >> 5) == pass) {
#pragma unroll 1
        for (int tau = 0; tau < 3; ++tau) {
          float a[16];
          tc5::tmem_ld16(tmem + lane_base + t5e::TACC + 64 * tau + 16 * cg, a);
          double* dst = part + CF::C_W2 + (tau * H2 + (jrow & 31)) * CF::PW_H + 16 * cg;
#pragma unroll
          for (int e = 0; e < 16; ++e) { atomicAdd(dst + e, (double)a[e]); a[e] = 0.f; }
          tc5::tmem_st16(tmem + lane_base + t5e::TACC + 64 * tau + 16 * cg, a);
        }
        tc5::tmem_st_wait();
      }
      tc5::fence_before_sync();
      __syncthreads();
      tc5::fence_after_sync();
    }
  };
  TileInfo ti, tn;
  int buf = 0;
  auto fetch = [&](const TileInfo& t_, float* xu, float* dq) {
    const int dmax = (t_.Leff + 2 < cm.T) ? t_.Leff + 2 : cm.T;
    prefetch_rows(xu, cm.x, C, t_.b, cm.T, t_.t0, cm.T);
    prefetch_rows(dq, cm.dq_dec, K, t_.b, cm.T, t_.t0, dmax);
    prefetch_rows(dq + 8 * PT, cm.dq_prior, K, t_.b, cm.T, t_.t0, dmax);
    prefetch_rows(dq + 4 * PT, cm.lg, K, t_.b, cm.T, t_.t0, dmax);   // rows 4.. of the dq_dec group are free (K <= 4)
  };
  int64_t tile = next_tile<3>(cm, blockIdx.x, ti);
  if (tile >= 0) fetch(ti, XU0, DQ0);
  cp_async_commit();
  int64_t nxt = (tile >= 0) ? next_tile<3>(cm, tile + gridDim.x, tn) : -1;
  while (tile >= 0) {
    if (nxt >= 0) fetch(tn, buf ? XU0 : XU1, buf ? DQ0 : DQ1);
    cp_async_commit();
    const TilePre pre = tile_pre(cm, nxt >= 0 ? nxt + gridDim.x : cm.ntiles);   // consumed after the tile
    cp_async_wait_prev();
    const float* XU = buf ? XU1 : XU0;
    const float* DQ = buf ? DQ1 : DQ0;
    const int t0 = ti.t0, T = cm.T, L = ti.L;
    if (tid == 0) { amax_sh = 0u; amax2_sh = 0u; }
    __syncthreads(); PROF(0);
    float sH;
    // d loss / d logits needs only the logits kernel A stored and the two dq streams: it runs while the tensor
    // core works on conv2 (the logits layer itself is not recomputed here)
    auto dlogits = [&]() {
      if (tid >= NTH - NC) {    // softmax backward of dq plus the entropy term (:134-137); the last 4 warps: warp 0 is issuing MMAs
        const int j = tid - (NTH - NC), t = t0 - HL + j;
        const bool inr = (j >= 2 && j < NC - 2 && t >= 0 && t < T);
        const bool mk = inr && t < L;
        float lg[K], m = -INFINITY, s = 0.f;
#pragma unroll
        for (int k = 0; k < K; ++k) { lg[k] = DQ[(4 + k) * PT + j]; m = fmaxf(m, lg[k]); }
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
    };
    encoder_forward_tc5<CF, false>(W1s, b1s, b2s, W3s, b3s, XU, H1, H2b, LQ, RA, sRA, sW2, tmem, &bar_f, par_f, &amax_sh, sW, sH, t0, T, dlogits PROF_ARGS); PROF(1);
    __syncthreads(); PROF(2);
    // to_logits gradients: dW3[k][c] += sum_t dl[k][t] h2[c][t] as FFMA (skinny.cuh; slice partials in the unused C_W2 block)
    static_assert(SkRows<K, H2>::FLOATS <= 3 * H2 * CF::PW_H, "to_logits gradient scratch must fit the unused C_W2 block");
    sk_wgrad_rows<K, H2>(LQ, H2b, acc + CF::C_W2);
    row_sums<J0, TT / 4, PT, NW>(LQ, K, acc + CF::C_B3);
    __syncthreads(); PROF(3);
    sk_wgrad_rows_reduce<K, H2>(acc + CF::C_W2, acc + CF::C_W3, CF::PW_H2);
    // dh2 = relu'(h2) * W3^T dl, in place
    {   // dh2 = relu'(h2) * W3^T dlogits as FFMA, with the tile maximum of |dh2| (scale of its fp16 parts)
      float m = sk_rowsT_mask<K, H2>(W3s, CF::PW_H2, LQ, H2b);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
      if (lane == 0) atomicMax(&amax2_sh, __float_as_uint(m));
    }
    __syncthreads(); PROF(4);
    row_sums<J0, TT / 4, PT, NW>(H2b, H2, acc + CF::C_B2);   // conv2 bias gradient
    const float sD = tc5::pow2_scale(__uint_as_float(amax2_sh));
    for (int i = tid; i < (H2 / 8) * NC; i += NTH) {   // dh2 parts -> RB (+ the tail chunk with columns >= J0+TT zeroed)
      const int pl = i / NC, j = i - pl * NC;
      uint4 hh[1], mm[1];
      {
        float v8[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) v8[e] = H2b[(pl * 8 + e) * PT + j];
        tc5::split8(v8, tc5::f2_pack(sD, sD), hh[0], mm[0]);
      }
      uint8_t* p = RB + pl * t5e::PLANE + (j + 1) * 16;
      *reinterpret_cast<uint4*>(p) = *reinterpret_cast<const uint4*>(hh);
      *reinterpret_cast<uint4*>(p + t5e::D2_PART) = *reinterpret_cast<const uint4*>(mm);
      if (j >= J0 + TT - 8 && j < J0 + TT) {
        uint8_t* sp = SBf + pl * t5e::SB_PLANE + (j - (J0 + TT - 8)) * 16;
        *reinterpret_cast<uint4*>(sp) = *reinterpret_cast<const uint4*>(hh);
        *reinterpret_cast<uint4*>(sp + 4 * t5e::SB_PLANE) = *reinterpret_cast<const uint4*>(mm);
      }
    }
    tc5::fence_async_smem();
    tc5::fence_before_sync();
    __syncthreads(); PROF(5);
    // ---- tensor core: dh1 pre-mask (input gradient of conv2) and the weight gradient of tap 0
    auto issue_wgrad = [&](int tau, bool lead, uint32_t acc_col, uint64_t* bar) {   // acc = [dh2_h; dh2_m; -] x [h1_h; h1_m](shifted by tau) over the 120 owned columns
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const uint64_t a = (i < 7) ? tc5::make_desc(sRB + (J0 + 1 + 16 * i) * 16, 128, t5e::PLANE)
                                   : tc5::make_desc(sSB, 128, t5e::SB_PLANE);
        const uint64_t bq = tc5::make_desc(sRA + (J0 + 16 * i + tau) * 16, 128, t5e::PLANE);
        if (lead) tc5::mma_f16(tmem + acc_col, a, bq, ID_W128, i ? 1u : 0u);
      }
      if (lead) tc5::commit(bar);
    };
    if (w == 0) {
      tc5::fence_after_sync();
      const bool lead = tc5::elect_one();
#pragma unroll 1
      for (int tau = 0; tau < 3; ++tau)
#pragma unroll
        for (int ch = 0; ch < 2; ++ch) {   // dh1[t][c] += sum_o dh2[o][t - tau + 1] W2[o][c][tau], o in [16 ch, 16 ch + 16)
          const uint64_t ah = tc5::make_desc(sRB + (2 * ch) * t5e::PLANE + (2 - tau) * 16, t5e::PLANE, 128);
          const uint64_t am = tc5::make_desc(sRB + t5e::D2_PART + (2 * ch) * t5e::PLANE + (2 - tau) * 16, t5e::PLANE, 128);
          const uint64_t bh = tc5::make_desc(sW2 + (tau * 8 * t5e::WROWS + 16 * ch) * 16, 128, t5e::WPLANE);
          const uint64_t bm = tc5::make_desc(sW2 + (tau * 8 * t5e::WROWS + H2 + 16 * ch) * 16, 128, t5e::WPLANE);
          if (lead) {
            tc5::mma_f16(tmem + t5e::ACC, ah, bh, ID_D64, (tau | ch) ? 1u : 0u);
            tc5::mma_f16(tmem + t5e::ACC + 64, ah, bm, ID_D64, (tau | ch) ? 1u : 0u);
            tc5::mma_f16(tmem + t5e::ACC + 64, am, bh, ID_D64, 1u);
          }
        }
      if (lead) tc5::commit(&bar_d);
      issue_wgrad(0, lead, t5e::WACC, &bar_w);
      __syncwarp();
    }
    tc5::mbar_wait(&bar_d, par_d); par_d ^= 1u;
    tc5::fence_after_sync();
    {   // epilogue: dh1 = relu'(h1) * acc/(sD sW), in place over H1
      float vm[16], vs[16];
      tc5::tmem_ld16(tmem + lane_base + t5e::ACC + 16 * cg, vm);
      tc5::tmem_ld16(tmem + lane_base + t5e::ACC + 64 + 16 * cg, vs);
      const float inv_d = 1.f / (sD * sW);
#pragma unroll
      for (int e = 0; e < 16; ++e) {
        float* p = H1 + (16 * cg + e) * PT + jrow;
        *p = (*p > 0.f) ? (vm[e] + vs[e]) * inv_d : 0.f;
      }
    }
    tc5::fence_before_sync();
    __syncthreads(); PROF(6);
    if (w == 0) {   // the input-gradient accumulator is free now: tap 1 runs there under the FFMA phase below
      tc5::fence_after_sync();
      issue_wgrad(1, tc5::elect_one(), t5e::ACC, &bar_w2);
      __syncwarp();
    }
    // conv1 gradients: dW1[tau][o][c] += sum_t dh1[o][t] x[c][t+tau-1], bias: register-blocked FFMA (skinny.cuh);
    // slice partials go to the C_W2 block of acc, which this build never uses (dW2 accumulates in TMEM)
    static_assert(SkTaps<C, H, 4>::FLOATS <= 3 * H2 * CF::PW_H, "conv1 gradient scratch must fit the unused C_W2 block");
    sk_wgrad_taps<C, H, 4>(H1, XU, acc + CF::C_W2);
    __syncthreads();
    sk_wgrad_taps_reduce<C, H, 4>(acc + CF::C_W2, acc + CF::C_W1, acc + CF::C_B1);
    PROF(7);
    {   // conv2 weight gradient: drain each tap's tile sum into the persistent TMEM accumulators.  Tap 0 (WACC) and
        // tap 1 (ACC) were issued under the previous phases; tap 2 reuses WACC as soon as tap 0 is drained.
      const float inv_w = 1.f / (sD * sH);
#pragma unroll 1
      for (int tau = 0; tau < 3; ++tau) {
        const uint32_t col = (tau == 1) ? t5e::ACC : t5e::WACC;
        if (tau == 1) { tc5::mbar_wait(&bar_w2, par_w2); par_w2 ^= 1u; }
        else          { tc5::mbar_wait(&bar_w, par_w); par_w ^= 1u; }
        tc5::fence_after_sync();
        if ((w & 3) < 2) {   // lanes 0..63 carry data
          float v0[16], v1[16], a[16];
          tc5::tmem_ld16(tmem + lane_base + col + 16 * cg, v0);        // x h1_h
          tc5::tmem_ld16(tmem + lane_base + col + 64 + 16 * cg, v1);   // x h1_m
          tc5::tmem_ld16(tmem + lane_base + t5e::TACC + 64 * tau + 16 * cg, a);
#pragma unroll
          for (int e = 0; e < 16; ++e) a[e] += (v0[e] + v1[e]) * inv_w;
          tc5::tmem_st16(tmem + lane_base + t5e::TACC + 64 * tau + 16 * cg, a);
          tc5::tmem_st_wait();
        }
        tc5::fence_before_sync();
        if (tau != 1) __syncthreads();
        if (tau == 0 && w == 0) {
          tc5::fence_after_sync();
          issue_wgrad(2, tc5::elect_one(), t5e::WACC, &bar_w);
          __syncwarp();
        }
      }
    }
    PROF(8);
    if (++ndone % FLUSH == 0) flush_all();
    tile = nxt; ti = tn; buf ^= 1;
    nxt = (tile >= 0) ? tile_resolve<3>(cm, pre, tn) : -1;
  }
  PROF_DUMP("enc_bwd_tc5", 19);
  flush_all();
  tc5::fence_before_sync();
  __syncthreads();
  if (w == 0) tc5::tmem_free512(tmem);
}
