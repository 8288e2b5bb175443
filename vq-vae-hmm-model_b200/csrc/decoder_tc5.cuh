// decoder_tc5.cuh -- kernel B of the fused ELBO path (decoder forward + backward, see elbo_fused.cu)
// with its three 64x64 k=3 GEMM families on the 5th-generation tensor core (tcgen05.mma, TMEM
// accumulators).  Included by elbo_fused.cu inside namespace vhmm::fz.
//
//   g2  = relu(conv3(g1; W2) + b2)          M = 128 time columns, N = 64 out channels, K = 3 x 64
//   dg1 = conv3^T(dg2; W2)                  M = 128 time columns, N = 64 in channels,  K = 3 x 64
//   dW2[tau] += dg2 . g1(shifted)^T         M = 64 (x2 parts),   N = 64 (x2 parts),   K = 120 owned columns
//
// Arithmetic.  fp32 parity (1e-5) on a 16-bit tensor-core path: every fp32 operand x is scaled by a
// power of two s and split into two fp16 parts, x*s = h + m (22 significant bits, the same as the
// hi/lo TF32 split of the mma.sync path); a product needs h*h + h*m + m*h (m*m ~ 2^-22 dropped).
// The scale keeps h, m inside fp16's range: a per-launch scale for the weights, a static bound for
// g1 (q is a probability vector, so |conv1'(q)| <= sum_tau max_k |W'|), a per-tile maximum for dg2.
// The tensor core's accumulate truncates toward zero, so the h*h terms go to their own TMEM columns
// and all small terms to others; they meet in an fp32 FADD in the epilogue.
//
// Operand layout.  Activations live in shared memory as fp16 act[part][ch/8][row][8] (row = window
// column + 1, one zero row each side).  The SAME buffer is the K-major A operand of the forward /
// input-gradient GEMMs (a conv tap = +16 bytes on the descriptor start address) and the MN-major
// operand of the weight-gradient GEMM (K = time).  The weights are W2s[tau][c/8][part*64+o][8]:
// K-major B of the forward GEMM with the two parts stacked along N (one N=128 MMA gives h*h | h*m),
// and, read MN-major, the B operand of the input-gradient GEMM.  In the weight-gradient GEMM both
// operands stack their parts ([dg2_h; dg2_m] x [g1_h; g1_m]): ONE M=128 N=128 MMA per (tap, 16
// columns) yields all four products.
//
// The small layers (folded conv1 3->64, to_params 64->10 and their gradients) stay on the mma.sync
// helpers of mma_tf32.cuh and work on one fp32 scratch array S[64][PT] that holds, in turn, g1, g2,
// dg2 and dg1.  Weight gradients: small ones in shared memory (as before); dW2 is drained from the
// MMA accumulator once per tile and tap and added (FADD, RN) into persistent TMEM columns, which are
// folded into the CTA's double partial every FLUSH tiles -- same numerics as the shared-memory
// accumulators of the mma.sync kernel, no shared memory spent.
#pragma once
#include "tc5.cuh"
#include "skinny.cuh"

namespace t5 {
constexpr int ROWS = 130;                 // rows of an activation plane: row r <-> window column r-1
constexpr int PLANE = ROWS * 16;          // bytes of one 8-channel plane
constexpr int NPL = 8;                    // 8-channel planes per part (H = 64)
constexpr int PART = NPL * PLANE;         // bytes of one part (h or m)
constexpr int ACT_BYTES = 2 * PART;       // h and m parts
constexpr int WROWS = 128;                // rows of a weight plane: part*64 + o
constexpr int WPLANE = WROWS * 16;
constexpr int W_BYTES = 3 * NPL * WPLANE; // 49152
constexpr int SB_ROWS = 16;               // side buffer: the last 16-column chunk of dg2 with the non-owned columns zeroed
constexpr int SB_PLANE = SB_ROWS * 16;
constexpr int SB_BYTES = 2 * NPL * SB_PLANE;
constexpr int OP_BYTES = W_BYTES + 2 * ACT_BYTES + SB_BYTES;
// TMEM columns
constexpr uint32_t ACC_F = 0;             // [0,64) h*h, [64,128) small terms of the forward / input-gradient GEMM
constexpr uint32_t WACC = 128;            // [128,256): weight-gradient tile accumulator of one tap
constexpr uint32_t TACC = 256;            // [256,448): persistent dW2 accumulators, 64 columns per tap
// floats after the operand buffers
template <typename CF>
struct Lay {
  static constexpr int WF = 0;                                  // [3][H][8] folded conv1
  static constexpr int WP = WF + 3 * CF::H * 8;                 // [16][PW_H]
  static constexpr int BD1 = WP + 16 * CF::PW_H;
  static constexpr int BD2 = BD1 + CF::H;
  static constexpr int BP = BD2 + CF::H;                        // [16]
  static constexpr int QX = (BP + 16 + 31) / 32 * 32;           // rows 0..7 q, 8..15 x; 128-byte aligned (TMA destination)
  static constexpr int S = QX + 16 * PT;                        // fp32 scratch [H][PT]
  static constexpr int PB = S + CF::H * PT;                     // [16][PT]
  static constexpr int ACC = PB + 16 * PT;                      // small gradient accumulators: layout of Cfg::B_* minus B_WF
  static constexpr int NACC = CF::ACC_B - CF::B_WF;
  static constexpr int QX1 = (ACC + NACC + 31) / 32 * 32;       // second input buffer (prefetch of the next tile), 128-byte aligned
  static constexpr int SCR = QX1 + 16 * PT;                     // slice partials of the folded conv1 weight gradient (skinny.cuh)
  static constexpr int WFT = SCR + SkTaps<CF::K, CF::H>::FLOATS;  // [3][K][H] folded conv1, transposed for the FFMA forward
  static constexpr int FLOATS = WFT + 3 * CF::K * CF::H;
  static constexpr size_t BYTES = (size_t)OP_BYTES + (size_t)FLOATS * 4;
};
}  // namespace t5

template <typename CF>
__global__ void __launch_bounds__(NTH, 1) decoder_tc5_kernel(Common cm, vhmm_params P, const __grid_constant__ TmaMaps tm) {
  constexpr int C = CF::C, H = CF::H, K = CF::K;
  static_assert(H == 64, "the tcgen05 decoder tiling is written for hidden_dim = 64");
  using L = t5::Lay<CF>;
  extern __shared__ __align__(1024) uint8_t smraw[];
  __shared__ double red[NW];
  __shared__ __align__(8) uint64_t bar_f, bar_d, bar_w, bar_w2, bar_in[2];
  __shared__ uint32_t tmem_sh, amax_sh;
  uint8_t* W2h = smraw;                                   // fp16 weights
  uint8_t* RA = W2h + t5::W_BYTES;                         // g1 parts
  uint8_t* RB = RA + t5::ACT_BYTES;                        // dg2 parts
  uint8_t* SBf = RB + t5::ACT_BYTES;                       // dg2 tail chunk
  float* fl = reinterpret_cast<float*>(smraw + t5::OP_BYTES);
  float* WFs = fl + L::WF;
  float* WPs = fl + L::WP;
  float* bd1s = fl + L::BD1;
  float* bd2s = fl + L::BD2;
  float* bps = fl + L::BP;
  float* QX0 = fl + L::QX;
  float* QX1 = fl + L::QX1;
  float* S = fl + L::S;
  float* Pb = fl + L::PB;
  float* acc = fl + L::ACC;                                // acc[i] <-> Cfg::B_WF + i
  float* accWF = acc;
  float* accWP = acc + (CF::B_WP - CF::B_WF);
  float* accBD2 = acc + (CF::B_BD2 - CF::B_WF);
  float* accBD1 = acc + (CF::B_BD1 - CF::B_WF);
  float* accBP = acc + (CF::B_BP - CF::B_WF);
  float* scrW = fl + L::SCR;
  float* WFT = fl + L::WFT;
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;

  for (int i = tid; i < (int)(L::BYTES / 4); i += NTH) reinterpret_cast<uint32_t*>(smraw)[i] = 0u;
  if (tid == 0) {
    tc5::mbar_init(&bar_f, 1); tc5::mbar_init(&bar_d, 1); tc5::mbar_init(&bar_w, 1); tc5::mbar_init(&bar_w2, 1);
    tc5::mbar_init(&bar_in[0], 1); tc5::mbar_init(&bar_in[1], 1);
    tc5::mbar_init_fence();
    amax_sh = 0u;
  }
  if (w == 0) tc5::tmem_alloc512(&tmem_sh);
  __syncthreads();
  stage_vec(WFs, cm.wfold, 3 * H * 8);
  for (int i = tid; i < 3 * K * H; i += NTH) {   // WFT[tau][k][o] = wfold[tau][o][k]
    const int o = i % H, k = (i / H) % K, tau = i / (H * K);
    WFT[i] = cm.wfold[(tau * H + o) * 8 + k];
  }
  stage_conv_w(WPs, P.p[P_DPW], 2 * C, H, 1, 16, CF::PW_H);
  stage_vec(bd1s, P.p[P_D1B], H); stage_vec(bd2s, P.p[P_D2B], H); stage_vec(bps, P.p[P_DPB], 2 * C);
  // weight scale: max |W2| -> power of two
  {
    float m = 0.f;
    for (int i = tid; i < H * H * 3; i += NTH) m = fmaxf(m, fabsf(P.p[P_D2W][i]));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) atomicMax(&amax_sh, __float_as_uint(m));
  }
  __syncthreads();
  const float sW = tc5::pow2_scale(__uint_as_float(amax_sh));
  // W2[o][c][tau] -> fp16 parts at W2h[((tau*8 + c/8)*128 + part*64 + o)*16 + (c%8)*2]
  for (int i = tid; i < H * H * 3; i += NTH) {
    const int tau = i % 3, c = (i / 3) % H, o = i / (3 * H);
    __half h, m;
    tc5::split_f16(P.p[P_D2W][i] * sW, h, m);
    uint8_t* p = W2h + ((tau * 8 + (c >> 3)) * t5::WROWS + o) * 16 + (c & 7) * 2;
    *reinterpret_cast<__half*>(p) = h;
    *reinterpret_cast<__half*>(p + 64 * 16) = m;
  }
  __syncthreads();
  // static bound of g1 = relu(conv1'(q) + b1), q a probability vector: max_o sum_tau max_k |W'| + |b1|
  if (tid == 0) amax_sh = 0u;
  __syncthreads();
  if (tid < H) {
    float bnd = fabsf(bd1s[tid]);
    for (int tau = 0; tau < 3; ++tau) {
      float mk = 0.f;
      for (int k = 0; k < K; ++k) mk = fmaxf(mk, fabsf(WFs[(tau * H + tid) * 8 + k]));
      bnd += mk;
    }
    atomicMax(&amax_sh, __float_as_uint(bnd));
  }
  tc5::fence_async_smem();
  tc5::fence_before_sync();
  __syncthreads();
  tc5::fence_after_sync();
  const float sA = tc5::pow2_scale(__uint_as_float(amax_sh));
  const uint32_t tmem = tmem_sh;
  const float inv_f = 1.f / (sA * sW);
  const float aR = (float)(1.0 / cm.norms[0]);
  double s_nll = 0.0;
  uint32_t par_f = 0, par_d = 0, par_w = 0, par_w2 = 0;   // mbarrier phase parities
  const uint32_t lane_base = (uint32_t)(32 * (w & 3)) << 16;   // this warp's TMEM lane quadrant
  const int jrow = 32 * (w & 3) + lane;                    // TMEM lane == window column (time GEMMs) or stacked channel (dW2)
  const int cg = w >> 2;                                    // 16-column group this warp reads
  const uint32_t sW2 = tc5::smem_u32(W2h), sRA = tc5::smem_u32(RA), sRB = tc5::smem_u32(RB), sSB = tc5::smem_u32(SBf);
  constexpr uint32_t ID_F128 = tc5::idesc_f16(128, 0, 0), ID_F64 = tc5::idesc_f16(64, 0, 0);
  constexpr uint32_t ID_D64 = tc5::idesc_f16(64, 0, 1), ID_W128 = tc5::idesc_f16(128, 1, 1);
  PROF_DECL

  double* part = cm.acc_part + (int64_t)blockIdx.x * CF::ACC_ALL + CF::ACC_A;
  int ndone = 0;
  // persistent dW2 accumulators start at zero
  {
    float z[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) z[e] = 0.f;
#pragma unroll
    for (int tau = 0; tau < 3; ++tau) tc5::tmem_st16(tmem + lane_base + t5::TACC + 64 * tau + 16 * cg, z);
    tc5::tmem_st_wait();
    tc5::fence_before_sync();
  }
  __syncthreads();   // also: everybody has read amax_sh (sA) before the first tile resets it
  tc5::fence_after_sync();
  // fold the fp32 accumulators (shared memory: small layers; TMEM: dW2) into the CTA's double partial
  auto flush_all = [&]() {
    flush_acc(part + CF::B_WF, acc, L::NACC);
    // TMEM lane l holds row o = l % 64 of dW2 (h part of dg2 in lanes 0..63, m part in 64..127): two ordered passes
    for (int pass = 0; pass < 2; ++pass) {
      if ((jrow >> 6) == pass) {
#pragma unroll 1
        for (int tau = 0; tau < 3; ++tau) {
          float a[16];
          tc5::tmem_ld16(tmem + lane_base + t5::TACC + 64 * tau + 16 * cg, a);
          double* dst = part + CF::B_W2 + (tau * H + (jrow & 63)) * CF::PW_H + 16 * cg;
#pragma unroll
          for (int e = 0; e < 16; ++e) { atomicAdd(dst + e, (double)a[e]); a[e] = 0.f; }
          tc5::tmem_st16(tmem + lane_base + t5::TACC + 64 * tau + 16 * cg, a);
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
  uint32_t par_in[2] = {0u, 0u};
  // q and x windows of a tile: two TMA boxes issued by one thread (tensor-map TMA, zeros outside [0,T) by the bounds fill),
  // or cp.async when the tensors do not qualify.  q beyond the dependency cone of the valid steps (t >= Leff + 6) was never
  // written by kernel A: the cp.async path zero-fills it, the TMA path clears those columns after the box has landed.
  auto fetch = [&](const TileInfo& t_, int which) {
    float* dst = which ? QX1 : QX0;
    if (cm.use_tma) {
      if (tid == 0) {
        tc5::fence_async_smem();
        ::vhmm::tma::expect_tx(&bar_in[which], (uint32_t)(PT * 4 * (K + C)));
        ::vhmm::tma::load_3d(dst, &tm.q, t_.t0 - HL, 0, (int)t_.b, &bar_in[which]);
        ::vhmm::tma::load_3d(dst + 8 * PT, &tm.x, t_.t0 - HL, 0, (int)t_.b, &bar_in[which]);
      }
      return;
    }
    const int qmax = (t_.Leff + 6 < cm.T) ? t_.Leff + 6 : cm.T;
    prefetch_rows(dst, cm.q, K, t_.b, cm.T, t_.t0, qmax);
    prefetch_rows(dst + 8 * PT, cm.x, C, t_.b, cm.T, t_.t0, cm.T);
  };
  int64_t tile = next_tile<2>(cm, blockIdx.x, ti);
  if (tile >= 0) fetch(ti, 0);
  cp_async_commit();
  int64_t nxt = (tile >= 0) ? next_tile<2>(cm, tile + gridDim.x, tn) : -1;
  while (tile >= 0) {
    if (nxt >= 0) fetch(tn, buf ^ 1);
    cp_async_commit();
    const TilePre pre = tile_pre(cm, nxt >= 0 ? nxt + gridDim.x : cm.ntiles);   // consumed after the tile
    if (cm.use_tma) {
      tc5::mbar_wait(&bar_in[buf], par_in[buf]); par_in[buf] ^= 1u;
      const int qmax = (ti.Leff + 6 < cm.T) ? ti.Leff + 6 : cm.T;
      const int j0 = qmax - (ti.t0 - HL);                        // first window column beyond the written part of q
      if (qmax < cm.T && j0 < NC) {
        float* qd = buf ? QX1 : QX0;
        for (int i = tid; i < K * NC; i += NTH) {
          const int r = i / NC, j = i - r * NC;
          if (j >= j0) qd[r * PT + j] = 0.f;
        }
      }
    } else {
      cp_async_wait_prev();
    }
    PROF(13);
    const float* QX = buf ? QX1 : QX0;
    const int t0 = ti.t0, T = cm.T, Lb = ti.L;
    const int64_t b = ti.b;
    if (tid == 0) amax_sh = 0u;
    __syncthreads(); PROF(0);
    // ---- g1 = relu(conv3'(q) + b1) -> S (fp32), then its fp16 parts -> RA   (:83-85)
    {   // FFMA (skinny.cuh) straight into the fp16 parts: the scale of g1 is static and nothing reads g1 in fp32
      float g1v[SkConv<H>::CS];
      int r, c0;
      (void)sk_conv_taps_regs<K, H, K>(WFT, bd1s, QX, t0, T, g1v, r, c0);   // edge columns 0, NC-1 are zero
      sk_store_parts(g1v, sA, RA, t5::PLANE, t5::PART, r, c0);
    }
    tc5::fence_async_smem();
    tc5::fence_before_sync();
    __syncthreads(); PROF(1);
    // ---- g2 pre-activation on the tensor core: [0,64) of an accumulator set = g1_h*W_h, [64,128) = g1_h*W_m + g1_m*W_h
    if (w == 0) {
      tc5::fence_after_sync();
      const bool lead = tc5::elect_one();
#pragma unroll 1
      for (int tau = 0; tau < 3; ++tau)
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          const uint64_t ah = tc5::make_desc(sRA + (2 * ch) * t5::PLANE + tau * 16, t5::PLANE, 128);
          const uint64_t am = tc5::make_desc(sRA + t5::PART + (2 * ch) * t5::PLANE + tau * 16, t5::PLANE, 128);
          const uint64_t bw = tc5::make_desc(sW2 + (tau * 8 + 2 * ch) * t5::WPLANE, t5::WPLANE, 128);
          // two accumulator sets (even / odd K chunks: ACC_F and the idle WACC columns) halve the length of the
          // h*h chains, whose truncating accumulate is the largest error of this path
          const uint32_t base = tmem + (((tau * 4 + ch) & 1) ? t5::WACC : t5::ACC_F);
          if (lead) {
            tc5::mma_f16(base, ah, bw, ID_F128, (tau * 4 + ch) >= 2 ? 1u : 0u);
            tc5::mma_f16(base + 64, am, bw, ID_F64, 1u);
          }
        }
      if (lead) tc5::commit(&bar_f);
      __syncwarp();
    }
    tc5::mbar_wait(&bar_f, par_f); par_f ^= 1u;
    tc5::fence_after_sync();
    {   // epilogue: g2 = relu(acc/(sA sW) + b2) inside the sequence, 0 outside -> S[o][j]
      float vm[16], vs[16], wm[16], ws[16];
      tc5::tmem_ld16(tmem + lane_base + t5::ACC_F + 16 * cg, vm);
      tc5::tmem_ld16(tmem + lane_base + t5::ACC_F + 64 + 16 * cg, vs);
      tc5::tmem_ld16(tmem + lane_base + t5::WACC + 16 * cg, wm);
      tc5::tmem_ld16(tmem + lane_base + t5::WACC + 64 + 16 * cg, ws);
#pragma unroll
      for (int e = 0; e < 16; ++e) { vm[e] += wm[e]; vs[e] += ws[e]; }
      const int t = t0 - HL + jrow;
      const bool in = (t >= 0 && t < T);
#pragma unroll
      for (int e = 0; e < 16; ++e) {
        const int o = 16 * cg + e;
        S[o * PT + jrow] = in ? fmaxf((vm[e] + vs[e]) * inv_f + bd2s[o], 0.f) : 0.f;
      }
    }
    tc5::fence_before_sync();
    __syncthreads(); PROF(2);
    // ---- p = Wp g2 + bp, Gaussian NLL and its gradient (:87-90, :118-120)
    cols_op<16, 1, 1, H / 8, 1, false>(WPs, 0, CF::PW_H, S, EpiBiasRows{Pb, bps, 2 * C});   // (an FFMA form split over 4 thread groups measured 12 % slower)
    __syncthreads(); PROF(3);
    {   // all 512 threads: thread (column j, channel group): channels c = tid/NC, tid/NC + 4, ...
      const int j = tid & (NC - 1), t = t0 - HL + j;
      const bool inr = (j >= 2 && j < NC - 2 && t >= 0 && t < T);
      const bool mk = inr && t < Lb;
      const bool own = mk && j >= J0 && j < J0 + TT;
      float nll = 0.f;
      for (int c = tid / NC; c < C; c += NTH / NC) {
        const float mu = Pb[c * PT + j], lv = Pb[(C + c) * PT + j], xv = QX[(8 + c) * PT + j];
        const float ev = expf(lv), v = fmaxf(ev, 1e-8f), d = mu - xv, r = d * d / v;
        nll += 0.5f * (logf(6.283185307179586f * v) + r);
        Pb[c * PT + j] = mk ? aR * d / v : 0.f;
        Pb[(C + c) * PT + j] = (mk && ev >= 1e-8f) ? aR * 0.5f * (1.f - r) : 0.f;
      }
      if (own) s_nll += (double)nll;
    }
    __syncthreads(); PROF(4);
    // ---- to_params gradients, then dg2 = relu'(g2) * Wp^T dp in place over S
    static_assert(SkRows<2 * C, H>::FLOATS <= SkTaps<CF::K, CF::H>::FLOATS, "to_params gradient partials share the scratch of phase 9");
    sk_wgrad_rows<2 * C, H>(Pb, S, scrW);             // dWp[p][o] += sum_t dp[p][t] g2[o][t] as FFMA (skinny.cuh)
    row_sums<J0, TT / 4, PT, NW>(Pb, 2 * C, accBP);
    __syncthreads(); PROF(5);
    sk_wgrad_rows_reduce<2 * C, H>(scrW, accWP, CF::PW_H);
    {   // dg2 in place over S (mma.sync: the FFMA form, sk_rowsT_mask, is slower here) + the tile maximum of |dg2|
      float m = 0.f;
      cols_op<H, 2, 2, 2, 1, true>(WPs, 0, CF::PW_H, Pb, EpiReluMaskMax{S, &m});
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
      if (lane == 0) atomicMax(&amax_sh, __float_as_uint(m));
    }
    __syncthreads(); PROF(6);
    row_sums<J0, TT / 4, PT, NW>(S, H, accBD2);   // conv2 bias gradient
    const float sB = tc5::pow2_scale(__uint_as_float(amax_sh));
    for (int i = tid; i < t5::NPL * NC; i += NTH) {   // dg2 parts -> RB (+ the tail chunk with columns >= J0+TT zeroed)
      const int pl = i / NC, j = i - pl * NC;
      uint4 hh[1], mm[1];
      {
        float v8[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) v8[e] = S[(pl * 8 + e) * PT + j];
        tc5::split8(v8, tc5::f2_pack(sB, sB), hh[0], mm[0]);
      }
      uint8_t* p = RB + pl * t5::PLANE + (j + 1) * 16;
      *reinterpret_cast<uint4*>(p) = *reinterpret_cast<const uint4*>(hh);
      *reinterpret_cast<uint4*>(p + t5::PART) = *reinterpret_cast<const uint4*>(mm);
      if (j >= J0 + TT - 8 && j < J0 + TT) {
        uint8_t* sp = SBf + pl * t5::SB_PLANE + (j - (J0 + TT - 8)) * 16;
        *reinterpret_cast<uint4*>(sp) = *reinterpret_cast<const uint4*>(hh);
        *reinterpret_cast<uint4*>(sp + t5::NPL * t5::SB_PLANE) = *reinterpret_cast<const uint4*>(mm);
      }
    }
    tc5::fence_async_smem();
    tc5::fence_before_sync();
    __syncthreads(); PROF(7);
    // ---- tensor core: dg1 pre-mask (input gradient of conv2) and the weight gradient of tap 0
    // weight-gradient tile sum of one tap: acc = [dg2_h; dg2_m] x [g1_h; g1_m](shifted by tau) over the 120 owned
    // columns, into 128 TMEM columns at `acc_col`; completion on `bar`
    auto issue_wgrad = [&](int tau, bool lead, uint32_t acc_col, uint64_t* bar) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const uint64_t a = (i < 7) ? tc5::make_desc(sRB + (J0 + 1 + 16 * i) * 16, 128, t5::PLANE)
                                   : tc5::make_desc(sSB, 128, t5::SB_PLANE);
        const uint64_t bq = tc5::make_desc(sRA + (J0 + 16 * i + tau) * 16, 128, t5::PLANE);
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
        for (int ch = 0; ch < 4; ++ch) {   // dg1[t][c] += sum_o dg2[o][t - tau + 1] W2[o][c][tau], o in [16 ch, 16 ch + 16)
          const uint64_t ah = tc5::make_desc(sRB + (2 * ch) * t5::PLANE + (2 - tau) * 16, t5::PLANE, 128);
          const uint64_t am = tc5::make_desc(sRB + t5::PART + (2 * ch) * t5::PLANE + (2 - tau) * 16, t5::PLANE, 128);
          const uint64_t bh = tc5::make_desc(sW2 + (tau * 8 * t5::WROWS + 16 * ch) * 16, 128, t5::WPLANE);
          const uint64_t bm = tc5::make_desc(sW2 + (tau * 8 * t5::WROWS + 64 + 16 * ch) * 16, 128, t5::WPLANE);
          if (lead) {
            tc5::mma_f16(tmem + t5::ACC_F, ah, bh, ID_D64, (tau | ch) ? 1u : 0u);
            tc5::mma_f16(tmem + t5::ACC_F + 64, ah, bm, ID_D64, (tau | ch) ? 1u : 0u);
            tc5::mma_f16(tmem + t5::ACC_F + 64, am, bh, ID_D64, 1u);
          }
        }
      if (lead) tc5::commit(&bar_d);
      issue_wgrad(0, lead, t5::WACC, &bar_w);
      __syncwarp();
    }
    tc5::mbar_wait(&bar_d, par_d); par_d ^= 1u;
    tc5::fence_after_sync();
    {   // epilogue: dg1 = relu'(g1) * acc/(sB sW) -> S[c][j]   (g1 > 0 <=> its h part is non-zero)
      float vm[16], vs[16];
      tc5::tmem_ld16(tmem + lane_base + t5::ACC_F + 16 * cg, vm);
      tc5::tmem_ld16(tmem + lane_base + t5::ACC_F + 64 + 16 * cg, vs);
      const float inv_d = 1.f / (sB * sW);
      const uint4 g0 = *reinterpret_cast<const uint4*>(RA + (2 * cg) * t5::PLANE + (jrow + 1) * 16);
      const uint4 g1v = *reinterpret_cast<const uint4*>(RA + (2 * cg + 1) * t5::PLANE + (jrow + 1) * 16);
      const uint32_t gw[8] = {g0.x, g0.y, g0.z, g0.w, g1v.x, g1v.y, g1v.z, g1v.w};
#pragma unroll
      for (int e = 0; e < 16; ++e) {
        const uint32_t hbits = (gw[e >> 1] >> ((e & 1) * 16)) & 0x7fffu;   // |g1_h| as fp16 bits
        S[(16 * cg + e) * PT + jrow] = hbits ? (vm[e] + vs[e]) * inv_d : 0.f;
      }
    }
    tc5::fence_before_sync();
    __syncthreads(); PROF(8);
    if (w == 0) {   // the forward / input-gradient accumulator is free now: tap 1 runs there under the FFMA phase below
      tc5::fence_after_sync();
      issue_wgrad(1, tc5::elect_one(), t5::ACC_F, &bar_w2);
      __syncwarp();
    }
    // ---- folded conv1 gradients dW'[tau][o][k] += sum_t dg1[o][t] q[k][t+tau-1], bias, d loss / d q -> HBM:
    // register-blocked FFMA (skinny.cuh): slice partials -> scrW / Pb, one barrier, ordered reduction
    sk_wgrad_taps<K, H>(S, QX, scrW);
    sk_dgrad_taps<K, H, 4>(WFs, S, Pb);
    __syncthreads();
    sk_wgrad_taps_reduce<K, H>(scrW, accWF, accBD1);
    sk_dgrad_taps_store<K, 4>(Pb, cm.dq_dec + b * K * (int64_t)T, t0, T);
    PROF(9);
    // ---- conv2 weight gradient: drain each tap's tile sum into the persistent TMEM accumulators.  Tap 0 (WACC) and
    // tap 1 (ACC_F) were issued under the previous phases; tap 2 reuses WACC as soon as tap 0 is drained.
    {
      const float inv_w = 1.f / (sA * sB);
#pragma unroll 1
      for (int tau = 0; tau < 3; ++tau) {
        const uint32_t col = (tau == 1) ? t5::ACC_F : t5::WACC;
        if (tau == 1) { tc5::mbar_wait(&bar_w2, par_w2); par_w2 ^= 1u; }
        else          { tc5::mbar_wait(&bar_w, par_w); par_w ^= 1u; }
        tc5::fence_after_sync();
        float v0[16], v1[16], a[16];
        tc5::tmem_ld16(tmem + lane_base + col + 16 * cg, v0);        // x g1_h
        tc5::tmem_ld16(tmem + lane_base + col + 64 + 16 * cg, v1);   // x g1_m
        tc5::tmem_ld16(tmem + lane_base + t5::TACC + 64 * tau + 16 * cg, a);
#pragma unroll
        for (int e = 0; e < 16; ++e) a[e] += (v0[e] + v1[e]) * inv_w;
        tc5::tmem_st16(tmem + lane_base + t5::TACC + 64 * tau + 16 * cg, a);
        tc5::tmem_st_wait();
        tc5::fence_before_sync();
        if (tau != 1) __syncthreads();
        if (tau == 0 && w == 0) {
          tc5::fence_after_sync();
          issue_wgrad(2, tc5::elect_one(), t5::WACC, &bar_w);
          __syncwarp();
        }
      }
    }
    PROF(10);
    if (++ndone % FLUSH == 0) flush_all();
    PROF(11);
    tile = nxt; ti = tn; buf ^= 1;
    nxt = (tile >= 0) ? tile_resolve<2>(cm, pre, tn) : -1;
    PROF(12);
  }
  PROF_DUMP("decoder_tc5", 14);
  flush_all();
  const double v = block_sum_d(s_nll, red);
  if (tid == 0) cm.loss_part[(int64_t)blockIdx.x * LP + 0] = v;
  tc5::fence_before_sync();
  __syncthreads();
  if (w == 0) tc5::tmem_free512(tmem);
}
