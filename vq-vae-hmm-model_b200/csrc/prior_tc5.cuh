// prior_tc5.cuh -- transition prior of the wide model (K = 32: BASELINE.json configs[4]) in ONE kernel:
//   logits = transition_net.2(a)              (128 -> K*K = 1024 channels, tcgen05 GEMM as in conv_tc5s_kernel<1>)
//   log A  = row log-softmax over j            (VQ_VAE_HMM_fixed.py:57)
//   pair term  sum_t sum_ij q_{t-1}(i) q_t(j) log A_t(i,j)  and its gradients  (VQ_VAE_HMM_fixed.py:125-131)
// The layer-wise path wrote the logits (4.3 GB per 1024 sequences), normalised them in place (read + write), read them twice
// more in the loss kernel and wrote d loss / d logits: five passes over the largest tensor of the model.  Here a thread of the
// GEMM epilogue owns one (time step, row i) of the transition matrix -- 32 accumulator columns = the 32 targets j -- so the
// log-softmax, the loss term and the gradient are formed in registers and only d loss / d logits is written.
//   dM_t(i,j)    = aP q_{t-1}(i) (q_t(j) - A_t(i,j))                        aP = -beta / norms[1]
//   dq_row_{t-1}(i) = R_t(i) = sum_j q_t(j) log A_t(i,j)                    (row i of the transition INTO t)
//   dq_col_t(j)  = sum_i q_{t-1}(i) log A_t(i,j)                            (summed over the CTA's 8 output tiles in registers,
//                                                                            over the 4 column quarters through shared memory)
// for 1 <= t < L, zero elsewhere (pointwise.cuh:loss_seed_rows4_kernel is the unfused statement of the same sums).
// Two TMEM accumulator buffers (one h*h set + one small-term set each: chains of K/16 = 8 accumulations): the MMAs of output
// tile ct+1 run under the epilogue of tile ct.
#pragma once
#include "conv_tc5.cuh"

namespace vhmm {
namespace ct5 {

constexpr int PK = 32;                                   // K of the fused prior kernel
// d loss / d logits in OPERAND form: the two consumers (input gradient and weight gradient of transition_net.2) are GEMMs over
// fp16 (h, m) parts; written here once, in their shared-memory image, they are fetched with bulk copies and never converted
// (the fp32 tensor cost both consumers a conversion pass over 10^9 elements per 1024 sequences).  |dM| <= |aP| is a static
// bound, so the scale is known before the first element exists.
//   parts[b][time tile][chunk of 16 channels (64)][part h/m][plane of 8 channels (2)][row = step in tile (128)][8] fp16
constexpr int PP_CHUNK = 2 * 2 * TM * 16;                // 8 KB: one chunk = the A operand of one K step of the input gradient
constexpr int PP_TILE = (PK * PK / KC) * PP_CHUNK;       // 512 KB per (sequence, time tile)
constexpr int PQ_P = 132;                                // q window pitch: columns t0-4 .. t0+127
constexpr int PCOL_P = 33;                               // column-sum tile pitch
struct SmP {
  using S = SmS<1>;
  static constexpr int Q_OFF = S::TOTAL;                 // [32 states][132] fp32
  static constexpr int COL_OFF = Q_OFF + PK * PQ_P * 4;  // [128 steps][33] fp32
  static constexpr int TOTAL = COL_OFF + TM * PCOL_P * 4;
};

struct PriorArgs {
  const float* a;          // (B,Cin,T) hidden layer of the transition net
  const uint8_t* Wg;       // prepared forward weights of transition_net.2
  const float* bias;       // (K*K)
  const uint32_t* amax_x;
  const uint32_t* amax_w;
  const float* q;          // (B,K,T)
  const int64_t* lengths;  // (B)
  const double* norms;
  float beta;
  float* dM;               // (B,K*K,T) or null (forward only)
  uint8_t* dM_parts;       // or: d loss / d logits as fp16 operand parts (layout below), scaled by pow2_scale(|aP|)
  float* dq_row;           // (B,K,T) or null: un-scaled (the consumer multiplies by aP)
  float* dq_col;           // (B,K,T) or null
  double* partial;         // [grid][4]: slot 1 receives the CTA's pair-term sum
  uint32_t* bound_dM;      // receives |aP| (float bits) = max |dM| bound: the operand scale of the consumers
  int64_t B;
  int Cin, T, ntt;
};

__global__ void __launch_bounds__(NT, 1) prior_tc5_kernel(PriorArgs a) {
  using S = SmS<1>;
  constexpr int NWS = S::NWS, GK = S::GK;
  constexpr int NCT = PK * PK / TN;                      // 8 output tiles of 128 logits = 4 rows i each
  extern __shared__ __align__(1024) uint8_t sm[];
  __shared__ __align__(8) uint64_t landed[NWS], empty[NWS], done[2];
  __shared__ uint32_t tmem_sh;
  __shared__ double red[32];
  uint8_t* parts = sm + S::WREGION;
  float* qs = reinterpret_cast<float*>(sm + SmP::Q_OFF);
  float* colred = reinterpret_cast<float*>(sm + SmP::COL_OFF);
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int i = 0; i < NWS; ++i) { tc5::mbar_init(&landed[i], 1); tc5::mbar_init(&empty[i], 1); }
    tc5::mbar_init(&done[0], 1); tc5::mbar_init(&done[1], 1);
    tc5::mbar_init_fence();
  }
  if (w == 0) tc5::tmem_alloc512(&tmem_sh);
  tc5::fence_before_sync();
  __syncthreads();
  tc5::fence_after_sync();
  const uint32_t tmem = tmem_sh;
  const float sx = tc5::pow2_scale(__uint_as_float(*a.amax_x));
  const float sw = tc5::pow2_scale(__uint_as_float(*a.amax_w));
  const float inv = 1.f / (sx * sw);
  const f2_t sx2 = f2_pack(sx, sx);
  const float aP = (float)(-(double)a.beta / a.norms[1]);
  const float sM = tc5::pow2_scale(fabsf(aP));
  const f2_t sM2 = f2_pack(sM, sM);
  if (blockIdx.x == 0 && tid == 0 && a.bound_dM) *a.bound_dM = __float_as_uint(fabsf(aP));
  const int nkc = a.Cin / KC;
  const int ngr = (nkc + GK - 1) / GK;                   // weight groups per output tile
  const int T = a.T;
  constexpr uint32_t IDESC = tc5::idesc_f16(TN, 0, 0);
  uint32_t g_load = 0, g_issue = 0;                      // weight groups requested (warp 1) / consumed (warp 0) so far
  uint32_t nuse0 = 0, nuse1 = 0;                         // completed uses of done[0] / done[1]
  const int64_t nbt = a.B * a.ntt;
  double trans = 0.0;

  auto load_tile_w = [&](int ct) {                       // warp 1: one TMA bulk copy per weight group
    for (int gr = 0; gr < ngr; ++gr, ++g_load) {
      const uint32_t g = g_load;
      if (g >= (uint32_t)NWS) tc5::mbar_wait(&empty[g % NWS], ((g / NWS) - 1u) & 1u);
      if (lane == 0) {
        const int k0 = gr * GK, nk = (nkc - k0 < GK) ? nkc - k0 : GK;
        mbar_expect_tx(&landed[g % NWS], (uint32_t)(nk * S::WBYTES));
        bulk_g2s(sm + (g % NWS) * (GK * S::WBYTES), a.Wg + ((int64_t)ct * nkc + k0) * S::WBYTES, (uint32_t)(nk * S::WBYTES),
                 &landed[g % NWS]);
      }
      __syncwarp();
    }
  };
  auto issue_tile = [&](int ct) {                        // warp 0: the 3 * nkc MMAs of output tile ct into TMEM buffer ct & 1
    const uint32_t acc = tmem + 256u * (uint32_t)(ct & 1);
    for (int gr = 0; gr < ngr; ++gr, ++g_issue) {
      const uint32_t g = g_issue;
      tc5::mbar_wait(&landed[g % NWS], (g / NWS) & 1u);
      tc5::fence_after_sync();
      const bool lead = tc5::elect_one();
      const int k0 = gr * GK, nk = (nkc - k0 < GK) ? nkc - k0 : GK;
      for (int kk = 0; kk < nk; ++kk) {
        const int kc = k0 + kk;
        const uint32_t sA = tc5::smem_u32(parts + kc * ABYTES);
        const uint32_t sW = tc5::smem_u32(sm + (g % NWS) * (GK * S::WBYTES) + kk * S::WBYTES);
        const uint64_t ah = tc5::make_desc(sA + 16, PLANE, 128);
        const uint64_t am = tc5::make_desc(sA + APART + 16, PLANE, 128);
        const uint64_t bh = tc5::make_desc(sW, WPLANE, 128);
        const uint64_t bm = tc5::make_desc(sW + TN * 16, WPLANE, 128);
        if (lead) {
          tc5::mma_f16(acc, ah, bh, IDESC, kc >= 1 ? 1u : 0u);
          tc5::mma_f16(acc + 128u, ah, bm, IDESC, kc >= 1 ? 1u : 0u);
          tc5::mma_f16(acc + 128u, am, bh, IDESC, 1u);
        }
      }
      if (lead) {
        tc5::commit(&empty[g % NWS]);
        if (gr == ngr - 1) tc5::commit(&done[ct & 1]);
      }
      __syncwarp();
    }
  };

  for (int64_t bt = blockIdx.x; bt < nbt; bt += gridDim.x) {
    const int tt = (int)(bt % a.ntt);
    const int64_t b = bt / a.ntt;
    const int t0 = tt * TM;
    const int L = (int)min((int64_t)T, max((int64_t)0, a.lengths[b]));
    // ---- phase A: stage the q window and the input tile, convert the input to fp16 parts (the weight region is free)
    {
      const float* qb = a.q + b * (int64_t)PK * T;
      for (int i = tid; i < PK * (PQ_P / 4); i += NT) {
        const int j = i / (PQ_P / 4), c = (i - j * (PQ_P / 4)) * 4, t = t0 - 4 + c;
        const bool ok = (t >= 0 && t < T);
        cp16(qs + j * PQ_P + c, qb + (int64_t)j * T + (ok ? t : 0), ok ? 16 : 0);
      }
    }
    const float* xb = a.a + b * (int64_t)a.Cin * T;
    for (int c0 = 0; c0 < nkc; c0 += S::XROUND) {
      const int nch = (nkc - c0 < S::XROUND) ? nkc - c0 : S::XROUND;
      for (int i = tid; i < nch * KC * (XW / 4); i += NT) {
        const int c = i / (XW / 4), j = (i - c * (XW / 4)) * 4, t = t0 - 4 + j;
        const bool ok = (t >= 0 && t < T);
        cp16(reinterpret_cast<float*>(sm) + c * XP + j, xb + (int64_t)(c0 * KC + c) * T + (ok ? t : 0), ok ? 16 : 0);
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      __syncthreads();
      const float* xs = reinterpret_cast<const float*>(sm);
      for (int i = tid; i < nch * 2 * ROWS; i += NT) {
        const int ch = i / (2 * ROWS), rem = i - ch * (2 * ROWS), pl = rem / ROWS, r = rem - pl * ROWS;
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = xs[(ch * KC + pl * 8 + e) * XP + r + 3];
        uint4 h, m;
        split8(v, sx2, h, m);
        uint8_t* p = parts + (c0 + ch) * ABYTES + pl * PLANE + r * 16;
        *reinterpret_cast<uint4*>(p) = h;
        *reinterpret_cast<uint4*>(p + APART) = m;
      }
      tc5::fence_async_smem();
      __syncthreads();
    }
    // ---- phase B: the 8 output tiles, MMAs one tile ahead of the epilogue
    if (w == 1) { load_tile_w(0); load_tile_w(1); }
    if (w == 0) { tc5::fence_after_sync(); issue_tile(0); }
    const int tl = 32 * (w & 3) + lane;                  // row of the time tile = TMEM lane
    const int t = t0 + tl;
    const int cq = w >> 2;
    const bool tm = (t >= 1 && t < L);
    const uint32_t lane_base = (uint32_t)(32 * (w & 3)) << 16;
    float colacc[PK];
#pragma unroll
    for (int j = 0; j < PK; ++j) colacc[j] = 0.f;
    float tr = 0.f;
    for (int ct = 0; ct < NCT; ++ct) {
      if (w == 0 && ct + 1 < NCT) issue_tile(ct + 1);    // its TMEM buffer was read out before the barrier that ended tile ct-1
      if (w == 1 && ct + 2 < NCT) load_tile_w(ct + 2);
      {
        uint32_t& nu = (ct & 1) ? nuse1 : nuse0;
        tc5::mbar_wait(&done[ct & 1], nu & 1u);
        ++nu;
      }
      tc5::fence_after_sync();
      const int irow = ct * 4 + cq;                      // row i of the transition matrix
      float z[PK];
#pragma unroll
      for (int cgp = 0; cgp < 2; ++cgp) {
        float v[16], s[16];
        tc5::tmem_ld16(tmem + lane_base + 256u * (uint32_t)(ct & 1) + 32 * cq + 16 * cgp, v);
        tc5::tmem_ld16(tmem + lane_base + 256u * (uint32_t)(ct & 1) + 128u + 32 * cq + 16 * cgp, s);
#pragma unroll
        for (int e = 0; e < 16; ++e) z[16 * cgp + e] = (v[e] + s[e]) * inv + __ldg(a.bias + irow * PK + 16 * cgp + e);
      }
      tc5::fence_before_sync();
      if (t < T) {
        float mx = z[0];
#pragma unroll
        for (int j = 1; j < PK; ++j) mx = fmaxf(mx, z[j]);
        float ssum = 0.f;
#pragma unroll
        for (int j = 0; j < PK; ++j) { z[j] -= mx; ssum += __expf(z[j]); }
        // (exp is evaluated again below instead of keeping 32 more registers beside z and the column sums)
        const float lsum = logf(ssum), rs = 1.f / ssum;
        const float qprev = tm ? qs[irow * PQ_P + tl + 3] : 0.f;      // q_{t-1}(i): window column of t-1
        const float gq = aP * qprev;
        float R = 0.f;
        float* dMp = a.dM ? a.dM + ((b * PK + irow) * (int64_t)PK) * T + t : nullptr;
        // chunk of this thread's first 16 logits: (ct * 128 + cq * 32) / 16
        uint8_t* dPp = a.dM_parts ? a.dM_parts + (bt * (PK * PK / KC) + ct * 8 + cq * 2) * (int64_t)PP_CHUNK + tl * 16 : nullptr;
#pragma unroll
        for (int j8 = 0; j8 < PK / 8; ++j8) {
          float d8[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const int j = j8 * 8 + e;
            const float la = z[j] - lsum;
            const float qc = qs[j * PQ_P + tl + 4];
            R = fmaf(qc, la, R);
            colacc[j] = fmaf(qprev, la, colacc[j]);
            d8[e] = gq * (qc - __expf(z[j]) * rs);
            if (dMp) dMp[(int64_t)j * T] = d8[e];
          }
          if (dPp) {
            uint4 h, m;
            split8(d8, sM2, h, m);
            uint8_t* pp = dPp + (j8 >> 1) * PP_CHUNK + (j8 & 1) * (TM * 16);     // chunk, plane; the m part follows 2 planes on
            *reinterpret_cast<uint4*>(pp) = h;
            *reinterpret_cast<uint4*>(pp + 2 * TM * 16) = m;
          }
        }
        if (tm) tr = fmaf(qprev, R, tr);
        if (a.dq_row) {
          float* dr = a.dq_row + (b * PK + irow) * (int64_t)T;
          if (t >= 1) dr[t - 1] = tm ? R : 0.f;
          if (t == T - 1) dr[t] = 0.f;
        }
      }
      else if (a.dM_parts) {                             // rows of the tile beyond T: zeros (0 x anything stays 0 in the consumers)
        uint8_t* dPp = a.dM_parts + (bt * (PK * PK / KC) + ct * 8 + cq * 2) * (int64_t)PP_CHUNK + tl * 16;
#pragma unroll
        for (int j8 = 0; j8 < PK / 8; ++j8) {
          uint8_t* pp = dPp + (j8 >> 1) * PP_CHUNK + (j8 & 1) * (TM * 16);
          *reinterpret_cast<uint4*>(pp) = make_uint4(0u, 0u, 0u, 0u);
          *reinterpret_cast<uint4*>(pp + 2 * TM * 16) = make_uint4(0u, 0u, 0u, 0u);
        }
      }
      __syncthreads();                                   // TMEM buffer ct & 1 is free for tile ct + 2
    }
    trans += (double)tr;
    // column sums over the four quarters (fixed order), the last quarter stores
    for (int r = 0; r < 4; ++r) {
      if (cq == r) {
        float* row = colred + tl * PCOL_P;
        if (r == 0) {
#pragma unroll
          for (int j = 0; j < PK; ++j) row[j] = colacc[j];
        } else if (r < 3) {
#pragma unroll
          for (int j = 0; j < PK; ++j) row[j] += colacc[j];
        } else if (a.dq_col && t < T) {
          float* dc = a.dq_col + b * (int64_t)PK * T + t;
#pragma unroll
          for (int j = 0; j < PK; ++j) dc[(int64_t)j * T] = row[j] + colacc[j];
        }
      }
      __syncthreads();
    }
  }
  {
    const double s = block_sum<double>(trans, red);
    if (tid == 0) {
      a.partial[blockIdx.x * 4 + 0] = 0.0; a.partial[blockIdx.x * 4 + 1] = s;
      a.partial[blockIdx.x * 4 + 2] = 0.0; a.partial[blockIdx.x * 4 + 3] = 0.0;
    }
  }
  tc5::fence_before_sync();
  __syncthreads();
  if (w == 0) tc5::tmem_free512(tmem);
}


// ---------------------------------------------------------------------------------------------------------
// Input gradient of transition_net.2 from the operand-form d loss / d logits:  da[t, c] = mask(a) * sum_o dM[t, o] W[o, c].
// No conversion anywhere: warp 1 streams (8 KB of dM parts + the 8 KB weight chunk) per K step into a ring of stages with
// one TMA bulk copy each, warp 0 issues the three MMAs of the step, all 16 warps run the epilogue.  HBM-bound by design
// (the 4.3 GB of dM are read once at the tensor pipe's full appetite of 16 KB per 390 cycles).
// ---------------------------------------------------------------------------------------------------------
constexpr int DP_NST = 8;                                // stages of 16 KB
constexpr int DP_STAGE = PP_CHUNK + 2 * WPLANE;          // dM parts chunk | weight chunk (kw = 1)
constexpr int DP_SMEM = DP_NST * DP_STAGE;
struct DgradPartsArgs {
  const uint8_t* dM_parts;
  const uint8_t* Wg;       // prepared input-gradient view of transition_net.2 (Cin = K*K, Cout = TH <= 128: one output tile)
  const float* act;        // (B,Cout,T): ReLU mask source
  float* y;                // (B,Cout,T)
  const uint32_t* amax_x;  // the |aP| word
  const uint32_t* amax_w;
  uint32_t* amax_y;
  int64_t B;
  int Cout, T, ntt;
};

__global__ void __launch_bounds__(NT, 1) dgrad_parts_tc5_kernel(DgradPartsArgs a) {
  constexpr int NKC = PK * PK / KC;                      // 64 K steps
  extern __shared__ __align__(1024) uint8_t sm[];
  __shared__ __align__(8) uint64_t landed[DP_NST], empty[DP_NST], done;
  __shared__ uint32_t tmem_sh;
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int i = 0; i < DP_NST; ++i) { tc5::mbar_init(&landed[i], 1); tc5::mbar_init(&empty[i], 1); }
    tc5::mbar_init(&done, 1);
    tc5::mbar_init_fence();
  }
  if (w == 0) tc5::tmem_alloc512(&tmem_sh);
  tc5::fence_before_sync();
  __syncthreads();
  tc5::fence_after_sync();
  const uint32_t tmem = tmem_sh;
  const float sx = tc5::pow2_scale(__uint_as_float(*a.amax_x));
  const float sw = tc5::pow2_scale(__uint_as_float(*a.amax_w));
  const float inv = 1.f / (sx * sw);
  const int T = a.T;
  constexpr uint32_t IDESC = tc5::idesc_f16(TN, 0, 0);
  const int64_t nbt = a.B * a.ntt;
  uint32_t g = 0;                                        // K steps loaded (warp 1) / issued (warp 0) so far: same sequence in both
  uint32_t ntile = 0;
  int pre = 0;                                           // warp 1: K steps of the current tile already requested during the previous one

  for (int64_t bt = blockIdx.x; bt < nbt; bt += gridDim.x) {
    const int tt = (int)(bt % a.ntt);
    const int64_t b = bt / a.ntt;
    const int t0 = tt * TM;
    if (w == 1) {
      auto load_steps = [&](int64_t tile, int k0, int k1) {
        const uint8_t* src = a.dM_parts + tile * (int64_t)PP_TILE;
        for (int kc = k0; kc < k1; ++kc, ++g) {
          if (g >= (uint32_t)DP_NST) tc5::mbar_wait(&empty[g % DP_NST], ((g / DP_NST) - 1u) & 1u);
          if (lane == 0) {
            uint8_t* st = sm + (g % DP_NST) * DP_STAGE;
            mbar_expect_tx(&landed[g % DP_NST], (uint32_t)DP_STAGE);
            bulk_g2s(st, src + (int64_t)kc * PP_CHUNK, (uint32_t)PP_CHUNK, &landed[g % DP_NST]);
            bulk_g2s(st + PP_CHUNK, a.Wg + (int64_t)kc * (2 * WPLANE), (uint32_t)(2 * WPLANE), &landed[g % DP_NST]);
          }
          __syncwarp();
        }
      };
      load_steps(bt, pre, NKC);
      // the first steps of the NEXT tile go out before this warp joins the epilogue (their stages free up as the last MMAs
      // of this tile complete): the ring stays full across the tile boundary
      pre = 0;
      if (bt + gridDim.x < nbt) { load_steps(bt + gridDim.x, 0, DP_NST); pre = DP_NST; }
    } else if (w == 0) {
      for (int kc = 0; kc < NKC; ++kc, ++g) {
        tc5::mbar_wait(&landed[g % DP_NST], (g / DP_NST) & 1u);
        tc5::fence_after_sync();
        const bool lead = tc5::elect_one();
        const uint32_t sA = tc5::smem_u32(sm + (g % DP_NST) * DP_STAGE);
        const uint32_t sW = sA + PP_CHUNK;
        const uint64_t ah = tc5::make_desc(sA, TM * 16, 128);
        const uint64_t am = tc5::make_desc(sA + 2 * TM * 16, TM * 16, 128);
        const uint64_t bh = tc5::make_desc(sW, WPLANE, 128);
        const uint64_t bm = tc5::make_desc(sW + TN * 16, WPLANE, 128);
        if (lead) {
          tc5::mma_f16(tmem + HH0 + 128u * (uint32_t)(kc % 3), ah, bh, IDESC, kc >= 3 ? 1u : 0u);
          tc5::mma_f16(tmem + SMALL, ah, bm, IDESC, kc >= 1 ? 1u : 0u);
          tc5::mma_f16(tmem + SMALL, am, bh, IDESC, 1u);
          tc5::commit(&empty[g % DP_NST]);
          if (kc == NKC - 1) tc5::commit(&done);
        }
        __syncwarp();
      }
    } else {
      g += (uint32_t)NKC;
    }
    tc5::mbar_wait(&done, ntile & 1u);
    ++ntile;
    tc5::fence_after_sync();
    {   // epilogue: TMEM lane = time row; warp w reads lane quadrant w & 3, columns [32 (w >> 2), +32)
      const uint32_t lane_base = (uint32_t)(32 * (w & 3)) << 16;
      const int t = t0 + 32 * (w & 3) + lane;
      float ymax = 0.f;
#pragma unroll 1
      for (int cgp = 0; cgp < 2; ++cgp) {
        const int col = 32 * (w >> 2) + 16 * cgp;
        float v[16], s[16], v1[16], v2[16];
        tc5::tmem_ld16(tmem + lane_base + SMALL + col, s);
        tc5::tmem_ld16(tmem + lane_base + HH0 + col, v);
        tc5::tmem_ld16(tmem + lane_base + HH0 + 128 + col, v1);
        tc5::tmem_ld16(tmem + lane_base + HH0 + 256 + col, v2);
#pragma unroll
        for (int e = 0; e < 16; ++e) v[e] += v1[e] + v2[e];
        if (t < T && col < a.Cout) {
          const int64_t off0 = (b * a.Cout + col) * (int64_t)T + t;
          float mk[16];
#pragma unroll
          for (int e = 0; e < 16; ++e) mk[e] = a.act ? __ldg(a.act + off0 + (int64_t)e * T) : 1.f;
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            const float r = mk[e] > 0.f ? (v[e] + s[e]) * inv : 0.f;
            a.y[off0 + (int64_t)e * T] = r;
            ymax = fmaxf(ymax, fabsf(r));
          }
        }
      }
      if (a.amax_y) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) ymax = fmaxf(ymax, __shfl_xor_sync(0xffffffffu, ymax, o));
        if (lane == 0 && ymax > 0.f) atomicMax(a.amax_y, __float_as_uint(ymax));
      }
    }
    tc5::fence_before_sync();
    __syncthreads();                                     // the accumulators are free for the next tile
    tc5::fence_after_sync();
  }
  tc5::fence_before_sync();
  __syncthreads();
  if (w == 0) tc5::tmem_free512(tmem);
}

// bias gradient of transition_net.2 from the operand-form tensor: partial[s][o] = sum over the tiles of split s and the
// rows of (h + m) / scale.  grid (64 chunks, nsplit), 256 threads: thread = (plane, row); 8 channels each.
__global__ void __launch_bounds__(256) bias_parts_kernel(const uint8_t* __restrict__ parts, int64_t ntiles, int nsplit,
                                                         const uint32_t* __restrict__ amax, float* __restrict__ partial) {
  __shared__ float red[256 * 8];
  const int kc = blockIdx.x, s = blockIdx.y, tid = threadIdx.x;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int64_t it = s; it < ntiles; it += nsplit) {
    const uint4* src = reinterpret_cast<const uint4*>(parts + (it * (PK * PK / KC) + kc) * (int64_t)PP_CHUNK);
    const uint4 h = src[tid], m = src[256 + tid];        // [part][plane][row]: tid = plane * 128 + row
    const __half2* hh = reinterpret_cast<const __half2*>(&h);
    const __half2* mm = reinterpret_cast<const __half2*>(&m);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 a2 = __half22float2(hh[e]), b2 = __half22float2(mm[e]);
      acc[2 * e] += a2.x + b2.x; acc[2 * e + 1] += a2.y + b2.y;
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) red[e * 256 + tid] = acc[e];
  __syncthreads();
  if (tid < 16) {                                        // channel tid of the chunk: plane tid / 8, element tid % 8
    const float* r = red + (tid & 7) * 256 + (tid >> 3) * 128;
    float sum = 0.f;
    for (int i = 0; i < 128; ++i) sum += r[i];
    const float inv = 1.f / tc5::pow2_scale(__uint_as_float(*amax));
    partial[(int64_t)s * (PK * PK) + kc * KC + tid] = sum * inv;
  }
}

}  // namespace ct5

// fused prior: usable when K == 32, the transition hidden width is a multiple of 16 up to 128, T % 4 == 0, 16-byte aligned tensors
static inline bool prior_tc5_ok(int K, int TH, int64_t T, const void* a, const void* q, const void* dM) {
  return K == ct5::PK && TH % ct5::KC == 0 && TH / ct5::KC <= ct5::SMAX_KC && (T & 3) == 0 && T >= 4 &&
         (((uintptr_t)a | (uintptr_t)q | (uintptr_t)dM) & 15) == 0;
}
static inline int64_t prior_tc5_grid(int64_t B, int64_t T) {
  int64_t grid = device_sm_count();
  const int64_t nbt = B * cdiv(T, ct5::TM);
  return grid < nbt ? grid : nbt;
}
// partial: 4 * prior_tc5_grid doubles; amax_x_known: |a| max word or null (measured into W.amax_x)
static inline int launch_prior_tc5(const float* a, const Tc5Weights& W, const float* bias, const float* q, const int64_t* lengths,
                                   const double* norms, float beta, int64_t B, int TH, int64_t T, float* dM, float* dq_row,
                                   float* dq_col, double* partial, uint32_t* bound_dM, cudaStream_t st,
                                   const uint32_t* amax_x_known = nullptr, bool dM_as_parts = false) {
  if (B <= 0 || T <= 0) return 0;
  const uint32_t* ax = amax_x_known;
  if (!ax) {
    VHMM_CUDA_CHECK(cudaMemsetAsync(W.amax_x, 0, 4, st));
    int rc = launch_absmax(a, B * (int64_t)TH * T, W.amax_x, st);
    if (rc) return rc;
    ax = W.amax_x;
  }
  ct5::PriorArgs p{};
  p.a = a; p.Wg = W.Wg; p.bias = bias; p.amax_x = ax; p.amax_w = W.amax_w; p.q = q; p.lengths = lengths; p.norms = norms;
  p.beta = beta; p.dM = dM_as_parts ? nullptr : dM; p.dM_parts = dM_as_parts ? reinterpret_cast<uint8_t*>(dM) : nullptr; p.dq_row = dq_row; p.dq_col = dq_col; p.partial = partial; p.bound_dM = bound_dM;
  p.B = B; p.Cin = TH; p.T = (int)T; p.ntt = (int)cdiv(T, ct5::TM);
  static PerDeviceOnce once;
  if (once.pending()) {
    VHMM_CUDA_CHECK(cudaFuncSetAttribute(ct5::prior_tc5_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ct5::SmP::TOTAL));
    once.mark();
  }
  ct5::prior_tc5_kernel<<<(unsigned)prior_tc5_grid(B, T), ct5::NT, ct5::SmP::TOTAL, st>>>(p);
  VHMM_LAUNCH_CHECK();
  return 0;
}

// da = mask(act) * (dM W) from the operand-form dM (TH == 128: one output tile); amax_dM: the |aP| word
static inline int launch_dgrad_parts_tc5(const uint8_t* dM_parts, const Tc5Weights& Wd, const float* act, float* y, int64_t B, int TH,
                                         int64_t T, const uint32_t* amax_dM, uint32_t* amax_y, cudaStream_t st) {
  if (B <= 0 || T <= 0) return 0;
  ct5::DgradPartsArgs a{};
  a.dM_parts = dM_parts; a.Wg = Wd.Wg; a.act = act; a.y = y; a.amax_x = amax_dM; a.amax_w = Wd.amax_w; a.amax_y = amax_y;
  a.B = B; a.Cout = TH; a.T = (int)T; a.ntt = (int)cdiv(T, ct5::TM);
  static PerDeviceOnce once;
  if (once.pending()) {
    VHMM_CUDA_CHECK(cudaFuncSetAttribute(ct5::dgrad_parts_tc5_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, ct5::DP_SMEM));
    once.mark();
  }
  ct5::dgrad_parts_tc5_kernel<<<(unsigned)prior_tc5_grid(B, T), ct5::NT, ct5::DP_SMEM, st>>>(a);
  VHMM_LAUNCH_CHECK();
  return 0;
}
// gB[o] += sum_{b,t} dM[b,o,t] from the operand-form dM; bpart: 32 * K*K floats
static inline int launch_bias_parts(const uint8_t* dM_parts, int64_t B, int64_t T, const uint32_t* amax_dM, float* bpart, float* gB,
                                    cudaStream_t st) {
  const int64_t ntiles = B * cdiv(T, ct5::TM);
  const int ns = ntiles < 32 ? (int)ntiles : 32;
  constexpr int K2 = ct5::PK * ct5::PK;
  ct5::bias_parts_kernel<<<dim3(K2 / ct5::KC, (unsigned)ns), 256, 0, st>>>(dM_parts, ntiles, ns, amax_dM, bpart);
  VHMM_LAUNCH_CHECK();
  reduce_partials_kernel<<<(unsigned)cdiv(K2, 256), 256, 0, st>>>(bpart, K2, ns, K2, gB);
  VHMM_LAUNCH_CHECK();
  return 0;
}

}  // namespace vhmm
