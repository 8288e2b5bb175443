// conv_tc5.cuh -- layer kernels on the 5th-generation tensor core (tcgen05.mma, TMEM accumulators) for the
// LAYER-WISE path at widths where the convolutions are compute-bound GEMMs (north_star: "tensor cores ... at
// hidden_dim >= 128"; BASELINE.json configs[4]: hidden_dim 256, hidden_dim2 128, K 32 -> K*K = 1024 transition
// logits).  Every layer of the model is a 1-D cross-correlation of width 3 (pad 1) or 1 over (B,C,T) tensors
// (VQ_VAE_HMM_fixed.py:34-36, :53-57, :77-79), i.e. per tap a GEMM  Y[t, o] += X[t + tau - pad, c] W[o, c, tau]:
//
//   conv_tc5_kernel   forward AND input gradient (the latter through the transposed / tap-flipped weight view, as in
//                     conv_generic.cuh):  M = 128 time steps, N = 128 output channels, K = Cin in chunks of 16
//   wgrad_tc5_kernel  weight gradient  dW[o, c, tau] = sum_{b,t} dY[b,o,t] X[b,c,t+tau-pad]:
//                     M = 64 out channels x 2 parts, N = 64 in channels x 2 parts, K = time
//
// Arithmetic: the fp32-parity scheme of the fused tile kernels (decoder_tc5.cuh): every fp32 operand is scaled by a
// power of two and split into two fp16 parts, x*s = h + m; a product is h*h + h*m + m*h with fp32 accumulation in
// TMEM.  The tensor core's accumulate truncates toward zero, so the h*h terms are spread round-robin over THREE
// accumulator sets (chains of K/48 instead of K/16 accumulations) and the small terms go to a fourth; the sets meet in
// fp32 FADDs in the epilogue.  Scales: one power of two per TENSOR (device-side abs-max, absmax_kernel) -- fp16 is a
// floating format, the scale only has to keep the largest element below 2^14.
//
// Operand layouts are the no-swizzle canonical ones of tc5.cuh: activations act[part][ch/8][row][8] (row = time,
// K-major; a conv tap is a +16-byte shift of the descriptor start), weights W[tau][ch/8][part*128 + o][8], prepared
// ONCE per call in global memory in exactly the shared-memory image of a (128-out-channel, 16-in-channel) chunk, so
// that a chunk is one contiguous block fetched with 16-byte cp.async.  Four-stage cp.async pipeline over the K
// chunks; the fp32 -> fp16-parts conversion of chunk k+1 runs under the MMAs of chunk k.
#pragma once
#include "conv_generic.cuh"
#include "tc5.cuh"
#include "tma.cuh"

namespace vhmm {
namespace ct5 {

constexpr int NT = 512;                 // threads per CTA (16 warps)
constexpr int TM = 128;                 // time steps per tile (MMA M)
constexpr int TN = 128;                 // output channels per tile (MMA N)
constexpr int KC = 16;                  // input channels per chunk (one MMA K step = 2 planes of 8)
constexpr int ROWS = 130;               // rows of an activation plane: row r <-> t = t0 - 1 + r
constexpr int PLANE = ROWS * 16;        // bytes of one 8-channel plane
constexpr int APART = 2 * PLANE;        // one part (h or m) of a chunk
constexpr int ABYTES = 2 * APART;       // h and m
constexpr int WPLANE = 2 * TN * 16;     // rows = part*128 + o
constexpr int XW = 136;                 // staged fp32 columns: t0-4 .. t0+131 (16-byte aligned start)
constexpr int XP = 140;                 // pitch (floats)
constexpr int XBYTES = KC * XP * 4;
constexpr int STAGES = 6;                // load stages of the streaming kernel: chunks g+1 .. g+3 in flight
constexpr int NPB = 3;                   // operand-parts buffers: chunk g+2 is converted while chunks g and g+1 feed the MMAs
constexpr uint32_t HH0 = 0, SMALL = 384;   // TMEM columns: three h*h sets at 0 / 128 / 256, the small terms at 384

template <int KW>
struct Sm {
  static constexpr int WBYTES = KW * 2 * WPLANE;          // weights of one chunk: [tau][plane][256 rows][16 B]
  static constexpr int STAGE = WBYTES + XBYTES;
  static constexpr int TOTAL = STAGES * STAGE + NPB * ABYTES;
};

// |x| max of a tensor -> *out (float bits; non-negative floats order like unsigned integers).  *out zeroed by the host.
__global__ void __launch_bounds__(256) absmax_kernel(const float* __restrict__ x, int64_t n, uint32_t* __restrict__ out) {
  float m = 0.f;
  const int64_t n4 = n >> 2;
  const float4* x4 = reinterpret_cast<const float4*>(x);
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += (int64_t)gridDim.x * 256) {
    const float4 v = x4[i];
    m = fmaxf(fmaxf(m, fmaxf(fabsf(v.x), fabsf(v.y))), fmaxf(fabsf(v.z), fabsf(v.w)));
  }
  for (int64_t i = (n4 << 2) + (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) m = fmaxf(m, fabsf(x[i]));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0 && m > 0.f) atomicMax(out, __float_as_uint(m));
}

// W(o,c,tau) of a weight view -> fp16 parts in the chunk image Wg[ct][kc][tau][plane][part*128 + o%128][c%8]
__global__ void __launch_bounds__(256) prep_weights_kernel(WView wv, int Cin, int Cout, int KW, const uint32_t* __restrict__ amax,
                                                           uint8_t* __restrict__ Wg) {
  const int64_t n = (int64_t)Cout * Cin * KW;
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  const int tau = (int)(i % KW), c = (int)((i / KW) % Cin), o = (int)(i / ((int64_t)KW * Cin));
  const float s = tc5::pow2_scale(__uint_as_float(*amax));
  const float v = wv.w[wv.base + (int64_t)o * wv.so + (int64_t)c * wv.sc + (int64_t)tau * wv.st] * s;
  __half h, m;
  tc5::split_f16(v, h, m);
  const int nkc = Cin / KC;
  const int ct = o / TN, ol = o % TN, kc = c / KC, pl = (c % KC) / 8, e = c % 8;
  uint8_t* p = Wg + ((((int64_t)(ct * nkc + kc) * KW + tau) * 2 + pl) * (2 * TN) + ol) * 16 + e * 2;
  *reinterpret_cast<__half*>(p) = h;
  *reinterpret_cast<__half*>(p + TN * 16) = m;
}

__device__ __forceinline__ void cp16(void* smem, const void* g, int bytes) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(d), "l"(g), "r"(bytes) : "memory");
}

using tc5::f2_t; using tc5::f2_pack; using tc5::f2_unpack; using tc5::f2_add; using tc5::f2_sub; using tc5::f2_mul; using tc5::split8;

struct ConvArgs {
  const float* x;          // (B,Cin,T)
  const uint8_t* Wg;       // prepared weights
  const float* bias;       // (Cout) or null
  const float* act;        // (B,Cout,T) ReLU mask source for EPI_MASK, or null
  float* y;                // (B,Cout,T)
  const uint32_t* amax_x;  // device: |x| max (float bits)
  const uint32_t* amax_w;  // device: |W| max
  uint32_t* amax_y;        // device, or null: receives |y| max (the next layer's operand scale comes for free)
  int64_t B;
  int Cin, Cout, T, epi;
  int ntt, nct;            // time tiles per sequence, out-channel tiles
  int64_t nitems;
  int use_tma;             // streaming kernel: the input chunk arrives as ONE TMA box (pitch XW) instead of 16-byte cp.async
};

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(tc5::smem_u32(bar)) : "memory");
}
// the barrier receives one arrival when all cp.async issued so far by this thread have landed (count pre-charged: .noinc)
__device__ __forceinline__ void cp_async_arrive(uint64_t* bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" :: "r"(tc5::smem_u32(bar)) : "memory");
}

// one arrival + `bytes` of expected transaction count (completed by the bulk copies below)
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(tc5::smem_u32(bar)), "r"(bytes) : "memory");
}
// TMA bulk copy (1-D, no tensor map): global -> shared, completion reported to the mbarrier as transaction bytes.  A weight
// chunk is ONE contiguous block in the prepared image, so one instruction by one thread replaces 1536 16-byte cp.async.
__device__ __forceinline__ void bulk_g2s(void* smem, const void* g, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"(tc5::smem_u32(smem)), "l"(g), "r"(bytes), "r"(tc5::smem_u32(bar)) : "memory");
}

// y = epi(bias + conv_KW(x, W)); requires Cin % 16 == 0, Cout % 128 == 0, T % 4 == 0, x 16-byte aligned.
// Warp-specialised main loop over the K chunks (global chunk counter g, continuing across the CTA's work items):
//   warps 1..15  load chunk g+2 (cp.async into stage (g+2) % 4, completion -> landed[stage]) and convert chunk g from its
//                fp32 staging to the fp16 parts buffer g % 3, then arrive on full[g % 3]
//   warp 0       waits for full[g % 3], issues the 3 KW MMAs of chunk g, commits them to empty[g % 3]
// (Two parts buffers left the issuer waiting ~600 cycles per chunk: from "MMAs of chunk g done" to "chunk g+2 converted"
//  takes ~2 k cycles, more than the 1.2-1.4 k of MMA time of chunk g+1.  With three, the conversion has two chunk times.)
// A parts buffer / stage is rewritten only after empty[] of the chunk that last read it; no CTA-wide barrier inside the
// loop (the first build had two per chunk: tensor pipe 23 % active, 0.4 eligible warps per cycle).
// (Measured and dropped, twice: A_h x [W_h | W_m] as one N = 256 MMA -- same kernel time as the three N = 128 MMAs.)
// (First build: every loader thread arrived on landed[] (cp.async) and on full[] -- ~1000 mbarrier arrivals per chunk, and
// 2.0 k cycles per chunk against 1.17 k of MMA time.  Now one TMA box per chunk and one arrival per warp.)
template <int KW>
__global__ void __launch_bounds__(NT, 1) conv_tc5_kernel(ConvArgs a, const __grid_constant__ CUtensorMap tm_x) {
  constexpr int PAD = KW / 2;
  constexpr int NLD = NT - 32;                       // loader / converter threads
  using S = Sm<KW>;
  extern __shared__ __align__(1024) uint8_t sm[];
  __shared__ __align__(8) uint64_t landed[STAGES], full[NPB], empty[NPB];
  __shared__ uint32_t tmem_sh;
  uint8_t* parts = sm + STAGES * S::STAGE;
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
  for (int i = tid; i < NPB * ABYTES / 4; i += NT) reinterpret_cast<uint32_t*>(parts)[i] = 0u;
  if (tid == 0) {
    // cp.async build: NLD cp.async arrivals (x) + the weight block's expect_tx; TMA build: the one expect_tx
    for (int i = 0; i < STAGES; ++i) tc5::mbar_init(&landed[i], a.use_tma ? 1 : NLD + 1);
    for (int i = 0; i < NPB; ++i) { tc5::mbar_init(&full[i], NLD / 32); tc5::mbar_init(&empty[i], 1); }
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
  const int nkc = a.Cin / KC;
  const int T = a.T;
  constexpr uint32_t IDESC = tc5::idesc_f16(TN, 0, 0);
  uint32_t g0 = 0;                                   // global index of this item's chunk 0
#ifdef CT5_PROF
  long long pq[4] = {0, 0, 0, 0}, pql = clock64();
#define CT5Q(i) do { if (tid == 0) { long long n_ = clock64(); pq[i] += n_ - pql; pql = n_; } } while (0)
#else
#define CT5Q(i)
#endif

  for (int64_t item = blockIdx.x; item < a.nitems; item += gridDim.x, g0 += (uint32_t)nkc) {
    const int ct = (int)(item % a.nct);
    const int64_t bt = item / a.nct;
    const int tt = (int)(bt % a.ntt);
    const int64_t b = bt / a.ntt;
    const int t0 = tt * TM;
    if (w == 0) {
      // ---- MMA issuer -------------------------------------------------------------------------------------
      for (int kc = 0; kc < nkc; ++kc) {
        const uint32_t g = g0 + (uint32_t)kc;
        CT5Q(3);
        tc5::mbar_wait(&full[g % NPB], (g / NPB) & 1u);
        CT5Q(kc == 0 ? 0 : 1);
        tc5::fence_after_sync();
        const bool lead = tc5::elect_one();
        const uint32_t sA = tc5::smem_u32(parts + (g % NPB) * ABYTES);
        const uint32_t sW = tc5::smem_u32(sm + (g % STAGES) * S::STAGE);
#pragma unroll
        for (int tau = 0; tau < KW; ++tau) {
          const int step = kc * KW + tau;
          const uint64_t ah = tc5::make_desc(sA + (tau - PAD + 1) * 16, PLANE, 128);
          const uint64_t am = tc5::make_desc(sA + APART + (tau - PAD + 1) * 16, PLANE, 128);
          const uint64_t bh = tc5::make_desc(sW + tau * 2 * WPLANE, WPLANE, 128);
          const uint64_t bm = tc5::make_desc(sW + tau * 2 * WPLANE + TN * 16, WPLANE, 128);
          if (lead) {
            tc5::mma_f16(tmem + HH0 + 128u * (uint32_t)(step % 3), ah, bh, IDESC, step >= 3 ? 1u : 0u);
            tc5::mma_f16(tmem + SMALL, ah, bm, IDESC, step >= 1 ? 1u : 0u);
            tc5::mma_f16(tmem + SMALL, am, bh, IDESC, 1u);
          }
        }
        if (lead) tc5::commit(&empty[g % NPB]);
        __syncwarp();
        CT5Q(2);
      }
    } else {
      // ---- loaders / converters ---------------------------------------------------------------------------
      const int lt = tid - 32;
      const float* xb = a.x + b * (int64_t)a.Cin * T;
      const uint8_t* wct = a.Wg + (int64_t)ct * nkc * S::WBYTES;
      auto issue_loads_at = [&](const float* xb, const uint8_t* wct, int64_t b, int t0, uint32_t g, int kc) {
        uint8_t* st = sm + (g % STAGES) * S::STAGE;
        const uint8_t* wsrc = wct + (int64_t)kc * S::WBYTES;
        if (a.use_tma) {                                 // weight chunk (bulk copy) + input chunk (tensor box, zero fill outside [0,T))
          if (lt == 0) {
            mbar_expect_tx(&landed[g % STAGES], (uint32_t)(S::WBYTES + KC * XW * 4));
            bulk_g2s(st, wsrc, (uint32_t)S::WBYTES, &landed[g % STAGES]);
            ::vhmm::tma::load_3d(st + S::WBYTES, &tm_x, t0 - 4, kc * KC, (int)b, &landed[g % STAGES]);
          }
          return;
        }
        if (lt == 0) {                                   // the weight chunk: one TMA bulk copy
          mbar_expect_tx(&landed[g % STAGES], (uint32_t)S::WBYTES);
          bulk_g2s(st, wsrc, (uint32_t)S::WBYTES, &landed[g % STAGES]);
        }
        float* xs = reinterpret_cast<float*>(st + S::WBYTES);
        for (int i = lt; i < KC * (XW / 4); i += NLD) {
          const int c = i / (XW / 4), j = (i - c * (XW / 4)) * 4, t = t0 - 4 + j;
          const bool ok = (t >= 0 && t < T);             // T % 4 == 0: a granule is inside or outside as a whole
          cp16(xs + c * XP + j, xb + (int64_t)(kc * KC + c) * T + (ok ? t : 0), ok ? 16 : 0);
        }
        cp_async_arrive(&landed[g % STAGES]);
      };
      auto issue_loads = [&](int kc) { issue_loads_at(xb, wct, b, t0, g0 + (uint32_t)kc, kc); };
      // the first three chunks of every item but the CTA's first were requested at the end of the previous item
      if (item == (int64_t)blockIdx.x) {
        issue_loads(0);
        if (nkc > 1) issue_loads(1);
        if (nkc > 2) issue_loads(2);
      }
      for (int kc = 0; kc < nkc; ++kc) {
        const uint32_t g = g0 + (uint32_t)kc;
        // parts buffer g % 3 and stage (g + 3) % 6 were last read by the MMAs of chunk g - 3: ~70 KB must be in flight to
        // cover the L2 latency at the tensor pipe's appetite of 24 KB per ~1000 cycles
        if (g >= (uint32_t)NPB) tc5::mbar_wait(&empty[g % NPB], ((g - NPB) / NPB) & 1u);
        if (kc + 3 < nkc) issue_loads(kc + 3);
        tc5::mbar_wait(&landed[g % STAGES], (g / STAGES) & 1u);
        const float* xs = reinterpret_cast<const float*>(sm + (g % STAGES) * S::STAGE + S::WBYTES);
        uint8_t* pb = parts + (g % NPB) * ABYTES;
        const int xp = a.use_tma ? XW : XP;
        for (int i = lt; i < 2 * ROWS; i += NLD) {
          const int pl = i / ROWS, r = i - pl * ROWS;
          float v[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) v[e] = xs[(pl * 8 + e) * xp + r + 3];
          uint4 h, m;
          split8(v, sx2, h, m);
          uint8_t* p = pb + pl * PLANE + r * 16;
          *reinterpret_cast<uint4*>(p) = h;
          *reinterpret_cast<uint4*>(p + APART) = m;
        }
        tc5::fence_async_smem();          // this thread's generic-proxy writes (parts) -> async proxy
        __syncwarp();
        if (lane == 0) mbar_arrive(&full[g % NPB]);    // one arrival per loader warp
      }
      // drain: the last two chunks' MMAs (idempotent waits: a completed phase answers at once)
      for (int kc = (nkc >= NPB ? nkc - NPB : 0); kc < nkc; ++kc) {
        const uint32_t g = g0 + (uint32_t)kc;
        tc5::mbar_wait(&empty[g % NPB], (g / NPB) & 1u);
      }
      // every stage is free: request the next item's first chunks now, so that they land under the epilogue
      const int64_t nitem = item + gridDim.x;
      if (nitem < a.nitems) {
        const int nct = (int)(nitem % a.nct);
        const int64_t nbt = nitem / a.nct;
        const int64_t nb = nbt / a.ntt;
        const int nt0 = (int)(nbt % a.ntt) * TM;
        const float* nxb = a.x + nb * (int64_t)a.Cin * T;
        const uint8_t* nwct = a.Wg + (int64_t)nct * nkc * S::WBYTES;
        for (int kc = 0; kc < 3 && kc < nkc; ++kc) issue_loads_at(nxb, nwct, nb, nt0, g0 + (uint32_t)(nkc + kc), kc);
      }
    }
    __syncthreads();                       // every MMA of the item has completed (the converters saw its commits)
    tc5::fence_after_sync();
    {   // epilogue: TMEM lane = time row; warp w reads lane quadrant w & 3, columns [32 (w >> 2), +32)
      const uint32_t lane_base = (uint32_t)(32 * (w & 3)) << 16;
      const int t = t0 + 32 * (w & 3) + lane;
      const int nsets = (nkc * KW < 3) ? nkc * KW : 3;
      float ymax = 0.f;
#pragma unroll 1
      for (int cgp = 0; cgp < 2; ++cgp) {
        const int col = 32 * (w >> 2) + 16 * cgp;
        float v[16], s[16];
        tc5::tmem_ld16(tmem + lane_base + SMALL + col, s);
        tc5::tmem_ld16(tmem + lane_base + HH0 + col, v);
        if (nsets > 1) {
          float v1[16];
          tc5::tmem_ld16(tmem + lane_base + HH0 + 128 + col, v1);
#pragma unroll
          for (int e = 0; e < 16; ++e) v[e] += v1[e];
        }
        if (nsets > 2) {
          float v2[16];
          tc5::tmem_ld16(tmem + lane_base + HH0 + 256 + col, v2);
#pragma unroll
          for (int e = 0; e < 16; ++e) v[e] += v2[e];
        }
        if (t < T && ct * TN + col < a.Cout) {      // (columns beyond Cout: zero padding rows of the weights)
          // every load (bias, ReLU mask) is issued BEFORE the first store: y, act and bias are plain pointers, so a load
          // placed after a store would have to wait for it (possible aliasing) and the 16 columns would serialise on
          // global-memory latency
          const int64_t off0 = (b * a.Cout + ct * TN + col) * (int64_t)T + t;
          float bz[16], mk[16] = {};
#pragma unroll
          for (int e = 0; e < 16; ++e) bz[e] = a.bias ? __ldg(a.bias + ct * TN + col + e) : 0.f;
          if (a.epi == EPI_MASK) {
#pragma unroll
            for (int e = 0; e < 16; ++e) mk[e] = __ldg(a.act + off0 + (int64_t)e * T);
          }
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            float r = (v[e] + s[e]) * inv + bz[e];
            if (a.epi == EPI_RELU) r = fmaxf(r, 0.f);
            else if (a.epi == EPI_MASK) r = mk[e] > 0.f ? r : 0.f;
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
    __syncthreads();
    tc5::fence_after_sync();
  }
#ifdef CT5_PROF
  if (tid == 0 && blockIdx.x == 0)
    printf("conv_tc5<%d> Cin=%d Cout=%d epi=%d cycles: wait first full %lld, wait full (steady) %lld, issue %lld, epilogue+sync %lld\n", KW,
           a.Cin, a.Cout, a.epi, pq[0], pq[1], pq[2], pq[3]);
#endif
  tc5::fence_before_sync();
  __syncthreads();
  if (w == 0) tc5::tmem_free512(tmem);
}

// ---------------------------------------------------------------------------------------------------------
// Input-stationary build (Cin <= 256, several output tiles): the input tile is staged and converted to fp16 parts ONCE
// per (sequence, time tile) and stays resident while the CTA walks the 128-channel output tiles; inside that loop only
// weights stream (cp.async, NWS stages) and the issuer warp is bound by the tensor pipe.  The streaming kernel above
// re-staged and re-converted the input once per output tile (8x for transition_net.2 with its 1024 outputs).
// ---------------------------------------------------------------------------------------------------------
constexpr int SMAX_KC = 8;                               // resident chunks: Cin <= 128
template <int KW>
struct SmS {
  static constexpr int GK = (KW == 3) ? 1 : 4;           // weight chunks per stage = per TMA bulk copy (24 KB / 32 KB)
  static constexpr int NWS = (KW == 3) ? 5 : 3;          // weight stages (~120 KB / ~100 KB in flight)
  static constexpr int WBYTES = KW * 2 * WPLANE;         // one chunk
  static constexpr int WREGION = NWS * GK * WBYTES;      // also the fp32 staging of the conversion phase
  static constexpr int XROUND = WREGION / XBYTES;        // input chunks staged per conversion round
  static constexpr int TOTAL = WREGION + SMAX_KC * ABYTES;
};

template <int KW>
__global__ void __launch_bounds__(NT, 1) conv_tc5s_kernel(ConvArgs a) {
  constexpr int PAD = KW / 2;
  using S = SmS<KW>;
  constexpr int NWS = S::NWS;
  extern __shared__ __align__(1024) uint8_t sm[];
  __shared__ __align__(8) uint64_t landed[NWS], empty[NWS], done;
  __shared__ uint32_t tmem_sh;
  uint8_t* parts = sm + S::WREGION;
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int i = 0; i < NWS; ++i) { tc5::mbar_init(&landed[i], 1); tc5::mbar_init(&empty[i], 1); }
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
  const int nkc = a.Cin / KC;
  const int T = a.T;
  constexpr uint32_t IDESC = tc5::idesc_f16(TN, 0, 0);
#ifdef CT5_PROF
  long long pf[6] = {0, 0, 0, 0, 0, 0}, pfl = clock64();
#define CT5P(i) do { if (tid == 0) { long long n_ = clock64(); pf[i] += n_ - pfl; pfl = n_; } } while (0)
#else
#define CT5P(i)
#endif
  uint32_t gw = 0;      // weight chunks loaded / consumed so far by this CTA (same sequence in the loaders and the issuer)
  uint32_t gct = 0;     // output tiles finished so far
  const int64_t nbt = a.B * a.ntt;

  constexpr int GK = S::GK;
  const int ngr = (nkc + GK - 1) / GK;                   // weight groups per output tile
  auto load_w = [&](int ct, int gr, uint32_t g) {        // warp 1 only: one TMA bulk copy per weight group
    if (g >= (uint32_t)NWS) tc5::mbar_wait(&empty[g % NWS], ((g / NWS) - 1u) & 1u);
    if (lane == 0) {
      const int k0 = gr * GK, nk = (nkc - k0 < GK) ? nkc - k0 : GK;
      uint8_t* st = sm + (g % NWS) * (GK * S::WBYTES);
      const uint8_t* wsrc = a.Wg + ((int64_t)ct * nkc + k0) * S::WBYTES;
      mbar_expect_tx(&landed[g % NWS], (uint32_t)(nk * S::WBYTES));
      bulk_g2s(st, wsrc, (uint32_t)(nk * S::WBYTES), &landed[g % NWS]);
    }
    __syncwarp();
  };

  for (int64_t bt = blockIdx.x; bt < nbt; bt += gridDim.x) {
    const int tt = (int)(bt % a.ntt);
    const int64_t b = bt / a.ntt;
    const int t0 = tt * TM;
    const float* xb = a.x + b * (int64_t)a.Cin * T;
    // ---- phase A: stage and convert the whole input tile (the weight region is free: the previous item is finished)
    for (int c0 = 0; c0 < nkc; c0 += S::XROUND) {
      const int nch = (nkc - c0 < S::XROUND) ? nkc - c0 : S::XROUND;
      for (int i = tid; i < nch * KC * (XW / 4); i += NT) {
        const int c = i / (XW / 4), j = (i - c * (XW / 4)) * 4, t = t0 - 4 + j;     // c: channel within the round
        const bool ok = (t >= 0 && t < T);
        cp16(reinterpret_cast<float*>(sm) + c * XP + j, xb + (int64_t)(c0 * KC + c) * T + (ok ? t : 0), ok ? 16 : 0);
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      __syncthreads();
      const float* xs = reinterpret_cast<const float*>(sm);
      for (int i = tid; i < nch * 2 * ROWS; i += NT) {
        const int ch = i / (2 * ROWS), rem = i - ch * (2 * ROWS), pl = rem / ROWS, r = rem - pl * ROWS;
        __align__(16) __half hh[8], mm[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) tc5::split_f16(xs[(ch * KC + pl * 8 + e) * XP + r + 3] * sx, hh[e], mm[e]);
        uint8_t* p = parts + (c0 + ch) * ABYTES + pl * PLANE + r * 16;
        *reinterpret_cast<uint4*>(p) = *reinterpret_cast<const uint4*>(hh);
        *reinterpret_cast<uint4*>(p + APART) = *reinterpret_cast<const uint4*>(mm);
      }
      tc5::fence_async_smem();
      __syncthreads();
    }
    CT5P(0);
    // ---- phase B: the output tiles
    if (w == 1) {   // first weight groups of tile 0
      const int npre = ngr < NWS ? ngr : NWS;
      for (int gr = 0; gr < npre; ++gr) load_w(0, gr, gw + (uint32_t)gr);
    }
    for (int ct = 0; ct < a.nct; ++ct) {
      const int npre = ngr < NWS ? ngr : NWS;      // groups of this tile already requested
      if (w == 0) {
        for (int gr = 0; gr < ngr; ++gr) {
          const uint32_t g = gw + (uint32_t)gr;
          tc5::mbar_wait(&landed[g % NWS], (g / NWS) & 1u);
          CT5P(1);
          tc5::fence_after_sync();
          const bool lead = tc5::elect_one();
          const int k0 = gr * GK, nk = (nkc - k0 < GK) ? nkc - k0 : GK;
          for (int kk = 0; kk < nk; ++kk) {
            const int kc = k0 + kk;
            const uint32_t sA = tc5::smem_u32(parts + kc * ABYTES);
            const uint32_t sW = tc5::smem_u32(sm + (g % NWS) * (GK * S::WBYTES) + kk * S::WBYTES);
#pragma unroll
            for (int tau = 0; tau < KW; ++tau) {
              const int step = kc * KW + tau;
              const uint64_t ah = tc5::make_desc(sA + (tau - PAD + 1) * 16, PLANE, 128);
              const uint64_t am = tc5::make_desc(sA + APART + (tau - PAD + 1) * 16, PLANE, 128);
              const uint64_t bh = tc5::make_desc(sW + tau * 2 * WPLANE, WPLANE, 128);
              const uint64_t bm = tc5::make_desc(sW + tau * 2 * WPLANE + TN * 16, WPLANE, 128);
              if (lead) {
                tc5::mma_f16(tmem + HH0 + 128u * (uint32_t)(step % 3), ah, bh, IDESC, step >= 3 ? 1u : 0u);
                tc5::mma_f16(tmem + SMALL, ah, bm, IDESC, step >= 1 ? 1u : 0u);
                tc5::mma_f16(tmem + SMALL, am, bh, IDESC, 1u);
              }
            }
          }
          if (lead) {
            tc5::commit(&empty[g % NWS]);
            if (gr == ngr - 1) tc5::commit(&done);
          }
          __syncwarp();
          CT5P(2);
        }
      } else if (w == 1) {
        for (int gr = npre; gr < ngr; ++gr) load_w(ct, gr, gw + (uint32_t)gr);
      }
      gw += (uint32_t)ngr;
      tc5::mbar_wait(&done, gct & 1u);
      CT5P(3);
      ++gct;
      tc5::fence_after_sync();
      if (w == 1 && ct + 1 < a.nct) {   // the stages are free: request the next tile's first groups under the epilogue
        for (int gr = 0; gr < npre; ++gr) load_w(ct + 1, gr, gw + (uint32_t)gr);
      }
      {   // epilogue: TMEM lane = time row; warp w reads lane quadrant w & 3, columns [32 (w >> 2), +32)
        const uint32_t lane_base = (uint32_t)(32 * (w & 3)) << 16;
        const int t = t0 + 32 * (w & 3) + lane;
        const int nsets = (nkc * KW < 3) ? nkc * KW : 3;
        float ymax = 0.f;
  #pragma unroll 1
        for (int cgp = 0; cgp < 2; ++cgp) {
          const int col = 32 * (w >> 2) + 16 * cgp;
          float v[16], s[16];
          tc5::tmem_ld16(tmem + lane_base + SMALL + col, s);
          tc5::tmem_ld16(tmem + lane_base + HH0 + col, v);
          if (nsets > 1) {
            float v1[16];
            tc5::tmem_ld16(tmem + lane_base + HH0 + 128 + col, v1);
  #pragma unroll
            for (int e = 0; e < 16; ++e) v[e] += v1[e];
          }
          if (nsets > 2) {
            float v2[16];
            tc5::tmem_ld16(tmem + lane_base + HH0 + 256 + col, v2);
  #pragma unroll
            for (int e = 0; e < 16; ++e) v[e] += v2[e];
          }
          if (t < T && ct * TN + col < a.Cout) {
            // every load (bias, ReLU mask) is issued BEFORE the first store: y, act and bias are plain pointers, so a load
            // placed after a store would have to wait for it (possible aliasing) and the 16 columns would serialise on
            // global-memory latency
            const int64_t off0 = (b * a.Cout + ct * TN + col) * (int64_t)T + t;
            float bz[16], mk[16] = {};
  #pragma unroll
            for (int e = 0; e < 16; ++e) bz[e] = a.bias ? __ldg(a.bias + ct * TN + col + e) : 0.f;
            if (a.epi == EPI_MASK) {
  #pragma unroll
              for (int e = 0; e < 16; ++e) mk[e] = __ldg(a.act + off0 + (int64_t)e * T);
            }
  #pragma unroll
            for (int e = 0; e < 16; ++e) {
              float r = (v[e] + s[e]) * inv + bz[e];
              if (a.epi == EPI_RELU) r = fmaxf(r, 0.f);
              else if (a.epi == EPI_MASK) r = mk[e] > 0.f ? r : 0.f;
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
      __syncthreads();
      tc5::fence_after_sync();
      CT5P(4);
    }
  }
#ifdef CT5_PROF
  if (tid == 0 && blockIdx.x == 0)
    printf("conv_tc5s<%d> Cin=%d Cout=%d cycles: phaseA %lld, wait landed %lld, issue %lld, wait done %lld, epilogue %lld\n", KW, a.Cin,
           a.Cout, pf[0], pf[1], pf[2], pf[3], pf[4]);
#endif
  tc5::fence_before_sync();
  __syncthreads();
  if (w == 0) tc5::tmem_free512(tmem);
}

}  // namespace ct5
}  // namespace vhmm

namespace vhmm {

// ---- host side -------------------------------------------------------------------------------------------
// 0 = auto (tensor cores when the shape allows), 1 = FP32 FFMA layer kernels only, 2 = tensor cores required
int conv_impl_mode();

static inline bool conv_tc5_ok(int Cin, int Cout, int64_t T, int kw, const void* x, const void* y) {
  return (kw == 1 || kw == 3) && Cin % ct5::KC == 0 && Cout % 16 == 0 && (T & 3) == 0 && T >= 4 &&
         (((uintptr_t)x | (uintptr_t)y) & 15) == 0;
}
// bytes of prepared weights (+ two abs-max words) for one (layer, view)
// Cout that is not a multiple of the 128-channel output tile is padded with zero weight rows (the MMA costs the same for any
// N <= 128); the epilogue stores the real channels only.  Needs Cout % 16 == 0 (the epilogue works in groups of 16 columns).
static inline int conv_tc5_cout_padded(int Cout) { return (Cout + ct5::TN - 1) / ct5::TN * ct5::TN; }
static inline size_t conv_tc5_weight_bytes(int Cin, int Cout, int kw) {
  return align_up((size_t)Cin * conv_tc5_cout_padded(Cout) * kw * 4, 256) + 256;
}

struct Tc5Weights {        // a weight view prepared for conv_tc5_kernel
  const uint8_t* Wg = nullptr;
  const uint32_t* amax_w = nullptr;
  uint32_t* amax_x = nullptr;      // scratch word for the activation abs-max of the launch that uses these weights
};

// prepare one weight view into `buf` (conv_tc5_weight_bytes); n_w = number of floats of the underlying tensor
static inline int conv_tc5_prepare(WView wv, int64_t n_w, int Cin, int Cout, int kw, void* buf, Tc5Weights* out, cudaStream_t st) {
  uint8_t* base = (uint8_t*)buf;
  const size_t wb = align_up((size_t)Cin * conv_tc5_cout_padded(Cout) * kw * 4, 256);
  uint32_t* words = (uint32_t*)(base + wb);
  VHMM_CUDA_CHECK(cudaMemsetAsync(words, 0, 8, st));
  if (Cout % ct5::TN) VHMM_CUDA_CHECK(cudaMemsetAsync(base, 0, wb, st));       // the padding rows
  int64_t blocks = cdiv(n_w, 1024);
  if (blocks > 1184) blocks = 1184;
  ct5::absmax_kernel<<<(unsigned)blocks, 256, 0, st>>>(wv.w, n_w, words);
  VHMM_LAUNCH_CHECK();
  const int64_t n = (int64_t)Cin * Cout * kw;
  ct5::prep_weights_kernel<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(wv, Cin, Cout, kw, words, base);
  VHMM_LAUNCH_CHECK();
  out->Wg = base; out->amax_w = words; out->amax_x = words + 1;
  return 0;
}

static inline int launch_absmax(const float* p, int64_t n, uint32_t* word, cudaStream_t st) {
  int64_t blocks = cdiv(n, 1024 * 8);
  if (blocks > 1184) blocks = 1184;
  if (blocks < 1) blocks = 1;
  ct5::absmax_kernel<<<(unsigned)blocks, 256, 0, st>>>(p, n, word);
  VHMM_LAUNCH_CHECK();
  return 0;
}

// amax_x_known: a device word that already holds |x| max (from the producer's epilogue or an earlier pass), or null;
// amax_y: a ZEROED device word that receives |y| max, or null
static inline int launch_conv_tc5(const float* x, const Tc5Weights& W, const float* bias, const float* act, float* y,
                                  int64_t B, int Cin, int Cout, int64_t T, int kw, int epi, cudaStream_t st,
                                  const uint32_t* amax_x_known = nullptr, uint32_t* amax_y = nullptr) {
  if (B <= 0 || T <= 0) return 0;
  const uint32_t* ax = amax_x_known;
  if (!ax) {
    VHMM_CUDA_CHECK(cudaMemsetAsync(W.amax_x, 0, 4, st));
    int rc = launch_absmax(x, B * (int64_t)Cin * T, W.amax_x, st);
    if (rc) return rc;
    ax = W.amax_x;
  }
  ct5::ConvArgs a{};
  a.x = x; a.Wg = W.Wg; a.bias = bias; a.act = act; a.y = y; a.amax_x = ax; a.amax_w = W.amax_w; a.amax_y = amax_y;
  a.B = B; a.Cin = Cin; a.Cout = Cout; a.T = (int)T; a.epi = epi;
  a.ntt = (int)cdiv(T, ct5::TM); a.nct = conv_tc5_cout_padded(Cout) / ct5::TN;
  a.nitems = B * a.ntt * a.nct;
  int64_t grid = device_sm_count();
  if (a.nct >= 2 && Cin / ct5::KC <= ct5::SMAX_KC) {     // input-stationary build: convert the input tile once for all output tiles
    const int64_t nbt = B * a.ntt;
    if (grid > nbt) grid = nbt;
    if (kw == 3) {
      static PerDeviceOnce once;
      if (once.pending()) {
        VHMM_CUDA_CHECK(cudaFuncSetAttribute(ct5::conv_tc5s_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, ct5::SmS<3>::TOTAL));
        once.mark();
      }
      ct5::conv_tc5s_kernel<3><<<(unsigned)grid, ct5::NT, ct5::SmS<3>::TOTAL, st>>>(a);
    } else {
      static PerDeviceOnce once;
      if (once.pending()) {
        VHMM_CUDA_CHECK(cudaFuncSetAttribute(ct5::conv_tc5s_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, ct5::SmS<1>::TOTAL));
        once.mark();
      }
      ct5::conv_tc5s_kernel<1><<<(unsigned)grid, ct5::NT, ct5::SmS<1>::TOTAL, st>>>(a);
    }
    VHMM_LAUNCH_CHECK();
    return 0;
  }
  if (grid > a.nitems) grid = a.nitems;
  CUtensorMap tm_x;
  a.use_tma = tma::make_map(&tm_x, x, B, Cin, T, ct5::XW, ct5::KC) ? 1 : 0;
  if (!a.use_tma) memset(&tm_x, 0, sizeof(tm_x));
  if (kw == 3) {
    static PerDeviceOnce once;
    if (once.pending()) {
      VHMM_CUDA_CHECK(cudaFuncSetAttribute(ct5::conv_tc5_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, ct5::Sm<3>::TOTAL));
      once.mark();
    }
    ct5::conv_tc5_kernel<3><<<(unsigned)grid, ct5::NT, ct5::Sm<3>::TOTAL, st>>>(a, tm_x);
  } else {
    static PerDeviceOnce once;
    if (once.pending()) {
      VHMM_CUDA_CHECK(cudaFuncSetAttribute(ct5::conv_tc5_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, ct5::Sm<1>::TOTAL));
      once.mark();
    }
    ct5::conv_tc5_kernel<1><<<(unsigned)grid, ct5::NT, ct5::Sm<1>::TOTAL, st>>>(a, tm_x);
  }
  VHMM_LAUNCH_CHECK();
  return 0;
}

}  // namespace vhmm

// =============================================================================================================
// weight gradient on the tensor core
// =============================================================================================================
namespace vhmm {
namespace ct5 {

constexpr int WNW = 16;                      // worker warps: 4 TMEM lane quadrants x 4 column quarters
constexpr int WNT = 32 + WNW * 32;           // threads per CTA: the issuer warp + the workers
constexpr int WB = 64;                       // channels per weight block side (x 2 parts = MMA M / N of 128)
constexpr int DY_ROWS = 128, X_ROWS = 130;   // dy row r <-> t0 + r; x row r <-> t0 - 1 + r
constexpr int DY_PLANE = DY_ROWS * 16, X_PLANE = X_ROWS * 16;
constexpr int DY_PARTS = 2 * 8 * DY_PLANE;   // [part][plane][row][8]: the m planes directly follow the h planes
constexpr int X_PARTS = 2 * 8 * X_PLANE;
constexpr int DY_P = 128, X_P = 136;         // fp32 staging pitches = TMA box widths; x window = t0-4 .. t0+131
// One buffer holds an item first as fp32 (a TMA box of dy and one of x: out-of-range columns and the padding channels of
// a skinny layer arrive as zeros) and then, converted IN PLACE, as fp16 operand parts (an element is 4 bytes either
// way).  Three buffers: item i feeds the MMAs, item i+1 is converted, item i+2 is in flight.
constexpr int WDY_BYTES = WB * DY_P * 4, WX_BYTES = WB * X_P * 4;
static_assert(WDY_BYTES >= DY_PARTS && WX_BYTES >= X_PARTS, "parts fit in their fp32 staging");
static_assert(WDY_BYTES % 128 == 0 && WX_BYTES % 128 == 0, "TMA destinations stay 128-byte aligned");
constexpr int WBUF = WDY_BYTES + WX_BYTES;
constexpr int WNBUF = 3;
constexpr int WSM_TOTAL = WNBUF * WBUF;
static_assert(WSM_TOTAL >= 128 * (3 * 64 + 1) * 4, "final reduction tile fits");

struct WgradArgs {
  const float* dy;         // (B,Cout,T)
  const float* x;          // (B,Cin,T)
  const uint8_t* dy_parts; // or (kw = 1, Cout = 1024): dy already as fp16 operand parts, prior_tc5.cuh layout -- one bulk copy per
                           // item lands [chunk (4)][part][plane (2)][row][8] in the dy region: nothing to convert; TMEM lane l <->
                           // channel 16 (l / 32) + l % 16, part (l / 16) % 2
  float* part;             // [nsplit][part_stride] partial weight gradients in the PARAMETER's layout
  int64_t part_stride;
  int64_t w_base, w_so, w_sc, w_st;
  const uint32_t* amax_dy;
  const uint32_t* amax_x;
  int64_t B;
  int Cin, Cout, T, ntt, nbc, nsplit;
};

// part[s][W(o,c,tau)] = sum over the (b, time tile) items of split s of dy[b,o,t] x[b,c,t+tau-pad].
// grid = (Cout/64)(Cin/64) nsplit; CTA = one 64 x 64 x KW weight block and one split of the (b, tile) items.
// Per item: [dy_h; dy_m] (M = 128) x [x_h; x_m] (N = 128) over K = 128 time steps per tap into TMEM (all four part
// products at once, the m*m quadrant is ignored), then drained into per-thread fp32 registers (RN adds): the TMEM
// accumulation chains stay 8 MMAs long, whatever the batch size.
// Warp-specialised: 16 worker warps + 1 issuer warp (544 threads).  tcgen05.mma issue BLOCKS while the tensor pipe's queue is
// full, so a warp that issues an item's 24 MMAs is gone for ~3.5 k cycles; as long as that warp also took part in the
// conversion (with its CTA-wide barriers) nothing overlapped: 5.2 k cycles per item for 3.6 k of MMA time.
//   workers, item i: thread 0 requests the two TMA boxes of item i+1; all convert item i IN PLACE (named barrier between the
//                    reads and the writes), arrive on ready[buf]; then drain item i-1 slot by slot (mma_done[slot] ->
//                    tcgen05.ld -> RN adds into registers -> tfree[slot])
//   issuer, item i:  waits ready[buf], then per slot: waits tfree[slot] (the previous tenant has been read), issues the 8 MMAs,
//                    commits to mma_done[slot]
// A slot is one 128-column accumulator: the three taps of an item (KW = 3) or the two alternating sets (KW = 1).  With one
// slot per tap the taps of item i+1 start as soon as the same tap of item i is drained -- the tensor pipe never waits for a
// whole-item drain.  (The first build of this kernel loaded with 16-byte cp.async: 825 instructions per thread and item, half
// of them address arithmetic, issue-bound at IPC 2.)
// (17 warps: registers are granted per warp in units that leave 96 per thread -- 112 and 120 fail to launch)
template <int KW>
__global__ void __launch_bounds__(WNT, 1) wgrad_tc5_kernel(WgradArgs a, const __grid_constant__ CUtensorMap tm_dy,
                                                          const __grid_constant__ CUtensorMap tm_x) {
  constexpr int PAD = KW / 2;
  constexpr int NSLOT = (KW == 1) ? 2 : KW;
  extern __shared__ __align__(1024) uint8_t sm[];
  __shared__ __align__(8) uint64_t mma_done[NSLOT], tfree[NSLOT];
  __shared__ __align__(8) uint64_t ld_bar[WNBUF], ready[WNBUF];
  __shared__ uint32_t tmem_sh;
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < NSLOT; ++i) { tc5::mbar_init(&mma_done[i], 1); tc5::mbar_init(&tfree[i], WNW); }
#pragma unroll
    for (int i = 0; i < WNBUF; ++i) { tc5::mbar_init(&ld_bar[i], 1); tc5::mbar_init(&ready[i], WNW); }
    tc5::mbar_init_fence();
  }
  if (w == 0) tc5::tmem_alloc512(&tmem_sh);
  tc5::fence_before_sync();
  __syncthreads();
  tc5::fence_after_sync();
  const uint32_t tmem = tmem_sh;
  const float sdy = tc5::pow2_scale(__uint_as_float(*a.amax_dy));
  const float sx = tc5::pow2_scale(__uint_as_float(*a.amax_x));
  const int s = blockIdx.x % a.nsplit;
  const int blk = blockIdx.x / a.nsplit;
  const int bc = blk % a.nbc, bo = blk / a.nbc;
  const int64_t nitems = a.B * a.ntt;
  constexpr uint32_t IDESC = tc5::idesc_f16(128, 1, 1);
  const int wt = tid - 32, ww = wt >> 5;                   // worker thread / warp index (issuer warp: negative)
  f2_t acc[KW][8];
#pragma unroll
  for (int k = 0; k < KW; ++k)
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[k][e] = 0ull;          // {0.f, 0.f}

  if (w == 0) {
    // ---- issuer -------------------------------------------------------------------------------------------------------
    int buf = 0;
    uint32_t n = 0;
    for (int64_t item = s; item < nitems; item += a.nsplit, ++n) {
      tc5::mbar_wait(&ready[buf], (n / WNBUF) & 1u);
      tc5::fence_after_sync();
      const bool lead = tc5::elect_one();
      const uint32_t sD = tc5::smem_u32(sm + buf * WBUF), sX = sD + WDY_BYTES;
#pragma unroll
      for (int k = 0; k < KW; ++k) {
        const int slot = (KW == 1) ? (int)(n & 1u) : k;
        const uint32_t use = (KW == 1) ? (n >> 1) : n;     // how often this slot has been filled before
        if (use >= 1) { tc5::mbar_wait(&tfree[slot], (use - 1u) & 1u); tc5::fence_after_sync(); }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint64_t da = tc5::make_desc(sD + (16 * i) * 16, 128, DY_PLANE);
          const uint64_t db = tc5::make_desc(sX + (16 * i + k - PAD + 1) * 16, 128, X_PLANE);
          if (lead) tc5::mma_f16(tmem + 128u * (uint32_t)slot, da, db, IDESC, i ? 1u : 0u);
        }
        if (lead) tc5::commit(&mma_done[slot]);
        __syncwarp();
      }
      buf = (buf == WNBUF - 1) ? 0 : buf + 1;
    }
  } else {
    // ---- workers ------------------------------------------------------------------------------------------------------
    // a warp reaches the TMEM lanes 32 (w % 4) .. of its HARDWARE warp index w = ww + 1: quadrants 1,2,3,0 per column quarter
    const uint32_t lane_base = (uint32_t)(32 * (w & 3)) << 16;
    const int cq = ww >> 2;
    const f2_t sdy2 = f2_pack(sdy, sdy), sx2 = f2_pack(sx, sx);
    // x rows needed by the MMAs: KW = 3 all 130, KW = 1 rows 1..128 only
    constexpr int XR0 = (KW == 1) ? 1 : 0, XNR = (KW == 1) ? 128 : X_ROWS;
    constexpr int NWT = WNW * 32;
    auto issue_loads = [&](int64_t item, int buf, uint32_t n) {          // worker thread 0 only
      const int64_t b = item / a.ntt;
      const int t0 = (int)(item - b * a.ntt) * TM;
      ::vhmm::tma::expect_tx(&ld_bar[buf], (uint32_t)WBUF);
      if (a.dy_parts) bulk_g2s(sm + buf * WBUF, a.dy_parts + (item * 64 + bo * 4) * (int64_t)8192, (uint32_t)WDY_BYTES, &ld_bar[buf]);
      else ::vhmm::tma::load_3d(sm + buf * WBUF, &tm_dy, t0, bo * WB, (int)b, &ld_bar[buf]);
      ::vhmm::tma::load_3d(sm + buf * WBUF + WDY_BYTES, &tm_x, t0 - 4, bc * WB, (int)b, &ld_bar[buf]);
    };
    auto drain_item = [&](uint32_t m) {
#pragma unroll
      for (int k = 0; k < KW; ++k) {
        const int slot = (KW == 1) ? (int)(m & 1u) : k;
        const uint32_t use = (KW == 1) ? (m >> 1) : m;
        tc5::mbar_wait(&mma_done[slot], use & 1u);
        tc5::fence_after_sync();
        float v0[16], v1[16];
        tc5::tmem_ld16(tmem + lane_base + 128u * (uint32_t)slot + 16 * cq, v0);
        tc5::tmem_ld16(tmem + lane_base + 128u * (uint32_t)slot + 64 + 16 * cq, v1);
#pragma unroll
        for (int e = 0; e < 8; ++e)
          acc[k][e] = f2_add(f2_add(acc[k][e], f2_pack(v0[2 * e], v0[2 * e + 1])), f2_pack(v1[2 * e], v1[2 * e + 1]));
        tc5::fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tfree[slot]);           // one arrival per worker warp
      }
    };

    if (wt == 0 && s < nitems) issue_loads(s, 0, 0);
    int buf = 0;
    uint32_t n = 0;                      // items done by this CTA
    for (int64_t item = s; item < nitems; item += a.nsplit, ++n) {
      const int64_t nxt = item + a.nsplit;
      const int nbuf = (buf == WNBUF - 1) ? 0 : buf + 1;
      // that buffer's last reader (the MMAs of item n-2) was drained by this thread in the previous iteration
      if (wt == 0 && nxt < nitems) issue_loads(nxt, nbuf, n + 1);
      tc5::mbar_wait(&ld_bar[buf], (n / WNBUF) & 1u);
      uint8_t* dyp = sm + buf * WBUF;
      uint8_t* xp = dyp + WDY_BYTES;
      {
        const float* ds = reinterpret_cast<const float*>(dyp);
        const float* xs = reinterpret_cast<const float*>(xp);
        // the dy region and the x region are converted in place one after the other (fewer live registers: 17 warps leave
        // 96 per thread): read -> named barrier -> write, twice
        const bool cdy = (a.dy_parts == nullptr);      // dy still to be converted
        if (cdy) {
          uint4 dh[2], dm[2];
#pragma unroll
          for (int q = 0; q < 2; ++q) {
            const int i = wt + q * NWT;
            const int pl = i / DY_ROWS, r = i - pl * DY_ROWS;
            float v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = ds[(pl * 8 + e) * DY_P + r];
            split8(v, sdy2, dh[q], dm[q]);
          }
          asm volatile("bar.sync 1, %0;" :: "n"(NWT) : "memory");
#pragma unroll
          for (int q = 0; q < 2; ++q) {
            const int i = wt + q * NWT;
            const int pl = i / DY_ROWS, r = i - pl * DY_ROWS;
            uint8_t* p = dyp + pl * DY_PLANE + r * 16;
            *reinterpret_cast<uint4*>(p) = dh[q];
            *reinterpret_cast<uint4*>(p + 8 * DY_PLANE) = dm[q];
          }
        }
        uint4 xh[3], xm[3];
#pragma unroll
        for (int q = 0; q < 3; ++q) {
          const int i = wt + q * NWT;
          if (i < 8 * XNR) {
            const int pl = i / XNR, r = XR0 + i - pl * XNR;
            float v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = xs[(pl * 8 + e) * X_P + r + 3];
            split8(v, sx2, xh[q], xm[q]);
          }
        }
        asm volatile("bar.sync 1, %0;" :: "n"(NWT) : "memory");
#pragma unroll
        for (int q = 0; q < 3; ++q) {
          const int i = wt + q * NWT;
          if (i < 8 * XNR) {
            const int pl = i / XNR, r = XR0 + i - pl * XNR;
            uint8_t* p = xp + pl * X_PLANE + r * 16;
            *reinterpret_cast<uint4*>(p) = xh[q];
            *reinterpret_cast<uint4*>(p + 8 * X_PLANE) = xm[q];
          }
        }
      }
      tc5::fence_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&ready[buf]);
      if (n > 0) drain_item(n - 1);        // under the MMAs of item n
      buf = nbuf;
    }
    if (n > 0) drain_item(n - 1);
  }
  __syncthreads();
  // lanes l and l + 64 hold the h / m rows of dy for out channel l: combine through shared memory, undo the scales
  float* red = reinterpret_cast<float*>(sm);               // [128 rows][KW * 64 + 1]
  constexpr int RP = KW * 64 + 1;
  if (w > 0) {
    const int row = 32 * (w & 3) + lane, cq = ww >> 2;
#pragma unroll
    for (int k = 0; k < KW; ++k)
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        float a0, a1;
        f2_unpack(acc[k][e], a0, a1);
        red[row * RP + k * 64 + 16 * cq + 2 * e] = a0;
        red[row * RP + k * 64 + 16 * cq + 2 * e + 1] = a1;
      }
  }
  __syncthreads();
  const float inv = 1.f / (sdy * sx);
  float* dst = a.part + (int64_t)s * a.part_stride;
  for (int i = tid; i < WB * WB * KW; i += WNT) {
    const int k = i % KW, c = (i / KW) % WB, o = i / (KW * WB);
    if (bo * WB + o >= a.Cout || bc * WB + c >= a.Cin) continue;     // padding channels of a skinny layer
    const int lh = a.dy_parts ? ((o >> 4) * 32 + (o & 15)) : o, lm = a.dy_parts ? lh + 16 : o + 64;
    const float v = (red[lh * RP + k * 64 + c] + red[lm * RP + k * 64 + c]) * inv;
    dst[a.w_base + (int64_t)(bo * WB + o) * a.w_so + (int64_t)(bc * WB + c) * a.w_sc + (int64_t)k * a.w_st] = v;
  }
  tc5::fence_before_sync();
  __syncthreads();
  if (w == 0) tc5::tmem_free512(tmem);
}

// partial[s][o] = sum over the sequences b = s (mod nsplit) and all t of dy[b,o,t]; grid (Cout, nsplit), 128-bit loads.
// The partials are added to gB in a fixed order by reduce_partials_kernel (deterministic).
__global__ void __launch_bounds__(256) bias_grad_kernel(const float* __restrict__ dy, int64_t B, int Cout, int T, int nsplit,
                                                        float* __restrict__ partial) {
  __shared__ double sh[32];
  const int o = blockIdx.x, s = blockIdx.y;
  const int T4 = T >> 2;
  double acc = 0.0;
  for (int64_t b = s; b < B; b += nsplit) {
    const float4* row = reinterpret_cast<const float4*>(dy + (b * Cout + o) * (int64_t)T);
    float a = 0.f;
    for (int i = threadIdx.x; i < T4; i += 256) {
      const float4 v = row[i];
      a += (v.x + v.y) + (v.z + v.w);
    }
    acc += (double)a;
  }
  acc = block_sum(acc, sh);
  if (threadIdx.x == 0) partial[(int64_t)s * Cout + o] = (float)acc;
}

}  // namespace ct5

static inline bool wgrad_tc5_ok(int Cin, int Cout, int64_t T, int kw, const void* dy, const void* x) {
  // channel counts are padded up to the 64-channel block (zero-filled): skinny layers (the embedding 256 x 32, to_logits
  // 32 x 128, to_params 10 x 256, transition_net.0 128 x 16) still beat their FFMA build by ~3x; below 8 channels FFMA stays
  // (measured: encoder.conv1's 256 x 5 gradient gains nothing over its FFMA kernel)
  return (kw == 1 || kw == 3) && Cin >= 8 && Cout >= 8 && (Cin >= 64 || Cout >= 64) && (T & 3) == 0 && T >= 4 &&
         (((uintptr_t)dy | (uintptr_t)x) & 15) == 0 && tma::encode_fn() != nullptr;
}
static inline int wgrad_tc5_nsplit(int Cin, int Cout, int64_t B, int64_t T) {
  const int64_t blocks = cdiv(Cout, ct5::WB) * cdiv(Cin, ct5::WB);
  // one CTA per SM (the three buffers take the whole shared memory): whole waves only -- a grid of 2 waves + 8 CTAs
  // costs 3.  Pick the wave count (1..3) that fills the machine best; ties go to fewer waves (fewer partials).
  const int64_t sms = device_sm_count();
  int64_t ns = 1;
  double best = -1.0;
  for (int64_t wv = 1; wv <= 3; ++wv) {
    const int64_t k = (wv * sms) / blocks;
    if (k < 1) continue;
    const double fill = (double)(k * blocks) / (double)(wv * sms);
    if (fill > best + 1e-9) { best = fill; ns = k; }
  }
  const int64_t items = B * cdiv(T, ct5::TM);
  if (ns > items) ns = items;
  return (int)(ns < 1 ? 1 : ns);
}
// floats of scratch: partials + two abs-max words
static inline int64_t wgrad_tc5_scratch_floats(int Cin, int Cout, int kw, int64_t B, int64_t T) {
  return (int64_t)wgrad_tc5_nsplit(Cin, Cout, B, T) * ((int64_t)Cin * Cout * kw) + 64 + 32 * (int64_t)Cout;
}

// gW (layer layout, +=), gB (+=, may be null)
static inline int launch_wgrad_tc5(const float* dy, const float* x, float* gW, float* gB, WView layout, int64_t B, int Cin,
                                   int Cout, int64_t T, int kw, float* scratch, cudaStream_t st,
                                   const uint32_t* amax_dy_known = nullptr, const uint32_t* amax_x_known = nullptr,
                                   const uint8_t* dy_parts = nullptr) {
  if (B <= 0 || T <= 0) return 0;
  const int ns = wgrad_tc5_nsplit(Cin, Cout, B, T);
  const int64_t nW = (int64_t)Cin * Cout * kw;
  uint32_t* words = reinterpret_cast<uint32_t*>(scratch + (int64_t)ns * nW);
  VHMM_CUDA_CHECK(cudaMemsetAsync(words, 0, 8, st));
  if (!amax_dy_known) { int rc = launch_absmax(dy, B * (int64_t)Cout * T, words, st); if (rc) return rc; }
  if (!amax_x_known) { int rc = launch_absmax(x, B * (int64_t)Cin * T, words + 1, st); if (rc) return rc; }
  ct5::WgradArgs a{};
  a.dy = dy; a.dy_parts = dy_parts; a.x = x; a.part = scratch; a.part_stride = nW;
  a.w_base = layout.base; a.w_so = layout.so; a.w_sc = layout.sc; a.w_st = layout.st;
  a.amax_dy = amax_dy_known ? amax_dy_known : words; a.amax_x = amax_x_known ? amax_x_known : words + 1;
  a.B = B; a.Cin = Cin; a.Cout = Cout; a.T = (int)T; a.ntt = (int)cdiv(T, ct5::TM); a.nbc = (int)cdiv(Cin, ct5::WB); a.nsplit = ns;
  const unsigned grid = (unsigned)(cdiv(Cout, ct5::WB) * cdiv(Cin, ct5::WB) * ns);
  CUtensorMap tm_dy, tm_x;
  if (dy_parts) dy = reinterpret_cast<const float*>(dy_parts);     // (the dy map is built but never used)
  if (!tma::make_map(&tm_dy, dy, B, Cout, T, ct5::DY_P, ct5::WB) || !tma::make_map(&tm_x, x, B, Cin, T, ct5::X_P, ct5::WB)) {
    set_error("wgrad_tc5: cuTensorMapEncodeTiled failed (B=%lld Cin=%d Cout=%d T=%lld)", (long long)B, Cin, Cout, (long long)T);
    return (int)cudaErrorInvalidValue;
  }
  if (kw == 3) {
    static PerDeviceOnce once;
    if (once.pending()) {
      VHMM_CUDA_CHECK(cudaFuncSetAttribute(ct5::wgrad_tc5_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, ct5::WSM_TOTAL));
      once.mark();
    }
    ct5::wgrad_tc5_kernel<3><<<grid, ct5::WNT, ct5::WSM_TOTAL, st>>>(a, tm_dy, tm_x);
  } else {
    static PerDeviceOnce once;
    if (once.pending()) {
      VHMM_CUDA_CHECK(cudaFuncSetAttribute(ct5::wgrad_tc5_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, ct5::WSM_TOTAL));
      once.mark();
    }
    ct5::wgrad_tc5_kernel<1><<<grid, ct5::WNT, ct5::WSM_TOTAL, st>>>(a, tm_dy, tm_x);
  }
  VHMM_LAUNCH_CHECK();
  reduce_partials_kernel<<<(unsigned)cdiv(nW, 256), 256, 0, st>>>(scratch, nW, ns, nW, gW);
  VHMM_LAUNCH_CHECK();
  if (gB && !dy_parts) {   // (T % 4 == 0 and 16-byte alignment were checked by wgrad_tc5_ok; operand-form dy: the caller's kernel)
    float* bpart = scratch + (int64_t)ns * nW + 64;
    int nb = B < 32 ? (int)B : 32;
    ct5::bias_grad_kernel<<<dim3((unsigned)Cout, (unsigned)nb), 256, 0, st>>>(dy, B, Cout, (int)T, nb, bpart);
    VHMM_LAUNCH_CHECK();
    reduce_partials_kernel<<<(unsigned)cdiv(Cout, 256), 256, 0, st>>>(bpart, Cout, nb, Cout, gB);
    VHMM_LAUNCH_CHECK();
  }
  return 0;
}

}  // namespace vhmm
