// Viterbi decode of FEW LONG sequences (K <= 4): one warp per sequence.
//
// The max-plus recursion is serial in time when the result has to be bit-exact (fp32 addition does not associate), so a
// sequence cannot be cut into segments like the sum-product recursions of hmm_scan.cuh.  What CAN be removed is everything
// around the dependent chain.  The thread-per-sequence sweep of hmm_small.cuh (hs::fwd_kernel<.., VIT>) is built for
// B >> #SM: at B = 64, T = 65536 it is one CTA of 64 threads waiting on its own 4-step slabs, 0.48 us per step (31.6 ms).
// Here:
//   forward    the warp's 32 lanes run the recursion redundantly (same cost as one lane) on inputs that lane 0 streams
//              into a shared-memory ring with bulk copies (cp.async.bulk + mbarrier, 128 steps = 6 KB per stage,
//              NS stages in flight), so a step costs its dependent chain (add, compare, select, add) and 12 broadcast
//              shared-memory words; every lane keeps the packed back-pointers of "its" 4 steps of the 128-step block
//              and the block is stored as one coalesced 128-byte line.
//   back-track path[t-1] = psi_t[path[t]] is a composition of maps {0..K-1} -> {0..K-1}: exact integer work, so it IS
//              time-parallel.  Per 128-step block every lane composes the maps of its 4 steps, a 5-round shuffle scan
//              gives each lane its incoming state, and the lane replays its 4 steps: one coalesced 128-byte store of
//              path per block, the next block's back-pointers prefetched meanwhile.
// Semantics are those of hs::fwd_kernel (reference: VQ_VAE_HMM_fixed.py viterbi): ties -> lowest
// index, padding steps -> VHMM_PAD_STATE, score 0 for an empty sequence.
#pragma once
#include "hmm_small.cuh"

namespace vhmm {
namespace hv {

constexpr int BLK = 128;     // steps per stage / per back-pointer block (32 lanes x 4 steps)
constexpr int NS = 4;        // stages of the ring

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(s32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(s32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* smem, const void* g, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"(s32(smem)), "l"(g), "r"(bytes), "r"(s32(bar)) : "memory");
}
// every lane polls (the value it needs is behind the barrier anyway); bounded, a lost copy traps instead of hanging the GPU
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t a = s32(bar);
  uint32_t ok = 0;
  for (int it = 0; it < (1 << 24) && !ok; ++it)
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(ok) : "r"(a), "r"(parity) : "memory");
  if (!ok) __trap();
}

template <int K>
struct Ring {
  static constexpr int A_FLOATS = BLK * K * K;
  static constexpr int B_FLOATS = BLK * K;
  static constexpr int STAGE = A_FLOATS + B_FLOATS;          // floats (a multiple of 4: 16-byte aligned stages)
  static constexpr size_t SMEM = (size_t)NS * STAGE * sizeof(float);
};

// one max-plus step of the CHAIN: dl[j] <- lb[j] + max_i (dl[i] + la[i][j]).  add -> max -> add, nothing else: the maximum is
// taken arithmetically (fmaxf; 3-input FMNMX) and the back-pointers are not computed here at all (vit_args below).
template <int K>
__device__ __forceinline__ void vit_chain(float (&dl)[K], const float* la, const float* lb) {
  float nd[K];
#pragma unroll
  for (int j = 0; j < K; ++j) {
    float best = __fadd_rn(dl[0], la[j]);
#pragma unroll
    for (int i = 1; i < K; ++i) best = fmaxf(best, __fadd_rn(dl[i], la[i * K + j]));
    nd[j] = __fadd_rn(lb[j], best);
  }
#pragma unroll
  for (int j = 0; j < K; ++j) dl[j] = nd[j];
}
// the back-pointers of one step from its INCOMING vector: lowest i with dl[i] + la[i][j] == max_i (the same additions as the
// chain, so the same maximum).  For inputs without NaN this is the strictly-greater scan of hs::fwd_kernel bit for bit (a
// +0 / -0 tie may pick the other zero: equal in every comparison, and x + (+-0) = x); a NaN score (only (+inf) + (-inf)
// makes one) is garbage in both kernels, not the same garbage.
template <int K>
__device__ __forceinline__ unsigned vit_args(const float (&dl)[K], const float* la) {
  unsigned packed = 0;
#pragma unroll
  for (int j = 0; j < K; ++j) {
    float c[K];
#pragma unroll
    for (int i = 0; i < K; ++i) c[i] = __fadd_rn(dl[i], la[i * K + j]);
    float best = c[0];
#pragma unroll
    for (int i = 1; i < K; ++i) best = fmaxf(best, c[i]);
    unsigned arg = (unsigned)(K - 1);
#pragma unroll
    for (int i = K - 2; i >= 0; --i) arg = (c[i] == best) ? (unsigned)i : arg;
    packed |= arg << (2 * j);
  }
  return packed;
}

// maps {0..3} -> {0..3} packed 2 bits per entry
__device__ __forceinline__ unsigned map_at(unsigned m, unsigned s) { return (m >> (2 * s)) & 3u; }
// (f o g)(s) = f(g(s))
__device__ __forceinline__ unsigned map_compose(unsigned f, unsigned g) {
  return map_at(f, map_at(g, 0)) | (map_at(f, map_at(g, 1)) << 2) | (map_at(f, map_at(g, 2)) << 4) | (map_at(f, map_at(g, 3)) << 6);
}
constexpr unsigned MAP_ID = 0xE4u;   // 0 -> 0, 1 -> 1, 2 -> 2, 3 -> 3

template <int K>
__device__ __forceinline__ float4 pack4(const float (&v)[K]) {
  return make_float4(v[0], v[1], K > 2 ? v[K > 2 ? 2 : 0] : 0.f, K > 3 ? v[K > 3 ? 3 : 0] : 0.f);
}
template <int K>
__device__ __forceinline__ void unpack4(const float4& h, float (&v)[K]) {
  v[0] = h.x; v[1] = h.y;
  if constexpr (K > 2) v[2] = h.z;
  if constexpr (K > 3) v[3] = h.w;
}

// grid = B CTAs of one warp.  Requires T % 4 == 0, log_A / log_b 16-byte aligned, path 4-byte aligned, psi rows of T bytes.
template <int K>
__global__ void __launch_bounds__(32) vit_long_kernel(hs::Args a) {
  using R = Ring<K>;
  extern __shared__ __align__(16) float ring[];
  __shared__ __align__(8) uint64_t full[NS];
  __shared__ float4 hist[BLK];                       // the vector every step of the current block starts from
  const int lane = threadIdx.x;
  const int64_t b = blockIdx.x;
  const int T = (int)a.T;
  int64_t L = a.lengths[b];
  const int n = (int)(L < 0 ? 0 : (L > a.T ? a.T : L));
  uint32_t* prow = reinterpret_cast<uint32_t*>(a.path + b * a.T);
  uint32_t* psirow = reinterpret_cast<uint32_t*>(a.psi + b * a.T);
  if (n == 0) {
    if (lane == 0 && a.score) a.score[b] = 0.f;
    for (int w = lane; w < T / 4; w += 32) prow[w] = 0xFFFFFFFFu;
    return;
  }
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < NS; ++i) mbar_init(&full[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  const int nld = (n + 3) & ~3;                      // steps loaded (<= T: T % 4 == 0)
  const int nblk = (nld + BLK - 1) / BLK;
  const float* gA = a.A + b * a.T * (K * K);
  const float* gB = a.Bm + b * a.T * K;
  auto issue = [&](int c) {                          // lane 0: block c into stage c % NS
    float* st = ring + (c % NS) * R::STAGE;
    const int cn = min(BLK, nld - c * BLK);
    const uint32_t ba = (uint32_t)cn * K * K * 4u, bb = (uint32_t)cn * K * 4u;
    mbar_expect_tx(&full[c % NS], ba + bb);
    bulk_g2s(st, gA + (int64_t)c * BLK * K * K, ba, &full[c % NS]);
    bulk_g2s(st + R::A_FLOATS, gB + (int64_t)c * BLK * K, bb, &full[c % NS]);
  };
  if (lane == 0)
    for (int c = 0; c < NS - 1 && c < nblk; ++c) issue(c);

  float dl[K];
#pragma unroll
  for (int j = 0; j < K; ++j) dl[j] = 0.f;
  for (int c = 0; c < nblk; ++c) {
    // stage (c-1) % NS was read by every lane in the previous trip: the warp barrier orders those reads before the refill
    __syncwarp();
    if (lane == 0 && c + NS - 1 < nblk) issue(c + NS - 1);
    mbar_wait(&full[c % NS], (uint32_t)(c / NS) & 1u);
    const float* sA = ring + (c % NS) * R::STAGE;
    const float* sB = sA + R::A_FLOATS;
    const int t0 = c * BLK;
    const int ng = (min(BLK, nld - t0)) >> 2;        // groups of 4 steps in this block
    auto load_group = [&](int g, float (&Av)[4 * K * K], float (&Bv)[4 * K]) {
      const float4* pa = reinterpret_cast<const float4*>(sA + g * 4 * K * K);
      const float4* pb = reinterpret_cast<const float4*>(sB + g * 4 * K);
#pragma unroll
      for (int q = 0; q < K * K; ++q) {
        const float4 v = pa[q];
        Av[4 * q] = v.x; Av[4 * q + 1] = v.y; Av[4 * q + 2] = v.z; Av[4 * q + 3] = v.w;
      }
#pragma unroll
      for (int q = 0; q < K; ++q) {
        const float4 v = pb[q];
        Bv[4 * q] = v.x; Bv[4 * q + 1] = v.y; Bv[4 * q + 2] = v.z; Bv[4 * q + 3] = v.w;
      }
    };
    auto edge_group = [&](int g) {                   // the group with step 0 and the group that the length cuts
      float Av[4 * K * K], Bv[4 * K];
      load_group(g, Av, Bv);
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        const int t = t0 + g * 4 + s;
        if (t == 0) {
#pragma unroll
          for (int j = 0; j < K; ++j) dl[j] = __fadd_rn(a.log_pi[j], Bv[j]);
        } else if (t < n) {
          if (lane == 0) hist[g * 4 + s] = pack4<K>(dl);
          vit_chain<K>(dl, Av + s * K * K, Bv + s * K);
        }
      }
    };
    // full groups [g_lo, g_hi): a branch-free body, so that the loads of the next groups are scheduled over the chain
    const int g_lo = (c == 0) ? 1 : 0;
    const int g_hi = max(g_lo, min(ng, (n - t0) >> 2));
    if (g_lo) edge_group(0);
#pragma unroll 4
    for (int g = g_lo; g < g_hi; ++g) {
      float Av[4 * K * K], Bv[4 * K];
      load_group(g, Av, Bv);
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        if (lane == 0) hist[g * 4 + s] = pack4<K>(dl);           // the vector step t0+4g+s starts from
        vit_chain<K>(dl, Av + s * K * K, Bv + s * K);
      }
    }
    for (int g = g_hi; g < ng; ++g) edge_group(g);
    __syncwarp();
    // back-pointers of the block, time-parallel: lane l takes steps 4l .. 4l+3 (one 32-bit word, one coalesced line per block)
    if (t0 + 4 * lane < nld) {
      const float4* pa = reinterpret_cast<const float4*>(sA + lane * 4 * K * K);
      float Av[4 * K * K];
#pragma unroll
      for (int q = 0; q < K * K; ++q) {
        const float4 v = pa[q];
        Av[4 * q] = v.x; Av[4 * q + 1] = v.y; Av[4 * q + 2] = v.z; Av[4 * q + 3] = v.w;
      }
      uint32_t word = 0;
#pragma unroll
      for (int s = 0; s < 4; ++s) {
        const int t = t0 + 4 * lane + s;
        if (t >= 1 && t < n) {
          const float4 h = hist[4 * lane + s];
          float in[K];
          unpack4<K>(h, in);
          word |= vit_args<K>(in, Av + s * K * K) << (8 * s);
        }
      }
      psirow[t0 / 4 + lane] = word;
    }
  }
  // final state: lowest j attaining the max
  float best = dl[0];
  unsigned state = 0;
#pragma unroll
  for (int j = 1; j < K; ++j)
    if (dl[j] > best) { best = dl[j]; state = (unsigned)j; }
  if (lane == 0 && a.score) a.score[b] = best;
  __syncwarp();                                      // the block's own back-pointer stores are visible to its lanes below

  // padding blocks beyond the last valid step
  for (int w = nblk * (BLK / 4) + lane; w < T / 4; w += 32) prow[w] = 0xFFFFFFFFu;
  // back-track, block by block from the end; `state` = path[t] at the block's last step (virtual steps >= n: identity)
  uint32_t nxt = 0;
  {
    const int w = (nblk - 1) * (BLK / 4) + lane;
    if (4 * w < nld) nxt = psirow[w];
  }
  for (int c = nblk - 1; c >= 0; --c) {
    const uint32_t word = nxt;
    if (c > 0) nxt = psirow[(c - 1) * (BLK / 4) + lane];       // prefetch (every word of a full block is valid)
    const int tl = c * BLK + 4 * lane;                         // this lane's steps tl .. tl+3
    unsigned m[4];
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      const int t = tl + s;
      m[s] = (t < n) ? ((word >> (8 * s)) & 0xFFu) : MAP_ID;   // path[t-1] = m[s](path[t])
    }
    // F = m0 o m1 o m2 o m3: path[tl+3] -> path[tl-1]
    unsigned F = map_compose(m[0], map_compose(m[1], map_compose(m[2], m[3])));
    // suffix scan over the lanes: G_l = F_{l+1} o F_{l+2} o ... o F_31 maps the block's incoming state to path[tl+3] of lane l
    unsigned S = F;                                            // inclusive: F_l o ... o F_31
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned o = __shfl_down_sync(0xffffffffu, S, d);
      if (lane + d < 32) S = map_compose(S, o);
    }
    unsigned G = __shfl_down_sync(0xffffffffu, S, 1);
    if (lane == 31) G = MAP_ID;
    unsigned s_in = map_at(G, state);                          // path[tl+3]
    uint32_t out = 0;
#pragma unroll
    for (int s = 3; s >= 0; --s) {
      const int t = tl + s;
      out |= ((t < n) ? s_in : 0xFFu) << (8 * s);
      s_in = map_at(m[s], s_in);
    }
    if (tl < T) prow[c * (BLK / 4) + lane] = out;
    state = map_at(__shfl_sync(0xffffffffu, S, 0), state);     // path[c*BLK - 1]
  }
}

}  // namespace hv
}  // namespace vhmm
