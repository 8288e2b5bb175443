// hmm.cu -- host side of the exact-HMM stages (extensions): dispatch
// between the tuned small-K kernels (hmm_small.cuh, hmm_scan.cuh) and the any-K warp kernels
// (hmm_generic.cuh), workspace carving, emission log-probabilities.
#include "hmm_generic.cuh"
#include "hmm_wide.cuh"
#include "hmm_scan.cuh"
#include "hmm_warp.cuh"
#include "hmm_vitlong.cuh"
#include "hmm_wide_scan.cuh"
#include <atomic>
#include <stdlib.h>

namespace vhmm {

// numerics of the sum-product recursions: 0 = scaled linear space (default: the tuned kernels), 1 = log space
// (logsumexp recursions of hmm_generic.cuh for every K <= 32; include/vhmm.h: vhmm_hmm_numerics_select)
static std::atomic<int> g_numerics{0};
int hmm_numerics_select(int mode) { g_numerics = mode; return 0; }
static inline bool small_k(int64_t K) { return K >= 2 && K <= 4 && g_numerics.load(std::memory_order_relaxed) == 0; }
static inline bool aligned16(const void* p) { return ((uintptr_t)p & 15) == 0; }

#ifndef VHMM_HS_TC
#define VHMM_HS_TC 4   /* 4 steps x 2 stages = 416 B of smem per row -> 512 rows resident per SM (measured best) */
#endif
#ifndef VHMM_HS_NT
#define VHMM_HS_NT 64
#endif
constexpr int HS_TC = VHMM_HS_TC;   // steps per staged slab
constexpr int HS_NT = VHMM_HS_NT;   // rows (threads) per CTA
#ifndef VHMM_SCAN_CH
#define VHMM_SCAN_CH 64   /* measured at cfg 4 (B=64, T=65536): 32 -> 0.160 ms, 64 -> 0.110 ms, 128 -> 0.117 ms (log-lik) */
#endif
constexpr int SCAN_CH = VHMM_SCAN_CH; // steps per scan chunk

static int64_t psi_bytes(int64_t K) { return small_k(K) ? 1 : K; }   // packed 2 bits/state only in the small-K kernels

size_t hmm_workspace_bytes(int op, int64_t B, int64_t T, int64_t K) {
  if (K <= 0) K = 32;
  const size_t stash = align_up((size_t)B * T * K * sizeof(float), 256);
  const size_t psi = align_up((size_t)B * T * psi_bytes(K), 256);
  const int64_t nck = cdiv(T, SCAN_CH);
  // few long sequences at K > 4 (hmm_wide_scan.cuh): segment products, exponents, emission maxima, vin / bin per 128-step row
  const int64_t wrows = B * cdiv(T, (int64_t)hx::WS_SEG);
  const size_t wide_rows = (K > 4 && T >= 512 && B <= 256)
      ? align_up((size_t)wrows * K * K * 4, 256) + 3 * align_up((size_t)wrows * K * 4, 256) + align_up((size_t)wrows * 4, 256) +
        align_up((size_t)B * 4, 256) + 256 : 0;
  switch (op) {
    case VHMM_OP_HMM_FORWARD: return 0;
    case VHMM_OP_HMM_POSTERIOR: return stash + 256 + wide_rows;
    case VHMM_OP_VITERBI: return psi + 256;
    case VHMM_OP_HMM_DECODE: return stash + psi + 512 + wide_rows;
    case VHMM_OP_HMM_SCAN:
      return align_up((size_t)B * nck * K * K * 4, 256) + 2 * align_up((size_t)B * nck * 8, 256) +
             align_up((size_t)B * nck * K * 4, 256) + align_up((size_t)B * 4, 256) + 1024;
    default: return 0;
  }
}

// ---- emission log-probabilities (B.1) -----------------------------------------------------
// log_b[b,t,k] = -0.5 sum_c [ log(2 pi v) + (x - mu)^2 / v ],  v = max(exp(logvar), 1e-8)
__global__ void __launch_bounds__(256)
emission_kernel(const float* __restrict__ x, const float* __restrict__ mu, const float* __restrict__ lv,
                int C, int T, int K, float* __restrict__ log_b) {
  const int t = blockIdx.x * 256 + threadIdx.x;
  if (t >= T) return;
  const int64_t b = blockIdx.y;
  for (int k = 0; k < K; ++k) {
    float acc = 0.f;
    for (int c = 0; c < C; ++c) {
      const float xv = x[(b * C + c) * (int64_t)T + t];
      const float m = mu[((int64_t)k * C + c) * T + t];
      const float v = fmaxf(expf(lv[((int64_t)k * C + c) * T + t]), 1e-8f);
      const float d = xv - m;
      acc += logf(6.283185307179586f * v) + d * d / v;
    }
    log_b[(b * T + t) * K + k] = -0.5f * acc;
  }
}

int hmm_emission_logprob(const float* x, const float* mu_tab, const float* logvar_tab, int64_t B,
                         int64_t C, int64_t T, int64_t K, float* log_b, cudaStream_t st) {
  for (int64_t b0 = 0; b0 < B; b0 += 32768) {
    const int64_t nb = B - b0 < 32768 ? B - b0 : 32768;
    dim3 g((unsigned)cdiv(T, 256), (unsigned)nb);
    emission_kernel<<<g, 256, 0, st>>>(x + b0 * C * T, mu_tab, logvar_tab, (int)C, (int)T, (int)K,
                                       log_b + b0 * T * K);
    VHMM_LAUNCH_CHECK();
  }
  return 0;
}

// ---- small-K launch helpers ---------------------------------------------------------------
template <int K, bool SP, bool VIT>
static int launch_fwd_small(const hs::Args& a, cudaStream_t st) {
  using TL = hs::Tile<K, HS_TC, HS_NT>;
  auto kern = hs::fwd_kernel<K, HS_TC, HS_NT, SP, VIT>;
  VHMM_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TL::SMEM));
  kern<<<(unsigned)cdiv(a.nrows, HS_NT), HS_NT, TL::SMEM, st>>>(a);
  VHMM_LAUNCH_CHECK();
  return 0;
}

// Viterbi of small K: few sequences (fewer than 16 warps per SM's worth) -> one warp per sequence (hmm_vitlong.cuh: the
// thread-per-sequence sweep leaves such a batch on a handful of CTAs, 0.48 us per step); environment VHMM_VITERBI=sweep keeps
// the sweep.  Same results bit for bit.
static bool vit_long_enabled() {
  static const bool on = [] { const char* e = getenv("VHMM_VITERBI"); return !(e && e[0] == 's'); }();
  return on;
}
static inline bool vit_long_fits(int64_t B, int64_t T, const hs::Args& a) {
  return vit_long_enabled() && B < 2368 && T >= 64 && (T & 3) == 0 && a.vec_in && a.vec_path && !a.ahat && !a.alpha_log;
}
template <int K>
static int launch_vit_small(const hs::Args& a, cudaStream_t st) {
  if (vit_long_fits(a.nrows, a.T, a)) {
    auto kern = hv::vit_long_kernel<K>;
    VHMM_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hv::Ring<K>::SMEM));
    kern<<<(unsigned)a.nrows, 32, hv::Ring<K>::SMEM, st>>>(a);
    VHMM_LAUNCH_CHECK();
    return 0;
  }
  return launch_fwd_small<K, false, true>(a, st);
}

template <int K>
static int launch_bwd_small(const hs::Args& a, cudaStream_t st) {
  using TL = hs::Tile<K, HS_TC, HS_NT>;
  auto kern = hs::bwd_kernel<K, HS_TC, HS_NT>;
  VHMM_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TL::SMEM));
  kern<<<(unsigned)cdiv(a.nrows, HS_NT), HS_NT, TL::SMEM, st>>>(a);
  VHMM_LAUNCH_CHECK();
  return 0;
}

#define DISPATCH_K(K, CALL)                                  \
  switch ((int)(K)) {                                        \
    case 2: { constexpr int KK = 2; CALL; break; }           \
    case 3: { constexpr int KK = 3; CALL; break; }           \
    case 4: { constexpr int KK = 4; CALL; break; }           \
    default: rc = VHMM_E_DIMS;                               \
  }


// ---- one-pass posterior (hmm_warp.cuh): whole sequence in shared memory, one warp each ------
static std::atomic<int> g_posterior_mode{-1};   // -1: from env VHMM_POSTERIOR (onepass|sweep), 0 auto, 1 sweep only
int hmm_posterior_select(int mode) { g_posterior_mode = mode; return 0; }
static bool onepass_allowed() {
  if (g_posterior_mode < 0) {
    const char* e = getenv("VHMM_POSTERIOR");
    g_posterior_mode = (e && e[0] == 's') ? 1 : 0;
  }
  return g_posterior_mode == 0;
}
constexpr size_t HW_MAX_WARP_BYTES = 28 * 1024;   // >= 8 resident warps per SM
static bool onepass_fits(int64_t T, int64_t K, const void* A, const void* Bm, const void* gamma, const void* xi) {
  if (!small_k(K) || T < 1 || ((T * K) & 3) != 0) return false;
  if (!aligned16(A) || !aligned16(Bm) || !aligned16(gamma) || (xi && !aligned16(xi))) return false;
  const int S = 4 * (int)cdiv(T, 128);
  const size_t wb = 32 * (size_t)(hw::pitch_odd16(S * (int)(K * K)) + hw::pitch_odd16(S * (int)K)) * 4;
  return wb <= HW_MAX_WARP_BYTES;
}
template <int K, bool XI, bool LL_ONLY = false, bool SEG = false, bool ALPHA = false>
static int launch_onepass(hw::Args a, cudaStream_t st) {
  auto kern = hw::posterior_kernel<K, XI, LL_ONLY, SEG, false, ALPHA>;
  const size_t wb = 32 * (size_t)(a.SA + a.SB) * 4;
  int wpb = (int)((56 * 1024) / wb);
  wpb = wpb < 1 ? 1 : (wpb > 4 ? 4 : wpb);
  const size_t smem = wb * wpb;
  const int sms = device_sm_count();   // current device
  VHMM_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  VHMM_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
  int per_sm = 0;
  VHMM_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, wpb * 32, smem));
  if (per_sm < 1) per_sm = 1;
  int64_t grid = (int64_t)sms * per_sm;
  const int64_t need = cdiv(a.B, wpb);
  if (grid > need) grid = need;
  kern<<<(unsigned)grid, wpb * 32, smem, st>>>(a);
  VHMM_LAUNCH_CHECK();
  return 0;
}
// cfg-3 decode from ONE read of the inputs (hmm_warp.cuh, VIT): `wpb` posterior warps + the Viterbi helper warp per CTA
// (measured SLOWER than the two kernels -- 14.5 vs 7.0 ms at cfg 3, see hmm_warp.cuh -- hence opt-in)
static std::atomic<int> g_decode_mode{-1};   // -1: from env VHMM_DECODE (one = the single kernel), 0 two kernels, 1 one kernel
int hmm_decode_select(int mode) { g_decode_mode = mode; return 0; }
static bool decode_onekernel() {
  if (g_decode_mode < 0) {
    const char* e = getenv("VHMM_DECODE");
    g_decode_mode = (e && e[0] == 'o') ? 1 : 0;
  }
  return g_decode_mode == 1;
}
template <int K>
static int launch_decode_onepass(hw::Args a, cudaStream_t st) {
  auto kern = hw::posterior_kernel<K, false, false, false, true>;
  const size_t tp = (size_t)((a.T + 15) & ~(int64_t)15);
  const size_t wb = (32 * (size_t)(a.SA + a.SB) + 4) * 4 + tp;       // skewed region + psi / path row
  int wpb = (int)((110 * 1024) / wb);                                 // two CTAs per SM
  wpb = wpb < 1 ? 1 : (wpb > 8 ? 8 : wpb);
  const size_t smem = wb * wpb;
  const int sms = device_sm_count();
  VHMM_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  VHMM_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
  int per_sm = 0;
  VHMM_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, (wpb + 1) * 32, smem));
  if (per_sm < 1) per_sm = 1;
  int64_t grid = (int64_t)sms * per_sm;
  const int64_t need = cdiv(a.B, wpb);
  if (grid > need) grid = need;
  kern<<<(unsigned)grid, (wpb + 1) * 32, smem, st>>>(a);
  VHMM_LAUNCH_CHECK();
  return 0;
}

static int hmm_posterior_onepass(const float* log_pi, const float* A, const float* Bm, const int64_t* lengths,
                                 int64_t B, int64_t T, int64_t K, float* loglik, float* gamma, float* xi,
                                 cudaStream_t st) {
  hw::Args a{};
  a.log_pi = log_pi; a.A = A; a.Bm = Bm; a.lengths = lengths; a.B = B; a.T = T;
  a.S = 4 * (int)cdiv(T, 128);
  a.SA = hw::pitch_odd16(a.S * (int)(K * K));
  a.SB = hw::pitch_odd16(a.S * (int)K);
  const uint64_t ga = (uint64_t)a.S * K * K / 4, gb = (uint64_t)a.S * K / 4;
  a.magicA = (unsigned)(((1ull << 32) + ga - 1) / ga);
  a.magicB = (unsigned)(((1ull << 32) + gb - 1) / gb);
  a.loglik = loglik; a.gamma = gamma; a.xi = xi;
  int rc = 0;
  if (!gamma) { DISPATCH_K(K, (rc = launch_onepass<KK, false, true>(a, st))); }   // log-likelihood only
  else if (xi) { DISPATCH_K(K, (rc = launch_onepass<KK, true>(a, st))); }
  else         { DISPATCH_K(K, (rc = launch_onepass<KK, false>(a, st))); }
  return rc;
}

// all-prefix log alpha by the one-pass kernel: rows of T steps (whole sequences: nseg = 1, vin = null; or segments of longer
// sequences with the incoming vectors of the carry scan)
static int hmm_alpha_onepass(const float* log_pi, const float* A, const float* Bm, const int64_t* lengths, int64_t rows,
                             int64_t T, int64_t K, int nseg, const float* vin, const double* ll_in, float* loglik,
                             float* alpha, cudaStream_t st, int64_t seqT = 0) {
  hw::Args a{};
  a.log_pi = log_pi; a.A = A; a.Bm = Bm; a.lengths = lengths; a.B = rows; a.T = T;
  a.S = 4 * (int)cdiv(T, 128);
  a.SA = hw::pitch_odd16(a.S * (int)(K * K));
  a.SB = hw::pitch_odd16(a.S * (int)K);
  const uint64_t ga = (uint64_t)a.S * K * K / 4, gb = (uint64_t)a.S * K / 4;
  a.magicA = (unsigned)(((1ull << 32) + ga - 1) / ga);
  a.magicB = (unsigned)(((1ull << 32) + gb - 1) / gb);
  a.loglik = loglik; a.gamma = alpha; a.xi = nullptr;
  a.nseg = nseg; a.vin = vin; a.ll_in = ll_in; a.seqT = seqT;
  int rc = 0;
  DISPATCH_K(K, (rc = launch_onepass<KK, false, false, false, true>(a, st)));
  return rc;
}

static hs::Args base_args(const float* log_pi, const float* A, const float* Bm, const int64_t* lengths,
                          int64_t B, int64_t T, int64_t K) {
  hs::Args a{};
  a.log_pi = log_pi; a.A = A; a.Bm = Bm; a.lengths = lengths;
  a.nrows = B; a.T = T; a.rows_per_seq = 1; a.CH = (int)T;
  const bool tk4 = ((T * K) & 3) == 0;
  a.vec_in = tk4 && aligned16(A) && aligned16(Bm);
  a.vec_out = tk4;
  a.vec_xi = tk4;
  a.full = 0;
  return a;
}

// K > 4: the warp-per-sequence kernels of hmm_wide.cuh, instantiated for lane counts 8 / 16 / 32
// (VHMM_WIDE=generic selects the first-round log-space kernels of hmm_generic.cuh, kept for comparison)
static bool wide_allowed() {
  static const int v = [] { const char* e = getenv("VHMM_WIDE"); return (e && e[0] == 'g') ? 0 : 1; }();
  return v != 0 && g_numerics.load(std::memory_order_relaxed) == 0;
}
#define DISPATCH_KP(K, expr)                                         \
  do {                                                               \
    if ((K) <= 8) { constexpr int KP = 8; expr; }                    \
    else if ((K) <= 16) { constexpr int KP = 16; expr; }             \
    else { constexpr int KP = 32; expr; }                            \
  } while (0)

static bool wide_scan_enabled() {            // environment VHMM_WIDE_SCAN=off keeps the warp-per-sequence forward for K > 4
  static const bool on = [] { const char* e = getenv("VHMM_WIDE_SCAN"); return !(e && e[0] == 'o'); }();
  return on;
}

static bool alpha_onepass_enabled() {        // environment VHMM_ALPHA=sweep keeps the thread-per-row sweeps for all-prefix alpha
  static const bool on = [] { const char* e = getenv("VHMM_ALPHA"); return !(e && e[0] == 's'); }();
  return on;
}

int hmm_forward(const float* log_pi, const float* A, const float* Bm, const int64_t* lengths, int64_t B,
                int64_t T, int64_t K, float* loglik, float* alpha, cudaStream_t st) {
  // log-likelihood only: the time-parallel warp kernel (hmm_warp.cuh) without its sweeps
  if (!alpha && loglik && onepass_allowed() && onepass_fits(T, K, A, Bm, A, nullptr))
    return hmm_posterior_onepass(log_pi, A, Bm, lengths, B, T, K, loglik, nullptr, nullptr, st);
  // all-prefix alpha of sequences that fit one warp's shared memory: the same kernel with its alpha sweep (inputs read once,
  // alpha written once, time-parallel over the lanes) instead of the thread-per-sequence sweep below
  if (alpha && onepass_allowed() && alpha_onepass_enabled() && onepass_fits(T, K, A, Bm, alpha, nullptr))
    return hmm_alpha_onepass(log_pi, A, Bm, lengths, B, T, K, 1, nullptr, nullptr, loglik, alpha, st);
  if (small_k(K)) {
    hs::Args a = base_args(log_pi, A, Bm, lengths, B, T, K);
    a.loglik = loglik; a.alpha_log = alpha;
    a.vec_out = a.vec_out && (!alpha || aligned16(alpha));
    a.full = alpha ? 1 : 0;
    int rc = 0;
    DISPATCH_K(K, (rc = launch_fwd_small<KK, true, false>(a, st)));
    return rc;
  }
  VHMM_REQUIRE(K <= 32, VHMM_E_DIMS, "hmm_forward: K=%lld > 32", (long long)K);
  const int wpb = 4;
  if (wide_allowed()) {
    DISPATCH_KP(K, (hx::fwd_kernel<KP><<<(unsigned)cdiv(B, hx::WPB), hx::WPB * 32, 0, st>>>(
                       log_pi, A, Bm, lengths, B, (int)T, (int)K, loglik, alpha, nullptr, hx::RowView{})));
    VHMM_LAUNCH_CHECK();
    return 0;
  }
  hg::fwd_kernel<false><<<(unsigned)cdiv(B, wpb), wpb * 32, 0, st>>>(log_pi, A, Bm, lengths, B, (int)T, (int)K,
                                                                     loglik, alpha, 0, nullptr, nullptr, nullptr);
  VHMM_LAUNCH_CHECK();
  return 0;
}

// gamma (and xi) of LONG sequences, time-parallel: 256-step segments -- segment products (one-pass kernel), forward carry scan
// (incoming alpha vector of every segment, log-likelihood), backward carry scan (beta after every segment), then the one-pass
// posterior kernel on the segment rows.  The sweep kernels below run ONE THREAD per sequence: at B = 64, T = 65536 they take
// 94 ms (3 GB/s); this path reads the inputs twice at HBM speed.
template <int K>
static int posterior_segments(const float* log_pi, const float* A, const float* Bm, const int64_t* lengths, int64_t B, int64_t T,
                              float* loglik, float* gamma, float* xi, void* ws, size_t ws_bytes, cudaStream_t st) {
  constexpr int64_t SEG_T = 256;
  const int64_t nseg = cdiv(T, SEG_T), rows = B * nseg;     // (the last row of a sequence may be shorter: Args.seqT)
  Arena ar(ws, ws_bytes);
  hs::ScanArgs sa{};
  sa.Pmat = ar.take<float>((size_t)rows * K * K);
  sa.Pscale = ar.take<double>((size_t)rows);
  sa.vin = ar.take<float>((size_t)rows * K);
  sa.ll_in = ar.take<double>((size_t)rows);
  sa.suspect = ar.take<int>((size_t)B);
  float* bin = ar.take<float>((size_t)rows * K);
  float* ll_tmp = ar.take<float>((size_t)B);
  VHMM_REQUIRE(ar.ok, VHMM_E_WORKSPACE, "hmm_posterior: workspace %zu < %zu", ws_bytes, ar.used);
  sa.loglik = loglik ? loglik : ll_tmp;
  hw::Args w{};
  w.log_pi = log_pi; w.A = A; w.Bm = Bm; w.lengths = lengths; w.B = rows; w.T = SEG_T;
  w.S = 4 * (int)cdiv(SEG_T, 128);
  w.SA = hw::pitch_odd16(w.S * (int)(K * K));
  w.SB = hw::pitch_odd16(w.S * (int)K);
  const uint64_t ga = (uint64_t)w.S * K * K / 4, gb = (uint64_t)w.S * K / 4;
  w.magicA = (unsigned)(((1ull << 32) + ga - 1) / ga);
  w.magicB = (unsigned)(((1ull << 32) + gb - 1) / gb);
  w.nseg = (int)nseg; w.seqT = T; w.Pmat = sa.Pmat; w.Pscale = sa.Pscale;
  int rc = launch_onepass<K, false, true, true>(w, st);                 // segment products
  if (rc) return rc;
  hs::Args c = base_args(log_pi, A, Bm, lengths, B, T, K);
  c.nrows = rows; c.rows_per_seq = (int)nseg; c.CH = (int)SEG_T;
  hs::scan_carry_kernel<K><<<(unsigned)B, 32 * hs::CARRY_W, 0, st>>>(c, sa);
  VHMM_LAUNCH_CHECK();
  hs::scan_fixup_kernel<K><<<(unsigned)B, 32, 0, st>>>(c, sa, nullptr);   // log-likelihood of sequences whose products vanished
  VHMM_LAUNCH_CHECK();
  hs::scan_carry_back_kernel<K><<<(unsigned)B, 32, 0, st>>>(c, sa, bin);
  VHMM_LAUNCH_CHECK();
  hw::Args p = w;
  p.Pmat = nullptr; p.Pscale = nullptr; p.loglik = nullptr;
  p.gamma = gamma; p.xi = xi; p.vin = sa.vin; p.ll_in = sa.ll_in; p.bin_in = bin;
  return xi ? launch_onepass<K, true>(p, st) : launch_onepass<K, false>(p, st);
}

int hmm_posterior(const float* log_pi, const float* A, const float* Bm, const int64_t* lengths, int64_t B,
                  int64_t T, int64_t K, float* loglik, float* gamma, float* xi, void* ws, size_t ws_bytes,
                  cudaStream_t st) {
  if (onepass_allowed() && onepass_fits(T, K, A, Bm, gamma, xi))
    return hmm_posterior_onepass(log_pi, A, Bm, lengths, B, T, K, loglik, gamma, xi, st);
  if (onepass_allowed() && T >= 512 && ((T * K) & 3) == 0 && B * cdiv(T, 256) < (1ll << 31) && onepass_fits(256, K, A, Bm, gamma, xi) &&
      ws_bytes >= (size_t)B * cdiv(T, 256) * (K * K + 2 * K + 6) * 4 + (size_t)B * 8 + 4096) {
    int rc = 0;
    DISPATCH_K(K, (rc = posterior_segments<KK>(log_pi, A, Bm, lengths, B, T, loglik, gamma, xi, ws, ws_bytes, st)));
    return rc;
  }
  Arena ar(ws, ws_bytes);
  float* stash = ar.take<float>((size_t)B * T * K);
  VHMM_REQUIRE(ar.ok, VHMM_E_WORKSPACE, "hmm_posterior: workspace %zu < %zu", ws_bytes, ar.used);
  if (small_k(K)) {
    hs::Args a = base_args(log_pi, A, Bm, lengths, B, T, K);
    a.loglik = loglik; a.ahat = stash; a.gamma = gamma; a.xi = xi;
    a.vec_out = a.vec_out && aligned16(stash) && aligned16(gamma);
    a.vec_xi = a.vec_xi && (!xi || aligned16(xi));
    int rc = 0;
    DISPATCH_K(K, (rc = launch_fwd_small<KK, true, false>(a, st)));
    if (rc) return rc;
    DISPATCH_K(K, (rc = launch_bwd_small<KK>(a, st)));
    return rc;
  }
  VHMM_REQUIRE(K <= 32, VHMM_E_DIMS, "hmm_posterior: K=%lld > 32", (long long)K);
  const int wpb = 4;
  if (wide_allowed() && wide_scan_enabled() && K > 4 && T >= 512 && B <= 256) {
    // few long sequences: segment products + forward / backward carries (hmm_wide_scan.cuh), then the same forward and
    // backward kernels on 128-step ROWS (each from its incoming alpha_hat / outgoing beta_hat) -- B*T/128 warps instead of B
    const int64_t nseg = cdiv(T, (int64_t)hx::WS_SEG), rows = B * nseg;
    float* Pmat = ar.take<float>((size_t)rows * K * K);
    int* Pex = ar.take<int>((size_t)rows * K);
    float* mbs = ar.take<float>((size_t)rows);
    float* vin = ar.take<float>((size_t)rows * K);
    float* bin = ar.take<float>((size_t)rows * K);
    int* suspect = ar.take<int>((size_t)B);
    if (ar.ok) {       // (a caller's workspace sized for the stash alone takes the warp-per-sequence kernels below)
      DISPATCH_KP(K, (hx::seg_product_kernel<KP><<<(unsigned)cdiv(rows, hx::WS_WPB), hx::WS_WPB * 32, 0, st>>>(
                         A, Bm, lengths, rows, (int)nseg, (int)T, (int)K, Pmat, Pex, mbs)));
      VHMM_LAUNCH_CHECK();
      hx::seg_carry_kernel<<<(unsigned)cdiv(B, hx::WPB), hx::WPB * 32, 0, st>>>(log_pi, Bm, lengths, B, (int)nseg, (int)T, (int)K,
                                                                              Pmat, Pex, mbs, loglik, suspect, vin, nullptr, nullptr);
      VHMM_LAUNCH_CHECK();
      hx::seg_carry_back_kernel<<<(unsigned)cdiv(B, hx::WPB), hx::WPB * 32, 0, st>>>(lengths, B, (int)nseg, (int)T, (int)K, Pmat, Pex,
                                                                                   bin, suspect);
      VHMM_LAUNCH_CHECK();
      const hx::RowView rv{(int)nseg, hx::WS_SEG, vin, bin, nullptr, nullptr, nullptr};
      const hx::RowView redo{0, 0, nullptr, nullptr, suspect, nullptr, nullptr};
      DISPATCH_KP(K, (hx::fwd_kernel<KP, true><<<(unsigned)cdiv(rows, hx::WPB), hx::WPB * 32, 0, st>>>(
                         log_pi, A, Bm, lengths, rows, (int)T, (int)K, nullptr, nullptr, stash, rv)));
      VHMM_LAUNCH_CHECK();
      if (xi) {
        DISPATCH_KP(K, (hx::bwd_kernel<KP, true, true><<<(unsigned)cdiv(rows, hx::WPB), hx::WPB * 32, 0, st>>>(
                           A, Bm, lengths, rows, (int)T, (int)K, stash, gamma, xi, rv)));
      } else {
        DISPATCH_KP(K, (hx::bwd_kernel<KP, false, true><<<(unsigned)cdiv(rows, hx::WPB), hx::WPB * 32, 0, st>>>(
                           A, Bm, lengths, rows, (int)T, (int)K, stash, gamma, nullptr, rv)));
      }
      VHMM_LAUNCH_CHECK();
      // sequences whose carries vanished (flagged): the warp-per-sequence kernels again, for those only
      DISPATCH_KP(K, (hx::fwd_kernel<KP><<<(unsigned)cdiv(B, hx::WPB), hx::WPB * 32, 0, st>>>(
                         log_pi, A, Bm, lengths, B, (int)T, (int)K, loglik, nullptr, stash, redo)));
      VHMM_LAUNCH_CHECK();
      if (xi) {
        DISPATCH_KP(K, (hx::bwd_kernel<KP, true><<<(unsigned)cdiv(B, hx::WPB), hx::WPB * 32, 0, st>>>(
                           A, Bm, lengths, B, (int)T, (int)K, stash, gamma, xi, redo)));
      } else {
        DISPATCH_KP(K, (hx::bwd_kernel<KP, false><<<(unsigned)cdiv(B, hx::WPB), hx::WPB * 32, 0, st>>>(
                           A, Bm, lengths, B, (int)T, (int)K, stash, gamma, nullptr, redo)));
      }
      VHMM_LAUNCH_CHECK();
      return 0;
    }
  }
  if (wide_allowed()) {
    DISPATCH_KP(K, (hx::fwd_kernel<KP><<<(unsigned)cdiv(B, hx::WPB), hx::WPB * 32, 0, st>>>(
                       log_pi, A, Bm, lengths, B, (int)T, (int)K, loglik, nullptr, stash, hx::RowView{})));
    VHMM_LAUNCH_CHECK();
    if (xi) {
      DISPATCH_KP(K, (hx::bwd_kernel<KP, true><<<(unsigned)cdiv(B, hx::WPB), hx::WPB * 32, 0, st>>>(
                         A, Bm, lengths, B, (int)T, (int)K, stash, gamma, xi, hx::RowView{})));
    } else {
      DISPATCH_KP(K, (hx::bwd_kernel<KP, false><<<(unsigned)cdiv(B, hx::WPB), hx::WPB * 32, 0, st>>>(
                         A, Bm, lengths, B, (int)T, (int)K, stash, gamma, nullptr, hx::RowView{})));
    }
    VHMM_LAUNCH_CHECK();
    return 0;
  }
  hg::fwd_kernel<false><<<(unsigned)cdiv(B, wpb), wpb * 32, 0, st>>>(log_pi, A, Bm, lengths, B, (int)T, (int)K,
                                                                     loglik, stash, 1, nullptr, nullptr, nullptr);
  VHMM_LAUNCH_CHECK();
  hg::bwd_kernel<<<(unsigned)cdiv(B, wpb), wpb * 32, 0, st>>>(A, Bm, lengths, B, (int)T, (int)K, stash, gamma, xi);
  VHMM_LAUNCH_CHECK();
  return 0;
}

template <int KP>
static int launch_vit_ring(const float* log_pi, const float* A, const float* Bm, const int64_t* lengths, int64_t B, int64_t T,
                           int64_t K, uint8_t* psi, uint8_t* path, float* score, cudaStream_t st) {
  auto kern = (K == KP) ? hx::vit_ring_kernel<KP, true> : hx::vit_ring_kernel<KP, false>;
  VHMM_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hx::vit_ring_smem<KP>()));
  kern<<<(unsigned)B, 32, hx::vit_ring_smem<KP>(), st>>>(log_pi, A, Bm, lengths, B, (int)T, (int)K, psi, path, score);
  VHMM_LAUNCH_CHECK();
  return 0;
}

int hmm_viterbi(const float* log_pi, const float* A, const float* Bm, const int64_t* lengths, int64_t B,
                int64_t T, int64_t K, uint8_t* path, float* score, void* ws, size_t ws_bytes,
                cudaStream_t st) {
  Arena ar(ws, ws_bytes);
  uint8_t* psi = ar.take<uint8_t>((size_t)B * T * psi_bytes(K));
  VHMM_REQUIRE(ar.ok, VHMM_E_WORKSPACE, "viterbi: workspace %zu < %zu", ws_bytes, ar.used);
  if (small_k(K)) {
    hs::Args a = base_args(log_pi, A, Bm, lengths, B, T, K);
    a.psi = psi; a.path = path; a.vec_path = ((T & 3) == 0) && (((uintptr_t)path & 3) == 0); a.score = score;
    int rc = 0;
    DISPATCH_K(K, (rc = launch_vit_small<KK>(a, st)));
    return rc;
  }
  VHMM_REQUIRE(K <= 32, VHMM_E_DIMS, "hmm_viterbi: K=%lld > 32", (long long)K);
  const int wpb = 4;
  if (wide_allowed() && wide_scan_enabled() && (K & 3) == 0 && B <= 148 && T >= 256 && aligned16(A) && aligned16(Bm)) {
    // few sequences: one warp per CTA, inputs through a bulk-copy ring (hx::vit_ring_kernel) -- bit-identical results
    int rc = 0;
    DISPATCH_KP(K, (rc = launch_vit_ring<KP>(log_pi, A, Bm, lengths, B, T, K, psi, path, score, st)));
    return rc;
  }
  if (wide_allowed()) {
    DISPATCH_KP(K, (hx::vit_kernel<KP><<<(unsigned)cdiv(B, hx::WPB), hx::WPB * 32, 0, st>>>(
                       log_pi, A, Bm, lengths, B, (int)T, (int)K, psi, path, score)));
    VHMM_LAUNCH_CHECK();
    return 0;
  }
  hg::fwd_kernel<true><<<(unsigned)cdiv(B, wpb), wpb * 32, 0, st>>>(log_pi, A, Bm, lengths, B, (int)T, (int)K,
                                                                    nullptr, nullptr, 0, psi, path, score);
  VHMM_LAUNCH_CHECK();
  return 0;
}

// ---- allocation-free Viterbi (include/vhmm.h: vhmm_viterbi_static) ------------------------------------------------------
// The back-pointer scratch of hmm_viterbi lives in a file-scope device array instead of a caller workspace: nothing is
// allocated by the call (usable where cudaMalloc is off limits).  Every GPU that loads the library holds the array until
// exit.  In HOST code the array's name is only a handle: the device address comes from cudaGetSymbolAddress, resolved once per
// device with that device current (below), and THAT pointer goes to the kernels.  All calls share the array -- calls that
// overlap on the GPU (two streams) race on it; nothing in it has to start at zero (every back-pointer that the back-track
// reads was written by the same call's forward sweep).
static __device__ __align__(256) uint8_t g_viterbi_scratch[VHMM_VITERBI_STATIC_MAX_BYTES];
static std::atomic<void*> g_viterbi_scratch_ptr[VHMM_MAX_DEVICES];

int hmm_viterbi_static(const float* log_pi, const float* A, const float* Bm, const int64_t* lengths, int64_t B,
                       int64_t T, int64_t K, uint8_t* path, float* score, cudaStream_t st) {
  VHMM_REQUIRE(K >= 1 && K <= 32, VHMM_E_DIMS, "viterbi_static: K=%lld outside [1, 32]", (long long)K);
  const int dev = current_device();
  void* ptr = g_viterbi_scratch_ptr[dev].load(std::memory_order_acquire);
  if (!ptr) {   // set-up, once per device (a call with B == 0 does only this: e.g. before a stream capture begins)
    VHMM_CUDA_CHECK(cudaGetSymbolAddress(&ptr, g_viterbi_scratch));
    VHMM_REQUIRE(ptr != nullptr, VHMM_E_WORKSPACE, "viterbi_static: no device address for the scratch array");
    g_viterbi_scratch_ptr[dev].store(ptr, std::memory_order_release);
  }
  if (B == 0) return 0;
  // (B, T <= 2^31 each, checked by the caller: the product below cannot overflow)
  const uint64_t need = (uint64_t)B * (uint64_t)T * (uint64_t)psi_bytes(K);
  VHMM_REQUIRE(need <= (uint64_t)VHMM_VITERBI_STATIC_MAX_BYTES, VHMM_E_WORKSPACE,
               "viterbi_static: B*T*%d = %llu bytes of back-pointers exceed the static scratch of %llu bytes "
               "(VHMM_VITERBI_STATIC_MAX_BYTES); use vhmm_viterbi with a caller workspace",
               (int)psi_bytes(K), (unsigned long long)need, (unsigned long long)VHMM_VITERBI_STATIC_MAX_BYTES);
  return hmm_viterbi(log_pi, A, Bm, lengths, B, T, K, path, score, ptr, (size_t)VHMM_VITERBI_STATIC_MAX_BYTES, st);
}

int hmm_decode(const float* log_pi, const float* A, const float* Bm, const int64_t* lengths, int64_t B,
               int64_t T, int64_t K, uint8_t* path, float* gamma, float* loglik, void* ws, size_t ws_bytes,
               cudaStream_t st) {
  Arena ar(ws, ws_bytes);
  uint8_t* psi = ar.take<uint8_t>((size_t)B * T * psi_bytes(K));
  float* stash = ar.take<float>((size_t)B * T * K);      // last: the posterior's region is the stash and whatever follows it
  VHMM_REQUIRE(ar.ok, VHMM_E_WORKSPACE, "hmm_decode: workspace %zu < %zu", ws_bytes, ar.used);
  if (small_k(K) && onepass_allowed() && decode_onekernel() && onepass_fits(T, K, A, Bm, gamma, nullptr)) {
    // ONE kernel, one read of the inputs (48 B/step read, 13 written): the one-pass posterior kernel with a Viterbi helper warp
    hw::Args w{};
    w.log_pi = log_pi; w.A = A; w.Bm = Bm; w.lengths = lengths; w.B = B; w.T = T;
    w.S = 4 * (int)cdiv(T, 128);
    w.SA = hw::pitch_odd16(w.S * (int)(K * K));
    w.SB = hw::pitch_odd16(w.S * (int)K);
    const uint64_t ga = (uint64_t)w.S * K * K / 4, gb = (uint64_t)w.S * K / 4;
    w.magicA = (unsigned)(((1ull << 32) + ga - 1) / ga);
    w.magicB = (unsigned)(((1ull << 32) + gb - 1) / gb);
    w.loglik = loglik; w.gamma = gamma; w.xi = nullptr;
    w.path = path; w.path_vec16 = ((T & 15) == 0) && (((uintptr_t)path & 15) == 0);
    int rc = 0;
    DISPATCH_K(K, (rc = launch_decode_onepass<KK>(w, st)));
    return rc;
  }
  if (small_k(K) && onepass_allowed() && onepass_fits(T, K, A, Bm, gamma, nullptr)) {
    // Viterbi sweep (49 B/step) + one-pass posterior (60 B/step): 109 B/step of traffic against
    // 181 B/step for the fused forward sweep + backward sweep below
    hs::Args a = base_args(log_pi, A, Bm, lengths, B, T, K);
    a.psi = psi; a.path = path; a.vec_path = ((T & 3) == 0) && (((uintptr_t)path & 3) == 0); a.score = nullptr;
    int rc = 0;
    DISPATCH_K(K, (rc = launch_vit_small<KK>(a, st)));
    if (rc) return rc;
    return hmm_posterior_onepass(log_pi, A, Bm, lengths, B, T, K, loglik, gamma, nullptr, st);
  }
  if (small_k(K) && onepass_allowed() && T >= 512 && ((T * K) & 3) == 0 && B * cdiv(T, 256) < (1ll << 31) &&
      onepass_fits(256, K, A, Bm, gamma, nullptr)) {
    // long sequences: the Viterbi sweep (serial in time by nature: bit-exact max-plus) + the time-parallel posterior
    hs::Args a = base_args(log_pi, A, Bm, lengths, B, T, K);
    a.psi = psi; a.path = path; a.vec_path = ((T & 3) == 0) && (((uintptr_t)path & 3) == 0); a.score = nullptr;
    int rc = 0;
    DISPATCH_K(K, (rc = launch_vit_small<KK>(a, st)));
    if (rc) return rc;
    DISPATCH_K(K, (rc = posterior_segments<KK>(log_pi, A, Bm, lengths, B, T, loglik, gamma, nullptr, stash,
                                               (size_t)B * T * K * sizeof(float), st)));
    return rc;
  }
  if (small_k(K)) {
    // one forward sweep computes alpha_hat AND the Viterbi back-pointers from the same staged
    // inputs; the backward sweep then produces gamma.
    hs::Args a = base_args(log_pi, A, Bm, lengths, B, T, K);
    a.loglik = loglik; a.ahat = stash; a.gamma = gamma; a.xi = nullptr;
    a.psi = psi; a.path = path; a.vec_path = ((T & 3) == 0) && (((uintptr_t)path & 3) == 0); a.score = nullptr;
    a.vec_out = a.vec_out && aligned16(stash) && aligned16(gamma);
    int rc = 0;
    DISPATCH_K(K, (rc = launch_fwd_small<KK, true, true>(a, st)));
    if (rc) return rc;
    DISPATCH_K(K, (rc = launch_bwd_small<KK>(a, st)));
    return rc;
  }
  int rc = hmm_viterbi(log_pi, A, Bm, lengths, B, T, K, path, nullptr, psi,
                       align_up((size_t)B * T * psi_bytes(K), 256), st);
  if (rc) return rc;
  return hmm_posterior(log_pi, A, Bm, lengths, B, T, K, loglik, gamma, nullptr, stash,
                       ws_bytes - ((size_t)((char*)stash - (char*)ws)), st);
}

template <int K>
static int scan_small(hs::Args a, hs::ScanArgs sa, float* alpha, int64_t B, cudaStream_t st) {
  using TL = hs::Tile<K, HS_TC, HS_NT>;
  auto k1 = hs::scan_local_kernel<K, HS_TC, HS_NT>;
  VHMM_CUDA_CHECK(cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TL::SMEM));
  k1<<<(unsigned)cdiv(a.nrows, HS_NT), HS_NT, TL::SMEM, st>>>(a, sa);
  VHMM_LAUNCH_CHECK();
  hs::scan_carry_kernel<K><<<(unsigned)B, 32 * hs::CARRY_W, 0, st>>>(a, sa);
  VHMM_LAUNCH_CHECK();
  if (alpha) {
    hs::Args a3 = a;
    a3.vin = sa.vin; a3.ll_in = sa.ll_in; a3.alpha_log = alpha; a3.full = 1; a3.loglik = nullptr;
    int rc = launch_fwd_small<K, true, false>(a3, st);
    if (rc) return rc;
  }
  hs::scan_fixup_kernel<K><<<(unsigned)B, 32, 0, st>>>(a, sa, alpha);   // sequences whose products vanished only
  VHMM_LAUNCH_CHECK();
  return 0;
}

int hmm_scan(const float* log_pi, const float* A, const float* Bm, const int64_t* lengths, int64_t B,
             int64_t T, int64_t K, float* loglik, float* alpha, void* ws, size_t ws_bytes, cudaStream_t st) {
  if (!small_k(K)) {
    // 4 < K <= 32, log-likelihood of few long sequences: segment products + carry (hmm_wide_scan.cuh); K^3 work per step
    // instead of K^2, but over B*T/128 warps -- it wins while the batch alone cannot fill the machine
    if (loglik && K > 4 && K <= 32 && wide_allowed() && wide_scan_enabled() && T >= 512 && B <= 256) {
      const int64_t nseg = cdiv(T, (int64_t)hx::WS_SEG), rows = B * nseg;
      Arena ar(ws, ws_bytes);
      float* Pmat = ar.take<float>((size_t)rows * K * K);
      int* Pex = ar.take<int>((size_t)rows * K);
      float* mbs = ar.take<float>((size_t)rows);
      int* suspect = ar.take<int>((size_t)B);
      float* vin = alpha ? ar.take<float>((size_t)rows * K) : nullptr;
      double* ll_in = alpha ? ar.take<double>((size_t)rows) : nullptr;
      float* carried = alpha ? ar.take<float>((size_t)B * K) : nullptr;
      VHMM_REQUIRE(ar.ok, VHMM_E_WORKSPACE, "hmm_scan: workspace %zu < %zu", ws_bytes, ar.used);
      DISPATCH_KP(K, (hx::seg_product_kernel<KP><<<(unsigned)cdiv(rows, hx::WS_WPB), hx::WS_WPB * 32, 0, st>>>(
                         A, Bm, lengths, rows, (int)nseg, (int)T, (int)K, Pmat, Pex, mbs)));
      VHMM_LAUNCH_CHECK();
      hx::seg_carry_kernel<<<(unsigned)cdiv(B, hx::WPB), hx::WPB * 32, 0, st>>>(log_pi, Bm, lengths, B, (int)nseg, (int)T, (int)K,
                                                                              Pmat, Pex, mbs, loglik, suspect, vin, ll_in, carried);
      VHMM_LAUNCH_CHECK();
      if (alpha) {   // all-prefix log alpha: the forward kernel on the 128-step rows, each from its incoming vector and log-scale
        const hx::RowView rv{(int)nseg, hx::WS_SEG, vin, nullptr, nullptr, ll_in, carried};
        DISPATCH_KP(K, (hx::fwd_kernel<KP, true><<<(unsigned)cdiv(rows, hx::WPB), hx::WPB * 32, 0, st>>>(
                           log_pi, A, Bm, lengths, rows, (int)T, (int)K, nullptr, alpha, nullptr, rv)));
        VHMM_LAUNCH_CHECK();
      }
      const hx::RowView redo{0, 0, nullptr, nullptr, suspect, nullptr, nullptr};
      DISPATCH_KP(K, (hx::fwd_kernel<KP><<<(unsigned)cdiv(B, hx::WPB), hx::WPB * 32, 0, st>>>(
                         log_pi, A, Bm, lengths, B, (int)T, (int)K, loglik, alpha, nullptr, redo)));   // flagged sequences only
      VHMM_LAUNCH_CHECK();
      return 0;
    }
    return hmm_forward(log_pi, A, Bm, lengths, B, T, K, loglik, alpha, st);
  }
  // a whole sequence fits one warp of the one-pass kernel (T up to ~580): that IS the time-parallel path, with no carry scan
  // (B = 2^20, T = 256 through the chunk kernels below: 11.8 ms; one-pass: 2.2 ms)
  if (onepass_allowed() && T < 512 && loglik && (!alpha || alpha_onepass_enabled()) &&
      onepass_fits(T, K, A, Bm, alpha ? (const void*)alpha : (const void*)A, nullptr))
    return hmm_forward(log_pi, A, Bm, lengths, B, T, K, loglik, alpha, st);
  const int64_t nck = cdiv(T, SCAN_CH);
  Arena ar(ws, ws_bytes);
  hs::ScanArgs sa{};
  sa.Pmat = ar.take<float>((size_t)B * nck * K * K);
  sa.Pscale = ar.take<double>((size_t)B * nck);
  sa.vin = ar.take<float>((size_t)B * nck * K);
  sa.ll_in = ar.take<double>((size_t)B * nck);
  sa.suspect = ar.take<int>((size_t)B);
  sa.loglik = loglik;
  VHMM_REQUIRE(ar.ok, VHMM_E_WORKSPACE, "hmm_scan: workspace %zu < %zu", ws_bytes, ar.used);
  // log-likelihood only, T a multiple of the segment length: segment products by the one-pass kernel (one warp per
  // 256-step segment, inputs resident in shared memory, ~HBM speed), then the per-sequence carry scan
  constexpr int64_t SEG_T = 256;
  if ((!alpha || (alpha_onepass_enabled() && aligned16(alpha))) && onepass_allowed() && T >= 2 * SEG_T && ((T * K) & 3) == 0 &&
      B * cdiv(T, SEG_T) < (1ll << 31) && cdiv(T, SEG_T) <= nck && onepass_fits(SEG_T, K, A, Bm, nullptr, nullptr)) {
    const int64_t nseg = cdiv(T, SEG_T);
    hw::Args w{};
    w.log_pi = log_pi; w.A = A; w.Bm = Bm; w.lengths = lengths; w.B = B * nseg; w.T = SEG_T;
    w.S = 4 * (int)cdiv(SEG_T, 128);
    w.SA = hw::pitch_odd16(w.S * (int)(K * K));
    w.SB = hw::pitch_odd16(w.S * (int)K);
    const uint64_t ga = (uint64_t)w.S * K * K / 4, gb = (uint64_t)w.S * K / 4;
    w.magicA = (unsigned)(((1ull << 32) + ga - 1) / ga);
    w.magicB = (unsigned)(((1ull << 32) + gb - 1) / gb);
    w.nseg = (int)nseg; w.seqT = T; w.Pmat = sa.Pmat; w.Pscale = sa.Pscale;
    int rc = 0;
    DISPATCH_K(K, (rc = launch_onepass<KK, false, true, true>(w, st)));
    if (rc) return rc;
    hs::Args c = base_args(log_pi, A, Bm, lengths, B, T, K);
    c.nrows = B * nseg; c.rows_per_seq = (int)nseg; c.CH = (int)SEG_T;
    DISPATCH_K(K, (hs::scan_carry_kernel<KK><<<(unsigned)B, 32 * hs::CARRY_W, 0, st>>>(c, sa), rc = 0));
    VHMM_LAUNCH_CHECK();
    if (alpha) {   // all-prefix alpha: the segments again, each from its incoming vector (second and last read of the inputs)
      rc = hmm_alpha_onepass(log_pi, A, Bm, lengths, B * nseg, SEG_T, K, (int)nseg, sa.vin, sa.ll_in, nullptr, alpha, st, T);
      if (rc) return rc;
    }
    DISPATCH_K(K, (hs::scan_fixup_kernel<KK><<<(unsigned)B, 32, 0, st>>>(c, sa, alpha), rc = 0));   // vanished products only
    VHMM_LAUNCH_CHECK();
    return rc;
  }
  hs::Args a = base_args(log_pi, A, Bm, lengths, B, T, K);
  a.nrows = B * nck; a.rows_per_seq = (int)nck; a.CH = SCAN_CH;
  a.vec_out = a.vec_out && (!alpha || aligned16(alpha));
  int rc = 0;
  DISPATCH_K(K, (rc = scan_small<KK>(a, sa, alpha, B, st)));
  return rc;
}

}  // namespace vhmm
