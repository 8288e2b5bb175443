// elbo_layerwise.cu -- the layer-wise (activation-stashing) implementation of the pinned path
// for ARBITRARY model dimensions: Encoder / Prior / Decoder forward and the mean-field ELBO
// with its full backward.  The batch is processed in chunks that fit the caller's workspace.
// Reference: VQ_VAE_HMM_fixed.py:31-143; maths in SURVEY.md Appendix A.2.
#include <atomic>
#include <stdlib.h>
#include "conv_generic.cuh"
#include "conv_tc5.cuh"
#include "prior_tc5.cuh"
#include "pointwise.cuh"

namespace vhmm {

static const int64_t kMaxChunk = 32768;  // gridDim.y/z limit is 65535

// layer kernels of the layer-wise path: 0 auto (tcgen05 where the layer shape allows: Cin % 16 == 0, Cout % 128 == 0,
// T % 4 == 0), 1 FP32 FFMA only, 2 tcgen05 required (error otherwise); environment VHMM_CONV=ffma|tc5 seeds the default
static std::atomic<int> g_conv_impl{-1};
int conv_impl_mode() {
  int v = g_conv_impl.load(std::memory_order_relaxed);
  if (v < 0) {
    const char* e = getenv("VHMM_CONV");
    v = (e && e[0] == 'f') ? 1 : ((e && e[0] == 't') ? 2 : 0);
    g_conv_impl.store(v, std::memory_order_relaxed);
  }
  return v;
}
int conv_impl_select(int mode) { g_conv_impl.store(mode, std::memory_order_relaxed); return 0; }

size_t conv1d_workspace_bytes(int Cin, int Cout, int kw) { return conv_tc5_weight_bytes(Cin, Cout, kw) + 256; }

// one Conv1d layer (kernel width 1 or 3, padding kw/2): y = [relu](bias + conv(x, w)); w (Cout,Cin,kw)
int conv1d_fwd(const float* x, const float* w, const float* bias, int64_t B, int Cin, int Cout, int64_t T, int kw, int relu,
               float* y, void* ws, size_t ws_bytes, cudaStream_t st) {
  const int mode = conv_impl_mode();
  const bool ok = conv_tc5_ok(Cin, Cout, T, kw, x, y);
  VHMM_REQUIRE(mode != 2 || ok, VHMM_E_DIMS, "conv1d: tensor-core layer kernel needs Cin %% 16 == 0, Cout %% 16 == 0, T %% 4 == 0");
  if (mode != 1 && ok) {
    VHMM_REQUIRE(ws && ws_bytes >= conv_tc5_weight_bytes(Cin, Cout, kw), VHMM_E_WORKSPACE, "conv1d: workspace too small");
    Tc5Weights W;
    int rc = conv_tc5_prepare(wview_fwd(w, Cin, kw), (int64_t)Cin * Cout * kw, Cin, Cout, kw, ws, &W, st);
    if (rc) return rc;
    return launch_conv_tc5(x, W, bias, nullptr, y, B, Cin, Cout, T, kw, relu ? EPI_RELU : EPI_NONE, st);
  }
  return launch_conv(x, wview_fwd(w, Cin, kw), bias, nullptr, y, B, Cin, Cout, T, kw, relu ? EPI_RELU : EPI_NONE, st);
}

// ---- tensor-core layer kernels inside the layer-wise path (conv_tc5.cuh) -------------------------------------
// The weights of the layers whose shape the tcgen05 kernels take are prepared once per call (fp16 parts in the
// kernel's chunk image): forward view and, where the input gradient is needed, the transposed tap-flipped view.
// transition prior of the wide model: 0 fused kernel + operand-form gradient, 1 fused kernel + fp32 gradient, 2 layer by layer
// (include/vhmm.h: vhmm_prior_select; environment VHMM_PRIOR_FUSED=0 / VHMM_PRIOR_PARTS=0 seed it)
static std::atomic<int> g_prior_mode{-1};
static int prior_mode() {
  int m = g_prior_mode.load();
  if (m < 0) {
    const char* f = getenv("VHMM_PRIOR_FUSED");
    const char* p = getenv("VHMM_PRIOR_PARTS");
    m = (f && f[0] == '0') ? 2 : ((p && p[0] == '0') ? 1 : 0);
    g_prior_mode = m;
  }
  return m;
}
int prior_impl_select(int mode) { g_prior_mode = mode; return 0; }
static bool prior_fused_enabled() { return prior_mode() != 2; }
static bool prior_parts_enabled() { return prior_mode() == 0; }
struct Tc5Ctx {
  struct Entry { const float* w; int dgrad; Tc5Weights tw; };
  Entry e[16];
  int n = 0;
  // |tensor| max words known for the CURRENT batch chunk (key = tensor pointer): filled by the tensor-core kernels' own
  // epilogues or by one absmax pass, reused by every later consumer of the same tensor (forward, input gradient, weight
  // gradient); reset when the chunk's buffers are rewritten
  static constexpr int NW = 64;
  uint32_t* pool = nullptr;
  mutable const void* key[NW];
  mutable int nkeys = 0;
  int reset(cudaStream_t st) const {
    nkeys = 0;
    if (pool) VHMM_CUDA_CHECK(cudaMemsetAsync(pool, 0, NW * 4, st));
    return 0;
  }
  uint32_t* known(const void* p) const {
    for (int i = 0; i < nkeys; ++i) if (key[i] == p) return pool + i;
    return nullptr;
  }
  uint32_t* fresh(const void* p) const {            // a zeroed word registered for tensor p (about to be produced / measured)
    for (int i = 0; i < nkeys; ++i) if (key[i] == p) key[i] = nullptr;   // p is being rewritten: its old word is stale
    if (!pool || nkeys >= NW) return nullptr;
    key[nkeys] = p;
    return pool + nkeys++;
  }
  // word holding |p| max, running the absmax pass if nobody produced it yet; null if the pool is exhausted
  uint32_t* measure(const float* p, int64_t n, cudaStream_t st, int* rc) const {
    *rc = 0;
    if (uint32_t* k = known(p)) return k;
    uint32_t* wd = fresh(p);
    if (wd) *rc = launch_absmax(p, n, wd, st);
    return wd;
  }
  const Tc5Weights* find(const float* w, int dgrad) const {
    for (int i = 0; i < n; ++i) if (e[i].w == w && e[i].dgrad == dgrad) return &e[i].tw;
    return nullptr;
  }
};
struct Tc5Layer { int param, Cin, Cout, kw, need_dgrad, emb; };   // emb: the (K,H) embedding read as a 1x1 conv K -> H
// (parameter, Cin, Cout, kw, dgrad needed) of the layers that can be GEMM-shaped at large widths
static int tc5_layers(const vhmm_dims& d, Tc5Layer* L) {
  int n = 0;
  L[n++] = Tc5Layer{P_E2W, d.H, d.H2, 3, 1, 0};
  L[n++] = Tc5Layer{P_E3W, d.H2, d.K, 1, 1, 0};          // forward 128 -> 32 on one zero-padded output tile
  L[n++] = Tc5Layer{P_T0W, d.U, d.TH, 1, 0, 0};
  L[n++] = Tc5Layer{P_T2W, d.TH, d.K * d.K, 1, 1, 0};
  L[n++] = Tc5Layer{P_EMB, d.K, d.H, 1, 0, 1};           // e = E^T q (VQ_VAE_HMM_fixed.py:83)
  L[n++] = Tc5Layer{P_D1W, d.H, d.H, 3, 1, 0};
  L[n++] = Tc5Layer{P_D2W, d.H, d.H, 3, 1, 0};
  return n;
}
static bool tc5_shape(int Cin, int Cout) { return Cin % ct5::KC == 0 && Cout % 16 == 0; }   // (Cout is zero-padded to 128-channel tiles)
static size_t tc5_ctx_bytes(const vhmm_dims& d) {
  Tc5Layer L[12];
  const int n = tc5_layers(d, L);
  size_t b = 0;
  for (int i = 0; i < n; ++i) {
    if (tc5_shape(L[i].Cin, L[i].Cout)) b += conv_tc5_weight_bytes(L[i].Cin, L[i].Cout, L[i].kw);
    if (L[i].need_dgrad && tc5_shape(L[i].Cout, L[i].Cin)) b += conv_tc5_weight_bytes(L[i].Cout, L[i].Cin, L[i].kw);
  }
  return b + 512;
}
static int tc5_ctx_prepare(const vhmm_dims& d, const vhmm_params& P, bool with_dgrad, Arena& A, Tc5Ctx* ctx, cudaStream_t st) {
  ctx->n = 0;
  ctx->nkeys = 0;
  ctx->pool = nullptr;
  if (conv_impl_mode() == 1) return 0;
  ctx->pool = A.take<uint32_t>(Tc5Ctx::NW);
  if (!A.ok) { ctx->pool = nullptr; return 0; }
  { int rc0 = ctx->reset(st); if (rc0) return rc0; }
  Tc5Layer L[12];
  const int n = tc5_layers(d, L);
  for (int i = 0; i < n; ++i) {
    const float* w = P.p[L[i].param];
    const int64_t nw = (int64_t)L[i].Cin * L[i].Cout * L[i].kw;
    if (tc5_shape(L[i].Cin, L[i].Cout)) {
      void* buf = A.take<uint8_t>(conv_tc5_weight_bytes(L[i].Cin, L[i].Cout, L[i].kw));
      if (!A.ok) return 0;     // the caller reports the workspace shortfall
      Tc5Ctx::Entry& en = ctx->e[ctx->n++];
      en.w = w; en.dgrad = 0;
      const WView fv = L[i].emb ? WView{w, 0, 1, L[i].Cout, 0} : wview_fwd(w, L[i].Cin, L[i].kw);   // emb: W(o=d, c=k) = E[k*H + d]
      int rc = conv_tc5_prepare(fv, nw, L[i].Cin, L[i].Cout, L[i].kw, buf, &en.tw, st);
      if (rc) return rc;
    }
    if (with_dgrad && L[i].need_dgrad && tc5_shape(L[i].Cout, L[i].Cin)) {
      void* buf = A.take<uint8_t>(conv_tc5_weight_bytes(L[i].Cout, L[i].Cin, L[i].kw));
      if (!A.ok) return 0;
      Tc5Ctx::Entry& en = ctx->e[ctx->n++];
      en.w = w; en.dgrad = 1;
      int rc = conv_tc5_prepare(wview_dgrad(w, L[i].Cin, L[i].kw), nw, L[i].Cout, L[i].Cin, L[i].kw, buf, &en.tw, st);
      if (rc) return rc;
    }
  }
  return 0;
}
// launch_conv with the tensor-core kernel where this (weight, view) was prepared and the tensors qualify
static int cconv(const Tc5Ctx* ctx, int dgrad, const float* x, WView wv, const float* bias, const float* act, float* y, int64_t B,
                 int Cin, int Cout, int64_t T, int kw, int epi, cudaStream_t st) {
  const Tc5Weights* tw = ctx ? ctx->find(wv.w, dgrad) : nullptr;
  if (tw && conv_tc5_ok(Cin, Cout, T, kw, x, y) && (!act || (((uintptr_t)act) & 15) == 0)) {
    int rc = 0;
    const uint32_t* ax = ctx->measure(x, B * (int64_t)Cin * T, st, &rc);
    if (rc) return rc;
    return launch_conv_tc5(x, *tw, bias, act, y, B, Cin, Cout, T, kw, epi, st, ax, ctx->fresh(y));
  }
  VHMM_REQUIRE(conv_impl_mode() != 2 || !ctx || !tc5_shape(Cin, Cout), VHMM_E_DIMS, "conv: VHMM_CONV_TC5 selected but T %% 4 != 0 or unaligned tensors");
  return launch_conv(x, wv, bias, act, y, B, Cin, Cout, T, kw, epi, st);
}
static int cwgrad(const Tc5Ctx* ctx, const float* dy, const float* x, float* gW, float* gB, WView layout, int64_t B, int Cin, int Cout,
                  int64_t T, int kw, float* scratch, cudaStream_t st) {
  if (conv_impl_mode() != 1 && wgrad_tc5_ok(Cin, Cout, T, kw, dy, x)) {
    int rc = 0;
    const uint32_t* ady = ctx ? ctx->measure(dy, B * (int64_t)Cout * T, st, &rc) : nullptr;
    if (rc) return rc;
    const uint32_t* ax = ctx ? ctx->measure(x, B * (int64_t)Cin * T, st, &rc) : nullptr;
    if (rc) return rc;
    return launch_wgrad_tc5(dy, x, gW, gB, layout, B, Cin, Cout, T, kw, scratch, st, ady, ax);
  }
  return launch_wgrad(dy, x, gW, gB, layout, B, Cin, Cout, T, kw, scratch, st);
}

// ---- decoder.embeddings folded into decoder.conv1 (ELBO of the wide model) ---------------------------------------------
// e = E^T q is linear and q is zero outside [0,T), so conv1(e) = conv(q, W') with W'[o,k,tau] = sum_d W1[o,d,tau] E[k,d]:
// a K -> H layer instead of H -> H (8x fewer MACs at H = 256, K = 32) in the forward pass, the input gradient (which now
// yields d loss / dq directly) and the weight gradient; the gradients of W1 and E follow from dW' by the chain rule after
// the last batch chunk.  Exact up to fp32 re-association (the fused tile kernels do the same: DESIGN.md 4.1).
__global__ void __launch_bounds__(256) fold_emb_kernel(const float* __restrict__ W1, const float* __restrict__ E, int H, int K,
                                                       float* __restrict__ Wf) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= H * K * 3) return;
  const int tau = i % 3, k = (i / 3) % K, o = i / (3 * K);
  float acc = 0.f;
  for (int dd = 0; dd < H; ++dd) acc = fmaf(W1[((int64_t)o * H + dd) * 3 + tau], E[(int64_t)k * H + dd], acc);
  Wf[i] = acc;
}
// gW1[o,d,tau] += sum_k gWf[o,k,tau] E[k,d]
__global__ void __launch_bounds__(256) unfold_w1_kernel(const float* __restrict__ gWf, const float* __restrict__ E, int H, int K,
                                                        float* __restrict__ gW1) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= H * H * 3) return;
  const int tau = i % 3, dd = (i / 3) % H, o = i / (3 * H);
  float acc = 0.f;
  for (int k = 0; k < K; ++k) acc = fmaf(gWf[((int64_t)o * K + k) * 3 + tau], E[(int64_t)k * H + dd], acc);
  gW1[i] += acc;
}
// gE[k,d] += sum_{o,tau} gWf[o,k,tau] W1[o,d,tau]
__global__ void __launch_bounds__(256) unfold_emb_kernel(const float* __restrict__ gWf, const float* __restrict__ W1, int H, int K,
                                                         float* __restrict__ gE) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= K * H) return;
  const int dd = i % H, k = i / H;
  float acc = 0.f;
  for (int o = 0; o < H; ++o)
#pragma unroll
    for (int tau = 0; tau < 3; ++tau) acc = fmaf(gWf[((int64_t)o * K + k) * 3 + tau], W1[((int64_t)o * H + dd) * 3 + tau], acc);
  gE[i] += acc;
}
static std::atomic<int> g_fold_mode{-1};    // environment VHMM_EMB_FOLD=0 keeps the embedding a layer of its own
static bool fold_enabled() {
  int m = g_fold_mode.load();
  if (m < 0) { const char* e = getenv("VHMM_EMB_FOLD"); m = (e && e[0] == '0') ? 0 : 1; g_fold_mode = m; }
  return m == 1;
}
int emb_fold_select(int on) { g_fold_mode = on ? 1 : 0; return 0; }
static bool fold_shape(const vhmm_dims& d) { return d.K % ct5::KC == 0 && d.H % ct5::KC == 0; }
static size_t fold_bytes(const vhmm_dims& d) {
  if (!fold_shape(d)) return 0;
  return 2 * align_up((size_t)d.H * d.K * 3 * 4, 256) + conv_tc5_weight_bytes(d.K, d.H, 3) + conv_tc5_weight_bytes(d.H, d.K, 3) + 1024;
}

struct LayerwiseSizes {
  int64_t per_step_fwd;   // floats per (sequence, step) kept by the forward pass
  int64_t per_step_bwd;   // extra floats per step for the backward pass
  int64_t scratch;        // wgrad partial scratch (floats), independent of the chunk size
};

static LayerwiseSizes layerwise_sizes(const vhmm_dims& d) {
  LayerwiseSizes s;
  const int64_t K2 = (int64_t)d.K * d.K;
  s.per_step_fwd = d.H + d.H2 + 3 * (int64_t)d.K + d.TH + K2 + 3 * (int64_t)d.H + 2 * d.C;
  s.per_step_bwd = 2 * d.C + d.K + K2 + 3 * (int64_t)d.H + 2 * (int64_t)d.K + d.H2 + d.H + d.TH + 2 * (int64_t)d.K;
  const int64_t BIG = 1 << 20;
  int64_t m = 0, v;
  v = wgrad_scratch_floats(d.C, d.H, 3, BIG); m = v > m ? v : m;
  v = wgrad_scratch_floats(d.H, d.H2, 3, BIG); m = v > m ? v : m;
  v = wgrad_scratch_floats(d.H2, d.K, 1, BIG); m = v > m ? v : m;
  v = wgrad_scratch_floats(d.U, d.TH, 1, BIG); m = v > m ? v : m;
  v = wgrad_scratch_floats(d.TH, d.K * d.K, 1, BIG); m = v > m ? v : m;
  v = wgrad_scratch_floats(d.K, d.H, 1, BIG); m = v > m ? v : m;
  v = wgrad_scratch_floats(d.K, d.H, 3, BIG); m = v > m ? v : m;      // folded embedding x conv1
  v = wgrad_scratch_floats(d.H, d.H, 3, BIG); m = v > m ? v : m;
  v = wgrad_scratch_floats(d.H, 2 * d.C, 1, BIG); m = v > m ? v : m;
  // tensor-core weight gradients keep one partial per (weight block, split): upper bound over the batch size
  auto tc5s = [&](int Cin, int Cout, int kw) {
    if (Cin < 8 || Cout < 8) return (int64_t)0;
    const int64_t blocks = cdiv(Cout, ct5::WB) * cdiv(Cin, ct5::WB);
    return ((3 * (int64_t)device_sm_count()) / blocks + 1) * ((int64_t)Cin * Cout * kw) + 64 + 32 * (int64_t)Cout;   // >= wgrad_tc5_nsplit
  };
  v = tc5s(d.H, d.H2, 3); m = v > m ? v : m;
  v = tc5s(d.TH, d.K * d.K, 1); m = v > m ? v : m;
  v = tc5s(d.H, d.H, 3); m = v > m ? v : m;
  v = tc5s(d.H2, d.K, 1); m = v > m ? v : m;
  v = tc5s(d.K, d.H, 1); m = v > m ? v : m;
  v = tc5s(d.K, d.H, 3); m = v > m ? v : m;
  v = tc5s(d.U, d.TH, 1); m = v > m ? v : m;
  v = tc5s(d.H, 2 * d.C, 1); m = v > m ? v : m;
  v = tc5s(d.C, d.H, 3); m = v > m ? v : m;
  s.scratch = m;
  return s;
}

static size_t fixed_bytes(const vhmm_dims& d, int64_t T) {
  LayerwiseSizes s = layerwise_sizes(d);
  size_t bytes = 0;
  bytes += align_up(sizeof(float) * KMAX, 256);                        // log_pi
  bytes += align_up(sizeof(double) * KMAX, 256);                       // qsum
  bytes += align_up(sizeof(float) * (size_t)s.scratch, 256);
  bytes += tc5_ctx_bytes(d);                                           // prepared tensor-core weights
  bytes += 64 * 1024;                                                  // per-CTA partial sums of the fused prior kernel
  bytes += fold_bytes(d);                                              // folded embedding x conv1 weights, their gradient
  return bytes + 4096;
}

size_t layerwise_workspace_bytes(int op, const vhmm_dims& d, int64_t B, int64_t T) {
  LayerwiseSizes s = layerwise_sizes(d);
  const int64_t Bc = B < 2048 ? B : 2048;   // sequences per chunk of the layer-wise passes (cfg 5: ~40 GB)
  const size_t slack = 64 * 256;
  switch (op) {
    case VHMM_OP_ENCODER_FWD:
      return (size_t)B * T * (d.H + d.H2 + d.K) * 4 + slack + tc5_ctx_bytes(d);
    case VHMM_OP_PRIOR_FWD:
      return (size_t)B * T * (d.TH + (int64_t)d.K * d.K) * 4 + slack + tc5_ctx_bytes(d);
    case VHMM_OP_DECODER_FWD:
      return (size_t)B * T * (3 * (int64_t)d.H) * 4 + slack + tc5_ctx_bytes(d);
    case VHMM_OP_EMISSION_TABLE:
      return (size_t)d.K * T * (3 * (int64_t)d.H + d.K + 2 * d.C) * 4 + slack + tc5_ctx_bytes(d) + 4096;
    case VHMM_OP_ELBO:
      return fixed_bytes(d, T) + (size_t)Bc * T * (s.per_step_fwd + s.per_step_bwd) * 4 +
             (size_t)Bc * cdiv(T, 32) * 32 + slack;
    default:
      return 0;
  }
}

// ---- forward building blocks (whole-tensor, caller provides scratch) -------------------
static int encoder_chain(const vhmm_dims& d, const vhmm_params& P, const float* x, int64_t B,
                         int64_t T, float* h1, float* h2, float* logits, cudaStream_t st, const Tc5Ctx* ctx = nullptr) {
  int rc;
  if ((rc = launch_conv(x, wview_fwd(P.p[P_E1W], d.C, 3), P.p[P_E1B], nullptr, h1, B, d.C, d.H, T, 3, EPI_RELU, st))) return rc;
  if ((rc = cconv(ctx, 0, h1, wview_fwd(P.p[P_E2W], d.H, 3), P.p[P_E2B], nullptr, h2, B, d.H, d.H2, T, 3, EPI_RELU, st))) return rc;
  return cconv(ctx, 0, h2, wview_fwd(P.p[P_E3W], d.H2, 1), P.p[P_E3B], nullptr, logits, B, d.H2, d.K, T, 1, EPI_NONE, st);
}

// e (B,H,T) = E^T q ; g1, g2 ; out rows [row0,row0+nrows) of to_params into `out`
static int decoder_chain(const vhmm_dims& d, const vhmm_params& P, const float* q, int64_t B,
                         int64_t T, float* e, float* g1, float* g2, cudaStream_t st, const Tc5Ctx* ctx = nullptr) {
  int rc;
  WView emb{P.p[P_EMB], 0, 1, d.H, 0};  // W(o=d, c=k) = E[k*H + d]   (VQ_VAE_HMM_fixed.py:83)
  if ((rc = cconv(ctx, 0, q, emb, nullptr, nullptr, e, B, d.K, d.H, T, 1, EPI_NONE, st))) return rc;
  if ((rc = cconv(ctx, 0, e, wview_fwd(P.p[P_D1W], d.H, 3), P.p[P_D1B], nullptr, g1, B, d.H, d.H, T, 3, EPI_RELU, st))) return rc;
  return cconv(ctx, 0, g1, wview_fwd(P.p[P_D2W], d.H, 3), P.p[P_D2B], nullptr, g2, B, d.H, d.H, T, 3, EPI_RELU, st);
}

int layerwise_encoder_fwd(const vhmm_dims& d, const vhmm_params& P, const float* x, int64_t B,
                          int64_t T, float* logits, float* q, void* ws, size_t ws_bytes,
                          cudaStream_t st) {
  Arena A(ws, ws_bytes);
  float* h1 = A.take<float>((size_t)B * d.H * T);
  float* h2 = A.take<float>((size_t)B * d.H2 * T);
  float* lg = logits ? logits : A.take<float>((size_t)B * d.K * T);
  Tc5Ctx ctx;
  { int rc0 = tc5_ctx_prepare(d, P, false, A, &ctx, st); if (rc0) return rc0; }
  VHMM_REQUIRE(A.ok, VHMM_E_WORKSPACE, "encoder_fwd: workspace %zu < %zu", ws_bytes, A.used);
  for (int64_t b0 = 0; b0 < B; b0 += kMaxChunk) {
    int64_t nb = B - b0 < kMaxChunk ? B - b0 : kMaxChunk;
    int rc = encoder_chain(d, P, x + b0 * d.C * T, nb, T, h1 + b0 * d.H * T, h2 + b0 * d.H2 * T,
                           lg + b0 * d.K * T, st, &ctx);
    if (rc) return rc;
    if (q) {
      dim3 g((unsigned)cdiv(T, 256), (unsigned)nb);
      softmax_kernel<<<g, 256, 0, st>>>(lg + b0 * d.K * T, q + b0 * d.K * T, nullptr, d.K, (int)T);
      VHMM_LAUNCH_CHECK();
    }
  }
  return 0;
}

bool hmm_transitions_supported(const vhmm_dims& d, int64_t T, const float* u, const float* out);   // hmm_inputs.cu
int hmm_transitions(const vhmm_dims& d, const vhmm_params& P, const float* u, int64_t B, int64_t T, float* out, int as_probs,
                    cudaStream_t st);

int layerwise_prior_fwd(const vhmm_dims& d, const vhmm_params& P, const float* u, int64_t B,
                        int64_t T, float* log_pi, float* log_A, void* ws, size_t ws_bytes,
                        cudaStream_t st, int as_probs, bool fused_transitions) {
  VHMM_REQUIRE(d.K <= KMAX, VHMM_E_DIMS, "prior_fwd: K=%d > %d", d.K, KMAX);
  if (fused_transitions && hmm_transitions_supported(d, T, u, log_A)) {   // base dims: one FFMA kernel, no hidden layer in HBM
    if (log_pi) {
      log_pi_kernel<<<1, 32, 0, st>>>(P.p[P_LOGPRIOR], log_pi, d.K);
      VHMM_LAUNCH_CHECK();
    }
    return hmm_transitions(d, P, u, B, T, log_A, as_probs, st);
  }
  Arena A(ws, ws_bytes);
  const int K2 = d.K * d.K;
  float* a = A.take<float>((size_t)B * d.TH * T);
  float* M = A.take<float>((size_t)B * K2 * T);
  Tc5Ctx ctx;
  { int rc0 = tc5_ctx_prepare(d, P, false, A, &ctx, st); if (rc0) return rc0; }
  VHMM_REQUIRE(A.ok, VHMM_E_WORKSPACE, "prior_fwd: workspace %zu < %zu", ws_bytes, A.used);
  if (log_pi) {
    log_pi_kernel<<<1, 32, 0, st>>>(P.p[P_LOGPRIOR], log_pi, d.K);
    VHMM_LAUNCH_CHECK();
  }
  int rc;
  if ((rc = cconv(&ctx, 0, u, wview_fwd(P.p[P_T0W], d.U, 1), P.p[P_T0B], nullptr, a, B, d.U, d.TH, T, 1, EPI_RELU, st))) return rc;
  if ((rc = cconv(&ctx, 0, a, wview_fwd(P.p[P_T2W], d.TH, 1), P.p[P_T2B], nullptr, M, B, d.TH, K2, T, 1, EPI_NONE, st))) return rc;
  for (int64_t b0 = 0; b0 < B; b0 += kMaxChunk) {
    int64_t nb = B - b0 < kMaxChunk ? B - b0 : kMaxChunk;
    dim3 g((unsigned)cdiv(T, 128), (unsigned)d.K, (unsigned)nb);
    logA_out_kernel<<<g, 128, 0, st>>>(M + b0 * K2 * T, log_A + b0 * T * K2, d.K, (int)T, as_probs);
    VHMM_LAUNCH_CHECK();
  }
  return 0;
}

int layerwise_decoder_fwd(const vhmm_dims& d, const vhmm_params& P, const float* q, int64_t B,
                          int64_t T, float* mu, float* logvar, void* ws, size_t ws_bytes,
                          cudaStream_t st) {
  Arena A(ws, ws_bytes);
  float* e = A.take<float>((size_t)B * d.H * T);
  float* g1 = A.take<float>((size_t)B * d.H * T);
  float* g2 = A.take<float>((size_t)B * d.H * T);
  Tc5Ctx ctx;
  { int rc0 = tc5_ctx_prepare(d, P, false, A, &ctx, st); if (rc0) return rc0; }
  VHMM_REQUIRE(A.ok, VHMM_E_WORKSPACE, "decoder_fwd: workspace %zu < %zu", ws_bytes, A.used);
  int rc;
  if ((rc = decoder_chain(d, P, q, B, T, e, g1, g2, st, &ctx))) return rc;
  // to_params rows [0,C) -> mu, rows [C,2C) -> logvar (VQ_VAE_HMM_fixed.py:87-90)
  if ((rc = launch_conv(g2, wview_fwd(P.p[P_DPW], d.H, 1), P.p[P_DPB], nullptr, mu, B, d.H, d.C, T, 1, EPI_NONE, st))) return rc;
  return launch_conv(g2, wview_fwd(P.p[P_DPW] + (int64_t)d.C * d.H, d.H, 1), P.p[P_DPB] + d.C, nullptr, logvar, B, d.H, d.C, T, 1, EPI_NONE, st);
}

// ---- ELBO forward + backward ------------------------------------------------------------
int layerwise_elbo(const vhmm_dims& d, const vhmm_params& P, const float* x, const float* u,
                   const int64_t* lengths, int64_t B, int64_t T, const double* norms, double beta,
                   double* sums, float* grads, void* ws, size_t ws_bytes, cudaStream_t st) {
  VHMM_REQUIRE(d.K <= KMAX, VHMM_E_DIMS, "elbo: K=%d > %d", d.K, KMAX);
  const LayerwiseSizes S = layerwise_sizes(d);
  const int K = d.K, K2 = d.K * d.K, C = d.C, H = d.H, H2 = d.H2, U = d.U, TH = d.TH;
  int64_t off[VHMM_NUM_PARAMS + 1];
  param_offsets(d, off);

  Arena A(ws, ws_bytes);
  float* log_pi = A.take<float>(KMAX);
  double* qsum = A.take<double>(KMAX);
  float* scratch = A.take<float>((size_t)S.scratch);
  double* prior_partial = A.take<double>(64 * 1024 / sizeof(double));
  Tc5Ctx ctx;
  {
    int rc0 = tc5_ctx_prepare(d, P, grads != nullptr, A, &ctx, st);
    if (rc0) return rc0;
  }
  // decoder.embeddings folded into decoder.conv1 (tensor-core layer kernels only)
  float *Wf = nullptr, *gWf = nullptr;
  const bool fold = ctx.pool && fold_enabled() && fold_shape(d) && (T & 3) == 0 && ctx.n + 2 <= 16;
  if (fold) {
    Wf = A.take<float>((size_t)H * K * 3);
    gWf = A.take<float>((size_t)H * K * 3);
    void* bf = A.take<uint8_t>(conv_tc5_weight_bytes(K, H, 3));
    void* bd = A.take<uint8_t>(conv_tc5_weight_bytes(H, K, 3));
    VHMM_REQUIRE(A.ok, VHMM_E_WORKSPACE, "elbo: workspace %zu too small (folded decoder weights)", ws_bytes);
    fold_emb_kernel<<<(unsigned)cdiv((int64_t)H * K * 3, 256), 256, 0, st>>>(P.p[P_D1W], P.p[P_EMB], H, K, Wf);
    VHMM_LAUNCH_CHECK();
    VHMM_CUDA_CHECK(cudaMemsetAsync(gWf, 0, (size_t)H * K * 3 * sizeof(float), st));
    Tc5Ctx::Entry& ef = ctx.e[ctx.n++];
    ef.w = Wf; ef.dgrad = 0;
    int rcf = conv_tc5_prepare(wview_fwd(Wf, K, 3), (int64_t)H * K * 3, K, H, 3, bf, &ef.tw, st);
    if (rcf) return rcf;
    if (grads) {
      Tc5Ctx::Entry& ed = ctx.e[ctx.n++];
      ed.w = Wf; ed.dgrad = 1;
      rcf = conv_tc5_prepare(wview_dgrad(Wf, K, 3), (int64_t)H * K * 3, H, K, 3, bd, &ed.tw, st);
      if (rcf) return rcf;
    }
  }
  VHMM_REQUIRE(A.ok, VHMM_E_WORKSPACE, "elbo: workspace %zu too small (fixed part %zu)", ws_bytes, A.used);
  const int64_t per_step = S.per_step_fwd + (grads ? S.per_step_bwd : 0);
  const size_t left = (ws_bytes - A.used > 64 * 256) ? ws_bytes - A.used - 64 * 256 : 0;
  int64_t Bc = (int64_t)(left / ((size_t)per_step * T * 4 + (size_t)cdiv(T, 32) * 32));
  if (Bc > B) Bc = B;
  if (Bc > kMaxChunk) Bc = kMaxChunk;
  VHMM_REQUIRE(Bc >= 1, VHMM_E_WORKSPACE, "elbo: workspace %zu cannot hold one sequence (T=%lld)", ws_bytes, (long long)T);
  const size_t n = (size_t)Bc * T;
  double* partial = A.take<double>(4 * (size_t)(Bc * cdiv(T, 32)));   // wide-K loss kernel: one block per 32 steps
  float* h1 = A.take<float>(n * H);   float* h2 = A.take<float>(n * H2);
  float* lg = A.take<float>(n * K);   float* q = A.take<float>(n * K);   float* lq = A.take<float>(n * K);
  float* a = A.take<float>(n * TH);   float* M = A.take<float>(n * K2);
  float* e = A.take<float>(n * H);    float* g1 = A.take<float>(n * H);  float* g2 = A.take<float>(n * H);
  float* p = A.take<float>(n * 2 * C);
  float *dp = nullptr, *dqp = nullptr, *dM = nullptr, *dg2 = nullptr, *dg1 = nullptr, *de = nullptr,
        *dqd = nullptr, *dl = nullptr, *dh2 = nullptr, *dh1 = nullptr, *da = nullptr, *dq_row = nullptr, *dq_col = nullptr;
  if (grads) {
    dq_row = A.take<float>(n * K); dq_col = A.take<float>(n * K);
    dp = A.take<float>(n * 2 * C); dqp = A.take<float>(n * K); dM = A.take<float>(n * K2);
    dg2 = A.take<float>(n * H); dg1 = A.take<float>(n * H); de = A.take<float>(n * H);
    dqd = A.take<float>(n * K); dl = A.take<float>(n * K);
    dh2 = A.take<float>(n * H2); dh1 = A.take<float>(n * H); da = A.take<float>(n * TH);
  }
  VHMM_REQUIRE(A.ok, VHMM_E_WORKSPACE, "elbo: internal workspace accounting error (%zu > %zu)", A.used, ws_bytes);

  VHMM_CUDA_CHECK(cudaMemsetAsync(sums, 0, 4 * sizeof(double), st));
  VHMM_CUDA_CHECK(cudaMemsetAsync(qsum, 0, KMAX * sizeof(double), st));
  if (grads) VHMM_CUDA_CHECK(cudaMemsetAsync(grads, 0, off[VHMM_NUM_PARAMS] * sizeof(float), st));
  log_pi_kernel<<<1, 32, 0, st>>>(P.p[P_LOGPRIOR], log_pi, K);
  VHMM_LAUNCH_CHECK();
  const float betaf = (float)beta;
  int rc;
#define RC(call) do { if ((rc = (call))) return rc; } while (0)
  for (int64_t b0 = 0; b0 < B; b0 += Bc) {
    const int64_t nb = B - b0 < Bc ? B - b0 : Bc;
    const float* xc = x + b0 * C * T;
    const float* uc = u + b0 * U * T;
    const int64_t* Lc = lengths + b0;
    RC(ctx.reset(st));                       // the chunk's activation buffers are about to be rewritten
    // ---------------- forward ----------------
    RC(encoder_chain(d, P, xc, nb, T, h1, h2, lg, st, &ctx));
    {
      dim3 g((unsigned)cdiv(T, 256), (unsigned)nb);
      softmax_kernel<<<g, 256, 0, st>>>(lg, q, lq, K, (int)T);
      VHMM_LAUNCH_CHECK();
    }
    RC(cconv(&ctx, 0, uc, wview_fwd(P.p[P_T0W], U, 1), P.p[P_T0B], nullptr, a, nb, U, TH, T, 1, EPI_RELU, st));
    // transition logits -> log A -> pair term and its gradients: ONE kernel when the shape allows (prior_tc5.cuh), else
    // the layer, the row log-softmax and the loops of the loss kernel
    const Tc5Weights* twT2 = (conv_impl_mode() != 1) ? ctx.find(P.p[P_T2W], 0) : nullptr;
    const bool al16 = ((((uintptr_t)xc | (uintptr_t)p | (uintptr_t)q | (uintptr_t)lq | (uintptr_t)M | (uintptr_t)dp | (uintptr_t)dqp |
                          (uintptr_t)dM) & 15) == 0);
    const bool fused_prior = twT2 && prior_fused_enabled() && al16 && prior_tc5_ok(K, TH, T, a, q, dM) &&
                             4 * prior_tc5_grid(nb, T) * sizeof(double) <= 64 * 1024;
    // d loss / d logits in operand form (fp16 parts in the dM buffer): needs both consumers on the tensor core
    const Tc5Weights* twT2d = (conv_impl_mode() != 1) ? ctx.find(P.p[P_T2W], 1) : nullptr;
    const bool dM_parts = fused_prior && grads && twT2d && TH == ct5::TN && prior_parts_enabled() && wgrad_tc5_ok(TH, K2, T, 1, dM, a) &&
                          (((uintptr_t)da) & 15) == 0;
    if (fused_prior) {
      int rcm = 0;
      const uint32_t* ax = ctx.measure(a, nb * (int64_t)TH * T, st, &rcm);
      RC(rcm);
      RC(launch_prior_tc5(a, *twT2, P.p[P_T2B], q, Lc, norms, betaf, nb, TH, T, grads ? dM : nullptr, dq_row, dq_col, prior_partial,
                          grads ? ctx.fresh(dM) : nullptr, st, ax, dM_parts));
      sum_partials_kernel<<<1, 1024, 0, st>>>(prior_partial, prior_tc5_grid(nb, T), sums);
      VHMM_LAUNCH_CHECK();
    } else {
      RC(cconv(&ctx, 0, a, wview_fwd(P.p[P_T2W], TH, 1), P.p[P_T2B], nullptr, M, nb, TH, K2, T, 1, EPI_NONE, st));
      dim3 g((unsigned)cdiv(T, 256), (unsigned)K, (unsigned)nb);
      row_logsoftmax_kernel<<<g, 256, 0, st>>>(M, K, (int)T);
      VHMM_LAUNCH_CHECK();
    }
    if (fold) {
      RC(cconv(&ctx, 0, q, wview_fwd(Wf, K, 3), P.p[P_D1B], nullptr, g1, nb, K, H, T, 3, EPI_RELU, st));
      RC(cconv(&ctx, 0, g1, wview_fwd(P.p[P_D2W], H, 3), P.p[P_D2B], nullptr, g2, nb, H, H, T, 3, EPI_RELU, st));
    } else {
      RC(decoder_chain(d, P, q, nb, T, e, g1, g2, st, &ctx));
    }
    RC(launch_conv(g2, wview_fwd(P.p[P_DPW], H, 1), P.p[P_DPB], nullptr, p, nb, H, 2 * C, T, 1, EPI_NONE, st));
    {
      dim3 g((unsigned)cdiv(T, 256), (unsigned)nb);
      int64_t nblk = (int64_t)g.x * g.y;
      if (K <= 4) loss_seed_kernel<4><<<g, 256, 0, st>>>(xc, p, q, lq, M, log_pi, Lc, norms, betaf, C, K, (int)T, partial, dp, dqp, dM);
      else if (K <= 8) loss_seed_kernel<8><<<g, 256, 0, st>>>(xc, p, q, lq, M, log_pi, Lc, norms, betaf, C, K, (int)T, partial, dp, dqp, dM);
      else {
        dim3 gr((unsigned)cdiv(T, 32), (unsigned)nb);
        nblk = (int64_t)gr.x * gr.y;
        if ((T & 3) == 0 && al16)     // four steps per thread, 128-bit accesses
          loss_seed_rows4_kernel<<<gr, 8 * K, 0, st>>>(xc, p, q, lq, fused_prior ? nullptr : M, log_pi, Lc, norms, betaf, C, K, (int)T,
                                                       partial, dp, dqp, fused_prior ? nullptr : dM);
        else
          loss_seed_rows_kernel<<<gr, 32 * K, 0, st>>>(xc, p, q, lq, M, log_pi, Lc, norms, betaf, C, K, (int)T, partial, dp, dqp, dM);
      }
      VHMM_LAUNCH_CHECK();
      sum_partials_kernel<<<1, 1024, 0, st>>>(partial, nblk, sums);
      VHMM_LAUNCH_CHECK();
      q0_sum_kernel<<<1, 1024, 0, st>>>(q, nb, K, (int)T, qsum);
      VHMM_LAUNCH_CHECK();
    }
    if (!grads) continue;
    // ---------------- backward: decoder ----------------
    RC(cwgrad(&ctx, dp, g2, grads + off[P_DPW], grads + off[P_DPB], wview_fwd(nullptr, H, 1), nb, H, 2 * C, T, 1, scratch, st));
    RC(launch_conv(dp, wview_dgrad(P.p[P_DPW], H, 1), nullptr, g2, dg2, nb, 2 * C, H, T, 1, EPI_MASK, st));
    RC(cwgrad(&ctx, dg2, g1, grads + off[P_D2W], grads + off[P_D2B], wview_fwd(nullptr, H, 3), nb, H, H, T, 3, scratch, st));
    RC(cconv(&ctx, 1, dg2, wview_dgrad(P.p[P_D2W], H, 3), nullptr, g1, dg1, nb, H, H, T, 3, EPI_MASK, st));
    if (fold) {     // d/dW' (K -> H weight gradient) and d/dq (H -> K input gradient, zero-padded to one 128-channel tile)
      RC(cwgrad(&ctx, dg1, q, gWf, grads + off[P_D1B], wview_fwd(nullptr, K, 3), nb, K, H, T, 3, scratch, st));
      RC(cconv(&ctx, 1, dg1, wview_dgrad(Wf, K, 3), nullptr, nullptr, dqd, nb, H, K, T, 3, EPI_NONE, st));
    } else {
    RC(cwgrad(&ctx, dg1, e, grads + off[P_D1W], grads + off[P_D1B], wview_fwd(nullptr, H, 3), nb, H, H, T, 3, scratch, st));
    RC(cconv(&ctx, 1, dg1, wview_dgrad(P.p[P_D1W], H, 3), nullptr, nullptr, de, nb, H, H, T, 3, EPI_NONE, st));
    {
      WView embL{nullptr, 0, 1, H, 0};  // dE[k,d] stored at k*H + d  with (o=d, c=k)
      RC(cwgrad(&ctx, de, q, grads + off[P_EMB], nullptr, embL, nb, K, H, T, 1, scratch, st));
      WView embT{P.p[P_EMB], 0, H, 1, 0};  // dq[k,t] = sum_d E[k,d] de[d,t]: W(o=k, c=d)
      RC(launch_conv(de, embT, nullptr, nullptr, dqd, nb, H, K, T, 1, EPI_NONE, st));
    }
    }
    {
      dim3 g((unsigned)cdiv(T, 256), (unsigned)nb);
      dlogits_kernel<<<g, 256, 0, st>>>(q, lq, dqd, dqp, Lc, norms, betaf, K, (int)T, dl, fused_prior ? dq_row : nullptr,
                                        fused_prior ? dq_col : nullptr);
      VHMM_LAUNCH_CHECK();
    }
    // ---------------- backward: encoder ----------------
    RC(cwgrad(&ctx, dl, h2, grads + off[P_E3W], grads + off[P_E3B], wview_fwd(nullptr, H2, 1), nb, H2, K, T, 1, scratch, st));
    RC(cconv(&ctx, 1, dl, wview_dgrad(P.p[P_E3W], H2, 1), nullptr, h2, dh2, nb, K, H2, T, 1, EPI_MASK, st));
    RC(cwgrad(&ctx, dh2, h1, grads + off[P_E2W], grads + off[P_E2B], wview_fwd(nullptr, H, 3), nb, H, H2, T, 3, scratch, st));
    RC(cconv(&ctx, 1, dh2, wview_dgrad(P.p[P_E2W], H, 3), nullptr, h1, dh1, nb, H2, H, T, 3, EPI_MASK, st));
    RC(cwgrad(&ctx, dh1, xc, grads + off[P_E1W], grads + off[P_E1B], wview_fwd(nullptr, C, 3), nb, C, H, T, 3, scratch, st));
    // ---------------- backward: prior ----------------
    if (dM_parts) {
      int rcm = 0;
      const uint32_t* adM = ctx.known(dM);
      const uint32_t* aa = ctx.measure(a, nb * (int64_t)TH * T, st, &rcm);
      RC(rcm);
      VHMM_REQUIRE(adM && aa, VHMM_E_WORKSPACE, "elbo: abs-max word pool exhausted");
      const uint8_t* dMp = reinterpret_cast<const uint8_t*>(dM);
      RC(launch_wgrad_tc5(dM, a, grads + off[P_T2W], nullptr, wview_fwd(nullptr, TH, 1), nb, TH, K2, T, 1, scratch, st, adM, aa, dMp));
      RC(launch_bias_parts(dMp, nb, T, adM, scratch, grads + off[P_T2B], st));     // (the wgrad partials in `scratch` are consumed)
      RC(launch_dgrad_parts_tc5(dMp, *twT2d, a, da, nb, TH, T, adM, ctx.fresh(da), st));
    } else {
      RC(cwgrad(&ctx, dM, a, grads + off[P_T2W], grads + off[P_T2B], wview_fwd(nullptr, TH, 1), nb, TH, K2, T, 1, scratch, st));
      RC(cconv(&ctx, 1, dM, wview_dgrad(P.p[P_T2W], TH, 1), nullptr, a, da, nb, K2, TH, T, 1, EPI_MASK, st));
    }
    RC(cwgrad(&ctx, da, uc, grads + off[P_T0W], grads + off[P_T0B], wview_fwd(nullptr, U, 1), nb, U, TH, T, 1, scratch, st));
  }
#undef RC
  if (fold && grads) {     // chain rule from dW' to decoder.conv1.weight and decoder.embeddings
    unfold_w1_kernel<<<(unsigned)cdiv((int64_t)H * H * 3, 256), 256, 0, st>>>(gWf, P.p[P_EMB], H, K, grads + off[P_D1W]);
    VHMM_LAUNCH_CHECK();
    unfold_emb_kernel<<<(unsigned)cdiv((int64_t)K * H, 256), 256, 0, st>>>(gWf, P.p[P_D1W], H, K, grads + off[P_EMB]);
    VHMM_LAUNCH_CHECK();
  }
  finalize_prior_kernel<<<1, 32, 0, st>>>(log_pi, qsum, norms, betaf, K, sums,
                                          grads ? grads + off[P_LOGPRIOR] : nullptr);
  VHMM_LAUNCH_CHECK();
  return 0;
}

// per-state emission table: decoder on the K one-hot sequences.
__global__ void onehot_kernel(float* __restrict__ q, int K, int T) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= (int64_t)K * K * T) return;
  const int64_t row = i / T;        // row = s*K + k
  q[i] = (row / K == row % K) ? 1.f : 0.f;
}

// The decoder's input here is CONSTANT over time (the one-hot of state k at every step) and its receptive field is +-2 steps
// (1x1, k=3, k=3, 1x1), so every position further than 2 steps from both ends sees the same window: the table of a long
// sequence is the table of a 16-step one with its middle column repeated.  (K*T positions through three H x H x 3 layers
// otherwise: 2 ms at K = 32, H = 256, T = 16384 -- most of hmm_inputs for few long sequences.)
constexpr int kTabShort = 16, kTabEdge = 4;
__global__ void __launch_bounds__(256) expand_table_kernel(const float* __restrict__ mu_s, const float* __restrict__ lv_s, int64_t rows,
                                                           int T, float* __restrict__ mu, float* __restrict__ lv) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= rows * T) return;
  const int64_t r = i / T;
  const int t = (int)(i - r * T);
  const int ts = t < kTabEdge ? t : (t >= T - kTabEdge ? kTabShort - (T - t) : kTabShort / 2);
  mu[i] = mu_s[r * kTabShort + ts];
  lv[i] = lv_s[r * kTabShort + ts];
}

int layerwise_emission_table(const vhmm_dims& d, const vhmm_params& P, int64_t T, float* mu_tab,
                             float* logvar_tab, void* ws, size_t ws_bytes, cudaStream_t st) {
  if (T > kTabShort) {
    Arena S(ws, ws_bytes);
    float* mu_s = S.take<float>((size_t)d.K * d.C * kTabShort);
    float* lv_s = S.take<float>((size_t)d.K * d.C * kTabShort);
    VHMM_REQUIRE(S.ok, VHMM_E_WORKSPACE, "emission_table: workspace too small");
    int rc = layerwise_emission_table(d, P, kTabShort, mu_s, lv_s, (char*)ws + S.used, ws_bytes - S.used, st);
    if (rc) return rc;
    const int64_t n = (int64_t)d.K * d.C * T;
    expand_table_kernel<<<(unsigned)cdiv(n, 256), 256, 0, st>>>(mu_s, lv_s, (int64_t)d.K * d.C, (int)T, mu_tab, logvar_tab);
    VHMM_LAUNCH_CHECK();
    return 0;
  }
  Arena A(ws, ws_bytes);
  float* q = A.take<float>((size_t)d.K * d.K * T);
  VHMM_REQUIRE(A.ok, VHMM_E_WORKSPACE, "emission_table: workspace too small");
  onehot_kernel<<<(unsigned)cdiv((int64_t)d.K * d.K * T, 256), 256, 0, st>>>(q, d.K, (int)T);
  VHMM_LAUNCH_CHECK();
  return layerwise_decoder_fwd(d, P, q, d.K, T, mu_tab, logvar_tab, (char*)ws + A.used,
                               ws_bytes - A.used, st);
}

#include "hmm_grad.cuh"

}  // namespace vhmm

namespace vhmm {
int launch_argmax(const float* q, int64_t B, int64_t K, int64_t T, int64_t* out, cudaStream_t st) {
  for (int64_t b0 = 0; b0 < B; b0 += kMaxChunk) {
    int64_t nb = B - b0 < kMaxChunk ? B - b0 : kMaxChunk;
    dim3 g((unsigned)cdiv(T, 256), (unsigned)nb);
    argmax_kernel<<<g, 256, 0, st>>>(q + b0 * K * T, (int)K, (int)T, out + b0 * T);
    VHMM_LAUNCH_CHECK();
  }
  return 0;
}
}  // namespace vhmm
