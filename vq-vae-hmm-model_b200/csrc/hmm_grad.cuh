// hmm_grad.cuh -- the beta-recursion gradient chained to the 18 parameters (last
// sentence).  Included by elbo_layerwise.cu (it reuses the layer kernels of conv_generic.cuh).
//
// loglik[b] = log p(x_b | u_b) of the exact HMM (B.2) depends on the parameters through
//   log_pi = log_softmax(prior.log_prior)                                     (VQ_VAE_HMM_fixed.py:54, :71)
//   log_A[b,t,i,:] = log_softmax_j( Wt2 relu(Wt0 u[b,:,t] + bt0) + bt2 )      (:68-69)
//   log_b[b,t,k]   = Gaussian log-density of x[b,:,t] under the decoder's output on the k-th one-hot sequence (B.1)
// and the recursion's own backward is the posterior pair (gamma, xi) of vhmm_hmm_posterior:
//   d loglik / d log_b = gamma,  d loglik / d log_A = xi,  d loglik / d log_pi = gamma[:,0].
// Given per-sequence upstream weights w[b] = d loss / d loglik[b] this file forms d loss / d (18 parameters):
//   1. dM[b,i*K+j,t] = w_b ( xi[b,t,i,j] - exp(log_A[b,t,i,j]) sum_j' xi[b,t,i,j'] )      (row log-softmax backward)
//   2. transition MLP backward with the layer kernels (a recomputed per batch chunk; never more than a chunk of
//      the hidden layer is alive)
//   3. d loss / d table: dmu[k,c,t] = sum_b w_b gamma[b,t,k] (x - mu)/v,
//      dlogvar[k,c,t] = sum_b w_b gamma[b,t,k] 0.5 ((x - mu)^2 / v - 1) [exp(logvar) >= 1e-8]   (clamp of :118)
//      -- deterministic two-level reduction over the batch
//   4. decoder backward on the K one-hot sequences (B = K, tiny) -> embeddings, conv1, conv2, to_params
//   5. d / d log_prior = g - softmax(log_prior) sum(g),  g[k] = sum_b w_b gamma[b,0,k]
// The encoder does not enter the exact likelihood: its six gradient tensors are exactly zero.
#pragma once

namespace hg_ {

// (B,T,K,K) xi, log_A -> dM in the layer layout (B, K*K, T).  grid (ceil(T/128), K, B), block 128.
__global__ void dM_kernel(const float* __restrict__ xi, const float* __restrict__ log_A, const float* __restrict__ w,
                          float* __restrict__ dM, int K, int T) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  const int i = blockIdx.y;
  const int64_t b = blockIdx.z;
  const float wb = w[b];
  const float* xr = xi + ((b * T + t) * K + i) * (int64_t)K;
  const float* ar = log_A + ((b * T + t) * K + i) * (int64_t)K;
  float s = 0.f;
  for (int j = 0; j < K; ++j) s += xr[j];
  float* o = dM + (b * K + i) * K * (int64_t)T + t;
  for (int j = 0; j < K; ++j) o[(int64_t)j * T] = (t >= 1) ? wb * (xr[j] - expf(ar[j]) * s) : 0.f;
}

// partial table gradients: block (t-tile, batch slice s) sums its sequences.  part[(s*2 + {0,1}) * K*C*T + (k*C+c)*T + t]
template <int KMAXT>
__global__ void __launch_bounds__(128) table_grad_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                         const float* __restrict__ w, const float* __restrict__ mu_tab,
                                                         const float* __restrict__ lv_tab, int64_t B, int C, int K, int T,
                                                         int nslice, float* __restrict__ part) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  const int s = blockIdx.y;
  const int64_t b0 = B * s / nslice, b1 = B * (s + 1) / nslice;
  const int64_t n = (int64_t)K * C * T;
  for (int c = 0; c < C; ++c) {
    float mu[KMAXT], iv[KMAXT], gate[KMAXT], am[KMAXT], av[KMAXT];
#pragma unroll
    for (int k = 0; k < KMAXT; ++k) {
      if (k < K) {
        const float lv = lv_tab[((int64_t)k * C + c) * T + t];
        const float ev = expf(lv);
        mu[k] = mu_tab[((int64_t)k * C + c) * T + t];
        iv[k] = 1.f / fmaxf(ev, 1e-8f);
        gate[k] = ev >= 1e-8f ? 1.f : 0.f;
      }
      am[k] = 0.f; av[k] = 0.f;
    }
    for (int64_t b = b0; b < b1; ++b) {
      const float xv = x[(b * C + c) * (int64_t)T + t];
      const float wb = w[b];
#pragma unroll
      for (int k = 0; k < KMAXT; ++k)
        if (k < K) {
          const float g = wb * gamma[(b * T + t) * (int64_t)K + k];
          const float d = xv - mu[k];
          am[k] = fmaf(g, d * iv[k], am[k]);
          av[k] = fmaf(g, 0.5f * (d * d * iv[k] - 1.f), av[k]);
        }
    }
#pragma unroll
    for (int k = 0; k < KMAXT; ++k)
      if (k < K) {
        part[((int64_t)s * 2 + 0) * n + ((int64_t)k * C + c) * T + t] = am[k];
        part[((int64_t)s * 2 + 1) * n + ((int64_t)k * C + c) * T + t] = av[k] * gate[k];
      }
  }
}

// dp (K, 2C, T): rows [0,C) = d/d mu, rows [C,2C) = d/d logvar, summed over the slices in a fixed order (double)
__global__ void table_reduce_kernel(const float* __restrict__ part, int nslice, int C, int K, int T, float* __restrict__ dp) {
  const int64_t n = (int64_t)K * C * T;
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= 2 * n) return;
  const int which = (int)(i / n);
  const int64_t r = i - which * n;
  double a = 0.0;
  for (int s = 0; s < nslice; ++s) a += (double)part[((int64_t)s * 2 + which) * n + r];
  const int64_t k = r / ((int64_t)C * T), ct = r - k * (int64_t)C * T;
  dp[(k * 2 * C + (int64_t)which * C) * T + ct] = (float)a;
}

// The decoder's input for the emission table is constant over time and the decoder's receptive field is +-2 steps, so every
// forward activation (and ReLU mask) at positions 2 .. T-3 is the same, the backward is linear in dp and shift-invariant
// there, and the parameter gradients are SUMS over t.  Hence the columns 8 .. T-9 of dp may be replaced by their sum in one
// column: the surrogate table gradient of 25 columns
//     [ dp[:, 0..7] | 0 0 0 0 | sum_{t=8}^{T-9} dp[:, t] | 0 0 0 0 | dp[:, T-8..T-1] ]
// run through the decoder backward of a 25-step one-hot sequence gives the same parameter gradients (the summed column's
// signal spreads over +-2 positions, pairs with activations +-1 further: all interior-class; the edge columns keep their
// distance to their end).  K*25 positions instead of K*T through three H x H x 3 layers, backward and recompute.
constexpr int kTabGradCols = 25, kTabGradEdge = 8;
__global__ void __launch_bounds__(256) collapse_table_grad_kernel(const float* __restrict__ dp, int64_t rows, int T, float* __restrict__ dps) {
  const int64_t r = blockIdx.x;                       // one CTA per row (k, channel)
  if (r >= rows) return;
  const float* src = dp + r * (int64_t)T;
  float* dst = dps + r * kTabGradCols;
  __shared__ double red[256];
  double a = 0.0;
  for (int t = kTabGradEdge + threadIdx.x; t < T - kTabGradEdge; t += 256) a += (double)src[t];
  red[threadIdx.x] = a;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x < kTabGradCols) {
    const int j = threadIdx.x;
    float v = 0.f;
    if (j < kTabGradEdge) v = src[j];
    else if (j >= kTabGradCols - kTabGradEdge) v = src[T - kTabGradCols + j];
    else if (j == kTabGradCols / 2) v = (float)red[0];
    dst[j] = v;
  }
}

// g[k] = sum_b w_b gamma[b,0,k]; grad_log_prior[k] += g[k] - softmax(log_prior)[k] sum(g).  one block of 256.
__global__ void log_prior_grad_kernel(const float* __restrict__ gamma, const float* __restrict__ w,
                                      const float* __restrict__ log_prior, int64_t B, int K, int T,
                                      float* __restrict__ grad) {
  __shared__ double red[256];
  __shared__ double g[256];
  for (int k = 0; k < K; ++k) {
    double a = 0.0;
    for (int64_t b = threadIdx.x; b < B; b += blockDim.x) a += (double)w[b] * (double)gamma[b * T * (int64_t)K + k];
    red[threadIdx.x] = a;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
      if ((int)threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
      __syncthreads();
    }
    if (threadIdx.x == 0) g[k] = red[0];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    double m = -1e300, s = 0.0, tot = 0.0;
    for (int k = 0; k < K; ++k) { m = fmax(m, (double)log_prior[k]); tot += g[k]; }
    for (int k = 0; k < K; ++k) s += exp((double)log_prior[k] - m);
    for (int k = 0; k < K; ++k) grad[k] += (float)(g[k] - exp((double)log_prior[k] - m) / s * tot);
  }
}

}  // namespace hg_

static const int kHmmGradSlices = 64;

size_t hmm_param_grads_workspace_bytes(const vhmm_dims& d, int64_t B, int64_t T) {
  const LayerwiseSizes S = layerwise_sizes(d);
  const int64_t K2 = (int64_t)d.K * d.K;
  const int64_t Bc = B < 2048 ? B : 2048;
  size_t s = 0;
  s += align_up((size_t)S.scratch * 4, 256);
  s += align_up((size_t)Bc * T * (2 * d.TH + K2) * 4, 256) * 1 + 3 * 256;
  s += align_up((size_t)kHmmGradSlices * 2 * d.K * d.C * T * 4, 256);
  s += align_up((size_t)d.K * T * (d.K + 2 * d.C + 3 * (int64_t)d.H + 2 * d.C + 3 * (int64_t)d.H) * 4, 256) + 16 * 256;
  s += tc5_ctx_bytes(d);          // prepared tensor-core weights (wide models)
  return s + 4096;
}

static bool hmm_table_collapse_enabled() {   // environment VHMM_TABLE_COLLAPSE=off: the decoder backward over all K*T positions
  static const bool on = [] { const char* e = getenv("VHMM_TABLE_COLLAPSE"); return !(e && e[0] == 'o'); }();
  return on;
}

// grads: flat float[vhmm_param_count], OVERWRITTEN with d loss / d params for loss = sum_b w[b] loglik[b].
int hmm_param_grads(const vhmm_dims& d, const vhmm_params& P, const float* x, const float* u, const float* log_A,
                    const float* gamma, const float* xi, const float* w, const float* mu_tab, const float* lv_tab,
                    int64_t B, int64_t T, float* grads, void* ws, size_t ws_bytes, cudaStream_t st) {
  VHMM_REQUIRE(d.K <= KMAX, VHMM_E_DIMS, "hmm_param_grads: K=%d > %d", d.K, KMAX);
  const LayerwiseSizes S = layerwise_sizes(d);
  const int K = d.K, K2 = d.K * d.K, C = d.C, H = d.H, U = d.U, TH = d.TH;
  int64_t off[VHMM_NUM_PARAMS + 1];
  param_offsets(d, off);
  Arena A(ws, ws_bytes);
  float* scratch = A.take<float>((size_t)S.scratch);
  float* part = A.take<float>((size_t)kHmmGradSlices * 2 * K * C * T);
  // decoder on the K one-hot sequences (B = K)
  const size_t nk = (size_t)K * T;
  float* qh = A.take<float>(nk * K);
  float* dp = A.take<float>(nk * 2 * C);
  float* e = A.take<float>(nk * H);  float* g1 = A.take<float>(nk * H);  float* g2 = A.take<float>(nk * H);
  float* dg2 = A.take<float>(nk * H); float* dg1 = A.take<float>(nk * H); float* de = A.take<float>(nk * H);
  // tensor-core layer kernels where a layer's shape allows (transition_net of a wide model: K*K = 1024 logits)
  Tc5Ctx ctx;
  { int rc0 = tc5_ctx_prepare(d, P, true, A, &ctx, st); if (rc0) return rc0; }
  VHMM_REQUIRE(A.ok, VHMM_E_WORKSPACE, "hmm_param_grads: workspace %zu too small (fixed part %zu)", ws_bytes, A.used);
  const size_t left = ws_bytes - A.used > 4 * 256 ? ws_bytes - A.used - 4 * 256 : 0;
  int64_t Bc = (int64_t)(left / ((size_t)T * (2 * TH + K2) * 4));
  if (Bc > B) Bc = B;
  if (Bc > kMaxChunk) Bc = kMaxChunk;
  VHMM_REQUIRE(Bc >= 1, VHMM_E_WORKSPACE, "hmm_param_grads: workspace %zu cannot hold one sequence (T=%lld)", ws_bytes, (long long)T);
  const size_t n = (size_t)Bc * T;
  float* a = A.take<float>(n * TH);
  float* da = A.take<float>(n * TH);
  float* dM = A.take<float>(n * K2);
  VHMM_REQUIRE(A.ok, VHMM_E_WORKSPACE, "hmm_param_grads: internal workspace accounting error");
  VHMM_CUDA_CHECK(cudaMemsetAsync(grads, 0, off[VHMM_NUM_PARAMS] * sizeof(float), st));
  int rc;
#define RC(call) do { if ((rc = (call))) return rc; } while (0)
  // ---- transition MLP (chunks of the batch) ----
  for (int64_t b0 = 0; b0 < B; b0 += Bc) {
    const int64_t nb = B - b0 < Bc ? B - b0 : Bc;
    const float* uc = u + b0 * U * T;
    RC(ctx.reset(st));                       // the chunk's buffers are about to be rewritten
    RC(cconv(&ctx, 0, uc, wview_fwd(P.p[P_T0W], U, 1), P.p[P_T0B], nullptr, a, nb, U, TH, T, 1, EPI_RELU, st));
    {
      dim3 g((unsigned)cdiv(T, 128), (unsigned)K, (unsigned)nb);
      hg_::dM_kernel<<<g, 128, 0, st>>>(xi + b0 * T * K2, log_A + b0 * T * K2, w + b0, dM, K, (int)T);
      VHMM_LAUNCH_CHECK();
    }
    // (TH x K*K weight gradient: on the tensor-core kernel where the shape allows -- 128 x 9 at the base dims, the 9 rows padded
    //  to a 64-channel block: 4.7 -> ~1 ms per 16 k sequences x 256 steps; the FFMA kernel otherwise)
    RC(cwgrad(&ctx, dM, a, grads + off[P_T2W], grads + off[P_T2B], wview_fwd(nullptr, TH, 1), nb, TH, K2, T, 1, scratch, st));
    RC(cconv(&ctx, 1, dM, wview_dgrad(P.p[P_T2W], TH, 1), nullptr, a, da, nb, K2, TH, T, 1, EPI_MASK, st));
    RC(cwgrad(&ctx, da, uc, grads + off[P_T0W], grads + off[P_T0B], wview_fwd(nullptr, U, 1), nb, U, TH, T, 1, scratch, st));
  }
  // ---- emission table gradient (reduction over the batch) ----
  {
    int ns = kHmmGradSlices;
    if (ns > B) ns = (int)B;
    dim3 g((unsigned)cdiv(T, 128), (unsigned)ns);
    if (K <= 4) hg_::table_grad_kernel<4><<<g, 128, 0, st>>>(x, gamma, w, mu_tab, lv_tab, B, C, K, (int)T, ns, part);
    else if (K <= 8) hg_::table_grad_kernel<8><<<g, 128, 0, st>>>(x, gamma, w, mu_tab, lv_tab, B, C, K, (int)T, ns, part);
    else hg_::table_grad_kernel<KMAX><<<g, 128, 0, st>>>(x, gamma, w, mu_tab, lv_tab, B, C, K, (int)T, ns, part);
    VHMM_LAUNCH_CHECK();
    hg_::table_reduce_kernel<<<(unsigned)cdiv((int64_t)2 * K * C * T, 256), 256, 0, st>>>(part, ns, C, K, (int)T, dp);
    VHMM_LAUNCH_CHECK();
  }
  // ---- decoder backward on the K one-hot sequences ----
  // a long table: its interior collapsed into one column (collapse_table_grad_kernel above): 25 steps instead of T
  int64_t Tb = T;
  const float* dpb = dp;
  if (T > 64 && hmm_table_collapse_enabled()) {
    float* dps = part;                                  // (the slices are reduced: the buffer is free; K*2C*25 <= its size)
    hg_::collapse_table_grad_kernel<<<(unsigned)(K * 2 * C), 256, 0, st>>>(dp, (int64_t)K * 2 * C, (int)T, dps);
    VHMM_LAUNCH_CHECK();
    Tb = hg_::kTabGradCols;
    dpb = dps;
  }
  onehot_kernel<<<(unsigned)cdiv((int64_t)K * K * Tb, 256), 256, 0, st>>>(qh, K, (int)Tb);
  VHMM_LAUNCH_CHECK();
  // (a long table that is NOT collapsed is real work: 3 H*H*3 layers over K*T positions forward, input gradient and weight
  //  gradient; the tensor-core layer kernels then, as in the layer-wise ELBO)
  const Tc5Ctx* dctx = ((int64_t)K * Tb >= 16384 && ctx.pool) ? &ctx : nullptr;
  if (dctx) RC(ctx.reset(st));
  RC(decoder_chain(d, P, qh, K, Tb, e, g1, g2, st, dctx));
#define WG(dy, xx, gw, gb, lay, cin, cout, kw) \
  (dctx ? cwgrad(dctx, dy, xx, gw, gb, lay, K, cin, cout, Tb, kw, scratch, st) : launch_wgrad(dy, xx, gw, gb, lay, K, cin, cout, Tb, kw, scratch, st))
  RC(WG(dpb, g2, grads + off[P_DPW], grads + off[P_DPB], wview_fwd(nullptr, H, 1), H, 2 * C, 1));
  RC(cconv(dctx, 1, dpb, wview_dgrad(P.p[P_DPW], H, 1), nullptr, g2, dg2, K, 2 * C, H, Tb, 1, EPI_MASK, st));
  RC(WG(dg2, g1, grads + off[P_D2W], grads + off[P_D2B], wview_fwd(nullptr, H, 3), H, H, 3));
  RC(cconv(dctx, 1, dg2, wview_dgrad(P.p[P_D2W], H, 3), nullptr, g1, dg1, K, H, H, Tb, 3, EPI_MASK, st));
  RC(WG(dg1, e, grads + off[P_D1W], grads + off[P_D1B], wview_fwd(nullptr, H, 3), H, H, 3));
  RC(cconv(dctx, 1, dg1, wview_dgrad(P.p[P_D1W], H, 3), nullptr, nullptr, de, K, H, H, Tb, 3, EPI_NONE, st));
  {
    WView embL{nullptr, 0, 1, H, 0};
    RC(launch_wgrad(de, qh, grads + off[P_EMB], nullptr, embL, K, K, H, Tb, 1, scratch, st));
  }
#undef WG
#undef RC
  // ---- initial distribution ----
  hg_::log_prior_grad_kernel<<<1, 256, 0, st>>>(gamma, w, P.p[P_LOGPRIOR], B, K, (int)T, grads + off[P_LOGPRIOR]);
  VHMM_LAUNCH_CHECK();
  return 0;
}
