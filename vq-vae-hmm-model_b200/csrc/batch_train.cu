// batch_train.cu -- the two callers either side of the ELBO path:
//
//  f1  gather_chunks : device-side batching.  The reference draws (sequence, start, length) per
//      item in Python and fills the zero-padded batch with 2B slice copies (RandomChunkDataset
//      VQ_VAE_HMM_fixed.py:20-29, collate_fn :164-179).  Here the corpus is resident in HBM as
//      x_cat (C, S_tot) / u_cat (U, S_tot) (sequences concatenated along time) and ONE kernel
//      writes x (B,C,T), u (B,U,T): out[b,r,t] = t < len[b] ? cat[r, col0[b] + t] : 0.
//      HBM bound: 4 (C+U) bytes read per valid step + 4 (C+U) bytes written per padded step.
//
//  f2  adam_step : the optimiser tail of train_model / Trainer.train_epoch (:146, :155-157;
//      src/training/trainer.py:24-34): optional clip_grad_norm_ (global L2 norm, coefficient
//      max_norm / (norm + 1e-6) clamped to 1) and torch.optim.Adam's update, over the flat
//      gradient buffer vhmm_elbo produces and the 18 parameter tensors, in two launches
//      (norm, update) instead of ~18 x 8 elementwise kernels.  Operation order follows
//      torch/optim/adam.py:_single_tensor_adam so that results agree with torch to rounding.
#include "common.cuh"

namespace vhmm {

// ---- f1 -------------------------------------------------------------------------------------
template <bool VEC>
__global__ void __launch_bounds__(128)
gather_chunks_kernel(const float* __restrict__ x_cat, const float* __restrict__ u_cat, int64_t S_tot, int C, int U,
                     const int64_t* __restrict__ col0, const int64_t* __restrict__ lengths, int64_t B, int T,
                     float* __restrict__ x_out, float* __restrict__ u_out) {
  const int64_t b = blockIdx.x;
  int64_t Ll = lengths[b];
  const int L = (int)(Ll < 0 ? 0 : (Ll > T ? T : Ll));
  const int64_t c0 = col0[b];
  const int R = C + U;
  if (VEC) {
    const int T4 = T >> 2;
    for (int i = threadIdx.x; i < R * T4; i += 128) {
      const int r = i / T4, t = (i - r * T4) * 4;
      const float* src = (r < C ? x_cat + (int64_t)r * S_tot : u_cat + (int64_t)(r - C) * S_tot) + c0 + t;
      float4 v;
      v.x = (t + 0 < L) ? __ldg(src + 0) : 0.f;
      v.y = (t + 1 < L) ? __ldg(src + 1) : 0.f;
      v.z = (t + 2 < L) ? __ldg(src + 2) : 0.f;
      v.w = (t + 3 < L) ? __ldg(src + 3) : 0.f;
      float* dst = (r < C ? x_out + (b * C + r) * (int64_t)T : u_out + (b * U + (r - C)) * (int64_t)T) + t;
      __stcs(reinterpret_cast<float4*>(dst), v);
    }
  } else {
    for (int i = threadIdx.x; i < R * T; i += 128) {
      const int r = i / T, t = i - r * T;
      const float* src = (r < C ? x_cat + (int64_t)r * S_tot : u_cat + (int64_t)(r - C) * S_tot) + c0 + t;
      float* dst = (r < C ? x_out + (b * C + r) * (int64_t)T : u_out + (b * U + (r - C)) * (int64_t)T) + t;
      *dst = (t < L) ? __ldg(src) : 0.f;
    }
  }
}

int gather_chunks(const float* x_cat, const float* u_cat, int64_t S_tot, int64_t C, int64_t U, const int64_t* col0,
                  const int64_t* lengths, int64_t B, int64_t T, float* x_out, float* u_out, cudaStream_t st) {
  if (B == 0 || T == 0) return 0;
  const bool vec = (T % 4 == 0) && (((uintptr_t)x_out | (uintptr_t)u_out) & 15) == 0;
  if (vec)
    gather_chunks_kernel<true><<<(unsigned)B, 128, 0, st>>>(x_cat, u_cat, S_tot, (int)C, (int)U, col0, lengths, B, (int)T, x_out, u_out);
  else
    gather_chunks_kernel<false><<<(unsigned)B, 128, 0, st>>>(x_cat, u_cat, S_tot, (int)C, (int)U, col0, lengths, B, (int)T, x_out, u_out);
  VHMM_LAUNCH_CHECK();
  return 0;
}

// ---- f2 -------------------------------------------------------------------------------------
struct ParamTable {
  float* p[VHMM_NUM_PARAMS];
  int64_t off[VHMM_NUM_PARAMS + 1];
};

// sum of squares of (scale * g) in double, fixed order: one CTA, strided, tree.  The buffer is
// 35 k (base) .. 645 k (scaled) floats, far too small for more than one CTA to matter.
__global__ void __launch_bounds__(1024)
grad_sqnorm_kernel(const float* __restrict__ g, int64_t n, const float* __restrict__ scale, double* __restrict__ out) {
  __shared__ double sh[32];
  const float s = scale ? *scale : 1.f;
  double acc = 0.0;
  for (int64_t i = threadIdx.x; i < n; i += 1024) {
    const float v = g[i] * s;
    acc += (double)v * (double)v;
  }
  acc = block_sum(acc, sh);
  if (threadIdx.x == 0) *out = acc;
}

__global__ void __launch_bounds__(256)
adam_kernel(ParamTable tab, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, int64_t n,
            float step_size /* lr / (1 - beta1^t) */, float w1 /* 1 - beta1 */, float beta2, float w2 /* 1 - beta2 */,
            float bc2_sqrt, float eps, const float* __restrict__ scale, const double* __restrict__ sqnorm,
            float max_norm, float* __restrict__ gnorm_out, const double* __restrict__ dyn /* train state or null */) {
  if (dyn) {   // graph-replayed step: lr / (1 - beta1^t) and sqrt(1 - beta2^t) of THIS step come from device memory
    step_size = (float)dyn[VHMM_TS_STEP_SIZE];
    bc2_sqrt = (float)dyn[VHMM_TS_BC2_SQRT];
  }
  const float s = scale ? *scale : 1.f;
  float coef = 1.f;
  if (sqnorm) {
    const float total = (float)sqrt(*sqnorm);
    if (gnorm_out && blockIdx.x == 0 && threadIdx.x == 0) *gnorm_out = total;
    if (max_norm > 0.f) coef = fminf(max_norm / (total + 1e-6f), 1.f);   // torch.nn.utils.clip_grad_norm_
  }
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256) {
    int t = 0;
#pragma unroll 1
    while (i >= tab.off[t + 1]) ++t;
    const float gi = __fmul_rn(__fmul_rn(g[i], s), coef);
    const float mi = __fmaf_rn(w1, __fsub_rn(gi, m[i]), m[i]);                                   // lerp_
    const float vi = __fmaf_rn(__fmul_rn(w2, gi), gi, __fmul_rn(v[i], beta2));                    // mul_, addcmul_
    m[i] = mi; v[i] = vi;
    const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(vi), bc2_sqrt), eps);
    float* p = tab.p[t] + (i - tab.off[t]);
    *p = __fmaf_rn(-step_size, __fdiv_rn(mi, denom), *p);                                         // addcdiv_
  }
}

int adam_step(const vhmm_dims& d, float* const* params, const float* grads, float* m, float* v, double lr,
              double beta1, double beta2, double eps, int64_t step, const float* grad_scale, double max_norm,
              float* gnorm_out, double* scratch /* 1 double on the device, or null when no norm is needed */,
              cudaStream_t st) {
  ParamTable tab;
  param_offsets(d, tab.off);
  for (int i = 0; i < VHMM_NUM_PARAMS; ++i) tab.p[i] = params[i];
  const int64_t n = tab.off[VHMM_NUM_PARAMS];
  const bool need_norm = (max_norm > 0.0) || gnorm_out;
  if (need_norm) {
    VHMM_REQUIRE(scratch, VHMM_E_WORKSPACE, "adam_step: clipping / norm output needs the 8-byte scratch");
    grad_sqnorm_kernel<<<1, 1024, 0, st>>>(grads, n, grad_scale, scratch);
    VHMM_LAUNCH_CHECK();
  }
  // bias corrections as torch computes them (python floats = double), then cast
  const double bc1 = 1.0 - pow(beta1, (double)step), bc2 = 1.0 - pow(beta2, (double)step);
  const float step_size = (float)(lr / bc1), bc2_sqrt = (float)sqrt(bc2);
  int64_t blocks = cdiv(n, 256);
  if (blocks > 1184) blocks = 1184;
  adam_kernel<<<(unsigned)blocks, 256, 0, st>>>(tab, grads, m, v, n, step_size, (float)(1.0 - beta1), (float)beta2,
                                               (float)(1.0 - beta2), bc2_sqrt, (float)eps, grad_scale,
                                               need_norm ? scratch : nullptr, (float)max_norm, gnorm_out, nullptr);
  VHMM_LAUNCH_CHECK();
  return 0;
}

// ---- f2: the whole small-batch training step with device-resident scalars (graph capture) -------------------
// one block: normalisers of this batch (VQ_VAE_HMM_fixed.py:120, :131, :135), step counter, Adam bias corrections
// computed as torch computes them (python floats = double)
__global__ void __launch_bounds__(256) train_prep_kernel(const int64_t* __restrict__ lengths, int64_t B, int C, int T,
                                                         double* __restrict__ st) {
  __shared__ double sh[32];
  double acc = 0.0;
  for (int64_t b = threadIdx.x; b < B; b += 256) {
    int64_t L = lengths[b];
    L = L < 0 ? 0 : (L > T ? T : L);
    acc += (double)L;
  }
  acc = block_sum(acc, sh);
  if (threadIdx.x == 0) {
    const double n0 = acc * (double)C;
    st[VHMM_TS_NORM0] = n0 < 1.0 ? 1.0 : n0;
    st[VHMM_TS_NORM1] = (double)B;
    const double step = st[VHMM_TS_STEP] + 1.0;
    st[VHMM_TS_STEP] = step;
    const double bc1 = 1.0 - pow(st[VHMM_TS_BETA1], step), bc2 = 1.0 - pow(st[VHMM_TS_BETA2], step);
    st[VHMM_TS_STEP_SIZE] = st[VHMM_TS_LR] / bc1;
    st[VHMM_TS_BC2_SQRT] = sqrt(bc2);
  }
}
// loss = recon + beta (prior - entropy) from the un-normalised sums (:137); fp32 like the reference's loss tensor
__global__ void train_loss_kernel(double* __restrict__ st) {
  if (threadIdx.x || blockIdx.x) return;
  const double* sums = st + VHMM_TS_SUMS;
  const double recon = sums[0] / st[VHMM_TS_NORM0], prior = -sums[1] / st[VHMM_TS_NORM1], ent = sums[2] / st[VHMM_TS_NORM1];
  const double loss = (double)(float)(recon + st[VHMM_TS_BETA] * (prior - ent));
  st[VHMM_TS_LOSS] = loss;
  st[VHMM_TS_LOSS_SUM] += loss;
  st[VHMM_TS_BATCHES] += 1.0;
}

int fused_elbo(const vhmm_dims&, const vhmm_params&, const float*, const float*, const int64_t*, int64_t, int64_t, const double*,
               double, double*, float*, void*, size_t, cudaStream_t, const double* beta_dev);
bool fused_supported(const vhmm_dims& d);
size_t fused_workspace_bytes(const vhmm_dims& d, int64_t B, int64_t T);

size_t train_step_workspace_bytes(const vhmm_dims& d, int64_t B, int64_t T) {
  return fused_supported(d) ? fused_workspace_bytes(d, B, T) + 256 : 0;
}

int train_step(const vhmm_dims& d, float* const* params, const float* x_cat, const float* u_cat, int64_t S_tot,
               const int64_t* col0, const int64_t* lengths, int64_t B, int64_t T, double* state, float* exp_avg,
               float* exp_avg_sq, double beta1, double beta2, double eps, double max_norm, float* x_buf, float* u_buf,
               float* grads, void* ws, size_t ws_bytes, cudaStream_t st) {
  VHMM_REQUIRE(fused_supported(d), VHMM_E_DIMS, "train_step: needs the dims of the fused ELBO path");
  int rc = gather_chunks(x_cat, u_cat, S_tot, d.C, d.U, col0, lengths, B, T, x_buf, u_buf, st);
  if (rc) return rc;
  train_prep_kernel<<<1, 256, 0, st>>>(lengths, B, d.C, (int)T, state);
  VHMM_LAUNCH_CHECK();
  vhmm_params P;
  for (int i = 0; i < VHMM_NUM_PARAMS; ++i) P.p[i] = params[i];
  Arena ar(ws, ws_bytes);
  double* sq = ar.take<double>(1);
  VHMM_REQUIRE(ar.ok, VHMM_E_WORKSPACE, "train_step: workspace too small");
  rc = fused_elbo(d, P, x_buf, u_buf, lengths, B, T, state + VHMM_TS_NORM0, 0.0, state + VHMM_TS_SUMS, grads,
                  (char*)ws + ar.used, ws_bytes - ar.used, st, state + VHMM_TS_BETA);
  if (rc) return rc;
  train_loss_kernel<<<1, 32, 0, st>>>(state);
  VHMM_LAUNCH_CHECK();
  ParamTable tab;
  param_offsets(d, tab.off);
  for (int i = 0; i < VHMM_NUM_PARAMS; ++i) tab.p[i] = params[i];
  const int64_t n = tab.off[VHMM_NUM_PARAMS];
  if (max_norm > 0.0) {
    grad_sqnorm_kernel<<<1, 1024, 0, st>>>(grads, n, nullptr, sq);
    VHMM_LAUNCH_CHECK();
  }
  int64_t blocks = cdiv(n, 256);
  if (blocks > 1184) blocks = 1184;
  // beta1 / beta2 are constants of a run (host arguments, also stored in the state for the bias corrections); the
  // step-dependent scalars (step size, sqrt of the second bias correction) are read from the state by the kernel
  adam_kernel<<<(unsigned)blocks, 256, 0, st>>>(tab, grads, exp_avg, exp_avg_sq, n, 0.f, (float)(1.0 - beta1), (float)beta2,
                                               (float)(1.0 - beta2), 1.f, (float)eps, nullptr,
                                               max_norm > 0.0 ? sq : nullptr, (float)max_norm, nullptr, state);
  VHMM_LAUNCH_CHECK();
  return 0;
}

}  // namespace vhmm
