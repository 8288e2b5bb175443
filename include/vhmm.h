/*
 * vhmm.h -- C ABI of libvhmm.so: the B200 (sm_100a) implementation of the VAE_HMM hot path
 * of yashnaray/VQ-VAE-HMM-model.
 *
 * The reference has no FFI / plugin layer: the path sits behind Python classes
 * (VQ_VAE_HMM_fixed.py:31-143).  Each entry point below states the reference lines it
 * replaces; the Python host (vq-vae-hmm-model_b200/) binds them with ctypes and wraps them in
 * torch.autograd.Function objects behind the reference's own class / method names.
 *
 * Conventions
 *   - Every pointer is a DEVICE pointer unless the name ends in _host.  The library allocates
 *     nothing and keeps no pointer after a call returns: the caller owns inputs, outputs and
 *     workspace (size from vhmm_workspace_bytes).  All work is enqueued on `stream`
 *     (a cudaStream_t passed as void*); no call synchronises the device.
 *   - Tensors are contiguous fp32 in the reference's layouts: x (B,C,T), u (B,U,T),
 *     logits/q (B,K,T), mu/logvar (B,C,T), log_A (B,T,K,K), log_b/gamma (B,T,K), xi (B,T,K,K);
 *     lengths is int64 (B,) on the device.
 *   - Return value: 0 ok; <0 invalid argument (VHMM_E_*); >0 a cudaError_t.  vhmm_last_error()
 *     returns a thread-local description of the last failure.  Nothing throws.
 *   - Parameters: 18 device pointers in state_dict order:
 *       0 encoder.conv1.weight (H,C,3)   1 encoder.conv1.bias (H)
 *       2 encoder.conv2.weight (H2,H,3)  3 encoder.conv2.bias (H2)
 *       4 encoder.to_logits.weight (K,H2,1) 5 encoder.to_logits.bias (K)
 *       6 prior.log_prior (K)
 *       7 prior.transition_net.0.weight (TH,U) 8 prior.transition_net.0.bias (TH)
 *       9 prior.transition_net.2.weight (K*K,TH) 10 prior.transition_net.2.bias (K*K)
 *      11 decoder.embeddings.weight (K,H)
 *      12 decoder.conv1.weight (H,H,3)  13 decoder.conv1.bias (H)
 *      14 decoder.conv2.weight (H,H,3)  15 decoder.conv2.bias (H)
 *      16 decoder.to_params.weight (2C,H,1) 17 decoder.to_params.bias (2C)
 *     The flat gradient buffer uses the same order, tensors back to back (vhmm_param_count).
 */
#ifndef VHMM_H_
#define VHMM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VHMM_VERSION 113 /* major*100 + minor */
#define VHMM_NUM_PARAMS 18
#define VHMM_PAD_STATE 255 /* Viterbi path value for t >= lengths[b] */

enum {
  VHMM_E_ARG = -1,       /* null pointer / non-positive size */
  VHMM_E_DIMS = -2,      /* unsupported model dimensions for this entry point */
  VHMM_E_WORKSPACE = -3, /* workspace too small */
  VHMM_E_ALIGN = -4      /* pointer not 16-byte aligned */
};

/* model dimensions: VAE_HMM(input_dim=C, hidden_dim=H, K, hidden_dim2=H2, u_dim=U,
 * trans_hidden=TH), VQ_VAE_HMM_fixed.py:93 */
typedef struct vhmm_dims {
  int32_t C, H, K, H2, U, TH;
} vhmm_dims;

typedef struct vhmm_params {
  const float* p[VHMM_NUM_PARAMS];
} vhmm_params;

/* ops for vhmm_workspace_bytes */
enum {
  VHMM_OP_ENCODER_FWD = 1,
  VHMM_OP_PRIOR_FWD = 2,
  VHMM_OP_DECODER_FWD = 3,
  VHMM_OP_ELBO = 4,
  VHMM_OP_HMM_FORWARD = 5,
  VHMM_OP_HMM_POSTERIOR = 6,
  VHMM_OP_VITERBI = 7,
  VHMM_OP_HMM_DECODE = 8,
  VHMM_OP_HMM_SCAN = 9,
  VHMM_OP_EMISSION_TABLE = 10,
  VHMM_OP_HMM_INPUTS = 11,
  VHMM_OP_HMM_PARAM_GRADS = 12,
  VHMM_OP_TRANSITION_PROBS = 13,
  VHMM_OP_FORWARD = 14,
  VHMM_OP_TRAIN_STEP = 15
};

int vhmm_version(void);
const char* vhmm_last_error(void);
/* kernels launched by this library in this process so far (bench.py gpu_launches) */
int64_t vhmm_launch_count(void);
/* number of floats in the flat parameter / gradient buffer for `dims` */
int64_t vhmm_param_count(const vhmm_dims* dims);
/* offsets (in floats) of the 18 tensors inside the flat buffer; out has 19 entries (last=total) */
int vhmm_param_offsets(const vhmm_dims* dims, int64_t* out_host);
/* minimum workspace for `op`; for VHMM_OP_ELBO any larger workspace lets the library process
 * more sequences per pass (it chunks the batch to fit). */
size_t vhmm_workspace_bytes(int op, const vhmm_dims* dims, int64_t B, int64_t T);

/* Encoder.forward, VQ_VAE_HMM_fixed.py:38-41 (+ softmax of :114/:141 when q != NULL).
 * x (B,C,T) -> logits (B,K,T) [, q (B,K,T)].  Either output may be NULL (not both). */
int vhmm_encoder_fwd(const vhmm_dims* dims, const vhmm_params* params, const float* x,
                     int64_t B, int64_t T, float* logits, float* q, void* workspace,
                     size_t workspace_bytes, void* stream);

/* VAE_HMM.forward, VQ_VAE_HMM_fixed.py:139-143: x (B,C,T) -> q = softmax(encoder(x)) (B,K,T), (mu, logvar) =
 * decoder(q) (B,C,T) each [, logits (B,K,T) when non-NULL].  For the dims of the FUSED path (see vhmm_elbo_select)
 * this -- and vhmm_encoder_fwd / vhmm_decoder_fwd -- is ONE persistent tile kernel (conv2 layers on tcgen05, every
 * activation on chip: 20 B/step read, 52 B/step written for forward()); other dims run the layer kernels.
 * Workspace: VHMM_OP_FORWARD. */
int vhmm_forward(const vhmm_dims* dims, const vhmm_params* params, const float* x, int64_t B, int64_t T,
                 float* logits, float* q, float* mu, float* logvar, void* workspace, size_t workspace_bytes,
                 void* stream);

/* Prior.forward, VQ_VAE_HMM_fixed.py:59-71.  u (B,U,T) (the layout collate_fn produces; the
 * Python wrapper applies the permute rule of :64-65) -> log_pi (K), log_A (B,T,K,K). */
int vhmm_prior_fwd(const vhmm_dims* dims, const vhmm_params* params, const float* u, int64_t B,
                   int64_t T, float* log_pi, float* log_A, void* workspace,
                   size_t workspace_bytes, void* stream);

/* Transition PROBABILITIES A = exp(log_A) = softmax rows of Prior.forward (VQ_VAE_HMM_fixed.py:68-69): the
 * (B,T,K,K) tensor the reference's consumers take as `transition_matrix` / `trans_matrix` / `transition_probs`
 * and index with [:, -1] (regime_utilities.py:42-53, delta_hedger.py:167-183, loss_functions.py:128-147;
 * VQ_VAE_HMM_fixed.py:214-228 regime_aware_sharpe_loss).  u (B,U,T) -> trans_probs (B,T,K,K), rows sum to 1.
 * Workspace: VHMM_OP_TRANSITION_PROBS. */
int vhmm_transition_probs(const vhmm_dims* dims, const vhmm_params* params, const float* u, int64_t B,
                          int64_t T, float* trans_probs, void* workspace, size_t workspace_bytes,
                          void* stream);

/* Decoder.forward, VQ_VAE_HMM_fixed.py:81-90.  q (B,K,T) -> mu (B,C,T), logvar (B,C,T). */
int vhmm_decoder_fwd(const vhmm_dims* dims, const vhmm_params* params, const float* q, int64_t B,
                     int64_t T, float* mu, float* logvar, void* workspace,
                     size_t workspace_bytes, void* stream);

/* VAE_HMM.compute_loss forward (+ its autograd backward when grads != NULL),
 * VQ_VAE_HMM_fixed.py:106-137 and loss.backward() at :156.
 *   norms      device double[2] = { max(N_valid_global*C, 1), B_global }: the reference's
 *              normalisers (:120, :131, :135); under batch sharding they are the GLOBAL values.
 *   beta       the KL weight (:137)
 *   sums       device double[4] out: { sum nll*mask, sum_b(init+trans), sum entropy*mask,
 *              valid steps } over THIS call's sequences (un-normalised).
 *              loss = sums[0]/norms[0] + beta*(-sums[1]/norms[1] - sums[2]/norms[1]).
 *   grads      device float[vhmm_param_count] out or NULL: d loss / d params for this call's
 *              sequences with the global normalisers applied (sum over ranks = full gradient).
 */
int vhmm_elbo(const vhmm_dims* dims, const vhmm_params* params, const float* x, const float* u,
              const int64_t* lengths, int64_t B, int64_t T, const double* norms, double beta,
              double* sums, float* grads, void* workspace, size_t workspace_bytes, void* stream);

/* Two implementations of vhmm_elbo share the signature above:
 *   VHMM_ELBO_LAYERWISE  one kernel per layer, activations stashed in the workspace; any dims.
 *   VHMM_ELBO_FUSED      three recompute-in-tile tensor-core kernels, no activation in HBM; needs
 *                        H%32==0, H2%16==0, TH%32==0, C,U,K<=8 (instantiated for the reference's
 *                        base config C=5,H=64,H2=32,K=3,U=4,TH=128).
 * VHMM_ELBO_AUTO (default) picks FUSED when the dims allow it; the environment variable
 * VHMM_ELBO_IMPL=layerwise|fused overrides AUTO.  vhmm_workspace_bytes(VHMM_OP_ELBO, ...) follows
 * the active choice; for FUSED it is an exact requirement (a larger workspace is not used). */
enum { VHMM_ELBO_AUTO = 0, VHMM_ELBO_LAYERWISE = 1, VHMM_ELBO_FUSED = 2 };
int vhmm_elbo_select(int impl);
/* the implementation vhmm_elbo would run for `dims` now (VHMM_ELBO_LAYERWISE / VHMM_ELBO_FUSED) */
int vhmm_elbo_active_impl(const vhmm_dims* dims);

/* The FUSED path has two builds of its decoder kernel (decoder forward + backward + NLL):
 *   VHMM_DECODER_TC5  the three 64x64 k=3 GEMM families on the 5th-generation tensor core
 *                     (tcgen05.mma, TMEM accumulators, 2 x fp16 split operands, fp32 accumulate)
 *   VHMM_DECODER_MMA  the same GEMMs on mma.sync with the 3 x TF32 split
 * Both meet the 1e-5 parity bar; AUTO (default) = TC5; environment VHMM_FUSED_DECODER=mma|tc5
 * overrides AUTO. */
enum { VHMM_DECODER_AUTO = 0, VHMM_DECODER_MMA = 1, VHMM_DECODER_TC5 = 2 };
int vhmm_elbo_decoder_select(int variant);
int vhmm_elbo_decoder_active(void);

/* Input windows of the tile kernels (FUSED ELBO kernels A and B, the fused forward kernel): 0 (default) = 16-byte cp.async
 * with zero-fill predicates, 1 = tensor-map TMA boxes (cp.async.bulk.tensor, out-of-bounds fill = the conv zero padding;
 * needs T % 4 == 0 and 16-byte aligned tensors, else the call falls back to cp.async).  Same results bit for bit; at cfg 2
 * the TMA build measured 0.8 % slower (85.4 vs 84.8 ms), hence the default.  Environment VHMM_TMA=1 seeds the choice. */
int vhmm_tma_select(int on);

/* Measurement hook for bench.py (roofline of the dominant kernel): while enabled, the FUSED path
 * records CUDA events on the caller's stream around its three tile kernels (up to 64 calls).
 * vhmm_elbo_timing_read waits for those events and returns the summed milliseconds of
 * {enc_prior, decoder, enc_bwd} and the number of calls recorded. */
int vhmm_elbo_timing(int enable);
int vhmm_elbo_timing_read(double* ms_sum_host /*[3]*/, int* ncalls_host);

/* Layer kernels of the LAYER-WISE path (any model dims; what runs BASELINE.json configs[4], hidden_dim 256).  Every layer
 * of the reference is an nn.Conv1d of width 3 (padding 1) or 1 -- or an nn.Linear applied per time step, the same thing
 * (VQ_VAE_HMM_fixed.py:34-36, :53-57, :77-79).  Two builds:
 *   VHMM_CONV_TC5   tcgen05.mma GEMM-convolutions (M = 128 time steps, N = 128 channels, fp32 parity through the
 *                   2 x fp16 split with fp32 TMEM accumulators): needs Cin % 16 == 0, Cout % 128 == 0, T % 4 == 0
 *   VHMM_CONV_FFMA  FP32 FFMA kernels, any shape
 * AUTO (default) picks TC5 per layer where the shape allows; environment VHMM_CONV=ffma|tc5 seeds the choice.
 * vhmm_conv1d_fwd is one such layer on its own: y (B,Cout,T) = [relu](bias + conv_kw(x (B,Cin,T), weight (Cout,Cin,kw))). */
enum { VHMM_CONV_AUTO = 0, VHMM_CONV_FFMA = 1, VHMM_CONV_TC5 = 2 };
int vhmm_conv_select(int impl);
size_t vhmm_conv1d_workspace_bytes(int64_t Cin, int64_t Cout, int64_t kw);
int vhmm_conv1d_fwd(const float* x, const float* weight, const float* bias, int64_t B, int64_t Cin, int64_t Cout,
                    int64_t T, int64_t kw, int relu, float* y, void* workspace, size_t workspace_bytes, void* stream);

/* Transition prior of the LAYER-WISE path at K = 32 with a 128-wide transition_net (BASELINE.json configs[4]):
 *   VHMM_PRIOR_FUSED   (AUTO) one kernel forms the K*K transition logits (tcgen05 GEMM), their row log-softmax, the pair term
 *                      sum_t sum_ij q_{t-1}(i) q_t(j) log A_t(i,j) and its gradients (VQ_VAE_HMM_fixed.py:53-57, :125-131);
 *                      d loss / d logits is written once as fp16 operand parts for the two backward GEMMs of transition_net.2
 *   VHMM_PRIOR_FUSED_F32  the same kernel, d loss / d logits as an fp32 tensor (the backward GEMMs convert it)
 *   VHMM_PRIOR_LAYERS  layer -> row log-softmax -> loss kernel, each a pass over the (B,K*K,T) tensor
 * All three meet the 1e-5 parity bar (tests run each); other shapes always take LAYERS.  Environment VHMM_PRIOR_FUSED=0 /
 * VHMM_PRIOR_PARTS=0 seed the choice. */
enum { VHMM_PRIOR_AUTO = 0, VHMM_PRIOR_FUSED_F32 = 1, VHMM_PRIOR_LAYERS = 2 };
int vhmm_prior_select(int mode);

/* ELBO of the LAYER-WISE path with tensor-core layer kernels (K % 16 == 0, hidden_dim % 16 == 0): decoder.embeddings folded into
 * decoder.conv1 -- e = E^T q (VQ_VAE_HMM_fixed.py:83) is linear, so conv1(e) is a K -> hidden_dim convolution of q with
 * W'[o,k,tau] = sum_d W1[o,d,tau] E[k,d]; forward, input gradient (d loss / dq directly) and weight gradient run on the
 * folded layer, the gradients of W1 and E follow by the chain rule.  1 (default) = on, 0 = the embedding stays a layer of its
 * own.  Same results up to fp32 re-association; environment VHMM_EMB_FOLD=0 seeds the choice. */
int vhmm_emb_fold_select(int on);

/* argmax regime decoding (backtesting.py:155, src/backtesting.py:107): q or logits (B,K,T) ->
 * regimes int64 (B,T), lowest index on ties. */
int vhmm_argmax_regimes(const float* q, int64_t B, int64_t K, int64_t T, int64_t* regimes,
                        void* stream);

/* ---- callers either side of the ELBO path ---- */

/* f1: device-side batching.  Replaces RandomChunkDataset.__getitem__'s slicing and collate_fn's
 * 2B slice copies (VQ_VAE_HMM_fixed.py:20-29, :164-179; src/utils/data.py:7-44) by one gather from
 * a corpus resident in HBM: x_cat (C,S_tot), u_cat (U,S_tot) = the sequences concatenated along
 * time.  col0[b] = absolute first column of item b (sequence offset + start), lengths[b] its
 * length; T = max length (the caller knows it: lengths are produced on the host, :165).
 * x_out (B,C,T), u_out (B,U,T): out[b,r,t] = t < lengths[b] ? cat[r, col0[b]+t] : 0. */
int vhmm_gather_chunks(const float* x_cat, const float* u_cat, int64_t S_tot, int64_t C, int64_t U,
                       const int64_t* col0, const int64_t* lengths, int64_t B, int64_t T,
                       float* x_out, float* u_out, void* stream);

/* f2: optimiser tail of train_model (VQ_VAE_HMM_fixed.py:146,:155-157) and Trainer.train_epoch
 * (src/training/trainer.py:24-34): torch.optim.Adam (no weight decay, no amsgrad) over the 18
 * parameter tensors from the FLAT gradient buffer of vhmm_elbo, optionally preceded by
 * clip_grad_norm_(max_norm) (trainer.py:31; max_norm <= 0: none).  grads are multiplied by
 * *grad_scale first when grad_scale != NULL (device float; the upstream d loss).  exp_avg /
 * exp_avg_sq: flat fp32 state, same layout as grads.  step = 1 for the first update.
 * grad_norm_out (device float, or NULL) receives the global L2 norm before clipping.  scratch:
 * 8 bytes of device memory (needed when clipping or grad_norm_out is requested). */
typedef struct vhmm_params_mut {
  float* p[VHMM_NUM_PARAMS];
} vhmm_params_mut;
int vhmm_adam_step(const vhmm_dims* dims, const vhmm_params_mut* params, const float* grads,
                   float* exp_avg, float* exp_avg_sq, double lr, double beta1, double beta2,
                   double eps, int64_t step, const float* grad_scale, double max_norm,
                   float* grad_norm_out, double* scratch, void* stream);

/* f2: ONE whole training step of train_model (VQ_VAE_HMM_fixed.py:153-157) / Trainer.train_epoch (trainer.py:24-34) for the
 * small-batch regime of BASELINE.json configs[0] (batch 64, T <= 200), where the step is launch-bound: batch assembly
 * from the resident corpus (vhmm_gather_chunks) + normalisers + ELBO forward/backward (FUSED path) + loss + optional
 * clip_grad_norm_ + Adam, enqueued by a single call in which EVERY scalar that changes from step to step lives in
 * DEVICE memory -- so the call can be captured once per (B, T) into a CUDA graph and replayed (the host then issues
 * one memcpy of (col0, lengths) and one graph launch per step).  `state`: device double[VHMM_TS_COUNT]:
 *   in : BETA (KL weight of this step), LR, BETA1, BETA2, STEP (number of updates done so far; incremented here)
 *   out: NORM0/NORM1 (the normalisers :120/:131), SUMS[4] (as vhmm_elbo), LOSS (this step, fp32-rounded like the
 *        reference's loss tensor), LOSS_SUM / BATCHES (running totals for the epoch mean the reference prints :160;
 *        the host resets them), STEP_SIZE / BC2_SQRT (Adam bias corrections of this step).
 * beta1 / beta2 / eps / max_norm are constants of a run (host arguments; beta1/beta2 must equal state[BETA1/2]).
 * x_buf (B,C,T), u_buf (B,U,T), grads (vhmm_param_count): caller-owned scratch.  Workspace: VHMM_OP_TRAIN_STEP. */
enum {
  VHMM_TS_NORM0 = 0, VHMM_TS_NORM1 = 1, VHMM_TS_BETA = 2, VHMM_TS_STEP = 3, VHMM_TS_LR = 4, VHMM_TS_BETA1 = 5,
  VHMM_TS_BETA2 = 6, VHMM_TS_LOSS = 7, VHMM_TS_LOSS_SUM = 8, VHMM_TS_BATCHES = 9, VHMM_TS_STEP_SIZE = 10,
  VHMM_TS_BC2_SQRT = 11, VHMM_TS_SUMS = 12, VHMM_TS_COUNT = 16
};
int vhmm_train_step(const vhmm_dims* dims, const vhmm_params_mut* params, const float* x_cat, const float* u_cat,
                    int64_t S_tot, const int64_t* col0, const int64_t* lengths, int64_t B, int64_t T, double* state,
                    float* exp_avg, float* exp_avg_sq, double beta1, double beta2, double eps, double max_norm,
                    float* x_buf, float* u_buf, float* grads, void* workspace, size_t workspace_bytes, void* stream);

/* ---- exact-HMM stages: extensions, absent from the reference ---- */

/* B.1: decoder on the K one-hot sequences of length T -> mu_tab (K,C,T), logvar_tab (K,C,T). */
int vhmm_emission_table(const vhmm_dims* dims, const vhmm_params* params, int64_t T,
                        float* mu_tab, float* logvar_tab, void* workspace, size_t workspace_bytes,
                        void* stream);
/* B.1: log_b[b,t,k] = -0.5 sum_c [log(2 pi v) + (x-mu)^2/v], v = max(exp(logvar),1e-8). */
int vhmm_emission_logprob(const float* x, const float* mu_tab, const float* logvar_tab, int64_t B,
                          int64_t C, int64_t T, int64_t K, float* log_b, void* stream);
/* Front end of the recursion stages, fused (one FFMA kernel, no activation traffic): log_A (B,T,K,K) =
 * Prior.forward(u) (VQ_VAE_HMM_fixed.py:59-71: Linear-ReLU-Linear on u[b,:,t], row log-softmax) and
 * log_b (B,T,K) = the per-state emission log-probabilities of B.1 from the emission table.  Same results
 * as vhmm_prior_fwd + vhmm_emission_logprob.  Instantiated for C=5, K=3, U=4, TH=128 and T % 4 == 0
 * (vhmm_hmm_inputs_supported); x, u, log_A, log_b 16-byte aligned. */
int vhmm_hmm_inputs_supported(const vhmm_dims* dims, int64_t T);
int vhmm_hmm_inputs(const vhmm_dims* dims, const vhmm_params* params, const float* x, const float* u,
                    int64_t B, int64_t T, const float* mu_tab, const float* logvar_tab, float* log_A,
                    float* log_b, void* workspace, size_t workspace_bytes, void* stream);
/* The whole log-likelihood path in one kernel: transition softmax + emission log-probabilities + forward
 * recursion, one warp per sequence, alpha in registers, log_A / log_b never materialised (HBM traffic = x and
 * u only).  loglik (B) = vhmm_hmm_forward(log_pi, vhmm_hmm_inputs(x, u)) to fp32 rounding.  log_pi: K
 * normalised log-probabilities on the device.  Same instantiation and workspace (VHMM_OP_HMM_INPUTS) as
 * vhmm_hmm_inputs.  Replaces the method api.py:18 / SURVEY B.2 reach through model.hmm_log_likelihood. */
int vhmm_hmm_loglik_fused(const vhmm_dims* dims, const vhmm_params* params, const float* log_pi, const float* x,
                          const float* u, const int64_t* lengths, int64_t B, int64_t T, const float* mu_tab,
                          const float* logvar_tab, float* loglik, void* workspace, size_t workspace_bytes,
                          void* stream);
/* Conventions of the recursion stages for ragged batches (`lengths`, clamped to [0, T]): steps t >= lengths[b] are padding --
 * alpha rows carry the last valid row unchanged, gamma / xi rows are zero, path bytes are VHMM_PAD_STATE (0xFF).  A sequence of
 * length 0 has NO steps: loglik = 0, score = 0, every output row is padding (the CPU oracle's forward() would return
 * logsumexp(alpha_0) for it, i.e. treats it as length 1; tests compare such rows through max(lengths, 1)). */
/* B.2: forward recursion.  loglik (B); alpha (B,T,K) log-space or NULL. */
int vhmm_hmm_forward(const float* log_pi, const float* log_A, const float* log_b,
                     const int64_t* lengths, int64_t B, int64_t T, int64_t K, float* loglik,
                     float* alpha, void* stream);
/* B.3: forward + backward recursion = gradient of loglik: gamma (B,T,K), xi (B,T,K,K) or NULL. */
int vhmm_hmm_posterior(const float* log_pi, const float* log_A, const float* log_b,
                       const int64_t* lengths, int64_t B, int64_t T, int64_t K, float* loglik,
                       float* gamma, float* xi, void* workspace, size_t workspace_bytes,
                       void* stream);
/* B.3, last sentence: the beta-recursion gradient chained to the 18 parameters.  For
 * loss = sum_b w[b] * loglik[b]  (w = d loss / d loglik, device float (B); the exact-HMM loss hmm_nll uses
 * w = -1/B) and the posterior pair (gamma, xi) of vhmm_hmm_posterior on the inputs vhmm_hmm_inputs /
 * vhmm_prior_fwd + vhmm_emission_logprob produced from (x, u):  grads (flat, vhmm_param_count floats, the layout
 * of vhmm_elbo's) is OVERWRITTEN with d loss / d params -- through the row log-softmax and the transition MLP
 * (VQ_VAE_HMM_fixed.py:68-69), through the per-state emission table and the decoder evaluated on the K one-hot
 * sequences (B.1; :81-90, :118-119), and through log_pi = log_softmax(log_prior) (:54, :71).  The six encoder
 * tensors do not enter the exact likelihood: their gradients are zero.  Any dims with K <= 32; the batch is
 * processed in chunks that fit the workspace (minimum: VHMM_OP_HMM_PARAM_GRADS).  This is what the backward of the
 * torch.autograd.Function behind model.hmm_nll / hmm_log_likelihood calls (train_model call shape :153-157). */
int vhmm_hmm_param_grads(const vhmm_dims* dims, const vhmm_params* params, const float* x, const float* u,
                         const float* log_A, const float* gamma, const float* xi, const float* w,
                         const float* mu_tab, const float* logvar_tab, int64_t B, int64_t T, float* grads,
                         void* workspace, size_t workspace_bytes, void* stream);
/* vhmm_hmm_posterior / vhmm_hmm_decode have two builds for K <= 4:
 *   VHMM_POSTERIOR_ONEPASS  one warp per sequence, the sequence resident in shared memory, time
 *                           parallelised over the lanes (chunk products + warp scan): inputs and
 *                           outputs cross HBM once (60 B/step for gamma, 96 with xi).  Needs
 *                           T*K % 4 == 0, 16-byte aligned tensors and T <= 512 (K=3).
 *   VHMM_POSTERIOR_SWEEP    forward sweep (alpha_hat stash) + backward sweep, one thread per
 *                           sequence: any T, 132 B/step of traffic.
 * AUTO (default) = ONEPASS when it applies; environment VHMM_POSTERIOR=sweep overrides AUTO.
 * Longer sequences (T >= 512, T*K % 4 == 0; AUTO, vhmm_hmm_posterior / _decode / _scan): time-parallel as well -- 256-step segments
 * (the last one of a sequence may be shorter): segment
 * products, a forward and a backward carry scan per sequence (incoming alpha / outgoing beta of every segment, the
 * log-likelihood), then the ONEPASS kernel on the segment rows; the inputs cross HBM twice.  (The sweep kernels run one
 * thread per sequence: 94 ms at B = 64, T = 65536 against 0.16 ms.)
 * 4 < K <= 32 with T >= 512 and B <= 256 (few long sequences; needs the workspace vhmm_workspace_bytes reports, a smaller one
 * takes the warp-per-sequence kernels): 128-step segment products (K x K), forward and backward carries, then the
 * warp-per-sequence forward / backward kernels on the segment rows; environment VHMM_WIDE_SCAN=off disables it. */
enum { VHMM_POSTERIOR_AUTO = 0, VHMM_POSTERIOR_SWEEP = 1 };
int vhmm_hmm_posterior_select(int variant);
/* Numerics of the sum-product recursions (vhmm_hmm_forward / _posterior / _decode / _scan and what is built on them).
 *   VHMM_NUMERICS_SCALED    (default) scaled linear space in fp32: alpha_hat / beta_hat normalised every step, transfer-matrix
 *                           products rescaled by powers of two.  Within 1e-5 of float64 as long as no state probability
 *                           that matters LATER has dropped below 2^-126 of the most likely state -- always the case for
 *                           dense transition rows (anything a softmax produced, e.g. Prior.forward) with per-step emission
 *                           log-ratios up to ~60.  NOT guaranteed with structural zeros (log_A = -inf: left-to-right
 *                           chains) under sharply discriminative emissions, where a state that cannot be re-entered can
 *                           underflow and be needed again.  vhmm_hmm_scan detects products that vanished entirely and
 *                           recomputes those sequences with the per-step recursion.
 *   VHMM_NUMERICS_LOGSPACE  logsumexp recursions for every K <= 32 (one warp per sequence): any dynamic range,
 *                           several times slower.  Viterbi is max-plus in log space in both modes (bit-exact paths). */
enum { VHMM_NUMERICS_SCALED = 0, VHMM_NUMERICS_LOGSPACE = 1 };
int vhmm_hmm_numerics_select(int mode);
/* B.4: Viterbi, uint8 back-pointers, lowest-index tie-break.  path (B,T) uint8, score (B).  The recursion is serial in time
 * (bit-exact max-plus); FEW sequences (K <= 4: B < 2368, T % 4 == 0; 4 < K <= 32: B <= 148, K % 4 == 0) take one warp each
 * with the inputs streamed through a shared-memory ring -- the same paths and scores as the batch kernels. */
int vhmm_viterbi(const float* log_pi, const float* log_A, const float* log_b,
                 const int64_t* lengths, int64_t B, int64_t T, int64_t K, uint8_t* path,
                 float* score, void* workspace, size_t workspace_bytes, void* stream);
/* Allocation-free variant of vhmm_viterbi.  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same
 * results as vhmm_viterbi, bit for bit; the back-pointer scratch is a static device array of
 * VHMM_VITERBI_STATIC_MAX_BYTES (4 MiB) inside the library, held by every GPU that loads it until exit.  Limit:
 * B * T * (K <= 4 ? 1 : K) bytes <= VHMM_VITERBI_STATIC_MAX_BYTES (K <= 4: 4,194,304 sequence-steps per call; K = 32:
 * 131,072) -- beyond it the call returns VHMM_E_WORKSPACE and vhmm_last_error() names the limit.  The call performs no
 * cudaMalloc / cudaFree and calls no allocating library routine; the first call on a device resolves the array's device
 * address (cudaGetSymbolAddress) -- a call with B == 0 does only that and may be used as set-up before a capture begins.
 * All calls share the one array: calls that may overlap on the GPU (different streams) race on it and must be ordered by the
 * caller; nothing in it needs clearing between calls. */
#define VHMM_VITERBI_STATIC_MAX_BYTES (4u << 20)
int vhmm_viterbi_static(const float* log_pi, const float* log_A, const float* log_b,
                        const int64_t* lengths, int64_t B, int64_t T, int64_t K, uint8_t* path,
                        float* score, void* stream);
/* cfg 3: Viterbi path + posterior gamma.  Default: the Viterbi sweep, then the one-pass posterior kernel (the inputs are read
 * twice: 109 B/step moved for 61 algorithmic).  VHMM_DECODE_ONE_KERNEL (environment VHMM_DECODE=one; K <= 4 with a sequence
 * that fits one warp's shared memory) reads them ONCE: the one-pass posterior kernel plus a Viterbi helper warp per CTA that
 * walks the raw log_A / log_b in shared memory before the posterior passes overwrite them.  Bit-identical paths, the same gamma
 * -- and measured 2x SLOWER at cfg 3 (14.5 vs 7.0 ms: the serial max-plus chain competes for issue slots with the issue-bound
 * posterior passes), hence opt-in. */
enum { VHMM_DECODE_AUTO = 0, VHMM_DECODE_ONE_KERNEL = 1 };
int vhmm_hmm_decode_select(int mode);
int vhmm_hmm_decode(const float* log_pi, const float* log_A, const float* log_b,
                    const int64_t* lengths, int64_t B, int64_t T, int64_t K, uint8_t* path,
                    float* gamma, float* loglik, void* workspace, size_t workspace_bytes,
                    void* stream);
/* B.5: log-semiring parallel scan over T (long sequences, small batch).  loglik (B);
 * alpha (B,T,K) all-prefix forward variables or NULL.  K <= 4: 256-step segments on the one-pass warp kernel + carry
 * scan (sequences that fit one warp, T < 512, are the one-pass kernel itself).  4 < K <= 32, T >= 512, B <= 256:
 * 128-step segment products (K x K, one warp each, row-wise power-of-two scaling) + per-sequence carry (+ the forward
 * kernel on the segment rows for alpha); otherwise the warp-per-sequence forward of vhmm_hmm_forward (same results). */
int vhmm_hmm_scan(const float* log_pi, const float* log_A, const float* log_b,
                  const int64_t* lengths, int64_t B, int64_t T, int64_t K, float* loglik,
                  float* alpha, void* workspace, size_t workspace_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VHMM_H_ */
