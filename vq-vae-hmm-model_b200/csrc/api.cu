// api.cu -- extern "C" entry points of libvhmm.so (include/vhmm.h).  Argument validation and
// dispatch only; kernels live in the other translation units.
#include <atomic>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace vhmm {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

int current_device() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0) dev = 0;
  return dev < VHMM_MAX_DEVICES ? dev : VHMM_MAX_DEVICES - 1;
}
int device_sm_count() {
  static std::atomic<int> sms[VHMM_MAX_DEVICES];
  const int dev = current_device();
  int n = sms[dev].load(std::memory_order_relaxed);
  if (n == 0) {
    int v = 0;
    n = (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0) ? v : 148;
    sms[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

// elbo_layerwise.cu
size_t layerwise_workspace_bytes(int op, const vhmm_dims& d, int64_t B, int64_t T);
int layerwise_encoder_fwd(const vhmm_dims&, const vhmm_params&, const float*, int64_t, int64_t, float*, float*, void*, size_t, cudaStream_t);
int layerwise_prior_fwd(const vhmm_dims&, const vhmm_params&, const float*, int64_t, int64_t, float*, float*, void*, size_t, cudaStream_t, int as_probs, bool fused_transitions);
size_t hmm_param_grads_workspace_bytes(const vhmm_dims& d, int64_t B, int64_t T);
int hmm_param_grads(const vhmm_dims&, const vhmm_params&, const float*, const float*, const float*, const float*, const float*,
                    const float*, const float*, const float*, int64_t, int64_t, float*, void*, size_t, cudaStream_t);
int layerwise_decoder_fwd(const vhmm_dims&, const vhmm_params&, const float*, int64_t, int64_t, float*, float*, void*, size_t, cudaStream_t);
int layerwise_elbo(const vhmm_dims&, const vhmm_params&, const float*, const float*, const int64_t*, int64_t, int64_t, const double*, double, double*, float*, void*, size_t, cudaStream_t);
int layerwise_emission_table(const vhmm_dims&, const vhmm_params&, int64_t, float*, float*, void*, size_t, cudaStream_t);
int launch_argmax(const float* q, int64_t B, int64_t K, int64_t T, int64_t* out, cudaStream_t st);
int conv_impl_select(int mode);
int prior_impl_select(int mode);
int emb_fold_select(int on);
size_t conv1d_workspace_bytes(int Cin, int Cout, int kw);
int conv1d_fwd(const float*, const float*, const float*, int64_t, int, int, int64_t, int, int, float*, void*, size_t, cudaStream_t);
// elbo_fused.cu
bool fused_supported(const vhmm_dims& d);
size_t fused_workspace_bytes(const vhmm_dims& d, int64_t B, int64_t T);
int fused_elbo(const vhmm_dims&, const vhmm_params&, const float*, const float*, const int64_t*, int64_t, int64_t, const double*, double, double*, float*, void*, size_t, cudaStream_t, const double* beta_dev);
size_t fused_forward_workspace_bytes(const vhmm_dims& d);
int fused_forward(const vhmm_dims&, const vhmm_params&, const float*, const float*, int64_t, int64_t, float*, float*, float*, float*,
                  void*, size_t, cudaStream_t);
int fused_timing_enable(int enable);
int fused_decoder_select(int variant);
int fused_decoder_active();
int fused_tma_select(int on);
int fused_timing_read(double* ms_sum, int* ncalls);
// hmm.cu
size_t hmm_workspace_bytes(int op, int64_t B, int64_t T, int64_t K);
int hmm_emission_logprob(const float*, const float*, const float*, int64_t, int64_t, int64_t, int64_t, float*, cudaStream_t);
int hmm_forward(const float*, const float*, const float*, const int64_t*, int64_t, int64_t, int64_t, float*, float*, cudaStream_t);
int hmm_posterior(const float*, const float*, const float*, const int64_t*, int64_t, int64_t, int64_t, float*, float*, float*, void*, size_t, cudaStream_t);
int hmm_viterbi(const float*, const float*, const float*, const int64_t*, int64_t, int64_t, int64_t, uint8_t*, float*, void*, size_t, cudaStream_t);
int hmm_decode_select(int mode);
int hmm_viterbi_static(const float*, const float*, const float*, const int64_t*, int64_t, int64_t, int64_t, uint8_t*, float*, cudaStream_t);
int hmm_decode(const float*, const float*, const float*, const int64_t*, int64_t, int64_t, int64_t, uint8_t*, float*, float*, void*, size_t, cudaStream_t);
int hmm_scan(const float*, const float*, const float*, const int64_t*, int64_t, int64_t, int64_t, float*, float*, void*, size_t, cudaStream_t);
int hmm_posterior_select(int mode);
int hmm_numerics_select(int mode);
// hmm_inputs.cu
bool hmm_inputs_supported(const vhmm_dims& d, int64_t T);
size_t hmm_inputs_workspace_bytes(const vhmm_dims& d, int64_t T);
int hmm_inputs(const vhmm_dims&, const vhmm_params&, const float*, const float*, int64_t, int64_t, const float*, const float*, float*, float*, void*, size_t, cudaStream_t);
int hmm_loglik_fused(const vhmm_dims&, const vhmm_params&, const float*, const float*, const float*, const int64_t*, int64_t, int64_t, const float*, const float*, float*, void*, size_t, cudaStream_t);
// batch_train.cu
int gather_chunks(const float*, const float*, int64_t, int64_t, int64_t, const int64_t*, const int64_t*, int64_t, int64_t, float*, float*, cudaStream_t);
int adam_step(const vhmm_dims&, float* const*, const float*, float*, float*, double, double, double, double, int64_t, const float*, double, float*, double*, cudaStream_t);
size_t train_step_workspace_bytes(const vhmm_dims& d, int64_t B, int64_t T);
int train_step(const vhmm_dims&, float* const*, const float*, const float*, int64_t, const int64_t*, const int64_t*, int64_t, int64_t,
               double*, float*, float*, double, double, double, double, float*, float*, float*, void*, size_t, cudaStream_t);

static std::atomic<int> g_elbo_impl{VHMM_ELBO_AUTO};
static int elbo_impl_for(const vhmm_dims& d) {
  int sel = g_elbo_impl.load();
  if (sel == VHMM_ELBO_AUTO) {
    const char* e = getenv("VHMM_ELBO_IMPL");
    if (e && !strcmp(e, "layerwise")) sel = VHMM_ELBO_LAYERWISE;
    if (e && !strcmp(e, "fused")) sel = VHMM_ELBO_FUSED;
  }
  if (sel == VHMM_ELBO_LAYERWISE) return VHMM_ELBO_LAYERWISE;
  return fused_supported(d) ? VHMM_ELBO_FUSED : VHMM_ELBO_LAYERWISE;
}

static bool params_ok(const vhmm_params* p) {
  if (!p) return false;
  for (int i = 0; i < VHMM_NUM_PARAMS; ++i)
    if (!p->p[i]) return false;
  return true;
}

}  // namespace vhmm

using namespace vhmm;

#define ARG(cond, what) VHMM_REQUIRE((cond), VHMM_E_ARG, "%s: invalid argument: %s", __func__, what)

extern "C" {

int vhmm_version(void) { return VHMM_VERSION; }

const char* vhmm_last_error(void) { return g_err; }

int64_t vhmm_launch_count(void) { return (int64_t)g_launches.load(); }

int64_t vhmm_param_count(const vhmm_dims* dims) {
  if (!dims_ok(dims)) return -1;
  int64_t off[VHMM_NUM_PARAMS + 1];
  param_offsets(*dims, off);
  return off[VHMM_NUM_PARAMS];
}

int vhmm_param_offsets(const vhmm_dims* dims, int64_t* out_host) {
  ARG(dims_ok(dims), "dims");
  ARG(out_host, "out_host");
  param_offsets(*dims, out_host);
  return 0;
}

size_t vhmm_workspace_bytes(int op, const vhmm_dims* dims, int64_t B, int64_t T) {
  if (B <= 0 || T <= 0) return 0;
  switch (op) {
    case VHMM_OP_HMM_FORWARD: case VHMM_OP_HMM_POSTERIOR: case VHMM_OP_VITERBI:
    case VHMM_OP_HMM_DECODE: case VHMM_OP_HMM_SCAN:
      return hmm_workspace_bytes(op, B, T, dims ? dims->K : 0);
    case VHMM_OP_HMM_INPUTS:
      return dims_ok(dims) ? hmm_inputs_workspace_bytes(*dims, T) : 0;
    case VHMM_OP_HMM_PARAM_GRADS:
      return dims_ok(dims) ? hmm_param_grads_workspace_bytes(*dims, B, T) : 0;
    case VHMM_OP_TRAIN_STEP:
      return dims_ok(dims) ? train_step_workspace_bytes(*dims, B, T) : 0;
    case VHMM_OP_TRANSITION_PROBS:
      return dims_ok(dims) ? layerwise_workspace_bytes(VHMM_OP_PRIOR_FWD, *dims, B, T) : 0;
    default:
      if (!dims_ok(dims)) return 0;
      if (op == VHMM_OP_ELBO && elbo_impl_for(*dims) == VHMM_ELBO_FUSED) return fused_workspace_bytes(*dims, B, T);
      if ((op == VHMM_OP_ENCODER_FWD || op == VHMM_OP_DECODER_FWD || op == VHMM_OP_FORWARD) &&
          elbo_impl_for(*dims) == VHMM_ELBO_FUSED)
        return fused_forward_workspace_bytes(*dims);   // the fused tile kernel keeps every activation on chip
      if (op == VHMM_OP_FORWARD)
        return layerwise_workspace_bytes(VHMM_OP_ENCODER_FWD, *dims, B, T) + layerwise_workspace_bytes(VHMM_OP_DECODER_FWD, *dims, B, T);
      return layerwise_workspace_bytes(op, *dims, B, T);
  }
}

int vhmm_encoder_fwd(const vhmm_dims* dims, const vhmm_params* params, const float* x, int64_t B,
                     int64_t T, float* logits, float* q, void* workspace, size_t workspace_bytes,
                     void* stream) {
  ARG(dims_ok(dims), "dims"); ARG(params_ok(params), "params"); ARG(x, "x");
  ARG(B > 0 && T > 0, "B,T"); ARG(logits || q, "logits and q both NULL"); ARG(workspace, "workspace");
  if (elbo_impl_for(*dims) == VHMM_ELBO_FUSED)
    return fused_forward(*dims, *params, x, nullptr, B, T, logits, q, nullptr, nullptr, workspace, workspace_bytes,
                         (cudaStream_t)stream);
  return layerwise_encoder_fwd(*dims, *params, x, B, T, logits, q, workspace, workspace_bytes, (cudaStream_t)stream);
}

int vhmm_forward(const vhmm_dims* dims, const vhmm_params* params, const float* x, int64_t B, int64_t T,
                 float* logits, float* q, float* mu, float* logvar, void* workspace, size_t workspace_bytes,
                 void* stream) {
  ARG(dims_ok(dims), "dims"); ARG(params_ok(params), "params"); ARG(x, "x");
  ARG(B > 0 && T > 0, "B,T"); ARG(q && mu && logvar, "q, mu, logvar"); ARG(workspace, "workspace");
  if (elbo_impl_for(*dims) == VHMM_ELBO_FUSED)
    return fused_forward(*dims, *params, x, nullptr, B, T, logits, q, mu, logvar, workspace, workspace_bytes,
                         (cudaStream_t)stream);
  const size_t we = layerwise_workspace_bytes(VHMM_OP_ENCODER_FWD, *dims, B, T);
  VHMM_REQUIRE(workspace_bytes >= we, VHMM_E_WORKSPACE, "vhmm_forward: workspace %zu < %zu", workspace_bytes, we);
  int rc = layerwise_encoder_fwd(*dims, *params, x, B, T, logits, q, workspace, we, (cudaStream_t)stream);
  if (rc) return rc;
  return layerwise_decoder_fwd(*dims, *params, q, B, T, mu, logvar, (char*)workspace + we, workspace_bytes - we,
                               (cudaStream_t)stream);
}

int vhmm_prior_fwd(const vhmm_dims* dims, const vhmm_params* params, const float* u, int64_t B,
                   int64_t T, float* log_pi, float* log_A, void* workspace, size_t workspace_bytes,
                   void* stream) {
  ARG(dims_ok(dims), "dims"); ARG(params_ok(params), "params"); ARG(u, "u");
  ARG(B > 0 && T > 0, "B,T"); ARG(log_pi && log_A, "outputs"); ARG(workspace, "workspace");
  return layerwise_prior_fwd(*dims, *params, u, B, T, log_pi, log_A, workspace, workspace_bytes, (cudaStream_t)stream, 0,
                             elbo_impl_for(*dims) == VHMM_ELBO_FUSED);
}

int vhmm_transition_probs(const vhmm_dims* dims, const vhmm_params* params, const float* u, int64_t B, int64_t T,
                          float* trans_probs, void* workspace, size_t workspace_bytes, void* stream) {
  ARG(dims_ok(dims), "dims"); ARG(params_ok(params), "params"); ARG(u, "u");
  ARG(B > 0 && T > 0, "B,T"); ARG(trans_probs, "trans_probs"); ARG(workspace, "workspace");
  return layerwise_prior_fwd(*dims, *params, u, B, T, nullptr, trans_probs, workspace, workspace_bytes, (cudaStream_t)stream, 1,
                             elbo_impl_for(*dims) == VHMM_ELBO_FUSED);
}

int vhmm_hmm_param_grads(const vhmm_dims* dims, const vhmm_params* params, const float* x, const float* u,
                         const float* log_A, const float* gamma, const float* xi, const float* w, const float* mu_tab,
                         const float* logvar_tab, int64_t B, int64_t T, float* grads, void* workspace,
                         size_t workspace_bytes, void* stream) {
  ARG(dims_ok(dims), "dims"); ARG(params_ok(params), "params"); ARG(x && u, "x,u");
  ARG(log_A && gamma && xi && w, "log_A,gamma,xi,w"); ARG(mu_tab && logvar_tab, "tables");
  ARG(B > 0 && T > 0, "B,T"); ARG(grads, "grads"); ARG(workspace, "workspace");
  return hmm_param_grads(*dims, *params, x, u, log_A, gamma, xi, w, mu_tab, logvar_tab, B, T, grads, workspace,
                         workspace_bytes, (cudaStream_t)stream);
}

int vhmm_decoder_fwd(const vhmm_dims* dims, const vhmm_params* params, const float* q, int64_t B,
                     int64_t T, float* mu, float* logvar, void* workspace, size_t workspace_bytes,
                     void* stream) {
  ARG(dims_ok(dims), "dims"); ARG(params_ok(params), "params"); ARG(q, "q");
  ARG(B > 0 && T > 0, "B,T"); ARG(mu && logvar, "outputs"); ARG(workspace, "workspace");
  if (elbo_impl_for(*dims) == VHMM_ELBO_FUSED)
    return fused_forward(*dims, *params, nullptr, q, B, T, nullptr, nullptr, mu, logvar, workspace, workspace_bytes,
                         (cudaStream_t)stream);
  return layerwise_decoder_fwd(*dims, *params, q, B, T, mu, logvar, workspace, workspace_bytes, (cudaStream_t)stream);
}

int vhmm_elbo(const vhmm_dims* dims, const vhmm_params* params, const float* x, const float* u,
              const int64_t* lengths, int64_t B, int64_t T, const double* norms, double beta,
              double* sums, float* grads, void* workspace, size_t workspace_bytes, void* stream) {
  ARG(dims_ok(dims), "dims"); ARG(params_ok(params), "params"); ARG(x && u && lengths, "x/u/lengths");
  ARG(B > 0 && T > 0, "B,T"); ARG(norms && sums, "norms/sums"); ARG(workspace, "workspace");
  if (elbo_impl_for(*dims) == VHMM_ELBO_FUSED)
    return fused_elbo(*dims, *params, x, u, lengths, B, T, norms, beta, sums, grads, workspace,
                      workspace_bytes, (cudaStream_t)stream, nullptr);
  return layerwise_elbo(*dims, *params, x, u, lengths, B, T, norms, beta, sums, grads, workspace,
                        workspace_bytes, (cudaStream_t)stream);
}

int vhmm_elbo_select(int impl) {
  ARG(impl == VHMM_ELBO_AUTO || impl == VHMM_ELBO_LAYERWISE || impl == VHMM_ELBO_FUSED, "impl");
  g_elbo_impl.store(impl);
  return 0;
}

int vhmm_elbo_decoder_select(int variant) {
  ARG(variant == VHMM_DECODER_AUTO || variant == VHMM_DECODER_MMA || variant == VHMM_DECODER_TC5, "variant");
  return fused_decoder_select(variant);
}
int vhmm_hmm_posterior_select(int variant) {
  ARG(variant == VHMM_POSTERIOR_AUTO || variant == VHMM_POSTERIOR_SWEEP, "variant");
  return hmm_posterior_select(variant);
}
int vhmm_conv_select(int impl) {
  ARG(impl == VHMM_CONV_AUTO || impl == VHMM_CONV_FFMA || impl == VHMM_CONV_TC5, "impl");
  return conv_impl_select(impl);
}
size_t vhmm_conv1d_workspace_bytes(int64_t Cin, int64_t Cout, int64_t kw) {
  return (Cin > 0 && Cout > 0 && (kw == 1 || kw == 3)) ? conv1d_workspace_bytes((int)Cin, (int)Cout, (int)kw) : 0;
}
int vhmm_conv1d_fwd(const float* x, const float* weight, const float* bias, int64_t B, int64_t Cin, int64_t Cout, int64_t T,
                    int64_t kw, int relu, float* y, void* workspace, size_t workspace_bytes, void* stream) {
  ARG(x && weight && y, "x/weight/y"); ARG(B > 0 && Cin > 0 && Cout > 0 && T > 0, "sizes"); ARG(kw == 1 || kw == 3, "kw");
  return conv1d_fwd(x, weight, bias, B, (int)Cin, (int)Cout, T, (int)kw, relu, y, workspace, workspace_bytes, (cudaStream_t)stream);
}
int vhmm_hmm_numerics_select(int mode) {
  ARG(mode == VHMM_NUMERICS_SCALED || mode == VHMM_NUMERICS_LOGSPACE, "mode");
  return hmm_numerics_select(mode);
}
int vhmm_prior_select(int mode) {
  ARG(mode == VHMM_PRIOR_AUTO || mode == VHMM_PRIOR_FUSED_F32 || mode == VHMM_PRIOR_LAYERS, "mode");
  return prior_impl_select(mode);
}
int vhmm_emb_fold_select(int on) { return emb_fold_select(on); }
int vhmm_hmm_decode_select(int mode) {
  ARG(mode == VHMM_DECODE_AUTO || mode == VHMM_DECODE_ONE_KERNEL, "mode");
  return hmm_decode_select(mode);
}
int vhmm_elbo_decoder_active(void) { return fused_decoder_active(); }
int vhmm_tma_select(int on) { return fused_tma_select(on); }

int vhmm_elbo_timing(int enable) { return fused_timing_enable(enable); }

int vhmm_elbo_timing_read(double* ms_sum_host, int* ncalls_host) {
  ARG(ms_sum_host && ncalls_host, "outputs");
  return fused_timing_read(ms_sum_host, ncalls_host);
}

int vhmm_elbo_active_impl(const vhmm_dims* dims) {
  if (!dims_ok(dims)) return VHMM_E_ARG;
  return elbo_impl_for(*dims);
}

int vhmm_argmax_regimes(const float* q, int64_t B, int64_t K, int64_t T, int64_t* regimes, void* stream) {
  ARG(q && regimes, "q/regimes"); ARG(B > 0 && K > 0 && T > 0, "B,K,T");
  return launch_argmax(q, B, K, T, regimes, (cudaStream_t)stream);
}

int vhmm_hmm_inputs_supported(const vhmm_dims* dims, int64_t T) { return dims_ok(dims) && T > 0 && hmm_inputs_supported(*dims, T) ? 1 : 0; }

int vhmm_hmm_inputs(const vhmm_dims* dims, const vhmm_params* params, const float* x, const float* u, int64_t B, int64_t T,
                    const float* mu_tab, const float* logvar_tab, float* log_A, float* log_b, void* workspace,
                    size_t workspace_bytes, void* stream) {
  ARG(dims_ok(dims), "dims"); ARG(params_ok(params), "params"); ARG(x && u && mu_tab && logvar_tab && log_A && log_b, "pointers");
  ARG(B > 0 && T > 0, "B,T"); ARG(workspace, "workspace");
  return hmm_inputs(*dims, *params, x, u, B, T, mu_tab, logvar_tab, log_A, log_b, workspace, workspace_bytes, (cudaStream_t)stream);
}

int vhmm_hmm_loglik_fused(const vhmm_dims* dims, const vhmm_params* params, const float* log_pi, const float* x, const float* u,
                          const int64_t* lengths, int64_t B, int64_t T, const float* mu_tab, const float* logvar_tab,
                          float* loglik, void* workspace, size_t workspace_bytes, void* stream) {
  ARG(dims_ok(dims), "dims"); ARG(params_ok(params), "params"); ARG(log_pi && x && u && mu_tab && logvar_tab && loglik, "pointers");
  ARG(B > 0 && T > 0, "B,T"); ARG(workspace, "workspace");
  return hmm_loglik_fused(*dims, *params, log_pi, x, u, lengths, B, T, mu_tab, logvar_tab, loglik, workspace, workspace_bytes,
                          (cudaStream_t)stream);
}

int vhmm_gather_chunks(const float* x_cat, const float* u_cat, int64_t S_tot, int64_t C, int64_t U,
                       const int64_t* col0, const int64_t* lengths, int64_t B, int64_t T, float* x_out,
                       float* u_out, void* stream) {
  ARG(x_cat && u_cat && col0 && lengths && x_out && u_out, "pointers");
  ARG(S_tot > 0 && C > 0 && U > 0 && B >= 0 && T >= 0, "sizes");
  return gather_chunks(x_cat, u_cat, S_tot, C, U, col0, lengths, B, T, x_out, u_out, (cudaStream_t)stream);
}

int vhmm_train_step(const vhmm_dims* dims, const vhmm_params_mut* params, const float* x_cat, const float* u_cat,
                    int64_t S_tot, const int64_t* col0, const int64_t* lengths, int64_t B, int64_t T, double* state,
                    float* exp_avg, float* exp_avg_sq, double beta1, double beta2, double eps, double max_norm,
                    float* x_buf, float* u_buf, float* grads, void* workspace, size_t workspace_bytes, void* stream) {
  ARG(dims_ok(dims), "dims"); ARG(params, "params");
  for (int i = 0; i < VHMM_NUM_PARAMS; ++i) ARG(params->p[i], "params->p[i]");
  ARG(x_cat && u_cat && S_tot > 0, "corpus"); ARG(col0 && lengths, "col0/lengths"); ARG(B > 0 && T > 0, "B,T");
  ARG(state && exp_avg && exp_avg_sq, "state"); ARG(x_buf && u_buf && grads, "buffers"); ARG(workspace, "workspace");
  return train_step(*dims, params->p, x_cat, u_cat, S_tot, col0, lengths, B, T, state, exp_avg, exp_avg_sq, beta1, beta2,
                    eps, max_norm, x_buf, u_buf, grads, workspace, workspace_bytes, (cudaStream_t)stream);
}

int vhmm_adam_step(const vhmm_dims* dims, const vhmm_params_mut* params, const float* grads, float* exp_avg,
                   float* exp_avg_sq, double lr, double beta1, double beta2, double eps, int64_t step,
                   const float* grad_scale, double max_norm, float* grad_norm_out, double* scratch, void* stream) {
  ARG(dims_ok(dims), "dims"); ARG(params && grads && exp_avg && exp_avg_sq, "pointers");
  for (int i = 0; i < VHMM_NUM_PARAMS; ++i) ARG(params->p[i], "params");
  ARG(step >= 1, "step"); ARG(beta1 >= 0 && beta1 < 1 && beta2 >= 0 && beta2 < 1 && eps >= 0, "hyper-parameters");
  return adam_step(*dims, params->p, grads, exp_avg, exp_avg_sq, lr, beta1, beta2, eps, step, grad_scale, max_norm,
                   grad_norm_out, scratch, (cudaStream_t)stream);
}

int vhmm_emission_table(const vhmm_dims* dims, const vhmm_params* params, int64_t T, float* mu_tab,
                        float* logvar_tab, void* workspace, size_t workspace_bytes, void* stream) {
  ARG(dims_ok(dims), "dims"); ARG(params_ok(params), "params"); ARG(T > 0, "T");
  ARG(mu_tab && logvar_tab, "outputs"); ARG(workspace, "workspace");
  return layerwise_emission_table(*dims, *params, T, mu_tab, logvar_tab, workspace, workspace_bytes, (cudaStream_t)stream);
}

int vhmm_emission_logprob(const float* x, const float* mu_tab, const float* logvar_tab, int64_t B,
                          int64_t C, int64_t T, int64_t K, float* log_b, void* stream) {
  ARG(x && mu_tab && logvar_tab && log_b, "pointers"); ARG(B > 0 && C > 0 && T > 0 && K > 0, "sizes");
  return hmm_emission_logprob(x, mu_tab, logvar_tab, B, C, T, K, log_b, (cudaStream_t)stream);
}

#define HMM_ARGS_OK()                                                                   \
  ARG(log_pi && log_A && log_b && lengths, "inputs");                                   \
  ARG(B > 0 && T > 0, "B,T");                                                           \
  VHMM_REQUIRE(K >= 1 && K <= 32, VHMM_E_DIMS, "%s: K=%lld outside [1,32]", __func__, (long long)K)

int vhmm_hmm_forward(const float* log_pi, const float* log_A, const float* log_b,
                     const int64_t* lengths, int64_t B, int64_t T, int64_t K, float* loglik,
                     float* alpha, void* stream) {
  HMM_ARGS_OK(); ARG(loglik, "loglik");
  return hmm_forward(log_pi, log_A, log_b, lengths, B, T, K, loglik, alpha, (cudaStream_t)stream);
}

int vhmm_hmm_posterior(const float* log_pi, const float* log_A, const float* log_b,
                       const int64_t* lengths, int64_t B, int64_t T, int64_t K, float* loglik,
                       float* gamma, float* xi, void* workspace, size_t workspace_bytes, void* stream) {
  HMM_ARGS_OK(); ARG(loglik && gamma, "loglik/gamma"); ARG(workspace, "workspace");
  return hmm_posterior(log_pi, log_A, log_b, lengths, B, T, K, loglik, gamma, xi, workspace, workspace_bytes, (cudaStream_t)stream);
}

int vhmm_viterbi(const float* log_pi, const float* log_A, const float* log_b, const int64_t* lengths,
                 int64_t B, int64_t T, int64_t K, uint8_t* path, float* score, void* workspace,
                 size_t workspace_bytes, void* stream) {
  HMM_ARGS_OK(); ARG(path, "path"); ARG(workspace, "workspace");
  return hmm_viterbi(log_pi, log_A, log_b, lengths, B, T, K, path, score, workspace, workspace_bytes, (cudaStream_t)stream);
}

int vhmm_viterbi_static(const float* log_pi, const float* log_A, const float* log_b, const int64_t* lengths,
                        int64_t B, int64_t T, int64_t K, uint8_t* path, float* score, void* stream) {
  if (B == 0) return hmm_viterbi_static(nullptr, nullptr, nullptr, nullptr, 0, 0, K >= 1 && K <= 32 ? K : 1, nullptr, nullptr,
                                        (cudaStream_t)stream);     // set-up only
  HMM_ARGS_OK(); ARG(path, "path"); ARG(B <= (int64_t)1 << 31 && T <= (int64_t)1 << 31, "B, T");
  return hmm_viterbi_static(log_pi, log_A, log_b, lengths, B, T, K, path, score, (cudaStream_t)stream);
}

int vhmm_hmm_decode(const float* log_pi, const float* log_A, const float* log_b,
                    const int64_t* lengths, int64_t B, int64_t T, int64_t K, uint8_t* path,
                    float* gamma, float* loglik, void* workspace, size_t workspace_bytes, void* stream) {
  HMM_ARGS_OK(); ARG(path && gamma && loglik, "outputs"); ARG(workspace, "workspace");
  return hmm_decode(log_pi, log_A, log_b, lengths, B, T, K, path, gamma, loglik, workspace, workspace_bytes, (cudaStream_t)stream);
}

int vhmm_hmm_scan(const float* log_pi, const float* log_A, const float* log_b, const int64_t* lengths,
                  int64_t B, int64_t T, int64_t K, float* loglik, float* alpha, void* workspace,
                  size_t workspace_bytes, void* stream) {
  HMM_ARGS_OK(); ARG(loglik, "loglik"); ARG(workspace, "workspace");
  return hmm_scan(log_pi, log_A, log_b, lengths, B, T, K, loglik, alpha, workspace, workspace_bytes, (cudaStream_t)stream);
}

}  // extern "C"
