// common.cuh -- shared host/device helpers for libvhmm (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/vhmm.h"

namespace vhmm {

void set_error(const char* fmt, ...);
void count_launch();  // process-wide kernel launch counter (vhmm_launch_count)

// Per-DEVICE cached state (api.cu).  Function attributes (cudaFuncSetAttribute) and SM counts belong to a
// device/context, not to the process: a process that drives several GPUs (model.to('cuda:1'), one thread per
// GPU) must see each device's own values.  Slots are indexed by cudaGetDevice(); entries are written with the
// same idempotent values by racing first callers.
constexpr int VHMM_MAX_DEVICES = 64;
int current_device();                 // cudaGetDevice(), 0 on failure; clamped to [0, VHMM_MAX_DEVICES)
int device_sm_count();                // multiprocessors of the CURRENT device (cached per device)
// once-per-device flag: returns true exactly when `slot` has not been marked for the current device yet
struct PerDeviceOnce {
  unsigned char done[VHMM_MAX_DEVICES] = {};
  bool pending() const { return !done[current_device()]; }
  void mark() { done[current_device()] = 1; }
};

#define VHMM_CUDA_CHECK(expr)                                                            \
  do {                                                                                   \
    cudaError_t e__ = (expr);                                                            \
    if (e__ != cudaSuccess) {                                                            \
      ::vhmm::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr,                    \
                        cudaGetErrorString(e__));                                        \
      return (int)e__;                                                                   \
    }                                                                                    \
  } while (0)

#define VHMM_LAUNCH_CHECK()                  \
  do {                                       \
    ::vhmm::count_launch();                  \
    VHMM_CUDA_CHECK(cudaPeekAtLastError());  \
  } while (0)

#define VHMM_REQUIRE(cond, code, ...)                                                    \
  do {                                                                                   \
    if (!(cond)) {                                                                       \
      ::vhmm::set_error(__VA_ARGS__);                                                    \
      return (code);                                                                     \
    }                                                                                    \
  } while (0)

static inline int64_t cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline size_t align_up(size_t a, size_t b) { return (a + b - 1) / b * b; }

// parameter indices (include/vhmm.h)
enum {
  P_E1W = 0, P_E1B, P_E2W, P_E2B, P_E3W, P_E3B, P_LOGPRIOR, P_T0W, P_T0B, P_T2W, P_T2B,
  P_EMB, P_D1W, P_D1B, P_D2W, P_D2B, P_DPW, P_DPB
};

static inline void param_sizes(const vhmm_dims& d, int64_t* n) {
  n[P_E1W] = (int64_t)d.H * d.C * 3;   n[P_E1B] = d.H;
  n[P_E2W] = (int64_t)d.H2 * d.H * 3;  n[P_E2B] = d.H2;
  n[P_E3W] = (int64_t)d.K * d.H2;      n[P_E3B] = d.K;
  n[P_LOGPRIOR] = d.K;
  n[P_T0W] = (int64_t)d.TH * d.U;      n[P_T0B] = d.TH;
  n[P_T2W] = (int64_t)d.K * d.K * d.TH; n[P_T2B] = (int64_t)d.K * d.K;
  n[P_EMB] = (int64_t)d.K * d.H;
  n[P_D1W] = (int64_t)d.H * d.H * 3;   n[P_D1B] = d.H;
  n[P_D2W] = (int64_t)d.H * d.H * 3;   n[P_D2B] = d.H;
  n[P_DPW] = (int64_t)2 * d.C * d.H;   n[P_DPB] = 2 * d.C;
}

static inline void param_offsets(const vhmm_dims& d, int64_t* off /*19*/) {
  int64_t n[VHMM_NUM_PARAMS];
  param_sizes(d, n);
  off[0] = 0;
  for (int i = 0; i < VHMM_NUM_PARAMS; ++i) off[i + 1] = off[i] + n[i];
}

static inline bool dims_ok(const vhmm_dims* d) {
  return d && d->C > 0 && d->H > 0 && d->K > 0 && d->H2 > 0 && d->U > 0 && d->TH > 0 && d->K <= 255;
}

// bump allocator over the caller's workspace (256-byte aligned slices)
struct Arena {
  char* base;
  size_t cap, used;
  bool ok;
  Arena(void* p, size_t n) : base((char*)p), cap(n), used(0), ok(true) {}
  template <typename T>
  T* take(size_t count) {
    size_t bytes = align_up(count * sizeof(T), 256);
    if (used + bytes > cap) { ok = false; used += bytes; return nullptr; }
    T* r = (T*)(base + used);
    used += bytes;
    return r;
  }
};

// ---- device helpers -----------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// deterministic block sum (blockDim.x multiple of 32, <= 1024); result valid in thread 0
template <typename T>
__device__ __forceinline__ T block_sum(T v, T* sh /*>=32*/) {
  v = warp_sum(v);
  int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) sh[w] = v;
  __syncthreads();
  int nw = (blockDim.x + 31) >> 5;
  T r = (threadIdx.x < nw) ? sh[threadIdx.x] : T(0);
  if (w == 0) r = warp_sum(r);
  return r;
}
#endif

}  // namespace vhmm
