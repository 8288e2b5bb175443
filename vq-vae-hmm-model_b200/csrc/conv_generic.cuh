// conv_generic.cuh -- layer-wise FP32 kernels with run-time dimensions: the "any VAE_HMM
// dims" path (smoke config, scaled config) and the building blocks of the fused kernels.
//
// Every layer of the reference model is a 1-D cross-correlation over (B,C,T) tensors with
// kernel width 3 (pad 1) or 1 (VQ_VAE_HMM_fixed.py:34-36, :53-57 on the (B,U,T) layout, :77-79,
// :83), so three kernels cover forward, input-gradient and weight-gradient of all of them:
//   conv1d_kernel   y = epi(bias + W * x)             (also dgrad, via weight strides)
//   wgrad_kernel    dW = sum_{b,t} dy (x) x, db = sum dy   (split over the batch, partials)
//   reduce_partials grads += sum_split partials        (fixed order -> deterministic)
#pragma once
#include "common.cuh"

namespace vhmm {

enum { EPI_NONE = 0, EPI_RELU = 1, EPI_MASK = 2 };

// Weight addressing: element (o, c, tau) lives at W[w_base + o*w_so + c*w_sc + tau*w_st].
// Forward conv of a (Cout,Cin,KW) tensor: {0, Cin*KW, KW, 1}.  dgrad of the same layer swaps
// the roles and flips the taps: {KW-1, KW, Cin*KW, -1}.
struct WView {
  const float* w;
  int64_t base, so, sc, st;
};

static inline WView wview_fwd(const float* w, int cin, int kw) {
  return WView{w, 0, (int64_t)cin * kw, kw, 1};
}
// dgrad: new "out" channel = original in channel, new "in" channel = original out channel
static inline WView wview_dgrad(const float* w, int cin_orig, int kw) {
  return WView{w, kw - 1, kw, (int64_t)cin_orig * kw, -1};
}

constexpr int CONV_TT = 128;   // time steps per CTA
constexpr int CONV_CC = 16;    // in-channel chunk staged in shared memory
constexpr int CONV_XP = 132;   // row pitch of the staged input tile (>= CONV_TT + 2, 16B multiple)

// grid (ceil(T/128), ceil(Cout/(8*TO)), B), block 256.
// thread (ty = tid/32, tx = tid%32) owns out channels [o0 + ty*TO, +TO) x steps [t0 + 4tx, +4).
template <int KW, int TO>
__global__ void __launch_bounds__(256)
conv1d_kernel(const float* __restrict__ x, WView wv, const float* __restrict__ bias,
              const float* __restrict__ act, float* __restrict__ y, int Cin, int Cout, int T,
              int epi) {
  constexpr int PAD = KW / 2;
  constexpr int OT = 8 * TO;
  __shared__ __align__(16) float xs[CONV_CC][CONV_XP];
  __shared__ __align__(16) float ws[CONV_CC][KW][OT];
  const int tid = threadIdx.x, tx = tid & 31, ty = tid >> 5;
  const int t0 = blockIdx.x * CONV_TT, o0 = blockIdx.y * OT;
  const int64_t b = blockIdx.z;
  const float* xb = x + b * (int64_t)Cin * T;

  float acc[TO][4];
#pragma unroll
  for (int i = 0; i < TO; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int c0 = 0; c0 < Cin; c0 += CONV_CC) {
    const int cc = min(CONV_CC, Cin - c0);
    __syncthreads();
    // stage x[c0..c0+cc)[t0-PAD .. t0+TT+PAD) ; index 0 <-> t0-PAD ; zero outside [0,T)
    for (int idx = tid; idx < cc * (CONV_TT + 2 * PAD); idx += 256) {
      int c = idx / (CONV_TT + 2 * PAD), j = idx - c * (CONV_TT + 2 * PAD);
      int t = t0 - PAD + j;
      xs[c][j] = (t >= 0 && t < T) ? xb[(int64_t)(c0 + c) * T + t] : 0.f;
    }
    for (int idx = tid; idx < cc * KW * OT; idx += 256) {
      int o = idx % OT, r = idx / OT, tau = r % KW, c = r / KW;
      float v = 0.f;
      if (o0 + o < Cout) v = wv.w[wv.base + (int64_t)(o0 + o) * wv.so + (int64_t)(c0 + c) * wv.sc + tau * wv.st];
      ws[c][tau][o] = v;
    }
    __syncthreads();
    for (int c = 0; c < cc; ++c) {
      float xv[4 + 2 * PAD];
      {
        float4 a = *reinterpret_cast<const float4*>(&xs[c][tx * 4]);
        xv[0] = a.x; xv[1] = a.y; xv[2] = a.z; xv[3] = a.w;
        if (PAD) {
          float2 e = *reinterpret_cast<const float2*>(&xs[c][tx * 4 + 4]);
          xv[4] = e.x; xv[4 + 2 * PAD - 1] = e.y;
        }
      }
#pragma unroll
      for (int tau = 0; tau < KW; ++tau) {
        float wr[TO];
        if (TO == 8) {
          float4 a = *reinterpret_cast<const float4*>(&ws[c][tau][ty * TO]);
          float4 e = *reinterpret_cast<const float4*>(&ws[c][tau][ty * TO + 4]);
          wr[0] = a.x; wr[1] = a.y; wr[2] = a.z; wr[3] = a.w;
          wr[TO - 4] = e.x; wr[TO - 3] = e.y; wr[TO - 2] = e.z; wr[TO - 1] = e.w;
        } else {
#pragma unroll
          for (int i = 0; i < TO; ++i) wr[i] = ws[c][tau][ty * TO + i];
        }
#pragma unroll
        for (int i = 0; i < TO; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(wr[i], xv[j + tau], acc[i][j]);
      }
    }
  }
  // epilogue
  const int tb = t0 + tx * 4;
  if (tb >= T) return;
  const bool vec = ((T & 3) == 0);
#pragma unroll
  for (int i = 0; i < TO; ++i) {
    const int o = o0 + ty * TO + i;
    if (o >= Cout) continue;
    const float bv = bias ? bias[o] : 0.f;
    const int64_t off = (b * Cout + o) * (int64_t)T + tb;
    float r[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) r[j] = acc[i][j] + bv;
    if (epi == EPI_RELU) {
#pragma unroll
      for (int j = 0; j < 4; ++j) r[j] = fmaxf(r[j], 0.f);
    }
    if (vec) {
      if (epi == EPI_MASK) {
        float4 m = *reinterpret_cast<const float4*>(act + off);
        r[0] = m.x > 0.f ? r[0] : 0.f; r[1] = m.y > 0.f ? r[1] : 0.f;
        r[2] = m.z > 0.f ? r[2] : 0.f; r[3] = m.w > 0.f ? r[3] : 0.f;
      }
      *reinterpret_cast<float4*>(y + off) = make_float4(r[0], r[1], r[2], r[3]);
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (tb + j < T) {
          float v = r[j];
          if (epi == EPI_MASK) v = act[off + j] > 0.f ? v : 0.f;
          y[off + j] = v;
        }
      }
    }
  }
}

// y (B,Cout,T) = epi(bias + conv(x (B,Cin,T), W))
static inline int launch_conv(const float* x, WView wv, const float* bias, const float* act, float* y,
                              int64_t B, int Cin, int Cout, int64_t T, int kw, int epi,
                              cudaStream_t st) {
  if (B <= 0 || T <= 0) return 0;
  const bool small = Cout <= 16;
  const int OT = small ? 16 : 64;
  // gridDim.z is limited to 65535: split B over several launches if needed
  for (int64_t b0 = 0; b0 < B; b0 += 65535) {
    int64_t nb = (B - b0 < 65535) ? (B - b0) : 65535;
    dim3 grid((unsigned)cdiv(T, CONV_TT), (unsigned)cdiv(Cout, OT), (unsigned)nb);
    const float* xp = x + b0 * (int64_t)Cin * T;
    const float* ap = act ? act + b0 * (int64_t)Cout * T : nullptr;
    float* yp = y + b0 * (int64_t)Cout * T;
    if (kw == 3) {
      if (small) conv1d_kernel<3, 2><<<grid, 256, 0, st>>>(xp, wv, bias, ap, yp, Cin, Cout, (int)T, epi);
      else conv1d_kernel<3, 8><<<grid, 256, 0, st>>>(xp, wv, bias, ap, yp, Cin, Cout, (int)T, epi);
    } else {
      if (small) conv1d_kernel<1, 2><<<grid, 256, 0, st>>>(xp, wv, bias, ap, yp, Cin, Cout, (int)T, epi);
      else conv1d_kernel<1, 8><<<grid, 256, 0, st>>>(xp, wv, bias, ap, yp, Cin, Cout, (int)T, epi);
    }
    VHMM_LAUNCH_CHECK();
  }
  return 0;
}

// ---------------------------------------------------------------------------------------
// weight gradient: part[split][...] holds, in the PARAMETER's own layout (via WView strides),
//   dW[o,c,tau] = sum over this split's (b,t) of dy[b,o,t] * x[b,c,t+tau-PAD]
// and (when dbias_part != null, from the blockIdx.y == 0 CTAs) db[o] = sum dy[b,o,t].
// grid (ceil(Cout/64), ceil(Cin/16), nsplit), block 256: c_l = tid%16, og = tid/16 owns
// out channels og*4..og*4+3.
// ---------------------------------------------------------------------------------------
constexpr int WG_OT = 64, WG_CT = 16, WG_TT = 128;

template <int KW>
__global__ void __launch_bounds__(256)
wgrad_kernel(const float* __restrict__ dy, const float* __restrict__ x, float* __restrict__ part,
             float* __restrict__ dbias_part, int64_t part_stride, int64_t dbias_stride,
             int64_t w_base, int64_t w_so, int64_t w_sc, int64_t w_st, int Cin, int Cout, int T,
             int64_t B) {
  constexpr int PAD = KW / 2;
  __shared__ __align__(16) float dys[WG_OT][WG_TT];
  __shared__ __align__(16) float xs[WG_CT][CONV_XP];
  const int tid = threadIdx.x, cl = tid & 15, og = tid >> 4;
  const int o0 = blockIdx.x * WG_OT, c0 = blockIdx.y * WG_CT;
  // the work units are (sequence, 128-step slab) pairs, split evenly over gridDim.z: a batch of FEW LONG sequences fills the
  // machine like many short ones (split over the batch alone, B = 4, T = 16384 ran 4 CTAs per tile for 2.5 ms a layer)
  const int nsplit = gridDim.z, split = blockIdx.z;
  const int64_t nslab = (T + WG_TT - 1) / WG_TT, units = B * nslab;
  const int64_t per = (units + nsplit - 1) / nsplit;
  const int64_t u_lo = split * per, u_hi = (u_lo + per < units) ? u_lo + per : units;

  float acc[4][KW];
  float dbs[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    dbs[i] = 0.f;
#pragma unroll
    for (int k = 0; k < KW; ++k) acc[i][k] = 0.f;
  }
  for (int64_t un = u_lo; un < u_hi; ++un) {
    const int64_t b = un / nslab;
    const float* dyb = dy + b * (int64_t)Cout * T;
    const float* xb = x + b * (int64_t)Cin * T;
    {
      const int t0 = (int)(un - b * nslab) * WG_TT;
      __syncthreads();
      for (int idx = tid; idx < WG_OT * WG_TT; idx += 256) {
        int o = idx / WG_TT, j = idx - o * WG_TT;
        int t = t0 + j;
        dys[o][j] = (o0 + o < Cout && t < T) ? dyb[(int64_t)(o0 + o) * T + t] : 0.f;
      }
      for (int idx = tid; idx < WG_CT * (WG_TT + 2 * PAD); idx += 256) {
        int c = idx / (WG_TT + 2 * PAD), j = idx - c * (WG_TT + 2 * PAD);
        int t = t0 - PAD + j;
        xs[c][j] = (c0 + c < Cin && t >= 0 && t < T) ? xb[(int64_t)(c0 + c) * T + t] : 0.f;
      }
      __syncthreads();
#pragma unroll 4
      for (int j = 0; j < WG_TT; j += 4) {
        float xv[4 + 2 * PAD];
        float4 a = *reinterpret_cast<const float4*>(&xs[cl][j]);
        xv[0] = a.x; xv[1] = a.y; xv[2] = a.z; xv[3] = a.w;
        if (PAD) {
          float2 e = *reinterpret_cast<const float2*>(&xs[cl][j + 4]);
          xv[4] = e.x; xv[4 + 2 * PAD - 1] = e.y;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float4 d = *reinterpret_cast<const float4*>(&dys[og * 4 + i][j]);
          dbs[i] += (d.x + d.y) + (d.z + d.w);
#pragma unroll
          for (int k = 0; k < KW; ++k) {
            acc[i][k] = fmaf(d.x, xv[0 + k], acc[i][k]);
            acc[i][k] = fmaf(d.y, xv[1 + k], acc[i][k]);
            acc[i][k] = fmaf(d.z, xv[2 + k], acc[i][k]);
            acc[i][k] = fmaf(d.w, xv[3 + k], acc[i][k]);
          }
        }
      }
    }
  }
  const int c = c0 + cl;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int o = o0 + og * 4 + i;
    if (o >= Cout) continue;
    if (c < Cin) {
#pragma unroll
      for (int k = 0; k < KW; ++k)
        part[split * part_stride + w_base + (int64_t)o * w_so + (int64_t)c * w_sc + k * w_st] = acc[i][k];
    }
    if (dbias_part && blockIdx.y == 0 && cl == 0) dbias_part[split * dbias_stride + o] = dbs[i];
  }
}

// out[i] += sum_{s<nsplit} part[s*stride + i]   (double accumulation, fixed order)
__global__ void reduce_partials_kernel(const float* __restrict__ part, int64_t stride, int nsplit,
                                       int64_t n, float* __restrict__ out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  double s = 0.0;
  for (int k = 0; k < nsplit; ++k) s += (double)part[k * stride + i];
  out[i] = (float)((double)out[i] + s);
}

static inline int wgrad_nsplit(int Cin, int Cout, int64_t units) {   // units = B * ceil(T / WG_TT)
  int64_t tiles = cdiv(Cout, WG_OT) * cdiv(Cin, WG_CT);
  int64_t ns = cdiv(592, tiles);
  if (ns > units) ns = units;
  if (ns < 1) ns = 1;
  return (int)ns;
}

// floats of scratch needed by launch_wgrad for one layer (B: an upper bound of the work units B * ceil(T / 128))
static inline int64_t wgrad_scratch_floats(int Cin, int Cout, int kw, int64_t B) {
  int ns = wgrad_nsplit(Cin, Cout, B);
  return (int64_t)ns * ((int64_t)Cin * Cout * kw + Cout);
}

// gW (layer layout, += ), gB (+=, may be null).  dy (B,Cout,T), x (B,Cin,T).
static inline int launch_wgrad(const float* dy, const float* x, float* gW, float* gB, WView layout,
                               int64_t B, int Cin, int Cout, int64_t T, int kw, float* scratch,
                               cudaStream_t st) {
  if (B <= 0 || T <= 0) return 0;
  const int ns = wgrad_nsplit(Cin, Cout, B * cdiv(T, (int64_t)WG_TT));
  const int64_t nW = (int64_t)Cin * Cout * kw;
  float* partW = scratch;
  float* partB = gB ? scratch + (int64_t)ns * nW : nullptr;
  dim3 grid((unsigned)cdiv(Cout, WG_OT), (unsigned)cdiv(Cin, WG_CT), (unsigned)ns);
  if (kw == 3)
    wgrad_kernel<3><<<grid, 256, 0, st>>>(dy, x, partW, partB, nW, Cout, layout.base, layout.so,
                                          layout.sc, layout.st, Cin, Cout, (int)T, B);
  else
    wgrad_kernel<1><<<grid, 256, 0, st>>>(dy, x, partW, partB, nW, Cout, layout.base, layout.so,
                                          layout.sc, layout.st, Cin, Cout, (int)T, B);
  VHMM_LAUNCH_CHECK();
  reduce_partials_kernel<<<(unsigned)cdiv(nW, 256), 256, 0, st>>>(partW, nW, ns, nW, gW);
  VHMM_LAUNCH_CHECK();
  if (gB) {
    reduce_partials_kernel<<<(unsigned)cdiv(Cout, 256), 256, 0, st>>>(partB, Cout, ns, Cout, gB);
    VHMM_LAUNCH_CHECK();
  }
  return 0;
}

}  // namespace vhmm
