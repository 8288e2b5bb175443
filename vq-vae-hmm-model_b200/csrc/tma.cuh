// tma.cuh -- tensor-map TMA loads (cp.async.bulk.tensor, SASS UTMALDG) for the input windows of the tile kernels.
// A window of the tile kernels is "rows r of sequence b, columns [t0-4, t0+128) of a (B, rows, T) fp32 tensor, zeros
// outside [0,T)" -- exactly a TMA box with out-of-bounds fill: one instruction by one thread replaces the per-thread
// cp.async address arithmetic and the zero-fill predicates of prefetch_rows.  Needs T % 4 == 0 (global strides are
// multiples of 16 bytes) and a 16-byte aligned base; the kernels keep the cp.async path for everything else.
// The driver entry point cuTensorMapEncodeTiled is fetched through the runtime (no link against libcuda).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace vhmm {
namespace tma {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static inline EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

// map of a (B, rows, T) fp32 tensor with box {box_w columns, box_rows rows, 1 sequence}; false if TMA cannot take it
static inline bool make_map(CUtensorMap* out, const float* base, int64_t B, int rows, int64_t T, int box_w, int box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (!fn || (T & 3) || (((uintptr_t)base) & 15) || box_w > 256 || box_rows > 256) return false;
  cuuint64_t dims[3] = {(cuuint64_t)T, (cuuint64_t)rows, (cuuint64_t)B};
  cuuint64_t strides[2] = {(cuuint64_t)T * 4, (cuuint64_t)T * 4 * (cuuint64_t)rows};
  cuuint32_t box[3] = {(cuuint32_t)box_w, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  return fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void*)base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

#ifdef __CUDACC__
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
// box at (column c0, row c1, sequence c2) -> shared memory (128-byte aligned); completion as transaction bytes on `bar`
__device__ __forceinline__ void load_3d(void* smem, const CUtensorMap* tm, int c0, int c1, int c2, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
      :: "r"(s32(smem)), "l"(tm), "r"(c0), "r"(c1), "r"(c2), "r"(s32(bar)) : "memory");
}
__device__ __forceinline__ void expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(s32(bar)), "r"(bytes) : "memory");
}
#endif

}  // namespace tma
}  // namespace vhmm
