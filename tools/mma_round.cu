// mma_round.cu -- rounding behaviour of the mma.sync TF32 accumulate on this GPU:
// d = c + a*b for c = 1 and a*b a fraction of ulp(1) = 2^-23, both signs; and chains.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
// A = diag-ish: only k=0 column nonzero (a0 for t4==0), B row k=0 = 1.  d[g][n] = c + A[g][0]*B[0][n]
__global__ void k(float* out, float cval, float aval, int reps) {
  int lane = threadIdx.x, t4 = lane & 3;
  uint32_t a[4] = {0, 0, 0, 0}, b[2] = {0, 0};
  if (t4 == 0) { a[0] = __float_as_uint(aval); a[1] = __float_as_uint(aval); b[0] = __float_as_uint(1.0f); }
  float d[4] = {cval, cval, cval, cval};
  for (int i = 0; i < reps; ++i) mma_tf32(d, a, b);
  out[lane] = d[0];
}
int main() {
  float* o; cudaMalloc(&o, 128);
  float h[32];
  const float ulp = 1.1920929e-07f;  // 2^-23
  float fr[] = {0.25f, 0.5f, 0.75f, 1.0f, 1.5f};
  for (float sgn : {1.f, -1.f})
    for (float f : fr) {
      k<<<1, 32>>>(o, 1.0f, sgn * f * ulp, 1); cudaMemcpy(h, o, 128, cudaMemcpyDeviceToHost);
      printf("c=1 + (%+.2f ulp): result-1 = %+.3f ulp\n", sgn * f, (h[0] - 1.0) / ulp);
    }
  for (float sgn : {1.f, -1.f}) {
    k<<<1, 32>>>(o, 1.0f, sgn * 0.75f * ulp, 1000); cudaMemcpy(h, o, 128, cudaMemcpyDeviceToHost);
    printf("chain of 1000 x (%+.2f ulp): result-1 = %+.1f ulp (exact %+.1f)\n", sgn * 0.75, (h[0] - 1.0) / ulp, sgn * 750.0);
  }
  // product-sum inside one mma: 8 products each 0.3 ulp (needs all k): use a full row
  return 0;
}
