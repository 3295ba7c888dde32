// Pipe-overlap microbenchmark: FFMA2 / FFMA / MUFU throughput alone and mixed.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 2000
template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, float seed) {
  float2 a[8]; float m[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { a[i] = make_float2(seed + i, seed - i); m[i] = seed * (i + 1) * 0.01f; }
  const float2 c = make_float2(0.999f, 1.001f), d = make_float2(1e-3f, -1e-3f);
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (MODE == 0 || MODE == 2 || MODE == 4) {  // 2 packed FMA per chain step
          a[i] = __ffma2_rn(a[i], c, d);
          a[i] = __ffma2_rn(a[i], c, d);
        }
        if (MODE == 3) {  // scalar FFMA x4 (same flops as 2 packed)
          a[i].x = fmaf(a[i].x, 0.999f, 1e-3f); a[i].y = fmaf(a[i].y, 1.001f, -1e-3f);
          a[i].x = fmaf(a[i].x, 0.999f, 1e-3f); a[i].y = fmaf(a[i].y, 1.001f, -1e-3f);
        }
        if (MODE == 5) {  // packed with register coefficients + scalar mix
          a[i] = __ffma2_rn(a[i], c, d);
          a[i].x = fmaf(a[i].x, c.x, d.x); a[i].y = fmaf(a[i].y, c.y, d.y);
        }
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if ((MODE == 1 || MODE == 2) && (r & 1) == 0) {  // 1 MUFU per 4 packed FMA
          float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(m[i])); m[i] = y * 0.5f;
        }
        if (MODE == 4 && (r & 1) == 0) {  // MUFU dependent on FMA chain (like ours)
          float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(a[i].x)); a[i].y = y;
        }
      }
    }
  }
  float s = 0; 
#pragma unroll
  for (int i = 0; i < 8; ++i) s += a[i].x + a[i].y + m[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE> void run(const char* name, int ctas_per_sm) {
  float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<148 * ctas_per_sm, 256>>>(out, 1.0f);
  cudaEventRecord(e0);
  k<MODE><<<148 * ctas_per_sm, 256>>>(out, 1.0f);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  // per SMSP: warps = ctas_per_sm*8/4 ; per warp: ITERS*4*8 chain steps
  double steps = (double)ITERS * 4 * 8 * (ctas_per_sm * 2);
  double cyc = ms * 1e-3 * 1.92e9;
  printf("%-44s ctas/sm=%d  %.3f ms  %.2f cycles per (chain-step, SMSP)\n", name, ctas_per_sm, ms, cyc / steps);
  cudaFree(out);
}
int main() {
  for (int c : {2, 4}) {
    run<0>("2xFFMA2 per step", c);
    run<3>("4xFFMA scalar-imm per step", c);
    run<5>("1xFFMA2 + 2xFFMA reg per step", c);
    run<1>("0.5 MUFU per step (MUFU only)", c);
    run<2>("2xFFMA2 + 0.5 MUFU independent", c);
    run<4>("2xFFMA2 + 0.5 MUFU dependent", c);
  }
  return 0;
}
