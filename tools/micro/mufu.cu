// MUFU throughput by opcode (ex2 / rcp / rsqrt / lg2 / sin) on sm_100a.
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 4000
template <int OP>
__global__ void __launch_bounds__(256) k(float* out, float seed) {
  float m[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) m[i] = seed + 0.01f * (i + 1) + threadIdx.x * 1e-6f;
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float y;
      if (OP == 0) asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(m[i]));
      if (OP == 1) asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(m[i]));
      if (OP == 2) asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(m[i]));
      if (OP == 3) asm volatile("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(m[i]));
      if (OP == 4) asm volatile("sin.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(m[i]));
      if (OP == 5) asm volatile("ex2.approx.f32 %0, %1;" : "=f"(y) : "f"(m[i]));
      if (OP == 6) asm volatile("rcp.approx.f32 %0, %1;" : "=f"(y) : "f"(m[i]));
      m[i] = y;
    }
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += m[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int OP> void run(const char* name) {
  float* out; cudaMalloc(&out, 148 * 4 * 256 * 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<OP><<<148 * 4, 256>>>(out, 1.0f);
  cudaEventRecord(e0);
  k<OP><<<148 * 4, 256>>>(out, 1.0f);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  double instr_per_smsp = (double)ITERS * 8 * (4 * 8 / 4);
  printf("%-22s %.3f ms  %.2f cycles per warp-instr per SMSP (at 1.92 GHz)\n", name, ms, ms * 1e-3 * 1.92e9 / instr_per_smsp);
  cudaFree(out);
}
int main() {
  run<0>("ex2.approx.ftz"); run<1>("rcp.approx.ftz"); run<2>("rsqrt.approx.ftz");
  run<3>("lg2.approx.ftz"); run<4>("sin.approx.ftz"); run<5>("ex2.approx"); run<6>("rcp.approx");
  return 0;
}
