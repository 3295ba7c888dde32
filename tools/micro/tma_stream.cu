// How fast can one persistent CTA per SM stream HBM through a TMA ring?
// sm_100a.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
//   -I csrc -o tools/micro/tma_stream tools/micro/tma_stream.cu -lcuda
//
// usage: tma_stream MODE ROWS STAGES CTAS_PER_SM [HOLD_NS] [GB] [HAMMER_WARPS] [DUTY]
//   MODE 0: 2-D tensor map, box = [64 bf16 = 128 B][ROWS], SWIZZLE_128B
//   MODE 1: 1-D bulk copy (cp.async.bulk) of ROWS*128 contiguous bytes
//   HOLD_NS: the consumer keeps every stage that long after it landed (models
//            the compute + UMMA phases of the real kernel) before freeing it
// Prints achieved GB/s.  The consumer does nothing with the data.
#include <cstdio>
#include <cstdlib>
#include <cuda.h>
#include <cuda_runtime.h>
#include "../../csrc/mgb_sm100.cuh"

using namespace sm100;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ void bulk_load_1d(void* dst, const void* src,
                                             uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes "
      "[%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes),
      "r"(smem_u32(bar)) : "memory");
}

__global__ void __launch_bounds__(64 + 32 * 8, 1)
stream_kernel(const __grid_constant__ CUtensorMap map, const char* base,
              long long ntiles, int mode, int rows, int stages, int hold_ns,
              unsigned long long* sink, int hammer_duty) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t full[16], empty[16];
  const int tid = threadIdx.x;
  const uint32_t bytes = (uint32_t)rows * 128u;
  if (tid == 0) {
    for (int s = 0; s < stages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    *reinterpret_cast<volatile int*>(&full[15]) = 0;
    mbar_fence_init();
  }
  __syncthreads();
  const long long G = gridDim.x;
  const long long mine = ntiles > blockIdx.x ? (ntiles - blockIdx.x + G - 1) / G : 0;
  if (tid == 0) {                       // producer
    for (long long it = 0; it < mine; ++it) {
      const int s = (int)(it % stages);
      const uint32_t ph = (uint32_t)((it / stages) & 1);
      mbar_wait(&empty[s], ph ^ 1u);
      const long long tile = blockIdx.x + it * G;
      mbar_arrive_expect_tx(&full[s], bytes);
      if (mode == 0)
        tma_load_2d(smem + (size_t)s * bytes, &map, &full[s], 0, (int)(tile * rows));
      else
        bulk_load_1d(smem + (size_t)s * bytes, base + tile * (long long)bytes, bytes, &full[s]);
    }
  } else if (tid == 32) {               // consumer
    unsigned long long acc = 0;
    for (long long it = 0; it < mine; ++it) {
      const int s = (int)(it % stages);
      const uint32_t ph = (uint32_t)((it / stages) & 1);
      mbar_wait(&full[s], ph);
      if (hold_ns > 0) {
        unsigned long long t0, t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        do { asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1)); } while (t1 - t0 < (unsigned)hold_ns);
      }
      acc += *reinterpret_cast<volatile unsigned long long*>(smem + (size_t)s * bytes);
      mbar_arrive(&empty[s]);
    }
    if (acc == 0x123456789ull) *sink = acc;
    *reinterpret_cast<volatile int*>(&full[15]) = 1;   // stop the hammers
  } else if (tid >= 64) {
    // "hammer" warps: conflict-free 16-byte shared loads (512 B per warp
    // instruction) for hammer_duty out of every 16 iterations: models the
    // UMMA operand fetches / G stores that share the shared-memory port
    // with the TMA fills.
    uint4 acc = make_uint4(0, 0, 0, 0);
    const uint32_t a0 = smem_u32(smem) + (tid - 64) * 16;
    const uint32_t span = (uint32_t)stages * bytes;
    uint32_t off = 0;
    int it = 0;
    while (*reinterpret_cast<volatile int*>(&full[15]) == 0) {
      if ((it++ & 15) < hammer_duty) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          uint4 v;
          asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                       : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                       : "r"(a0 + off));
          acc.x ^= v.x; acc.y ^= v.y; acc.z ^= v.z; acc.w ^= v.w;
          off += 4096; if (off >= span) off = 0;
        }
      } else {
        __nanosleep(0);
      }
    }
    if ((acc.x ^ acc.y ^ acc.z ^ acc.w) == 0x12345u) *sink = acc.x;
  }
}

int main(int argc, char** argv) {
  const int mode = argc > 1 ? atoi(argv[1]) : 0;
  const int rows = argc > 2 ? atoi(argv[2]) : 128;
  const int stages = argc > 3 ? atoi(argv[3]) : 8;
  const int cps = argc > 4 ? atoi(argv[4]) : 1;
  const int hold = argc > 5 ? atoi(argv[5]) : 0;
  const double gb = argc > 6 ? atof(argv[6]) : 2.56;
  const int hwarps = argc > 7 ? atoi(argv[7]) : 0;   // hammer warps (0..8)
  const int hduty = argc > 8 ? atoi(argv[8]) : 16;   // busy iterations of 16
  const size_t bytes = (size_t)(gb * 1e9) / (rows * 128) * (rows * 128);
  const long long ntiles = bytes / (rows * 128);
  char* d; CK(cudaMalloc(&d, bytes)); CK(cudaMemset(d, 1, bytes));
  unsigned long long* sink; CK(cudaMalloc(&sink, 8));
  CUtensorMap map;
  {
    typedef CUresult (*enc_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                              const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                              const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                              CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr; cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
    cuuint64_t gdim[2] = {64, (cuuint64_t)(bytes / 128)};
    cuuint64_t gstr[1] = {128};
    cuuint32_t box[2] = {64, (cuuint32_t)(rows > 256 ? 256 : rows)};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = ((enc_t)fn)(&map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, d, gdim, gstr, box,
                             estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                             CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
  }
  const size_t smem = (size_t)stages * rows * 128 + 1024;
  CK(cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int nsm = 0; CK(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, 0));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  for (int rep = 0; rep < 3; ++rep) {
    CK(cudaEventRecord(e0));
    stream_kernel<<<nsm * cps, 64 + 32 * hwarps, smem>>>(map, d, ntiles, mode, rows, stages,
                                                         hold, sink, hduty);
    CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    if (rep == 2)
      printf("mode %d rows %4d (%3d KiB/stage) stages %2d ctas/sm %d hold %5d ns hammer %d warps duty %2d/16: %.3f ms  %.0f GB/s\n",
             mode, rows, rows * 128 / 1024, stages, cps, hold, hwarps, hduty, ms, bytes / ms * 1e-6);
  }
  CK(cudaGetLastError());
  return 0;
}
