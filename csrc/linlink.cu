// Linear-link soft-binned statistic on the 5th-generation tensor cores.
//
// Model:  mu_h = x_h + sum_j W[h,j] theta_j   (W: per-object features, bf16)
//         sigma = link(p_sigma)               (global)
// so that  d s_k / d theta_j = c_k * sum_h (g_{k+1,h} - g_{k,h}) W[h,j]
// is a tall-skinny GEMM with the *object axis as the MMA K dimension*:
//
//   MMA #1 (forward)   D1[128 objects, 16]  = Wtile[128 x P]  . Theta[P x 16]
//   MMA #2 (Jacobian)  D2[32 rows, P]      += G[32 x 128 obj] . Wtile[128 x P]
//
// The same TMA-loaded, 128B-swizzled W tile is operand A (K-major) of #1 and
// operand B (MN-major) of #2; theta is split into three bf16 planes and g
// into two, so both products carry ~fp32 / ~2^-17 accuracy with bf16 inputs.
// Accumulators live in TMEM; CUDA cores only evaluate erfc / exp per
// (object, edge) and convert g to bf16.  W streams through HBM exactly once
// per step (the reference's forward + VJP would stream it twice).
//
// Reference counterpart: jax.vjp of a user sumstats function with a
// per-object linear predictor (OnePointModel._vjp, multigrad.py:508-538).
#include <cuda.h>
#include <cuda_bf16.h>
#include <math.h>
#include <string.h>
#include "mgb_common.cuh"
#include "mgb_sm100.cuh"

using namespace sm100;

namespace {

constexpr int LL_TILE = 128;     // objects per tile (UMMA M of #1, K of #2)
constexpr int LL_FB = 64;        // features per 128-byte swizzle block
constexpr int LL_GROWS = 32;     // rows of G: 16 x g_hi, 16 x g_lo

// cuTensorMapEncodeTiled through the runtime (no link-time libcuda symbol)
typedef CUresult (*encode_fn_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t,
                                void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*,
                                CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion,
                                CUtensorMapFloatOOBfill);
encode_fn_t g_encode = nullptr;

int get_encode(encode_fn_t* out) {
  if (!g_encode) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    MGB_CUDA_CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn,
                                           cudaEnableDefault, &q));
    MGB_REQUIRE(fn != nullptr && q == cudaDriverEntryPointSuccess, MGB_ESTATE,
                "cuTensorMapEncodeTiled not available in this driver");
    g_encode = (encode_fn_t)fn;
  }
  *out = g_encode;
  return MGB_OK;
}

// W is [H][P] row-major bf16; box = [LL_TILE objects][64 features], SW128.
int make_w_map(CUtensorMap* map, const void* w, uint64_t H, uint64_t P) {
  encode_fn_t enc;
  const int rc = get_encode(&enc);
  if (rc != MGB_OK) return rc;
  cuuint64_t gdim[2] = {P, H};
  cuuint64_t gstride[1] = {P * 2};
  cuuint32_t box[2] = {LL_FB, LL_TILE};
  cuuint32_t estr[2] = {1, 1};
  const CUresult r =
      enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(w), gdim,
          gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  MGB_REQUIRE(r == CUDA_SUCCESS, MGB_ESTATE,
              "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return MGB_OK;
}

__device__ __forceinline__ uint16_t bf16_bits(float v) {
  return __bfloat16_as_ushort(__float2bfloat16_rn(v));
}
__device__ __forceinline__ float bf16_val(float v) {
  return __bfloat162float(__float2bfloat16_rn(v));
}

// Write the Theta operand of MMA #1: 16 rows (n) x P (k), K-major SW128, one
// 2 KiB block per 64 features.  Rows 0..2 = the three bf16 planes of theta.
__device__ void write_theta_tile(uint8_t* base, const float* theta, int P,
                                 int tid, int nthreads) {
  const int nfb = P / LL_FB;
  for (int i = tid; i < nfb * 16 * LL_FB; i += nthreads) {
    const int fb = i / (16 * LL_FB), rem = i - fb * 16 * LL_FB;
    const int n = rem / LL_FB, k = rem - n * LL_FB;
    float v = 0.f;
    if (n < 3) {
      const float t = theta[fb * LL_FB + k];
      const float hi = bf16_val(t);
      const float mid = bf16_val(t - hi);
      v = n == 0 ? hi : (n == 1 ? mid : bf16_val((t - hi) - mid));
    }
    *reinterpret_cast<uint16_t*>(base + fb * 2048 + sw128_offset_16b(n, k)) =
        bf16_bits(v);
  }
}

// ---------------------------------------------------------------------------
// Self-test: one tile through both MMAs (validates the tensor map, the smem /
// instruction descriptors, the TMEM lane mapping and the manual swizzle).
//   out_mu[o]   = sum_j W[o,j] theta_j                       o < 128
//   out_j[r][j] = sum_o bf16(G[r][o]) W[o,j]                 r < 32, j < P
// ---------------------------------------------------------------------------
template <int P>
__global__ void __launch_bounds__(128, 1)
linlink_selftest_kernel(const __grid_constant__ CUtensorMap wmap,
                        const float* __restrict__ theta,
                        const float* __restrict__ gmat, float* out_mu,
                        float* out_j, int* status) {
  constexpr int NFB = P / LL_FB;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* s_w = smem;                              // NFB x 16 KiB
  uint8_t* s_th = s_w + NFB * LL_TILE * 128;        // NFB x 2 KiB
  uint8_t* s_g = s_th + NFB * 2048;                 // 2 x 4 KiB (+ read slack)
  __shared__ uint64_t bar_w, bar_mu, bar_j;
  __shared__ uint32_t s_tmem;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
    mbar_init(&bar_w, 1);
    mbar_init(&bar_mu, 1);
    mbar_init(&bar_j, 1);
    mbar_fence_init();
    tma_prefetch_desc(&wmap);
  }
  if (warp == 0) {
    tmem_alloc(&s_tmem, 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  const uint32_t tm_d1 = tmem;        // 16 columns
  const uint32_t tm_d2 = tmem + 32;   // P columns

  if (tid == 0) {
    mbar_arrive_expect_tx(&bar_w, NFB * LL_TILE * 128);
    for (int fb = 0; fb < NFB; ++fb)
      tma_load_2d(s_w + fb * LL_TILE * 128, &wmap, &bar_w, fb * LL_FB, 0);
  }
  write_theta_tile(s_th, theta, P, tid, blockDim.x);
  for (int i = tid; i < LL_GROWS * LL_TILE; i += blockDim.x) {
    const int r = i / LL_TILE, o = i - r * LL_TILE;
    const int kb = o >> 6, ko = o & 63;
    *reinterpret_cast<uint16_t*>(s_g + kb * 4096 + sw128_offset_16b(r, ko)) =
        bf16_bits(gmat[i]);
  }
  fence_proxy_async_smem();
  __syncthreads();

  int ok = 1;
  if (tid == 0) {
    ok &= mbar_wait(&bar_w, 0);
    tc_fence_after();
    const uint32_t id1 = idesc_bf16_f32(128, 16, false, false);
    for (int ks = 0; ks < P / 16; ++ks) {
      const int fb = ks >> 2, kk = ks & 3;
      const uint64_t a = smem_desc_sw128(
          smem_u32(s_w + fb * LL_TILE * 128) + kk * 32, 16, 1024);
      const uint64_t b =
          smem_desc_sw128(smem_u32(s_th + fb * 2048) + kk * 32, 16, 1024);
      umma_bf16(tm_d1, a, b, id1, ks > 0);
    }
    umma_commit(&bar_mu);
    const uint32_t id2 = idesc_bf16_f32(128, P, false, true);
    for (int ks = 0; ks < LL_TILE / 16; ++ks) {
      const int kb = ks >> 2, kk = ks & 3;
      const uint64_t a =
          smem_desc_sw128(smem_u32(s_g + kb * 4096) + kk * 32, 16, 1024);
      // B = W tile viewed MN-major: 16 object rows of 128 B per K step;
      // LBO = distance between 64-feature blocks, SBO = 8-row groups
      const uint64_t b = smem_desc_sw128(smem_u32(s_w) + ks * 2048,
                                         LL_TILE * 128, 1024);
      umma_bf16(tm_d2, a, b, id2, ks > 0);
    }
    umma_commit(&bar_j);
  }
  ok &= mbar_wait(&bar_mu, 0);
  tc_fence_after();
  {
    uint32_t v[16];
    tmem_ld_x16(tm_d1 + ((uint32_t)(warp * 32) << 16), v);
    tmem_ld_wait();
    out_mu[warp * 32 + lane] = __uint_as_float(v[0]) + __uint_as_float(v[1]) +
                               __uint_as_float(v[2]);
  }
  ok &= mbar_wait(&bar_j, 0);
  tc_fence_after();
  if (warp == 0) {
    for (int c = 0; c < P; c += 16) {
      uint32_t v[16];
      tmem_ld_x16(tm_d2 + c, v);
      tmem_ld_wait();
      for (int i = 0; i < 16; ++i)
        out_j[lane * P + c + i] = __uint_as_float(v[i]);
    }
  }
  if (!ok) atomicExch(status, 1);
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}

}  // namespace

extern "C" int mgb_linlink_selftest(const void* w_bf16, const float* theta,
                                    const float* gmat, float* out_mu,
                                    float* out_j, int* status_dev, int P,
                                    void* stream) {
  MGB_REQUIRE(w_bf16 && theta && gmat && out_mu && out_j && status_dev,
              MGB_EINVAL, "mgb_linlink_selftest: null argument");
  MGB_REQUIRE(P == 64 || P == 128, MGB_EINVAL,
              "mgb_linlink_selftest: P must be 64 or 128");
  CUtensorMap map;
  const int rc = make_w_map(&map, w_bf16, LL_TILE, (uint64_t)P);
  if (rc != MGB_OK) return rc;
  const int nfb = P / LL_FB;
  const size_t smem = (size_t)nfb * LL_TILE * 128 + nfb * 2048 + 2 * 4096 +
                      16384 + 1024;
  cudaStream_t st = (cudaStream_t)stream;
  if (P == 64) {
    MGB_CUDA_CHECK(cudaFuncSetAttribute(
        linlink_selftest_kernel<64>,
        cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    linlink_selftest_kernel<64><<<1, 128, smem, st>>>(map, theta, gmat, out_mu,
                                                      out_j, status_dev);
  } else {
    MGB_CUDA_CHECK(cudaFuncSetAttribute(
        linlink_selftest_kernel<128>,
        cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    linlink_selftest_kernel<128><<<1, 128, smem, st>>>(map, theta, gmat,
                                                       out_mu, out_j,
                                                       status_dev);
  }
  MGB_CUDA_CHECK(cudaGetLastError());
  return MGB_OK;
}
