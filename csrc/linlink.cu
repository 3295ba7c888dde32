// Linear-link soft-binned statistic on the 5th-generation tensor cores.
//
// Model:  mu_h = x_h + sum_j W[h,j] theta_j   (W: per-object features, bf16)
//         sigma = link(p_sigma)               (global)
// so that  d s_k / d theta_j = c_k * sum_h (g_{k+1,h} - g_{k,h}) W[h,j]
// is a tall-skinny GEMM with the *object axis as the MMA K dimension*:
//
//   forward   mu[128 objects]      = Wtile[128 x P] . theta        (3 columns)
//   Jacobian  D2[P, 32 rows]      += Wtile^T[P x 128 obj] . G^T[128 obj x 32]
//
// The Jacobian product runs on tcgen05 (UMMA, accumulators in TMEM): the
// TMA-loaded, 128B-swizzled W tile is its MN-major A operand, the Gaussian
// weights G (two bf16 planes per edge, ~2^-17 accuracy) its K-major B operand.
// The forward product has only three useful columns (theta as three bf16
// planes, fp32-grade); as a UMMA it re-read the whole tile through the tensor
// cores' operand path, which then bounded the kernel, so the compute warps do
// it themselves with ldmatrix + mma.sync on the same shared-memory tile.
// CUDA cores evaluate erfc / exp per (object, edge) and convert g to bf16.
// W streams through HBM exactly once per step (the reference's forward + VJP
// would stream it twice).
//
// Reference counterpart: jax.vjp of a user sumstats function with a
// per-object linear predictor (OnePointModel._vjp, multigrad.py:508-538).
#include <cuda.h>
#include <cuda_bf16.h>
#include <math.h>
#include <string.h>
#include "mgb_common.cuh"
#include "mgb_opt.cuh"
#include "mgb_sm100.cuh"

using namespace sm100;

#ifdef LL_TRACE
// Debug-only pipeline trace (tools/linlink_trace.py): SM-clock stamps of the
// pipeline events of CTA 0 for LL_TRACE_N tiles starting at LL_TRACE_T0.
#define LL_TRACE_N 256
#define LL_TRACE_T0 300
__device__ long long g_ll_trace[8][LL_TRACE_N];
#define LL_TR(ev, tile)                                                       \
  do {                                                                        \
    const long long _t = (tile) - LL_TRACE_T0;                                \
    if (blockIdx.x == 0 && _t >= 0 && _t < LL_TRACE_N)                        \
      g_ll_trace[ev][_t] = clock64();                                         \
  } while (0)
__device__ long long g_ll_trace_w[32][2][LL_TRACE_N];   // per compute warp
#define LL_TRW(cw, ev, tile)                                                  \
  do {                                                                        \
    const long long _t = (tile) - LL_TRACE_T0;                                \
    if (blockIdx.x == 0 && _t >= 0 && _t < LL_TRACE_N)                        \
      g_ll_trace_w[cw][ev][_t] = clock64();                                   \
  } while (0)
extern "C" int mgb_linlink_trace_read(long long* host) {
  return (int)cudaMemcpyFromSymbol(host, g_ll_trace, sizeof(g_ll_trace));
}
extern "C" int mgb_linlink_trace_read_warps(long long* host) {
  return (int)cudaMemcpyFromSymbol(host, g_ll_trace_w, sizeof(g_ll_trace_w));
}
#else
#define LL_TR(ev, tile) do {} while (0)
#define LL_TRW(cw, ev, tile) do {} while (0)
#endif

namespace {

constexpr int LL_TILE = 128;     // objects per tile (UMMA M of #1, K of #2)
constexpr int LL_FB = 64;        // features per 128-byte swizzle block
constexpr int LL_GROWS = 32;     // rows of G: 16 x g_hi, 16 x g_lo

#ifndef LL_SWAP
#define LL_SWAP 1
#endif
#ifndef LL_ABL
#define LL_ABL 0   // ablation bits for tuning experiments (0 in production)
#endif
#ifndef LL_L2_PROMOTION
#define LL_L2_PROMOTION CU_TENSOR_MAP_L2_PROMOTION_L2_256B
#endif

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
          CU_TENSOR_MAP_SWIZZLE_128B, LL_L2_PROMOTION,
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
// Entries at or beyond `nfeat` are zero: the parameter vector is only nfeat+1
// long, and a stray NaN there would poison the product even though the
// matching (padding) columns of W are zero.
__device__ void write_theta_tile(uint8_t* base, const float* theta, int P,
                                 int nfeat, int tid, int nthreads) {
  const int nfb = P / LL_FB;
  for (int i = tid; i < nfb * 16 * LL_FB; i += nthreads) {
    const int fb = i / (16 * LL_FB), rem = i - fb * 16 * LL_FB;
    const int n = rem / LL_FB, k = rem - n * LL_FB;
    float v = 0.f;
    if (n < 3 && fb * LL_FB + k < nfeat) {
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
                        float* out_j, int* status, int m64) {
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
  write_theta_tile(s_th, theta, P, P, tid, blockDim.x);
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
    const uint32_t id2 = idesc_bf16_f32(m64 ? 64 : 128, P, false, true);
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
  if (warp == 0 || m64) {
    // m64: dump every TMEM lane (row = lane) so the M=64 data-path layout
    // can be read off; otherwise rows 0..31 only
    for (int c = 0; c < P; c += 16) {
      uint32_t v[16];
      tmem_ld_x16(tm_d2 + c + ((uint32_t)(warp * 32) << 16), v);
      tmem_ld_wait();
      for (int i = 0; i < 16; ++i)
        out_j[(warp * 32 + lane) * P + c + i] = __uint_as_float(v[i]);
    }
  }
  if (!ok) atomicExch(status, 1);
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 256);
}


// ===========================================================================
// The fused step.
//   warp 0        producer: TMA of the W tile (128 objects x P features,
//                 SW128) + bulk copies of the tile's x / weights
//   warp 1        UMMA issuer (Jacobian product) + TMEM owner
//   warps 2..     compute: `sets` x 4 warps; a warp = 32 objects of a tile and
//                 all of its <= 12 edges
// Pipeline per tile (W stage s of LL_STAGES, G buffer gb of the G ring):
//   TMA -> full_w[s] -> compute (ldmatrix + mma.sync for mu, erfc, g -> bf16
//   hi/lo into G[gb]) -> g_full[gb] -> Jacobian UMMAs (accumulate D2) ->
//   empty[s], g_free[gb] -> TMA ...
// ===========================================================================
// pipeline depth: the W tile of a stage lives from its TMA until the Jacobian
// product has consumed it, so the depth has to cover the HBM latency
#ifndef LL_STAGES64
#define LL_STAGES64 8
#endif
__host__ __device__ constexpr int ll_stages(int P) {
  return P <= 64 ? LL_STAGES64 : 5;
}
constexpr int LL_MAX_STAGES = 12;
// independent TMEM accumulators of the Jacobian product: its UMMAs are tiny
// (64 x P x 16), so a single accumulator chain is bound by the ~95-cycle
// accumulate latency; K steps are dealt round-robin to LL_NACC chains that
// are summed at read-out
#if LL_SWAP
// (swapped form: an accumulator is only 32 columns wide, so every K step of a
// tile gets its own one: 8 x 32 columns from column 128)
#ifndef LL_NACC_SWAP
#define LL_NACC_SWAP 8
#endif
__host__ __device__ constexpr int ll_nacc(int) { return LL_NACC_SWAP; }
#else
__host__ __device__ constexpr int ll_nacc(int P) { return P <= 64 ? 4 : 2; }
#endif
// Compute-warp layout.  A tile is worked on by 4 * groups warps (one per TMEM
// lane quadrant and edge group); a warp owns LL_CH * nch edges of its 32
// objects; `sets` tiles are in the compute phase at the same time (set k owns
// tiles it % sets == k).  Few warps with many edges each amortise the
// per-tile hand-shakes (two barrier waits, tcgen05.ld, proxy fence, arrive)
// and give every warp nch * LL_CH independent chains.
#ifndef LL_CH
#define LL_CH 4
#endif
#ifndef LL_NCH64
#define LL_NCH64 3
#endif
#ifndef LL_SETS64
#define LL_SETS64 3
#endif
#ifndef LL_NCH128
#define LL_NCH128 3
#endif
#ifndef LL_SETS128
#define LL_SETS128 3
#endif
constexpr int LL_MAXE = MGB_LIN_MAX_BINS + 1;
__host__ __device__ constexpr int ll_nch(int P) {
  return P <= 64 ? LL_NCH64 : LL_NCH128;
}
__host__ __device__ constexpr int ll_sets(int P) {
  return P <= 64 ? LL_SETS64 : LL_SETS128;
}
__host__ __device__ constexpr int ll_groups(int P) {
  return (LL_MAXE + LL_CH * ll_nch(P) - 1) / (LL_CH * ll_nch(P));
}
__host__ __device__ constexpr int ll_threads(int P) {
  return 32 * (2 + ll_sets(P) * 4 * ll_groups(P));
}
constexpr int LL_MAX_CWARPS = 30;
// The Gaussian operand G of a tile (8 KiB) lives only from the compute phase
// to the Jacobian product, so it has its own, shorter ring: shared memory goes
// to W stages instead (the stream is bound by bytes in flight: a stage cycle is
// HBM latency under load + both products + the compute phase).
#ifndef LL_GEXTRA64
#define LL_GEXTRA64 1
#endif
#ifndef LL_GEXTRA128
#define LL_GEXTRA128 0
#endif
__host__ __device__ constexpr int ll_gextra(int P) {
  return P <= 64 ? LL_GEXTRA64 : LL_GEXTRA128;
}
__host__ __device__ constexpr int ll_gbufs(int P) {
  return (ll_sets(P) + ll_gextra(P)) < ll_stages(P)
             ? ll_sets(P) + ll_gextra(P)
             : ll_stages(P);
}
constexpr float LL_KAPPA = 1.2011224087864498f;
constexpr float LL_HSCALE = MGB_ERFCX_CSCALE / LL_KAPPA;

struct LinSmemLayout {
  int w_bytes, th_off, g_off, x_off, total;
};
// per stage: 128 floats of x, then 128 floats of per-object weights
constexpr int LL_XSTAGE = 2 * LL_TILE * 4;
__host__ __device__ inline LinSmemLayout lin_smem_layout(int P) {
  LinSmemLayout L;
  L.w_bytes = (P / LL_FB) * LL_TILE * 128;          // per stage
  L.th_off = 0;
  L.g_off = ll_stages(P) * L.w_bytes;
  L.x_off = L.g_off + ll_gbufs(P) * 8192;
  L.total = L.x_off + ll_stages(P) * LL_XSTAGE;
  return L;
}

template <int P, bool WEIGHTED>
__global__ void __launch_bounds__(ll_threads(P), 1)
linlink_step_kernel(const __grid_constant__ CUtensorMap wmap,
                    const MgbLinModel md, const MgbLinArgs a,
                    const MgbPeerTable pt, double* __restrict__ partials,
                    unsigned* __restrict__ counter) {
  constexpr int NFB = P / LL_FB;
  extern __shared__ __align__(1024) uint8_t smem[];
  const LinSmemLayout L = lin_smem_layout(P);
  uint8_t* s_w = smem;
  uint8_t* s_g = smem + L.g_off;
  // x (and weights) of a tile travel with its W tile: a 512-byte bulk copy on
  // the same mbarrier.  (Loading them with LDG from the compute warps put a
  // global-load latency in front of every G hand-over: the releasing
  // mbarrier.arrive drains the thread's outstanding loads first.)
  const float* s_xf = reinterpret_cast<const float*>(smem + L.x_off);

  constexpr int LL_STAGES = ll_stages(P);
  constexpr int NACC = ll_nacc(P);
  __shared__ uint64_t full_w[LL_MAX_STAGES],
      g_full[LL_MAX_STAGES], g_free[LL_MAX_STAGES], empty[LL_MAX_STAGES],
      d2_done;
  __shared__ uint32_t s_tmem;
  __shared__ float s_cs[LL_MAXE];
  constexpr int LL_THREADS = ll_threads(P);
  constexpr int SETS = ll_sets(P), NCH = ll_nch(P), GROUPS = ll_groups(P);
  constexpr int EPW = LL_CH * NCH;        // edges per compute warp
  constexpr int CWS = 4 * GROUPS;         // compute warps per set (= per tile)
  static_assert(SETS * CWS <= LL_MAX_CWARPS && LL_THREADS <= 1024, "layout");
  static_assert(SETS <= ll_stages(P) - 1, "a stage must be left for the TMA");
  __shared__ double s_warp[SETS * CWS][3 * EPW];
  __shared__ double s_edge[LL_MAXE][3];
  __shared__ int s_is_last, s_fail;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int G = gridDim.x;
  const int K = md.nbins, E = K + 1;
  const int NP = md.nfeat + 1;
  const int M = 2 * K + K * P;     // doubles per CTA partial
  const long long ntiles = (md.n + LL_TILE - 1) / LL_TILE;
#ifdef LL_CONTIG
  const long long per_cta = (ntiles + G - 1) / G;
  const long long tile0 = (long long)blockIdx.x * per_cta, tstride = 1;
  const long long my_tiles =
      ntiles > tile0 ? (ntiles - tile0 < per_cta ? ntiles - tile0 : per_cta) : 0;
#else
  const long long tile0 = blockIdx.x, tstride = G;
  const long long my_tiles =
      ntiles > blockIdx.x ? (ntiles - blockIdx.x + G - 1) / G : 0;
#endif

  const float* pv = (a.mode >= MGB_MODE_ADAM) ? a.opt_p : a.params;
  const float p_sg = pv[md.nfeat];
  const double sigma_d =
      md.sigma_link == 1 ? exp10((double)p_sg) : (double)p_sg;
  const double ks_d = 1.2011224087864498 * 0.7071067811865476 / sigma_d;
  const float ks = (float)ks_d;
  const float x0 = md.x0;

  if (tid == 0) {
    for (int s = 0; s < LL_STAGES; ++s) {
      mbar_init(&full_w[s], 1);
      mbar_init(&g_full[s], CWS);
      mbar_init(&g_free[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(&d2_done, 1);
    mbar_fence_init();
    tma_prefetch_desc(&wmap);
    s_fail = 0;
  }
  if (warp == 1) {
    tmem_alloc(&s_tmem, 512);
    tmem_relinquish();
  }
  if (tid < E)
    s_cs[tid] = (float)(((double)md.edges[tid] - (double)x0) * ks_d);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  constexpr int NGB = ll_gbufs(P);
  // (the parity waits of the compute warps rely on both)
  static_assert(NGB >= SETS && NGB <= LL_STAGES, "G ring length");
  const uint32_t tm_d2 = tmem;          // the D2 accumulators

  int fail = 0;
  if (warp == 0) {
    // ------------------------------------------------------ TMA producer
    if (lane == 0) {
      for (long long it = 0; it < my_tiles; ++it) {
        const int s = (int)(it % LL_STAGES);
        const uint32_t ph = (uint32_t)((it / LL_STAGES) & 1);
        if (!mbar_wait(&empty[s], ph ^ 1u)) { fail = 1; break; }
        LL_TR(0, it);
        const long long tile = tile0 + it * tstride;
        // a ragged last tile reads x / weights with plain loads instead
        const bool whole = (tile + 1) * LL_TILE <= md.n;
        const uint32_t xb = whole ? (WEIGHTED ? 2u : 1u) * (LL_TILE * 4) : 0u;
        mbar_arrive_expect_tx(&full_w[s], (uint32_t)L.w_bytes + xb);
        for (int fb = 0; fb < NFB; ++fb)
          tma_load_2d(s_w + s * L.w_bytes + fb * LL_TILE * 128, &wmap,
                      &full_w[s], fb * LL_FB, (int)(tile * LL_TILE));
        if (whole) {
          uint8_t* xs = smem + L.x_off + s * LL_XSTAGE;
          bulk_load(xs, md.x + tile * LL_TILE, LL_TILE * 4, &full_w[s]);
          if (WEIGHTED)
            bulk_load(xs + LL_TILE * 4, md.ow + tile * LL_TILE, LL_TILE * 4,
                      &full_w[s]);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------ UMMA issuer
    // The whole warp runs the (warp-uniform) polling loop and one elected
    // lane issues.  Descriptors are advanced by adding to their low word
    // (start address >> 4): building them from scratch cost ~17 dependent
    // uniform-datapath instructions per UMMA, which made this single thread
    // the bottleneck of the kernel (~95 cycles per 32-cycle UMMA).
    {
      // M = 64: only 32 rows of G are meaningful; the 64-row form halves the
      // operand fetch.  Its accumulator rows r land in TMEM lane
      // 32*(r/16) + r%16 (probed with tools/probe_m64.py).
#if LL_SWAP
      // Jacobian product with the operands swapped: D2[P x 32] += Wt . Gt,
      // A = W tile read MN-major (M = features), B = G read K-major
      // (N = 32 rows): 3 KiB (P = 64) / 5 KiB (P = 128) of operand fetch per
      // instruction instead of 4 / 6 KiB, and no zero-padded half of G.
      const uint32_t id2 = idesc_bf16_f32(P, LL_GROWS, true, false);
#else
      const uint32_t id2 = idesc_bf16_f32(64, P, false, true);
#endif
      const uint64_t dk = smem_desc_sw128(0, 16, 1024);                // K-major
      const uint64_t dmn = smem_desc_sw128(0, LL_TILE * 128, 1024);    // MN-major
      // Jacobian products are issued as soon as the compute warps have
      // published G for a tile.
      // One event stream (G published, in tile order), so the warp blocks in
      // mbarrier.try_wait instead of spinning over test_wait: a spinning
      // issuer takes issue slots from the three compute warps of its SM
      // sub-partition, and the slowest of a tile's four warps sets the time
      // its stage stays occupied (tools/linlink_trace.py: the quadrant-1
      // warps finished 0.2-0.75 us after the others).
      for (long long n2 = 0; n2 < my_tiles; ++n2) {
        const int s = (int)(n2 % LL_STAGES);
        const int gb = (int)(n2 % NGB);
        if (!mbar_wait(&g_full[gb], (uint32_t)((n2 / NGB) & 1))) {
          fail = 1;   // bounded: a broken pipeline must not hang the GPU
          break;
        }
        __syncwarp();
        LL_TR(4, n2);
        tc_fence_after();
        if (elect_one()) {
          const uint64_t b0 = dmn + (smem_u32(s_w + s * L.w_bytes) >> 4);
          const uint64_t a0 = dk + (smem_u32(s_g + gb * 8192) >> 4);
          const bool first = n2 == 0;
#pragma unroll
          for (int kstep = 0; kstep < ((LL_ABL & 1) ? 0 : LL_TILE / 16);
               ++kstep) {
#if LL_SWAP
            umma_bf16(tm_d2 + (kstep % NACC) * LL_GROWS,
                      b0 + kstep * (2048 / 16),
                      a0 + ((kstep >> 2) * (4096 / 16) + (kstep & 3) * 2), id2,
                      !first || kstep >= NACC);
#else
            umma_bf16(tm_d2 + (kstep % NACC) * P,
                      a0 + ((kstep >> 2) * (4096 / 16) + (kstep & 3) * 2),
                      b0 + kstep * (2048 / 16), id2, !first || kstep >= NACC);
#endif
          }
          umma_commit(&empty[s]);
          umma_commit(&g_free[gb]);
          LL_TR(7, n2);
        }
        __syncwarp();
      }
      // Only when somebody will wait for it: a commit is an *asynchronous*
      // arrive, and a CTA without tiles would exit before it lands -- the
      // late arrive then hits whatever the next CTA keeps at that shared
      // memory address (seen as sporadic NaNs in the following launch).
      if (my_tiles > 0 && elect_one()) umma_commit(&d2_done);
      __syncwarp();
    }
  } else {
    // ------------------------------------------------------ compute warps
    const int cw = warp - 2;
    const int set = cw / CWS;            // which tiles (it % SETS == set)
    const int grp = (cw % CWS) >> 2;     // edge group
    const int quad = warp & 3;           // TMEM lane quadrant this warp owns
    const int o = quad * 32 + lane;      // object within the tile
    const int e_first = grp * EPW;
    float cs[EPW];
    float accT[EPW], accZ[EPW];
    float accO[EPW];   // weighted: weight below the edge; else an int32 count
    uint32_t goff[EPW];
    // G word of this lane: even lanes store the g_hi pair of columns
    // (ko, ko+1) into row `edge`, odd lanes the g_lo pair into row
    // 16 + (edge ^ 4) (conflict-free: the ^4 moves the lo word four 16-byte
    // chunks away from the hi word inside the SW128 pattern)
    const bool odd = lane & 1;
    const uint32_t ko = (uint32_t)o & 63u;
    const uint32_t perm = odd ? 0x1054u : 0x7632u;
#pragma unroll
    for (int e = 0; e < EPW; ++e) {
      const uint32_t r = (uint32_t)(e_first + e);
      cs[e] = (e_first + e < E) ? s_cs[e_first + e] : 3.0e18f;
      accT[e] = 0.f; accZ[e] = 0.f;
      accO[e] = WEIGHTED ? 0.f : __int_as_float(0);
      goff[e] = sw128_offset_16b(odd ? 16u + (r ^ 4u) : r, ko & ~1u) +
                (uint32_t)(o >> 6) * 4096u;
    }
    // Forward product mu = W . theta of this warp's 32 objects on the
    // warp-level tensor path: ldmatrix of the SW128 tile + mma.sync
    // m16n8k16 (bf16 x bf16 -> fp32) against a register-resident B fragment
    // whose columns 0..2 are three bf16 planes of theta (fp32-grade product).
    // A tcgen05 product for it (M128 N16) re-read the whole 16 KiB tile through
    // the UMMA operand path, which -- not HBM -- then bounded the kernel
    // (measured: ~45 B/clk of operand fetch per SM for these tiny-N shapes).
    uint32_t bfr[P / 16][2];
    {
      const int pl = lane >> 2, q = lane & 3;      // plane (column), k pair
#pragma unroll
      for (int kk = 0; kk < P / 16; ++kk) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t word = 0;
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            const int f = kk * 16 + h * 8 + 2 * q + i;
            float th = (pl < 3 && f < md.nfeat) ? pv[f] : 0.0f;
            const float p0 = bf16_val(th);
            const float p1 = bf16_val(th - p0);
            const float vv = pl == 0 ? p0 : pl == 1 ? p1 : (th - p0) - p1;
            word |= (uint32_t)bf16_bits(vv) << (16 * i);
          }
          bfr[kk][h] = word;
        }
      }
    }
    // ldmatrix row address of this lane: matrix lane>>3 = (row half, k half),
    // row lane&7; the two 16-object blocks are 2 KiB apart
    uint32_t aoff[4];
    {
      const int mi = lane >> 3;
      const uint32_t R = (uint32_t)(quad * 32 + (mi & 1) * 8 + (lane & 7));
#pragma unroll
      for (int kc = 0; kc < 4; ++kc)
        aoff[kc] = R * 128u + (((uint32_t)(2 * kc + (mi >> 1)) ^ (R & 7u)) << 4);
    }
    const int src_lane = 4 * (lane & 7) + ((lane >> 3) & 1);
    long long obj = (tile0 + (long long)set * tstride) * LL_TILE + o;
    for (long long it = set; it < my_tiles; it += SETS) {
      const int s = (int)(it % LL_STAGES);
      const uint32_t ph = (uint32_t)((it / LL_STAGES) & 1);
      const bool valid = obj < md.n;
      const bool whole = obj - o + LL_TILE <= md.n;     // warp-uniform
      if (grp == 0 && quad == 0 && lane == 0) LL_TR(5, it);
      // Parity waits are only safe on a barrier that is at most one phase
      // behind or ahead.  TMA completions are not ordered between tiles and a
      // set skips the tiles of the other sets, so on its own this warp could
      // reach stage s while the PREVIOUS tile of that stage (it - LL_STAGES,
      // another set's) has not landed yet -- the parity test would then pass
      // on the phase before that (seen as wrong sums, barrier time-outs and
      // launch failures at P = 128).  The G buffer wait closes that: Jacobian
      // products are issued in tile order, so "product it - NGB complete"
      // implies every tile up to it - NGB >= it - LL_STAGES has landed; the
      // g_free parity itself is safe because NGB >= SETS.
      const int gb = (int)(it % NGB);
      if (it >= NGB &&
          !mbar_wait(&g_free[gb], (uint32_t)(((it / NGB) + 1) & 1))) {
        fail = 1;
        break;
      }
      // the tile's own barrier: it also makes the bulk-copied x visible
      if (!mbar_wait(&full_w[s], ph)) { fail = 1; break; }
      float xv = 1.0e30f, wv = 0.0f;
      if (!(LL_ABL & 16)) {
        if (whole) {
          xv = s_xf[s * (LL_XSTAGE / 4) + o];
          if (WEIGHTED) wv = s_xf[s * (LL_XSTAGE / 4) + LL_TILE + o];
        } else if (valid) {
          xv = __ldg(md.x + obj);
          if (WEIGHTED) wv = __ldg(md.ow + obj);
        }
      }
      obj += (long long)SETS * tstride * LL_TILE;
      if (grp == 0 && quad == 0 && lane == 0) LL_TR(2, it);
      if (lane == 0) LL_TRW(cw, 0, it);
      float muf = 0.0f;
      // lanes leave the polling loop of the barrier wait one by one; the
      // .aligned warp-level instructions below need them back together
      __syncwarp();
      if (!(LL_ABL & 32)) {
        const uint32_t wbase = smem_u32(s_w + s * L.w_bytes);
        float d[2][4];
#pragma unroll
        for (int b = 0; b < 2; ++b)
#pragma unroll
          for (int i = 0; i < 4; ++i) d[b][i] = 0.0f;
#pragma unroll
        for (int kk = 0; kk < P / 16; ++kk) {
#pragma unroll
          for (int b = 0; b < 2; ++b) {
            uint32_t a0, a1, a2, a3;
            const uint32_t addr = wbase + aoff[kk & 3] +
                                  (uint32_t)((kk >> 2) * (LL_TILE * 128) +
                                             b * 2048);
            asm volatile(
                "ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                : "=r"(a0), "=r"(a1), "=r"(a2), "=r"(a3)
                : "r"(addr));
            asm volatile(
                "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 "
                "{%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                : "+f"(d[b][0]), "+f"(d[b][1]), "+f"(d[b][2]), "+f"(d[b][3])
                : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(bfr[kk][0]),
                  "r"(bfr[kk][1]));
          }
        }
        // d[b][0..1]: row g = lane>>2 of block b, columns 2q, 2q+1;
        // d[b][2..3]: row g + 8.  Columns 0..2 are the planes: lanes q = 0, 1
        // hold them, their sum is mu of the row.
        float v00 = d[0][0] + d[0][1], v01 = d[0][2] + d[0][3];
        float v10 = d[1][0] + d[1][1], v11 = d[1][2] + d[1][3];
        v00 += __shfl_xor_sync(0xffffffffu, v00, 1);
        v01 += __shfl_xor_sync(0xffffffffu, v01, 1);
        v10 += __shfl_xor_sync(0xffffffffu, v10, 1);
        v11 += __shfl_xor_sync(0xffffffffu, v11, 1);
        // object o = 16 b + 8 half + g is served by lane 4 g + half, which
        // offers (b = 0, its half) and (b = 1, its half)
        const bool hsel = lane & 1;
        const float m0 = __shfl_sync(0xffffffffu, hsel ? v01 : v00, src_lane);
        const float m1 = __shfl_sync(0xffffffffu, hsel ? v11 : v10, src_lane);
        muf = (lane & 16) ? m1 : m0;
      }
      // u = (x - x0) + W.theta ; zs_e = cs_e - ks*u
      const float u = valid ? (xv - x0) + muf : 1.0e30f;
      uint8_t* gs = s_g + gb * 8192;
      if (LL_ABL & 4) {            // no math, no G stores: hand the tile on
        accT[0] += u;
        if (!(LL_ABL & 64)) fence_proxy_async_smem();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&g_full[gb]);
        continue;
      }
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        // chunks past the last edge are skipped (warp-uniform); their G rows
        // feed accumulator columns nobody reads
        if (e_first + c * LL_CH >= E) break;
        float zs[LL_CH], t[LL_CH], g[LL_CH], pp[LL_CH];
#pragma unroll
        for (int k = 0; k < LL_CH; ++k)
          zs[k] = fmaf(u, -ks, cs[c * LL_CH + k]);
#pragma unroll
        for (int k = 0; k < LL_CH; ++k)
          t[k] = mgb_rcp(fmaf(fabsf(zs[k]), LL_HSCALE, 1.0f));
#pragma unroll
        for (int k = 0; k < LL_CH; ++k) {
          g[k] = mgb_ex2(-(zs[k] * zs[k]));
          if (WEIGHTED) g[k] *= wv;      // the weight rides on the Gaussian
        }
        constexpr float kC[MGB_ERFCX_DEG + 1] = MGB_ERFCX_COEFFS;
#pragma unroll
        for (int k = 0; k < LL_CH; ++k)
          pp[k] = fmaf(0.5f * kC[MGB_ERFCX_DEG], t[k],
                       0.5f * kC[MGB_ERFCX_DEG - 1]);
#pragma unroll
        for (int d = MGB_ERFCX_DEG - 2; d >= 0; --d) {
#pragma unroll
          for (int k = 0; k < LL_CH; ++k)
            pp[k] = fmaf(pp[k], t[k], 0.5f * kC[d]);
        }
#pragma unroll
        for (int k = 0; k < LL_CH; ++k) {
          const int e = c * LL_CH + k;
          // T += +q below the edge, -q above, q = pp t g: the sign goes onto
          // g (sign bit from ~sign(zs)), the product into the FMA
          const float gsg = __uint_as_float(
              __float_as_uint(g[k]) | (~__float_as_uint(zs[k]) & 0x80000000u));
          accT[e] = fmaf(pp[k] * t[k], gsg, accT[e]);
          if (WEIGHTED) {
            accO[e] += zs[k] < 0.0f ? wv : 0.0f;
          } else {
            accO[e] = __int_as_float(__float_as_int(accO[e]) +
                                     (int)(__float_as_uint(zs[k]) >> 31));
          }
          accZ[e] = fmaf(zs[k], g[k], accZ[e]);
          // g -> two bf16 planes: hi = g truncated to bf16 (exact), lo =
          // bf16(g - hi); one cvt packs both.  Lanes pair up so that every
          // lane issues one full-word store (2-byte stores from all 32 lanes
          // cost 3x the wavefronts; ncu).
          const float ghi =
              __uint_as_float(__float_as_uint(g[k]) & 0xffff0000u);
          uint32_t mine;
          asm("cvt.rn.bf16x2.f32 %0, %1, %2;"
              : "=r"(mine) : "f"(ghi), "f"(g[k] - ghi));
          const uint32_t other = __shfl_xor_sync(0xffffffffu, mine, 1);
          if (!(LL_ABL & 2))
            *reinterpret_cast<uint32_t*>(gs + goff[e]) =
                __byte_perm(mine, other, perm);
        }
      }
      fence_proxy_async_smem();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&g_full[gb]);
      if (grp == 0 && quad == 0 && lane == 0) LL_TR(3, it);
      if (lane == 0) LL_TRW(cw, 1, it);
    }
#pragma unroll
    for (int e = 0; e < EPW; ++e) {
      const float tt = mgb_warp_sum(accT[e]);
      const float zz = mgb_warp_sum(accZ[e]);
      const double oo = WEIGHTED
          ? (double)mgb_warp_sum(accO[e])
          : (double)__reduce_add_sync(0xffffffffu, __float_as_int(accO[e]));
      if (lane == 0) {
        s_warp[cw][3 * e + 0] = (double)tt;
        s_warp[cw][3 * e + 1] = -oo;           // objects below -> sign flip
        s_warp[cw][3 * e + 2] = (double)zz;
      }
    }
  }
  if (fail) atomicExch(&s_fail, 1);
  __syncthreads();

  // ---- per-edge totals of T, O, Z (fp64) ---------------------------------
  for (int j = tid; j < 3 * E; j += LL_THREADS) {
    const int e = j / 3, q = j - 3 * e;
    const int grp = e / EPW, el = e - grp * EPW;
    double sacc = 0.0;
    for (int st = 0; st < SETS; ++st)
      for (int w = 0; w < 4; ++w)
        sacc += s_warp[st * CWS + grp * 4 + w][3 * el + q];
    s_edge[e][q] = sacc;
  }
  // ---- D2 (TMEM) -> shared: rows 0..15 g_hi, 16..31 g_lo, P columns -------
  float* s_gw = reinterpret_cast<float*>(smem);   // pipeline buffers are dead
#if LL_SWAP
  if (warp >= 4 && warp < 8) {
    // D2 is [feature -> TMEM lane][G row -> column].  Warp 4+q owns lanes
    // 32q..32q+31.  M = 128 (P = 128): lane = feature.  M = 64 (P = 64):
    // feature f sits in lane 32*(f/16) + f%16, lanes 16..31 of a quadrant
    // are unused.
    const int q = warp - 4;
    int okw = 1;
    if (my_tiles > 0) okw = mbar_wait(&d2_done, 0);
    tc_fence_after();
    const int f = (P == 64) ? 16 * q + lane : 32 * q + lane;
    const bool has_f = (P == 64) ? lane < 16 : true;
    for (int c = 0; c < LL_GROWS; c += 16) {
      float accv[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) accv[i] = 0.f;
      if (my_tiles > 0) {
        for (int ac = 0; ac < NACC; ++ac) {
          uint32_t v[16];
          tmem_ld_x16(tm_d2 + ac * LL_GROWS + c + ((uint32_t)(q * 32) << 16),
                      v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) accv[i] += __uint_as_float(v[i]);
        }
      }
      if (has_f) {
#pragma unroll
        for (int i = 0; i < 16; ++i) s_gw[(c + i) * P + f] = accv[i];
      }
    }
    if (!okw) atomicExch(&s_fail, 1);
  }
#else
  if (warp == 4 || warp == 5) {
    // warp 4 owns TMEM lanes 0..31 (rows 0..15 = g_hi in lanes 0..15),
    // warp 5 lanes 32..63 (rows 16..31 = g_lo in lanes 32..47)
    const int half = warp - 4;
    int okw = 1;
    if (my_tiles > 0) okw = mbar_wait(&d2_done, 0);
    tc_fence_after();
    for (int c = 0; c < P; c += 16) {
      float accv[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) accv[i] = 0.f;
      if (my_tiles > 0) {
        for (int ac = 0; ac < NACC; ++ac) {
          uint32_t v[16];
          tmem_ld_x16(tm_d2 + ac * P + c + ((uint32_t)(half * 32) << 16), v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 16; ++i) accv[i] += __uint_as_float(v[i]);
        }
      }
      if (lane < 16) {
#pragma unroll
        for (int i = 0; i < 16; ++i)
          s_gw[(half * 16 + lane) * P + c + i] = accv[i];
      }
    }
    if (!okw) atomicExch(&s_fail, 1);
  }
#endif
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);

  // ---- per-CTA partial: [D_k][Zd_k][GWd_k,j] -------------------------------
  for (int j = tid; j < M; j += LL_THREADS) {
    double val;
    if (j < K) {
      const int k = j;
      val = (s_edge[k + 1][1] - s_edge[k][1]) + (s_edge[k + 1][0] - s_edge[k][0]);
    } else if (j < 2 * K) {
      const int k = j - K;
      val = s_edge[k + 1][2] - s_edge[k][2];
    } else {
      const int k = (j - 2 * K) / P, f = (j - 2 * K) - k * P;
      const double hi1 = s_gw[(k + 1) * P + f];
      const double lo1 = s_gw[(16 + ((k + 1) ^ 4)) * P + f];
      const double hi0 = s_gw[k * P + f], lo0 = s_gw[(16 + (k ^ 4)) * P + f];
      val = (hi1 - hi0) + (lo1 - lo0);
    }
    partials[(size_t)blockIdx.x * M + j] = val;
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    const unsigned ticket = atomicAdd(counter, 1u);
    s_is_last = (ticket == (unsigned)(G - 1));
    if (s_fail) atomicExch(counter + 1, 1u);   // sticky pipeline-timeout flag
  }
  __syncthreads();
  if (!s_is_last) return;
  __threadfence();

  // ---- fixed-order grid reduction (8 lanes per value) ---------------------
  double* s_tot = reinterpret_cast<double*>(smem);            // M doubles
  double* s_stage = s_tot + ((M + 15) & ~15);                  // W * M doubles
  mgb_grid_reduce_f64(partials, G, M, s_tot, tid, LL_THREADS);
  if (tid == 0) *counter = 0u;
  __syncthreads();
  if (a.reduce && pt.world > 1) mgb_ll_allreduce_f64(pt, s_tot, s_stage, M);
  __syncthreads();

  // ---- epilogue: sumstats, loss, dL/ds, gradient, optimiser ---------------
  __shared__ double s_dl[MGB_LIN_MAX_BINS], s_ck[MGB_LIN_MAX_BINS],
      s_loss, s_gsig;
  const double inv = 0.7071067811865476 / sigma_d;
  const double rsqrtpi = 0.5641895835477563;
  const double dsdp = md.sigma_link == 1 ? 2.302585092994046 * sigma_d : 1.0;
  if (warp == 0) {
    double s_k = 0.0, l_k = 0.0, dl_k = 0.0, dsg_k = 0.0;
    if (lane < K) {
      const double nrm = (double)md.norm[lane];
      s_k = nrm * s_tot[lane];
      dsg_k = -nrm * (rsqrtpi * 0.8325546111576977 / sigma_d) * dsdp *
              s_tot[K + lane];
      s_ck[lane] = -nrm * inv * rsqrtpi;     // d s_k / d theta_j = ck * GWd
      if (a.out_sumstats) a.out_sumstats[lane] = (float)s_k;
      if (a.out_jac) a.out_jac[(size_t)lane * NP + md.nfeat] = (float)dsg_k;
      if (a.mode >= MGB_MODE_LOSS_GRAD) {
        const double tgt = (double)md.target[lane];
        if (md.loss_kind == 0) {
          const double eps = (double)md.loss_eps;
          const double r = log10(s_k + eps) - log10(tgt + eps);
          l_k = r * r / K;
          dl_k = 2.0 * r / (K * (s_k + eps) * 2.302585092994046);
        } else if (md.loss_kind == 1) {
          const double r = s_k - tgt;
          l_k = r * r / K;
          dl_k = 2.0 * r / K;
        } else {
          const double ie = (double)md.inv_err[lane];
          const double r = (s_k - tgt) * ie;
          l_k = r * r;
          dl_k = 2.0 * r * ie;
        }
      }
      s_dl[lane] = dl_k;
    }
    const double loss = mgb_warp_sum(l_k);
    const double gsig = mgb_warp_sum(dl_k * dsg_k);
    if (lane == 0) { s_loss = loss; s_gsig = gsig; }
  }
  __syncthreads();
  if (a.out_jac)
    for (int j = tid; j < K * md.nfeat; j += LL_THREADS) {
      const int k = j / md.nfeat, f = j - k * md.nfeat;
      a.out_jac[(size_t)k * NP + f] = (float)(s_ck[k] * s_tot[2 * K + k * P + f]);
    }
  if (a.mode < MGB_MODE_LOSS_GRAD) return;
  if (tid == 0 && a.out_loss) a.out_loss[0] = (float)s_loss;
  const int step = (a.mode >= MGB_MODE_ADAM) ? *a.opt_step : 0;
  const int row = step - a.step_base;
  if (tid < NP) {
    double gd = 0.0;
    if (tid < md.nfeat) {
      for (int k = 0; k < K; ++k)
        gd += s_dl[k] * s_ck[k] * s_tot[2 * K + k * P + tid];
    } else {
      gd = s_gsig;
    }
    const float grad = (float)gd;
    if (a.out_grad) a.out_grad[tid] = grad;
    if (a.mode >= MGB_MODE_ADAM)
      mgb_opt_update_param(a, tid, NP, step, row, grad,
                           a.bound_kind ? a.bound_kind[tid] : 0,
                           a.bound_lo ? a.bound_lo[tid] : 0.f,
                           a.bound_hi ? a.bound_hi[tid] : 0.f);
  }
  if (a.mode >= MGB_MODE_ADAM) {
    __syncthreads();
    if (tid == 0) {
      if (a.loss_hist && row >= 0 && row < a.traj_rows)
        a.loss_hist[row] = (float)s_loss;
      *a.opt_step = step + 1;
    }
  }
}

int g_ll_sms[64] = {0};
int ll_sm_count() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (dev >= 0 && dev < 64 && g_ll_sms[dev]) return g_ll_sms[dev];
  int sms = 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) !=
      cudaSuccess)
    return 0;
  if (dev >= 0 && dev < 64) g_ll_sms[dev] = sms;
  return sms;
}

}  // namespace

extern "C" uint64_t mgb_linlink_scratch_bytes(int nbins, int pfeat) {
  if (nbins < 1 || nbins > MGB_LIN_MAX_BINS || (pfeat != 64 && pfeat != 128))
    return 0;
  const uint64_t m = 2ull * nbins + (uint64_t)nbins * pfeat;
  return 256ull + (uint64_t)ll_sm_count() * m * 8ull;
}

extern "C" int mgb_linlink_step(const MgbLinModel* model,
                                const MgbLinArgs* args, const MgbPeerTable* pt,
                                void* scratch, uint64_t scratch_bytes,
                                void* stream) {
  MGB_REQUIRE(model && args && scratch, MGB_EINVAL,
              "mgb_linlink_step: null argument");
  const MgbLinModel& md = *model;
  const MgbLinArgs& a = *args;
  MGB_REQUIRE(md.nbins >= 1 && md.nbins <= MGB_LIN_MAX_BINS, MGB_ETOOBIG,
              "mgb_linlink_step: nbins=%d outside [1,%d]", md.nbins,
              MGB_LIN_MAX_BINS);
  MGB_REQUIRE(md.pfeat == 64 || md.pfeat == 128, MGB_EINVAL,
              "mgb_linlink_step: pfeat must be 64 or 128 (got %d)", md.pfeat);
  MGB_REQUIRE(md.nfeat >= 1 && md.nfeat <= md.pfeat, MGB_EINVAL,
              "mgb_linlink_step: nfeat=%d outside [1,%d]", md.nfeat, md.pfeat);
  MGB_REQUIRE(md.n >= 0 && md.w && (md.n == 0 || md.x), MGB_EINVAL,
              "mgb_linlink_step: x / w is null");
  MGB_REQUIRE(((uintptr_t)md.w & 15u) == 0 && ((uintptr_t)md.x & 15u) == 0 &&
                  ((uintptr_t)md.ow & 15u) == 0,
              MGB_EINVAL,
              "mgb_linlink_step: x, w and ow must be 16-byte aligned");
  MGB_REQUIRE(md.n < (1ll << 31), MGB_ETOOBIG,
              "mgb_linlink_step: n must be below 2^31 objects per GPU");
  MGB_REQUIRE(a.mode >= MGB_MODE_SUMSTATS_JAC && a.mode <= MGB_MODE_GD,
              MGB_EINVAL, "mgb_linlink_step: bad mode %d", a.mode);
  if (a.mode >= MGB_MODE_ADAM) {
    MGB_REQUIRE(a.opt_p && a.opt_step, MGB_EINVAL,
                "mgb_linlink_step: optimiser state missing");
    if (a.mode == MGB_MODE_ADAM)
      MGB_REQUIRE(a.opt_u && a.opt_m && a.opt_v, MGB_EINVAL,
                  "mgb_linlink_step: Adam state missing");
  } else {
    MGB_REQUIRE(a.params != nullptr, MGB_EINVAL,
                "mgb_linlink_step: params is null");
  }
  MgbPeerTable ptv;
  if (pt) {
    ptv = *pt;
  } else {
    memset(&ptv, 0, sizeof(ptv));
    ptv.world = 1;
  }
  const int mvals = 2 * md.nbins + md.nbins * md.pfeat;
  if (a.reduce && ptv.world > 1)
    MGB_REQUIRE(ptv.world <= MGB_MAX_WORLD &&
                    2u * (unsigned)mvals <= ptv.ll_units,
                MGB_ETOOBIG,
                "mgb_linlink_step: %d LL units needed, peer table has %u",
                2 * mvals, ptv.ll_units);
  const uint64_t need = mgb_linlink_scratch_bytes(md.nbins, md.pfeat);
  MGB_REQUIRE(need > 0 && scratch_bytes >= need, MGB_ETOOBIG,
              "mgb_linlink_step: scratch %llu < %llu bytes",
              (unsigned long long)scratch_bytes, (unsigned long long)need);
  CUtensorMap map;
  // at least one full box so that the encoder accepts tiny catalogues
  const uint64_t rows = md.n > 0 ? (uint64_t)md.n : 1;
  int rc = make_w_map(&map, md.w, rows, (uint64_t)md.pfeat);
  if (rc != MGB_OK) return rc;
  const LinSmemLayout L = lin_smem_layout(md.pfeat);
  // the epilogue reuses the pipeline buffers for the reduction staging
  size_t smem = (size_t)L.total;
  const size_t epi = ((size_t)((mvals + 15) & ~15) +
                      (size_t)(ptv.world > 1 ? ptv.world : 0) * mvals) * 8 + 64;
  if (epi > smem) smem = epi;
  MGB_REQUIRE(smem <= 227 * 1024, MGB_ETOOBIG,
              "mgb_linlink_step: %zu bytes of shared memory needed", smem);
  unsigned* counter = reinterpret_cast<unsigned*>(scratch);
  double* partials =
      reinterpret_cast<double*>(reinterpret_cast<char*>(scratch) + 256);
  const int grid = ll_sm_count();
  MGB_REQUIRE(grid > 0, MGB_ESTATE, "mgb_linlink_step: no CUDA device");
  cudaStream_t st = (cudaStream_t)stream;
  auto kern = md.pfeat == 64
      ? (md.ow ? linlink_step_kernel<64, true> : linlink_step_kernel<64, false>)
      : (md.ow ? linlink_step_kernel<128, true>
               : linlink_step_kernel<128, false>);
  MGB_CUDA_CHECK(cudaFuncSetAttribute(
      kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<grid, ll_threads(md.pfeat), smem, st>>>(map, md, a, ptv, partials,
                                                 counter);
  MGB_CUDA_CHECK(cudaGetLastError());
  return MGB_OK;
}

extern "C" int mgb_linlink_selftest(const void* w_bf16, const float* theta,
                                    const float* gmat, float* out_mu,
                                    float* out_j, int* status_dev, int P,
                                    void* stream) {
  // P + 1000 selects the M = 64 layout probe (out_j then holds 128 rows)
  const int m64 = P >= 1000 ? 1 : 0;
  if (m64) P -= 1000;
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
                                                      out_j, status_dev, m64);
  } else {
    MGB_CUDA_CHECK(cudaFuncSetAttribute(
        linlink_selftest_kernel<128>,
        cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    linlink_selftest_kernel<128><<<1, 128, smem, st>>>(map, theta, gmat,
                                                       out_mu, out_j,
                                                       status_dev, m64);
  }
  MGB_CUDA_CHECK(cudaGetLastError());
  return MGB_OK;
}
