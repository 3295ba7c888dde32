// K1/K2/K3 for USER-DEFINED links: fused soft-binned statistic + Jacobian +
// reductions + loss / optimiser epilogue for a model whose per-object mean and
// width are arbitrary differentiable functions of the object's coordinate, its
// features and the parameters,
//     mu_h = x_h + shift(x_h, f_h; p),   sigma_h = sigma(x_h, f_h; p).
// sm_100a.
//
// This file is a TEMPLATE: it is compiled once per model by
// multigrad_b200/ops/linkbin.py (nvcc, in tree) with -DUL_LINK_HEADER=<file>,
// a header generated from the user's Python link function by the expression
// tracer of multigrad_b200/link.py.  That header defines
//     UL_NP            number of parameters (<= MGB_MAX_PARAMS)
//     UL_NF            number of per-object feature columns
//     UL_MU_MASK       bit k set: d mu / d p_k is not identically zero
//     UL_SG_MASK       bit k set: d sigma / d p_k is not identically zero
//     UL_STOCHASTIC    the link draws random numbers
//     ul_eval(...)     shift, sigma and the non-zero derivatives for a pair
//                      of objects (packed float2), derivatives by symbolic
//                      differentiation of the traced expression
// The reference evaluates such a model by jax.vjp of the user's function
// (multigrad/multigrad.py:508-538: forward, Allreduce, VJP pass, Allreduce);
// here it is ONE pass: per object and bin edge
//     zs = kappa (edge - mu)/(sqrt2 sigma),  g = 2^(-zs^2),
//     T += -+ 1/2 erfc|z|,  O += [z >= 0],
//     J_k += g U_k + zs g V_k,   U_k = -dmu_k/(sqrt(2 pi) sigma),
//                                V_k = -dsigma_k/(sqrt(pi) kappa sigma)
// followed by the same CTA / grid / NVLink reductions and epilogue as
// csrc/erfbin.cu (one ordered exchange of (1+P) K doubles).
#include <math.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include "mgb_common.cuh"
#include "mgb_opt.cuh"

#ifndef UL_LINK_HEADER
#error "linkbin.cu is compiled per model: -DUL_LINK_HEADER=\"generated.h\""
#endif

namespace {

__device__ __forceinline__ float2 f2(float a, float b) {
  return make_float2(a, b);
}
__device__ __forceinline__ float2 f2s(float a) { return make_float2(a, a); }
__device__ __forceinline__ float2 vadd(float2 a, float2 b) {
  return __fadd2_rn(a, b);
}
__device__ __forceinline__ float2 vsub(float2 a, float2 b) {
  return __fadd2_rn(a, make_float2(-b.x, -b.y));
}
__device__ __forceinline__ float2 vmul(float2 a, float2 b) {
  return __fmul2_rn(a, b);
}
__device__ __forceinline__ float2 vfma(float2 a, float2 b, float2 c) {
  return __ffma2_rn(a, b, c);
}
__device__ __forceinline__ float2 vneg(float2 a) {
  return make_float2(-a.x, -a.y);
}
__device__ __forceinline__ float2 vdiv(float2 a, float2 b) {
  return make_float2(__fdividef(a.x, b.x), __fdividef(a.y, b.y));
}
#define UL_MAP1(name, expr)                                        \
  __device__ __forceinline__ float2 name(float2 a) {               \
    float v;                                                       \
    float2 r;                                                      \
    v = a.x; r.x = (expr);                                         \
    v = a.y; r.y = (expr);                                         \
    return r;                                                      \
  }
UL_MAP1(vexp, __expf(v))
UL_MAP1(vlog, __logf(v))
UL_MAP1(vlog10, __log10f(v))
UL_MAP1(vexp10, exp10f(v))
UL_MAP1(vsqrt, sqrtf(v))
UL_MAP1(vrcp, mgb_rcp(v))
UL_MAP1(vtanh, tanhf(v))
UL_MAP1(vabs, fabsf(v))
__device__ __forceinline__ float2 vpowc(float2 a, float c) {
  return make_float2(powf(a.x, c), powf(a.y, c));
}

// Counter-based random numbers for stochastic links: every draw is a pure
// function of (seed, step, global object index, draw number), so all ranks
// derive the same per-step key (the reference splits a jax
// key per step on the root, multigrad/adam.py:60-62) and a fit can be
// replayed.  SplitMix64 finaliser; the same arithmetic in
// multigrad_b200/link.py gives the PyTorch oracle.
struct UlRng {
  unsigned long long key;   // mixes seed and step
  long long i0, i1;         // global indices of the two objects of the pair
};
__device__ __forceinline__ unsigned long long ul_mix(unsigned long long z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
__device__ __forceinline__ float ul_u01(unsigned long long key, long long i,
                                        int draw) {
  const unsigned long long h =
      ul_mix(key ^ ul_mix((unsigned long long)i * 0x100ull +
                          (unsigned long long)draw));
  return ((float)(h >> 40) + 0.5f) * (1.0f / 16777216.0f);   // (0, 1)
}
__device__ __forceinline__ float2 ul_uniform(const UlRng& r, int draw) {
  return make_float2(ul_u01(r.key, r.i0, draw), ul_u01(r.key, r.i1, draw));
}
__device__ __forceinline__ float2 ul_normal(const UlRng& r, int draw) {
  // Box-Muller on two uniforms of the same object
  const float2 u = ul_uniform(r, 2 * draw), v = ul_uniform(r, 2 * draw + 1);
  return make_float2(
      sqrtf(-2.0f * __logf(u.x)) * __cosf(6.283185307179586f * v.x),
      sqrtf(-2.0f * __logf(u.y)) * __cosf(6.283185307179586f * v.y));
}

#include UL_LINK_HEADER

static_assert(UL_NP >= 1 && UL_NP <= MGB_MAX_PARAMS, "UL_NP out of range");

#define UL_KAPPA 1.2011224087864498f
#define UL_HSCALE (MGB_ERFCX_CSCALE / UL_KAPPA)
#define UL_ZS_SKIP 12.4f
#define UL_SENTINEL_X 1.0e30f
// Edges per warp (every warp of a group re-evaluates the link for its
// objects, so fewer, wider groups are better) limited by the accumulator
// registers, (2 + P) per edge: one group of up to 11 edges for P <= 3.
#if UL_NP <= 3
#define UL_ET 11
#elif UL_NP <= 5
#define UL_ET 6
#else
#define UL_ET 4
#endif
#define UL_CHUNK 3                   // edges per software-pipeline stage
#define UL_MAX_WARPS 12
#define UL_WIDE (UL_ET > 4)          // wide: one CTA of 12 warps per SM
#define UL_MAX_EDGES (MGB_MAX_BINS + 1)
#define UL_MAX_GROUPS ((UL_MAX_EDGES + UL_ET - 1) / UL_ET)
#define UL_NQ (2 + UL_NP)            // T, O, J_0 .. J_{P-1} per edge
#define UL_MAXVEC (3 * MGB_MAX_BINS) // (1+P) K doubles must fit

struct UlLayout {
  int ng, nwarps;
  int e0[UL_MAX_GROUPS + 1];
  int w0[UL_MAX_GROUPS + 1];
};

struct UlAcc {
  float T, O;
  float J[UL_NP];
};

// Stage A of a chunk of edges [e0, e0+CH): scaled coordinate, t = 1/(1+h|zs|)
// and Gaussian of every edge -- the MUFU work.
template <int ET, int CH>
__device__ __forceinline__ void ul_stage_a(const float* __restrict__ ec,
                                           float2 inv, float2 b, float2 wt,
                                           int e0, float2* zs, float2* g,
                                           float2* t) {
#pragma unroll
  for (int j = 0; j < CH; ++j)
    if (e0 + j < ET) zs[j] = vfma(f2s(ec[e0 + j]), inv, b);
#pragma unroll
  for (int j = 0; j < CH; ++j) {
    if (e0 + j < ET) {
      const float2 d = vfma(f2(fabsf(zs[j].x), fabsf(zs[j].y)),
                            f2s(UL_HSCALE), f2s(1.0f));
      t[j] = f2(mgb_rcp(d.x), mgb_rcp(d.y));
    }
  }
#pragma unroll
  for (int j = 0; j < CH; ++j) {
    if (e0 + j < ET) {
      const float2 zz = vmul(zs[j], zs[j]);
      g[j] = vmul(f2(mgb_ex2(-zz.x), mgb_ex2(-zz.y)), wt);
    }
  }
}

// Stage B: Horner polynomial and the accumulations -- the FMA work.
//   T += -+ 1/2 erfc|z| w,  O += [z >= 0] w,  J_k += g U_k + (zs g) V_k
template <int ET, int CH>
__device__ __forceinline__ void ul_stage_b(UlAcc* __restrict__ a, int e0,
                                           const float2* zs, const float2* g,
                                           const float2* t, float2 wt,
                                           const float2* U, const float2* V) {
  float2 p[CH];
  constexpr float kC[MGB_ERFCX_DEG + 1] = MGB_ERFCX_COEFFS;
#pragma unroll
  for (int j = 0; j < CH; ++j)
    if (e0 + j < ET)
      p[j] = vfma(f2s(0.5f * kC[MGB_ERFCX_DEG]), t[j],
                  f2s(0.5f * kC[MGB_ERFCX_DEG - 1]));
#pragma unroll
  for (int d = MGB_ERFCX_DEG - 2; d >= 0; --d) {
#pragma unroll
    for (int j = 0; j < CH; ++j)
      if (e0 + j < ET) p[j] = vfma(p[j], t[j], f2s(0.5f * kC[d]));
  }
#pragma unroll
  for (int j = 0; j < CH; ++j)
    if (e0 + j < ET) p[j] = vmul(p[j], t[j]);
#pragma unroll
  for (int j = 0; j < CH; ++j) {
    if (e0 + j < ET) {
      UlAcc& ac = a[e0 + j];
      const float gx = __uint_as_float(
          __float_as_uint(g[j].x) | (~__float_as_uint(zs[j].x) & 0x80000000u));
      const float gy = __uint_as_float(
          __float_as_uint(g[j].y) | (~__float_as_uint(zs[j].y) & 0x80000000u));
      ac.T = fmaf(p[j].x, gx, ac.T);
      ac.T = fmaf(p[j].y, gy, ac.T);
      ac.O += zs[j].x >= 0.0f ? wt.x : 0.0f;
      ac.O += zs[j].y >= 0.0f ? wt.y : 0.0f;
    }
  }
#pragma unroll
  for (int j = 0; j < CH; ++j) {
    if (e0 + j < ET) {
      UlAcc& ac = a[e0 + j];
      const float2 zg = vmul(zs[j], g[j]);
#pragma unroll
      for (int k = 0; k < UL_NP; ++k) {
        if ((UL_MU_MASK >> k) & 1) {
          ac.J[k] = fmaf(g[j].x, U[k].x, ac.J[k]);
          ac.J[k] = fmaf(g[j].y, U[k].y, ac.J[k]);
        }
        if ((UL_SG_MASK >> k) & 1) {
          ac.J[k] = fmaf(zg.x, V[k].x, ac.J[k]);
          ac.J[k] = fmaf(zg.y, V[k].y, ac.J[k]);
        }
      }
    }
  }
}

// one pair of objects against the ET edges of a group; the MUFU stage of the
// next chunk is issued ahead of the FMA stage of the current one (see
// csrc/erfbin.cu: with three warps per scheduler a warp has to keep both
// pipes busy by itself)
template <int ET>
__device__ __forceinline__ void ul_pair(UlAcc* __restrict__ a,
                                        const float* __restrict__ ec,
                                        float2 muc, float2 inv, float2 wt,
                                        const float2* U, const float2* V) {
  constexpr int CH = UL_CHUNK < ET ? UL_CHUNK : ET;
  constexpr int NCH = (ET + CH - 1) / CH;
  // zs_e = (ec_e - muc) inv = ec_e inv + b
  const float2 b = vneg(vmul(muc, inv));
  float2 zs[2][CH], g[2][CH], t[2][CH];
  ul_stage_a<ET, CH>(ec, inv, b, wt, 0, zs[0], g[0], t[0]);
#pragma unroll
  for (int n = 0; n < NCH; ++n) {
    if (n + 1 < NCH)
      ul_stage_a<ET, CH>(ec, inv, b, wt, (n + 1) * CH, zs[(n + 1) & 1],
                         g[(n + 1) & 1], t[(n + 1) & 1]);
    ul_stage_b<ET, CH>(a, n * CH, zs[n & 1], g[n & 1], t[n & 1], wt, U, V);
  }
}

struct UlShared {
  float ec[UL_MAX_EDGES];              // edge - x0
  float pv[MGB_MAX_PARAMS];
  double warp[UL_MAX_WARPS][UL_NQ * UL_ET];
  double edge[UL_MAX_EDGES][UL_NQ];
  double tot[UL_MAXVEC];
  double stage[MGB_MAX_WORLD * UL_MAXVEC];
  int is_last;
};

// four consecutive objects of one lane: x, weight, features
struct UlTile {
  float4 x, w;
  float4 f[UL_NF > 0 ? UL_NF : 1];
};

template <bool WEIGHTED>
__device__ __forceinline__ void ul_fetch(const MgbLinkModel& md, long long i,
                                         UlTile& t) {
  const long long n4 = md.n >> 2;
  if (i < n4) {
    t.x = __ldg(reinterpret_cast<const float4*>(md.x) + i);
    if (WEIGHTED) t.w = __ldg(reinterpret_cast<const float4*>(md.w) + i);
#pragma unroll
    for (int j = 0; j < UL_NF; ++j)
      t.f[j] = __ldg(reinterpret_cast<const float4*>(md.feat[j]) + i);
  } else {
    t.x = make_float4(UL_SENTINEL_X, UL_SENTINEL_X, UL_SENTINEL_X,
                      UL_SENTINEL_X);
    t.w = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int j = 0; j < UL_NF; ++j) t.f[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i == n4) {  // ragged tail: the last n % 4 objects
      const long long b4 = n4 << 2;
      const int rem = (int)(md.n - b4);
      if (rem > 0) t.x.x = md.x[b4];
      if (rem > 1) t.x.y = md.x[b4 + 1];
      if (rem > 2) t.x.z = md.x[b4 + 2];
      if (WEIGHTED) {
        if (rem > 0) t.w.x = md.w[b4];
        if (rem > 1) t.w.y = md.w[b4 + 1];
        if (rem > 2) t.w.z = md.w[b4 + 2];
      }
#pragma unroll
      for (int j = 0; j < UL_NF; ++j) {
        if (rem > 0) t.f[j].x = md.feat[j][b4];
        if (rem > 1) t.f[j].y = md.feat[j][b4 + 1];
        if (rem > 2) t.f[j].z = md.feat[j][b4 + 2];
      }
    }
  }
}

// The objects beyond the end of the shard (padding) must contribute nothing
// whatever the user's link makes of their coordinate: they get weight 0 and a
// harmless coordinate, and the unweighted hard count is corrected for them.
template <int ET, bool WEIGHTED>
__device__ __forceinline__ void ul_accumulate(
    const MgbLinkModel& md, const float* __restrict__ ec_all, int e_first,
    const float* __restrict__ pv, unsigned long long key, int wj, int wpg,
    int lane, double* s_out) {
  float ec[ET];
  UlAcc acc[ET];
#pragma unroll
  for (int e = 0; e < ET; ++e) {
    ec[e] = ec_all[e_first + e];
    acc[e].T = 0.f;
    acc[e].O = 0.f;
#pragma unroll
    for (int k = 0; k < UL_NP; ++k) acc[e].J[k] = 0.f;
  }
  float p[UL_NP];
#pragma unroll
  for (int k = 0; k < UL_NP; ++k) p[k] = pv[k];
  const long long n4 = md.n >> 2;
  const long long tstride = (long long)gridDim.x * wpg * 32;
  long long i0 = ((long long)blockIdx.x + (long long)wj * gridDim.x) * 32;
  UlTile nxt;
  if (i0 < n4 + 1) ul_fetch<WEIGHTED>(md, i0 + lane, nxt);
  for (; i0 < n4 + 1; i0 += tstride) {
    const UlTile tl = nxt;
    if (i0 + tstride < n4 + 1)
      ul_fetch<WEIGHTED>(md, i0 + tstride + lane, nxt);
    const long long obj0 = (i0 + lane) * 4;
#pragma unroll 1
    for (int h = 0; h < 2; ++h) {
      float2 x = h ? f2(tl.x.z, tl.x.w) : f2(tl.x.x, tl.x.y);
      float2 wt = h ? f2(tl.w.z, tl.w.w) : f2(tl.w.x, tl.w.y);
      float2 fe[UL_NF > 0 ? UL_NF : 1];
#pragma unroll
      for (int j = 0; j < UL_NF; ++j)
        fe[j] = h ? f2(tl.f[j].z, tl.f[j].w) : f2(tl.f[j].x, tl.f[j].y);
      // padding objects: evaluate the link at the centring coordinate, with
      // weight zero
      const bool padx = x.x == UL_SENTINEL_X, pady = x.y == UL_SENTINEL_X;
      if (padx) x.x = md.x0;
      if (pady) x.y = md.x0;
      if (!WEIGHTED) wt = f2(padx ? 0.f : 1.f, pady ? 0.f : 1.f);
      UlRng rng;
      rng.key = key;
      rng.i0 = md.index0 + obj0 + 2 * h;
      rng.i1 = rng.i0 + 1;
      const float2 xc = vadd(x, f2s(-md.x0));
      float2 shift, sigma, dmu[UL_NP], dsg[UL_NP];
      ul_eval(x, xc, fe, p, rng, shift, sigma, dmu, dsg);
      const float2 rs = f2(mgb_rcp(sigma.x), mgb_rcp(sigma.y));
      const float2 inv = vmul(rs, f2s(UL_KAPPA * 0.70710678118654752f));
      const float2 muc = vadd(xc, shift);
      float2 U[UL_NP], V[UL_NP];
#pragma unroll
      for (int k = 0; k < UL_NP; ++k) {
        // U_k = -dmu_k/(sqrt(2 pi) sigma), V_k = -dsg_k/(sqrt(pi) kappa sigma)
        U[k] = ((UL_MU_MASK >> k) & 1)
                   ? vmul(vmul(dmu[k], rs), f2s(-0.3989422804014327f))
                   : f2s(0.f);
        V[k] = ((UL_SG_MASK >> k) & 1)
                   ? vmul(vmul(dsg[k], rs),
                          f2s(-0.5641895835477563f / UL_KAPPA))
                   : f2s(0.f);
      }
      // every object carries a weight here (1 or 0 without user weights)
      ul_pair<ET>(acc, ec, muc, inv, wt, U, V);
    }
  }
#pragma unroll
  for (int e = 0; e < ET; ++e) {
    const float t = mgb_warp_sum(acc[e].T);
    const float o = mgb_warp_sum(acc[e].O);
    if (lane == 0) {
      s_out[UL_NQ * e + 0] = (double)t;
      s_out[UL_NQ * e + 1] = (double)o;
    }
#pragma unroll
    for (int k = 0; k < UL_NP; ++k) {
      const float j = mgb_warp_sum(acc[e].J[k]);
      if (lane == 0) s_out[UL_NQ * e + 2 + k] = (double)j;
    }
  }
}

// models with few edges: groups narrower than UL_ET - 1
template <bool WEIGHTED>
__device__ __noinline__ void ul_accumulate_small(
    const MgbLinkModel& md, const float* __restrict__ ec_all, int e_first,
    int et, const float* __restrict__ pv, unsigned long long key, int wj,
    int wpg, int lane, double* s_out) {
  switch (et) {
#define UL_CASE(n)                                                          \
  case n:                                                                   \
    if (n < UL_ET - 1)                                                      \
      ul_accumulate<(n < UL_ET - 1 ? n : 1), WEIGHTED>(                     \
          md, ec_all, e_first, pv, key, wj, wpg, lane, s_out);              \
    break;
    UL_CASE(1) UL_CASE(2) UL_CASE(3) UL_CASE(4) UL_CASE(5) UL_CASE(6)
    UL_CASE(7) UL_CASE(8) UL_CASE(9)
#undef UL_CASE
    default: break;
  }
}

template <bool WEIGHTED>
__global__ void __launch_bounds__(UL_MAX_WARPS * 32, UL_WIDE ? 1 : 2)
linkbin_step_kernel(const __grid_constant__ MgbLinkModel md,
                    const __grid_constant__ MgbStepArgs a,
                    const __grid_constant__ MgbPeerTable pt,
                    const __grid_constant__ UlLayout lay,
                    double* __restrict__ partials,
                    unsigned* __restrict__ counters) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nthreads = blockDim.x;
  const int G = gridDim.x;
  const int K = md.nbins, E = K + 1;
  constexpr int P = UL_NP;
  const int MP = (1 + P) * K;
  __shared__ UlShared sh;

  asm volatile("griddepcontrol.wait;" ::: "memory");
  const float* pv = (a.mode >= MGB_MODE_ADAM) ? a.opt_p : a.params;
  if (tid < P) sh.pv[tid] = pv[tid];
  if (tid < E) sh.ec[tid] = (float)(md.edges[tid] - (double)md.x0);
  __syncthreads();
  // per-step key of the stochastic links: the same on every rank
  const unsigned step =
      (a.mode >= MGB_MODE_ADAM && !(a.flags & MGB_STEP_CONST_KEY))
          ? (unsigned)__ldcg(a.opt_step)
          : a.rng_step;
  const unsigned long long key =
      ul_mix(md.seed ^ ul_mix(0x5851F42D4C957F2Dull + step));

  {
    int g = 0;
    while (g + 1 < lay.ng && warp >= lay.w0[g + 1]) ++g;
    const int e_first = lay.e0[g];
    const int et = lay.e0[g + 1] - e_first;
    const int wpg = lay.w0[g + 1] - lay.w0[g];
    const int wj = warp - lay.w0[g];
    // groups hold UL_ET or UL_ET - 1 edges (make_layout), fewer only when the
    // model has fewer edges than that
    if (et == UL_ET)
      ul_accumulate<UL_ET, WEIGHTED>(md, sh.ec, e_first, sh.pv, key, wj, wpg,
                                     lane, sh.warp[warp]);
    else if (et == UL_ET - 1 && UL_ET > 1)
      ul_accumulate<(UL_ET > 1 ? UL_ET - 1 : 1), WEIGHTED>(
          md, sh.ec, e_first, sh.pv, key, wj, wpg, lane, sh.warp[warp]);
    else
      ul_accumulate_small<WEIGHTED>(md, sh.ec, e_first, et, sh.pv, key, wj,
                                    wpg, lane, sh.warp[warp]);
  }
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  __syncthreads();

  // ---- CTA reduction in fp64: per-edge totals, then bin differences --------
  for (int j = tid; j < UL_NQ * E; j += nthreads) {
    const int e = j / UL_NQ, q = j - e * UL_NQ;
    int g = 0;
    while (g + 1 < lay.ng && e >= lay.e0[g + 1]) ++g;
    double sum = 0.0;
    for (int w = lay.w0[g]; w < lay.w0[g + 1]; ++w)
      sum += sh.warp[w][UL_NQ * (e - lay.e0[g]) + q];
    sh.edge[e][q] = sum;
  }
  __syncthreads();
  for (int j = tid; j < MP; j += nthreads) {
    const int q = j / K, k = j - q * K;   // q = 0: Phi differences, 1+p: J_p
    double val;
    if (q == 0)
      val = (sh.edge[k + 1][1] - sh.edge[k][1]) +
            (sh.edge[k + 1][0] - sh.edge[k][0]);
    else
      val = sh.edge[k + 1][1 + q] - sh.edge[k][1 + q];
    partials[(size_t)blockIdx.x * MP + j] = val;
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    const unsigned ticket = atomicAdd(&counters[0], 1u);
    sh.is_last = (ticket == (unsigned)(G - 1));
  }
  __syncthreads();
  if (!sh.is_last) return;
  __threadfence();
  mgb_grid_reduce_f64(partials, G, MP, sh.tot, tid, nthreads);
  if (tid == 0) counters[0] = 0u;
  __syncthreads();
  if (a.reduce && pt.world > 1) mgb_ll_allreduce_f64(pt, sh.tot, sh.stage, MP);

  // ---- epilogue (warp 0; lane k owns bin k) ----------------------------------
  if (warp != 0) return;
  double s_k = 0.0;
  double jk[P];
#pragma unroll
  for (int i = 0; i < P; ++i) jk[i] = 0.0;
  if (lane < K) {
    const double nrm = (double)md.norm[lane];
    s_k = nrm * sh.tot[lane];
#pragma unroll
    for (int i = 0; i < P; ++i) jk[i] = nrm * sh.tot[(1 + i) * K + lane];
    if (a.out_sumstats) a.out_sumstats[lane] = (float)s_k;
    if (a.out_jac && a.mode >= MGB_MODE_SUMSTATS_JAC) {
#pragma unroll
      for (int i = 0; i < P; ++i) a.out_jac[lane * P + i] = (float)jk[i];
    }
  }
  if (a.mode < MGB_MODE_LOSS_GRAD) return;
  double l_k = 0.0, dl_k = 0.0;
  if (lane < K) {
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
  const double loss = mgb_warp_sum(l_k);
  float grad = 0.f;
#pragma unroll
  for (int i = 0; i < P; ++i) {
    const double gi = mgb_warp_sum(dl_k * jk[i]);
    if (lane == i) grad = (float)gi;
  }
  const bool mine = lane < P;
  if (lane == 0 && a.out_loss) a.out_loss[0] = (float)loss;
  if (mine && a.out_grad) a.out_grad[lane] = grad;
  if (a.mode < MGB_MODE_ADAM) return;
  const int stp = *a.opt_step;
  const int row = stp - a.step_base;
  if (lane == 0 && a.loss_hist && row >= 0 && row < a.traj_rows)
    a.loss_hist[row] = (float)loss;
  if (mine)
    mgb_opt_update_param(a, lane, P, stp, row, grad, a.bound_kind[lane],
                         a.bound_lo[lane], a.bound_hi[lane]);
  __syncwarp();
  if (lane == 0) *a.opt_step = stp + 1;
}

UlLayout ul_make_layout(int nbins) {
  UlLayout L;
  memset(&L, 0, sizeof(L));
  const int E = nbins + 1;
  L.ng = (E + UL_ET - 1) / UL_ET;
  const int lo = E / L.ng, extra = E % L.ng;
  const int even = UL_MAX_WARPS / L.ng > 0 ? UL_MAX_WARPS / L.ng : 1;
  int e = 0, w = 0;
  for (int g = 0; g < L.ng; ++g) {
    const int et = lo + (g < extra ? 1 : 0);
    L.e0[g] = e;
    L.w0[g] = w;
    e += et;
    w += (L.ng == 1) ? UL_MAX_WARPS : ((E <= UL_MAX_WARPS) ? et : even);
  }
  L.e0[L.ng] = e;
  L.w0[L.ng] = w;
  L.nwarps = w;
  return L;
}

int ul_grid(int nbins) {
  int dev = 0, sms = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) !=
      cudaSuccess)
    return 0;
  return sms * (UL_WIDE ? 1 : 2);
}

}  // namespace

// Minimal error plumbing of a per-model library (the main library's
// workspace.cu is not linked into it).
static thread_local char g_ul_err[512] = "";
void mgb_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_ul_err, sizeof(g_ul_err), fmt, ap);
  va_end(ap);
}
extern "C" const char* mgb_linkbin_last_error(void) { return g_ul_err; }
extern "C" int mgb_linkbin_nparams(void) { return UL_NP; }
extern "C" int mgb_linkbin_nfeat(void) { return UL_NF; }
extern "C" int mgb_linkbin_stochastic(void) { return UL_STOCHASTIC; }
extern "C" uint32_t mgb_linkbin_sizeof_model(void) {
  return (uint32_t)sizeof(MgbLinkModel);
}

extern "C" uint64_t mgb_linkbin_scratch_bytes(int nbins) {
  if (nbins < 1 || nbins > MGB_MAX_BINS) return 0;
  return 256ull + (uint64_t)ul_grid(nbins) * (1 + UL_NP) * nbins * 8ull;
}

extern "C" int mgb_linkbin_step(const MgbLinkModel* model,
                                const MgbStepArgs* args,
                                const MgbPeerTable* pt, void* scratch,
                                uint64_t scratch_bytes, void* stream) {
  MGB_REQUIRE(model && args && scratch, MGB_EINVAL,
              "mgb_linkbin_step: null argument");
  const MgbLinkModel& md = *model;
  const MgbStepArgs& a = *args;
  MGB_REQUIRE(md.nbins >= 1 && md.nbins <= MGB_MAX_BINS, MGB_ETOOBIG,
              "mgb_linkbin_step: nbins=%d outside [1,%d]", md.nbins,
              MGB_MAX_BINS);
  MGB_REQUIRE(md.nparams == UL_NP && md.nfeat == UL_NF, MGB_EINVAL,
              "mgb_linkbin_step: this library was generated for %d parameters "
              "and %d features, the model has %d and %d", UL_NP, UL_NF,
              md.nparams, md.nfeat);
  MGB_REQUIRE((1 + UL_NP) * md.nbins <= UL_MAXVEC, MGB_ETOOBIG,
              "mgb_linkbin_step: (1 + nparams) * nbins = %d exceeds %d",
              (1 + UL_NP) * md.nbins, UL_MAXVEC);
  MGB_REQUIRE(md.n >= 0 && (md.n == 0 || md.x != nullptr), MGB_EINVAL,
              "mgb_linkbin_step: x is null");
  MGB_REQUIRE(((uintptr_t)md.x & 15u) == 0 && ((uintptr_t)md.w & 15u) == 0,
              MGB_EINVAL, "mgb_linkbin_step: x / w must be 16-byte aligned");
  for (int j = 0; j < UL_NF; ++j)
    MGB_REQUIRE(md.feat[j] != nullptr && ((uintptr_t)md.feat[j] & 15u) == 0,
                MGB_EINVAL,
                "mgb_linkbin_step: feature column %d missing or unaligned", j);
  MGB_REQUIRE(a.batch == 1, MGB_EINVAL, "mgb_linkbin_step: batch must be 1");
  MGB_REQUIRE(a.mode >= MGB_MODE_SUMSTATS && a.mode <= MGB_MODE_GD, MGB_EINVAL,
              "mgb_linkbin_step: bad mode %d", a.mode);
  if (a.mode >= MGB_MODE_ADAM) {
    MGB_REQUIRE(a.opt_p && a.opt_step, MGB_EINVAL,
                "mgb_linkbin_step: optimiser state missing");
    if (a.mode == MGB_MODE_ADAM)
      MGB_REQUIRE(a.opt_u && a.opt_m && a.opt_v, MGB_EINVAL,
                  "mgb_linkbin_step: Adam state missing");
  } else {
    MGB_REQUIRE(a.params != nullptr, MGB_EINVAL,
                "mgb_linkbin_step: params is null");
  }
  MgbPeerTable ptv;
  if (pt) {
    ptv = *pt;
  } else {
    memset(&ptv, 0, sizeof(ptv));
    ptv.world = 1;
  }
  if (a.reduce && ptv.world > 1)
    MGB_REQUIRE(ptv.world <= MGB_MAX_WORLD &&
                    2u * (1 + UL_NP) * md.nbins <= ptv.ll_units,
                MGB_ETOOBIG, "mgb_linkbin_step: peer table too small");
  const uint64_t need = mgb_linkbin_scratch_bytes(md.nbins);
  MGB_REQUIRE(need > 0 && scratch_bytes >= need, MGB_ETOOBIG,
              "mgb_linkbin_step: scratch %llu < %llu bytes",
              (unsigned long long)scratch_bytes, (unsigned long long)need);
  const UlLayout lay = ul_make_layout(md.nbins);
  unsigned* counters = reinterpret_cast<unsigned*>(scratch);
  double* partials =
      reinterpret_cast<double*>(reinterpret_cast<char*>(scratch) + 256);
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((unsigned)ul_grid(md.nbins), 1, 1);
  cfg.blockDim = dim3((unsigned)lay.nwarps * 32, 1, 1);
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (md.w)
    MGB_CUDA_CHECK(cudaLaunchKernelEx(&cfg, linkbin_step_kernel<true>, md, a,
                                      ptv, lay, partials, counters));
  else
    MGB_CUDA_CHECK(cudaLaunchKernelEx(&cfg, linkbin_step_kernel<false>, md, a,
                                      ptv, lay, partials, counters));
  return MGB_OK;
}
