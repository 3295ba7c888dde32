// K1/K2/K3: fused erf-soft-binned statistic + Jacobian + cross-CTA and
// cross-GPU reduction + loss / chain rule / optimiser epilogue.  sm_100a.
//
// What one launch does (see include/multigrad_b200.h for the contract):
//   main loop   every thread streams objects (float4 loads) and, for each of
//               the edges of its warp's edge group, evaluates on packed fp32x2
//               lanes (FFMA2/FMUL2, two objects per instruction)
//                   zs  = kappa*(edge - mu)/(sqrt2*sigma)      (kappa^2 = log2 e)
//                   g   = 2^(-zs^2) = exp(-z^2)
//                   q   = 1/2 erfc(|z|) = g * t*P(t)/2, t = 1/(1+|z|/2) (MUFU.RCP)
//               and accumulates per edge  T = sum -+q (tail part of the CDF),
//               O = sum [z>=0] (hard count), G = sum g, Z = sum zs*g.
//               Phi = O + T exactly, so bin sums never subtract numbers near 1.
//               Uniformly spaced edges: only the central edge of a group gets
//               its Gaussian from MUFU.EX2; the others follow by a
//               multiplicative recurrence
//                   g_{e+1} = g_e * rho_e,   rho_{e+1} = rho_e * 2^(-2 d^2)
//               Tiles with an object too far from the central edge for that
//               to be accurate, and non-uniform edges, take the per-edge
//               MUFU.EX2 path.
//   CTA reduce  warp shuffles in fp32, cross-warp and everything after in fp64.
//   cluster     the CTAs of a thread-block cluster push their 3K partial sums
//               into the leader CTA's shared memory (DSMEM, st.shared::cluster)
//               and the leader adds them in rank order: one global partial
//               per cluster instead of one per CTA.
//   grid reduce the last cluster leader (atomic ticket) sums the partials in a
//               fixed order -> run-to-run deterministic.
//   GPU reduce  the same CTA pushes the 3K doubles to every peer with the LL
//               protocol over NVLink and sums in rank order (mgb_common.cuh).
//   epilogue    sumstats, Jacobian, loss, dL/ds, gradient, bound-transform
//               chain rule and the Adam / GD update, trajectory + loss rows.
//   persistent  with args.nsteps > 1 all of the above repeats inside ONE
//               launch: the other CTAs wait on the device for the step counter
//               and re-read the parameters (their first tile already in
//               registers).
//
// Reference counterparts: OnePointModel._vjp (multigrad/multigrad.py:508-538),
// calc_smf_bin (tests/smf_example/smf_grad_descent.py:38-48), _adam_optimizer
// (multigrad/adam.py:52-68), transforms (adam.py:192-239), simple_grad_descent
// (multigrad/util.py:100-114), and for the persistent form the in-scan
// collectives of multigrad/mpi4jax/multigrad.py:38-58.
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <mutex>
#include "mgb_common.cuh"
#include "mgb_opt.cuh"

#define EB_SENTINEL_X 1.0e30f // padding object: contributes exactly zero

// words of the counter block (the block is 64 words; tickets are words
// [0, batch), the others are only used with batch == 1)
#define EB_RELAY_WORD 32      // host hand-shake relay (token + parameters)
#define EB_ABORT_WORD 24      // persistent fit: a step failed, everybody leave
#define EB_DYN_WORD 64        // + group*EB_NPOOL + pool: tile claim counters
#define EB_NPOOL 1            // claim pools per group (8 pools spread the
                              // atomics but balance worse: measured no gain)
#define EB_COUNTER_WORDS 256  // minimum size of the counter block
#ifndef EB_HOST_POLL_NS
#define EB_HOST_POLL_NS 100
#endif

namespace {

__device__ __forceinline__ float2 f2(float a, float b) {
  return make_float2(a, b);
}
__device__ __forceinline__ float2 f2s(float a) { return make_float2(a, a); }
#if defined(MGB_NO_F32X2)
__device__ __forceinline__ float2 vadd(float2 a, float2 b) {
  return make_float2(a.x + b.x, a.y + b.y);
}
__device__ __forceinline__ float2 vmul(float2 a, float2 b) {
  return make_float2(a.x * b.x, a.y * b.y);
}
__device__ __forceinline__ float2 vfma(float2 a, float2 b, float2 c) {
  return make_float2(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y));
}
#else
__device__ __forceinline__ float2 vadd(float2 a, float2 b) {
  return __fadd2_rn(a, b);
}
__device__ __forceinline__ float2 vmul(float2 a, float2 b) {
  return __fmul2_rn(a, b);
}
__device__ __forceinline__ float2 vfma(float2 a, float2 b, float2 c) {
  return __ffma2_rn(a, b, c);
}
#endif

// kappa = sqrt(log2(e)): exp(-z^2) == 2^(-(kappa z)^2)
#define EB_KAPPA 1.2011224087864498f
#define EB_INV_KAPPA 0.8325546111576977
// t = 1/(1 + c |z|) = 1/(1 + (c/kappa) |zs|), c = MGB_ERFCX_CSCALE
#define EB_HSCALE (MGB_ERFCX_CSCALE / EB_KAPPA)
#define EB_ZS_SKIP 12.4f   // |zs| beyond which g and q flush to zero
// Gaussian recurrence: used for a tile while every object is within za_lim of
// the group's central edge (in zs units), za_lim = min(EB_ZA_MAX,
// (EB_CHAIN_EXP_MAX - J^2 d^2) / (2 J d)): the exponent of every ratio then
// stays below EB_CHAIN_EXP_MAX in magnitude, which bounds the relative error
// a ratio inherits from the fp32 rounding of its exponent to
// EB_CHAIN_EXP_MAX * 2^-24 * ln 2 = 1e-6 (and keeps everything finite).
#define EB_ZA_MAX 10.0f
#define EB_CHAIN_EXP_MAX 24.0

// ------------------------------------------------------------ cluster / DSMEM
__device__ __forceinline__ unsigned eb_cluster_ctarank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ unsigned eb_cluster_nctarank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void eb_cluster_arrive() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
}
__device__ __forceinline__ void eb_cluster_wait() {
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `p` (a shared-memory object of this kernel) in
// the CTA of rank `rank` of this cluster
__device__ __forceinline__ unsigned eb_mapa(const void* p, unsigned rank) {
  unsigned a = (unsigned)__cvta_generic_to_shared(p), r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r)
               : "r"(a), "r"(rank));
  return r;
}
__device__ __forceinline__ void eb_st_cluster_f64(unsigned addr, double v) {
  asm volatile("st.shared::cluster.f64 [%0], %1;" ::"r"(addr), "d"(v)
               : "memory");
}
__device__ __forceinline__ int eb_ld_acquire_gpu_s32(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void eb_st_release_gpu_s32(int* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v)
               : "memory");
}

// Per-edge accumulators are scalar: a packed FADD2 and two FADDs cost the
// same FMA-pipe time and four registers per edge are what lets ptxas keep the
// edges' Horner chains in flight together.
// EB_ACC2 (off): T, G, Z accumulate the two objects of a pair separately
// (packed FFMA2 / FADD2 instead of two scalar FFMA / FADD each), the halves
// added once at the end.  Measured SLOWER (714 vs 663 us per step at 1e8
// objects with 12 warps, 748 with 8): a packed instruction holds the issue
// port for two cycles, so it saves nothing over two scalar ones, and the 33
// extra registers cost scheduling freedom.
#ifndef EB_ACC2
#define EB_ACC2 0
#endif
struct EdgeAcc {
#if EB_ACC2
  float2 T, G, Z;
  float O;
#else
  float T, O, G, Z;
#endif
};

// Gaussian recurrence (uniform edges).  With A0 = ET/2 the central edge of a
// group, zs_{A0+k} = za + k d and
//     g_{A0+k} = 2^(-zs_{A0+k}^2) = g_{A0+k-1} rho_k,
//     rho_k = 2^(-2 za d - (2k-1) d^2) = rho_1 kap^(k-1),  kap = 2^(-2 d^2)
// (downward: d -> -d).  MUFU.EX2 supplies g_A0, rho_1 of both directions and,
// EB_SKIP edges away from the centre, the direct ratio g_{A0+-J} / g_A0 =
// 2^(-+2 J za d - J^2 d^2), from which the recurrence restarts: no Gaussian
// is more than EB_SKIP - 1 multiplications by a MUFU-derived ratio away from
// one that MUFU.EX2 delivered, and all of them inherit the rounding of g_A0,
// so that differences between neighbouring edges (the mu-Jacobian) stay
// accurate for wide Gaussians.  7 to 10 MUFU.EX2 per pair of objects and
// group instead of 2 per edge.
#ifndef EB_SKIP
#define EB_SKIP 3
#endif
#ifndef EB_CHUNK
#define EB_CHUNK 3
#endif
#ifndef EB_RCP_BATCH
#define EB_RCP_BATCH 0
#endif


// warp-uniform constants of one edge group for the current parameters
template <int ET>
struct GroupConst {
  float cs[ET];     // zs_e = cs[e] - ks*xc
  float ks;         // kappa / (sqrt2 sigma)
  float two_d;      // 2 d,  d = ks * bin width
  float nd2;        // -d^2
  float kap;        // 2^(-2 d^2)
  float two_jd;     // 2 J d          (J = EB_SKIP)
  float nj2d2;      // -J^2 d^2
  float kapj;       // kap^J: rho_{J+1} = rho_1 kap^J
  float za_lim;     // recurrence allowed while |zs_A0| <= za_lim (<= 0: never)
};

// Processing order of the ET edges of a group: the central edge A0 first,
// then upward to the last edge, then downward from A0-1 to edge 0 (the order
// in which the recurrence produces the Gaussians).
template <int ET>
struct EbOrd {
  static constexpr int A0 = ET / 2;
  static constexpr int KU = ET - 1 - A0;  // edges above the centre
  static constexpr int KD = A0;           // edges below it
  __host__ __device__ static constexpr int edge(int i) {
    return i <= KU ? A0 + i : A0 - (i - KU);
  }
  // signed distance from the centre in units of the bin width
  __host__ __device__ static constexpr int k(int i) {
    return i <= KU ? i : -(i - KU);
  }
};

// per-object state of the Gaussian recurrence
struct EbChain {
  float2 ga;        // Gaussian of the central edge (times the weight)
  float2 ru, rd;    // rho_1 upward / downward
  float2 rju, rjd;  // direct ratio g_{A0+-J} / g_A0
  float2 gc, r;     // Gaussian of the previous edge, next ratio
};

// Stage A of a chunk (edges ord(i0) .. ord(i0+CH-1)): scaled coordinate,
// Gaussian and t = 1/(1 + h |zs|) of every edge -- all the MUFU work.
template <int ET, int CH, bool WEIGHTED, bool CHAIN>
__device__ __forceinline__ void eb_stage_a(const GroupConst<ET>& c, float2 xc,
                                           float2 wt, int i0, EbChain& ch,
                                           float2* zs, float2* g, float2* t) {
  typedef EbOrd<ET> OD;
#pragma unroll
  for (int j = 0; j < CH; ++j)
    if (i0 + j < ET)
      zs[j] = vfma(xc, f2s(-c.ks), f2s(c.cs[OD::edge(i0 + j)]));
  // t = 1/d, d = 1 + h |zs| (the |.| is an operand modifier of the FFMA2).
  // EB_RCP_BATCH (off): two reciprocals from one MUFU.RCP, r = 1/(d0 d1),
  // t0 = r d1, t1 = r d0.  Measured SLOWER (734 vs 660 us per step at 1e8
  // objects): the kernel is bound by issue slots (a packed FFMA2/FMUL2 holds
  // the port for two cycles), not by the XU pipe, so trading one MUFU pair for
  // three FMUL2 loses.
  {
    float2 d[CH];
#pragma unroll
    for (int j = 0; j < CH; ++j)
      if (i0 + j < ET)
        d[j] = vfma(f2(fabsf(zs[j].x), fabsf(zs[j].y)), f2s(EB_HSCALE),
                    f2s(1.0f));
#pragma unroll
    for (int j = 0; j < CH; ++j) {
      if (i0 + j < ET) {
#if EB_RCP_BATCH
        if ((j & 1) == 0 && j + 1 < CH && i0 + j + 1 < ET) {
          const float2 dd = vmul(d[j], d[j + 1]);
          const float2 r = f2(mgb_rcp(dd.x), mgb_rcp(dd.y));
          t[j] = vmul(r, d[j + 1]);
          t[j + 1] = vmul(r, d[j]);
        } else if ((j & 1) == 0) {
          t[j] = f2(mgb_rcp(d[j].x), mgb_rcp(d[j].y));
        }
#else
        t[j] = f2(mgb_rcp(d[j].x), mgb_rcp(d[j].y));
#endif
      }
    }
  }
#pragma unroll
  for (int j = 0; j < CH; ++j) {
    if (i0 + j < ET) {
      if (CHAIN) {
        const int k = OD::k(i0 + j);
        const int ak = k < 0 ? -k : k;
        const int kmax = k < 0 ? OD::KD : OD::KU;
        if (k == 0) {
          ch.gc = ch.ga;
          ch.r = ch.ru;
        } else if (k == -1) {          // the downward run starts at the centre
          ch.gc = vmul(ch.ga, ch.rd);
          ch.r = vmul(ch.rd, f2s(c.kap));
        } else if (ak == EB_SKIP) {    // direct ratio from the centre
          ch.gc = vmul(ch.ga, k > 0 ? ch.rju : ch.rjd);
          if (ak < kmax) ch.r = vmul(k > 0 ? ch.ru : ch.rd, f2s(c.kapj));
        } else {
          ch.gc = vmul(ch.gc, ch.r);
          if (ak < kmax && ak + 1 != EB_SKIP) ch.r = vmul(ch.r, f2s(c.kap));
        }
        g[j] = ch.gc;
      } else {
        const float2 zz = vmul(zs[j], zs[j]);
        g[j] = f2(mgb_ex2(-zz.x), mgb_ex2(-zz.y));
        if (WEIGHTED) g[j] = vmul(g[j], wt);
      }
    }
  }
}

// Stage B of a chunk: Horner polynomial and the four accumulations -- all
// FMA-pipe work.  The chains of the chunk are written *stage by stage* (all
// edges do step i before any edge does step i+1): ptxas keeps that order, so
// every warp has CH independent FFMA2 streams in flight and the Horner
// recurrence never exposes its latency.
template <int ET, int CH, bool WEIGHTED>
__device__ __forceinline__ void eb_stage_b(EdgeAcc* __restrict__ a, int i0,
                                           const float2* zs, const float2* g,
                                           const float2* t, float2 wt) {
  typedef EbOrd<ET> OD;
  float2 p[CH];
  // coefficients pre-halved so that p = erfcx/2 / t
  constexpr float kC[MGB_ERFCX_DEG + 1] = MGB_ERFCX_COEFFS;
#pragma unroll
  for (int j = 0; j < CH; ++j)
    if (i0 + j < ET)
      p[j] = vfma(f2s(0.5f * kC[MGB_ERFCX_DEG]), t[j],
                  f2s(0.5f * kC[MGB_ERFCX_DEG - 1]));
#pragma unroll
  for (int d = MGB_ERFCX_DEG - 2; d >= 0; --d) {
#pragma unroll
    for (int j = 0; j < CH; ++j)
      if (i0 + j < ET) p[j] = vfma(p[j], t[j], f2s(0.5f * kC[d]));
  }
#pragma unroll
  for (int j = 0; j < CH; ++j)
    if (i0 + j < ET) p[j] = vmul(p[j], t[j]);
#pragma unroll
  for (int j = 0; j < CH; ++j) {
    if (i0 + j < ET) {
      EdgeAcc& ac = a[OD::edge(i0 + j)];
      // T += +q for z < 0, -q for z >= 0, q = p*g = (weight) 1/2 erfc(|z|):
      // the sign goes onto g (sign bit from ~sign(zs)), the product into the
      // FMA
      const float gx = __uint_as_float(
          __float_as_uint(g[j].x) | (~__float_as_uint(zs[j].x) & 0x80000000u));
      const float gy = __uint_as_float(
          __float_as_uint(g[j].y) | (~__float_as_uint(zs[j].y) & 0x80000000u));
#if EB_ACC2
      ac.T = vfma(p[j], f2(gx, gy), ac.T);
#else
      ac.T = fmaf(p[j].x, gx, ac.T);
      ac.T = fmaf(p[j].y, gy, ac.T);
#endif
    }
  }
#pragma unroll
  for (int j = 0; j < CH; ++j) {
    if (i0 + j < ET) {
      EdgeAcc& ac = a[OD::edge(i0 + j)];
      if (WEIGHTED) {
        ac.O += zs[j].x >= 0.0f ? wt.x : 0.0f;
        ac.O += zs[j].y >= 0.0f ? wt.y : 0.0f;
      } else {
        // unweighted: count the objects *below* the edge on the integer pipe
        // (sign bit of zs; the bit pattern of O is an int32 counter)
        int cnt = __float_as_int(ac.O);
        cnt += (int)(__float_as_uint(zs[j].x) >> 31);
        cnt += (int)(__float_as_uint(zs[j].y) >> 31);
        ac.O = __int_as_float(cnt);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < CH; ++j) {
    if (i0 + j < ET) {
      EdgeAcc& ac = a[OD::edge(i0 + j)];
#if EB_ACC2
      ac.G = vadd(ac.G, g[j]);
      ac.Z = vfma(zs[j], g[j], ac.Z);
#else
      ac.G += g[j].x;
      ac.G += g[j].y;
      ac.Z = fmaf(zs[j].x, g[j].x, ac.Z);
      ac.Z = fmaf(zs[j].y, g[j].y, ac.Z);
#endif
    }
  }
}

// MUFU.EX2 part of the recurrence for one pair of objects: Gaussian of the
// central edge and the base / skip ratios.
template <int ET, bool WEIGHTED, bool CHAIN>
__device__ __forceinline__ void eb_anchor(const GroupConst<ET>& c, float2 xc,
                                          float2 wt, EbChain& ch) {
  typedef EbOrd<ET> OD;
  ch.ga = ch.ru = ch.rd = ch.rju = ch.rjd = ch.gc = ch.r = f2s(0.f);
  if (CHAIN) {
    // (the same rounded zs as the one the central edge's accumulations use:
    // all quantities of an object derive from ONE coordinate per edge)
    const float2 za = vfma(xc, f2s(-c.ks), f2s(c.cs[OD::A0]));
    const float2 zz = vmul(za, za);
    ch.ga = f2(mgb_ex2(-zz.x), mgb_ex2(-zz.y));
    if (WEIGHTED) ch.ga = vmul(ch.ga, wt);
    if (OD::KU >= 1) {
      const float2 e = vfma(za, f2s(-c.two_d), f2s(c.nd2));
      ch.ru = f2(mgb_ex2(e.x), mgb_ex2(e.y));
    }
    if (OD::KD >= 1) {
      const float2 e = vfma(za, f2s(c.two_d), f2s(c.nd2));
      ch.rd = f2(mgb_ex2(e.x), mgb_ex2(e.y));
    }
    if (OD::KU >= EB_SKIP) {
      const float2 e = vfma(za, f2s(-c.two_jd), f2s(c.nj2d2));
      ch.rju = f2(mgb_ex2(e.x), mgb_ex2(e.y));
    }
    if (OD::KD >= EB_SKIP) {
      const float2 e = vfma(za, f2s(c.two_jd), f2s(c.nj2d2));
      ch.rjd = f2(mgb_ex2(e.x), mgb_ex2(e.y));
    }
  }
}

// Contribution of NP pairs of objects (centred coordinates xc[], weights
// wt[]) to the accumulators of the ET edges of a group, in chunks of EB_CHUNK
// edges in the order of EbOrd.  The MUFU stage of the next chunk (of this pair
// or the next one) is issued in front of the FMA stage of the current chunk, so
// that a warp keeps both pipes busy by itself (with three warps per scheduler
// the bursts of MUFU would otherwise collide and stall the FMA pipe).  CHAIN
// selects the recurrence; otherwise every edge gets its own MUFU.EX2.
template <int ET, int NP, bool WEIGHTED, bool CHAIN>
__device__ __forceinline__ void edges_pairs(EdgeAcc* __restrict__ a,
                                            const GroupConst<ET>& c,
                                            const float2* xc,
                                            const float2* wt) {
  constexpr int CH = EB_CHUNK < ET ? EB_CHUNK : ET;
  constexpr int NCH = (ET + CH - 1) / CH;
  EbChain ch;
  float2 zs[2][CH], g[2][CH], t[2][CH];
  eb_anchor<ET, WEIGHTED, CHAIN>(c, xc[0], wt[0], ch);
  eb_stage_a<ET, CH, WEIGHTED, CHAIN>(c, xc[0], wt[0], 0, ch, zs[0], g[0],
                                      t[0]);
#pragma unroll
  for (int st = 0; st < NP * NCH; ++st) {
    const int h = st / NCH, n = st - h * NCH;
    const int nx = st + 1, hx = nx / NCH, nn = nx - hx * NCH;
    if (nx < NP * NCH) {
      if (nn == 0) eb_anchor<ET, WEIGHTED, CHAIN>(c, xc[hx], wt[hx], ch);
      eb_stage_a<ET, CH, WEIGHTED, CHAIN>(c, xc[hx], wt[hx], nn * CH, ch,
                                          zs[nx & 1], g[nx & 1], t[nx & 1]);
    }
    eb_stage_b<ET, CH, WEIGHTED>(a, n * CH, zs[st & 1], g[st & 1], t[st & 1],
                                 wt[h]);
  }
}

// ---------------------------------------------------------------------------
// Work decomposition.  The E = K+1 edges are split into NG groups of at most
// EB_ET_MAX consecutive edges; a warp belongs to one group and only carries the
// accumulators of that group's edges.  Every group has to see every object of
// the CTA, so a group with more edges gets proportionally more warps: all
// warps do the same amount of work.  Per-edge sums are independent, so no edge
// is ever evaluated twice; bin differences are taken after the CTA-level
// reduction.
// ---------------------------------------------------------------------------
#ifndef EB_ET_MAX
#define EB_ET_MAX 11
#endif
#ifndef EB_MIN_CTAS
#define EB_MIN_CTAS 2
#endif
#ifndef EB_TARGET_WARPS
#define EB_TARGET_WARPS 16
#endif
#ifndef EB_MAX_WARPS
#define EB_MAX_WARPS 8
#endif
#ifndef EB_WIDE_WARPS
#define EB_WIDE_WARPS 12
#endif
#ifndef EB_WIDE_CTAS
#define EB_WIDE_CTAS 1
#endif
// groups of more than EB_ET_NARROW edges run in EB_WIDE_CTAS CTA(s) per SM of
// EB_WIDE_WARPS warps (12 warps at up to 168 registers: measured 8% faster
// than 2 x 8 warps at 128), the narrow ones in CTAs of up to EB_MAX_WARPS warps
// at 80 registers
#ifndef EB_ET_NARROW
#define EB_ET_NARROW 4
#endif
#define EB_MAX_GROUPS ((MGB_MAX_BINS + 1 + EB_ET_MAX - 1) / EB_ET_MAX)
#define EB_MAX_EDGES (MGB_MAX_BINS + 1)
#define EB_MAX_CLUSTER 8

constexpr int eb_block_threads(int et_hi) {
  return et_hi > EB_ET_NARROW ? EB_WIDE_WARPS * 32 : EB_MAX_WARPS * 32;
}
constexpr int eb_min_ctas(int et_hi) {
  return et_hi > EB_ET_NARROW ? EB_WIDE_CTAS : EB_MIN_CTAS;
}
constexpr int eb_smem_warps() {
  return EB_WIDE_WARPS > EB_MAX_WARPS ? EB_WIDE_WARPS : EB_MAX_WARPS;
}

struct EbLayout {
  int ng;                       // number of edge groups
  int nwarps;                   // warps per CTA
  int uniform;                  // edges uniformly spaced (exactly, in fp64)
  int pad_;
  double width;                 // the common bin width when uniform
  int e0[EB_MAX_GROUPS + 1];    // first edge of group g (e0[ng] = E)
  int w0[EB_MAX_GROUPS + 1];    // first warp of group g (w0[ng] = nwarps)
};

// one tile = EB_TILE_F4 runs of 32 consecutive float4 (128 objects for 1):
// lane `lane` of the warp holds one float4 of every run.  (Two runs per
// iteration would halve the per-tile bookkeeping -- 10% of the instructions
// -- but cost 10% through register pressure: measured 723 vs 653 us.)
#ifndef EB_TILE_F4
#define EB_TILE_F4 1
#endif
#define EB_TILE_OBJS (128 * EB_TILE_F4)
struct EbTile {
  float4 x[EB_TILE_F4], w[EB_TILE_F4];
};

// fetch tile `tile` (padding objects beyond the end of the shard)
template <bool WEIGHTED>
__device__ __forceinline__ void eb_fetch(const MgbErfBinModel& md, int tile,
                                         int lane, EbTile& t) {
  const float4* __restrict__ x4 = reinterpret_cast<const float4*>(md.x);
  const float4* __restrict__ w4 = reinterpret_cast<const float4*>(md.w);
  const long long n4 = md.n >> 2;
#pragma unroll
  for (int q = 0; q < EB_TILE_F4; ++q) {
    const long long i = (long long)tile * (32 * EB_TILE_F4) + q * 32 + lane;
    if (i < n4) {
      t.x[q] = __ldg(x4 + i);
      if (WEIGHTED) t.w[q] = __ldg(w4 + i);
    } else {
      t.x[q] = make_float4(EB_SENTINEL_X, EB_SENTINEL_X, EB_SENTINEL_X,
                           EB_SENTINEL_X);
      if (WEIGHTED) t.w[q] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (i == n4) {  // ragged tail: the last n % 4 objects
        const long long b4 = n4 << 2;
        const int rem = (int)(md.n - b4);
        if (rem > 0) t.x[q].x = md.x[b4];
        if (rem > 1) t.x[q].y = md.x[b4 + 1];
        if (rem > 2) t.x[q].z = md.x[b4 + 2];
        if (WEIGHTED) {
          if (rem > 0) t.w[q].x = md.w[b4];
          if (rem > 1) t.w[q].y = md.w[b4 + 1];
          if (rem > 2) t.w[q].z = md.w[b4 + 2];
        }
      }
    }
  }
}

// Tile schedule of a warp (one tile = EB_TILE_OBJS consecutive objects).  A group of
// `wpg` warps per CTA sees every tile once: statically, warp wj of CTA c takes
// the tiles c + wj*G + j*G*wpg, j = 0, 1, ...  With the dynamic tail only the
// first ~90% of the tiles are dealt out that way; the rest is claimed two
// tiles at a time from a per-group counter in device memory, which evens out
// the few per cent by which the SMs differ in speed (measured: the first and
// the last CTA of a static split finish 3.3% of the loop time apart).  The
// price: which warp adds which tail tile varies from run to run, so the fp32
// partial sums -- not the reduction order, which stays fixed -- and with them
// the last bits of the results do.  Off = bit-reproducible.
struct EbSched {
  int ntiles;              // tiles in the shard (incl. the ragged last one)
  int per_iter;            // G * wpg: tiles the group consumes per iteration
  int static_end;          // tiles [0, static_end) are dealt out statically
  int first;               // this warp's first static tile
  int pool_lo, pool_hi;    // this CTA's pool of dynamic tiles
  unsigned* counter;       // claim counter of (group, pool) or NULL
};

__device__ __forceinline__ EbSched eb_make_sched(const MgbErfBinModel& md,
                                                 int wj, int wpg, bool dynamic,
                                                 unsigned* counter) {
  EbSched sc;
  sc.ntiles = (int)((md.n >> 2) / (32 * EB_TILE_F4)) + 1;
  sc.per_iter = (int)gridDim.x * wpg;
  sc.first = (int)blockIdx.x + wj * (int)gridDim.x;
  sc.static_end = sc.ntiles;
  sc.counter = nullptr;
  sc.pool_lo = sc.pool_hi = sc.ntiles;
  if (dynamic) {
    sc.static_end =
        (int)(((long long)sc.ntiles * 9 / 10) / sc.per_iter) * sc.per_iter;
    // the dynamic tiles are split into EB_NPOOL contiguous pools with a
    // counter each (one counter for all would see an atomic every 1.7 ns);
    // CTA c claims from pool c % EB_NPOOL
    const int pool = (int)(blockIdx.x % EB_NPOOL);
    const int dyn = sc.ntiles - sc.static_end;
    const int per = ((dyn + EB_NPOOL - 1) / EB_NPOOL + 1) & ~1;
    sc.pool_lo = min(sc.static_end + pool * per, sc.ntiles);
    sc.pool_hi = min(sc.pool_lo + per, sc.ntiles);
    sc.counter = counter + pool;
  }
  return sc;
}

// Accumulate the ET edges of a group over this warp's tiles (EbSched); leave
// the warp-reduced sums (T, O, G, Z per edge) in s_out[4*ET] (lane 0 writes).
// `first` holds the warp's first static tile, fetched by the caller (in the
// persistent fit before it waits for the parameters of the step).
template <int ET, bool WEIGHTED>
__device__ __forceinline__ void accumulate_group(
    const MgbErfBinModel& md, const GroupConst<ET>& c, float x0,
    const EbSched& sc, int lane, const EbTile& first, double* s_out) {
  EdgeAcc acc[ET];
#pragma unroll
  for (int e = 0; e < ET; ++e) {
#if EB_ACC2
    acc[e].T = f2s(0.f); acc[e].G = f2s(0.f); acc[e].Z = f2s(0.f);
#else
    acc[e].T = 0.f; acc[e].G = 0.f; acc[e].Z = 0.f;
#endif
    acc[e].O = 0.f;
  }
  // next tile after `t` (or -1): static stride first, then claims of two
  // consecutive tiles from the group's counter (warp-uniform)
  int second = -1;
  auto advance = [&](int t) -> int {
    const int t2 = t + sc.per_iter;
    if (t < sc.static_end && t2 < sc.static_end) return t2;
    if (sc.counter == nullptr) return -1;
    if (second >= 0) {
      const int t3 = second;
      second = -1;
      return t3;
    }
    unsigned cl = 0;
    if (lane == 0) cl = atomicAdd(sc.counter, 1u);
    cl = __shfl_sync(0xffffffffu, cl, 0);
    const unsigned t3 = (unsigned)sc.pool_lo + 2u * cl;
    if (t3 >= (unsigned)sc.pool_hi) return -1;
    second = (t3 + 1u < (unsigned)sc.pool_hi) ? (int)t3 + 1 : -1;
    return (int)t3;
  };
  int cur = sc.first;
  EbTile nxt = first;
  if (sc.first >= sc.static_end) {      // no static share: start with a claim
    cur = advance(sc.static_end);
    if (cur >= 0) eb_fetch<WEIGHTED>(md, cur, lane, nxt);
  }
  // The trip count is warp-uniform (tile indices do not depend on the lane)
  // because the body votes with __all_sync.  The next tile is fetched before
  // the current one is consumed (register double buffering hides the DRAM
  // latency, and that of the claim).
  while (cur >= 0) {
    const EbTile tl = nxt;
    cur = advance(cur);
    if (cur >= 0) eb_fetch<WEIGHTED>(md, cur, lane, nxt);
    float2 xcA[EB_TILE_F4], xcB[EB_TILE_F4];
    float cmin = 3.0e38f, cmax = -3.0e38f;
#pragma unroll
    for (int q = 0; q < EB_TILE_F4; ++q) {
      xcA[q] = vadd(f2(tl.x[q].x, tl.x[q].y), f2s(-x0));
      xcB[q] = vadd(f2(tl.x[q].z, tl.x[q].w), f2s(-x0));
      cmin = fminf(fminf(cmin, fminf(xcA[q].x, xcA[q].y)),
                   fminf(xcB[q].x, xcB[q].y));
      cmax = fmaxf(fmaxf(cmax, fmaxf(xcA[q].x, xcA[q].y)),
                   fmaxf(xcB[q].x, xcB[q].y));
    }
    // objects for which every edge of this group is saturated contribute
    // nothing to any difference between edges: skip them warp-uniformly
    const bool far = (fmaf(-c.ks, cmax, c.cs[0]) > EB_ZS_SKIP) ||
                     (fmaf(-c.ks, cmin, c.cs[ET - 1]) < -EB_ZS_SKIP);
    if (__all_sync(0xffffffffu, far)) {
      // A skipped lane still has to keep the hard counts consistent with the
      // other edge groups, for which the same objects may not be saturated.
      if (WEIGHTED) {
        // O = weight of objects at or above the edge: all objects of a lane
        // that is saturated on the high side (padding has weight 0)
        if (fmaf(-c.ks, cmax, c.cs[0]) > EB_ZS_SKIP) {
          float cnt = 0.f;
#pragma unroll
          for (int q = 0; q < EB_TILE_F4; ++q)
            cnt += (tl.w[q].x + tl.w[q].y) + (tl.w[q].z + tl.w[q].w);
#pragma unroll
          for (int e = 0; e < ET; ++e) acc[e].O += cnt;
        }
      } else {
        // O = number of objects below the edge: all objects of a lane
        // saturated on the low side (padding objects sit there too, in every
        // group alike, so they cancel in the differences between edges)
        if (fmaf(-c.ks, cmin, c.cs[ET - 1]) < -EB_ZS_SKIP) {
#pragma unroll
          for (int e = 0; e < ET; ++e)
            acc[e].O =
                __int_as_float(__float_as_int(acc[e].O) + 4 * EB_TILE_F4);
        }
      }
      continue;
    }
    // the recurrence needs every object of the tile within za_lim of the
    // central edge (in zs units); zs is monotone in xc
    const bool near = fmaxf(fabsf(fmaf(-c.ks, cmax, c.cs[ET / 2])),
                            fabsf(fmaf(-c.ks, cmin, c.cs[ET / 2]))) <= c.za_lim;
    const bool chain = __all_sync(0xffffffffu, near);
    // the 2*EB_TILE_F4 pairs of a lane go through the same code one after the
    // other (not inlined several times: the wide kernels would not fit the
    // instruction caches)
#pragma unroll 1
    for (int h = 0; h < 2 * EB_TILE_F4; ++h) {
      float2 xc1 = (h & 1) ? xcB[0] : xcA[0];
      float2 wt1 = (h & 1) ? f2(tl.w[0].z, tl.w[0].w) : f2(tl.w[0].x, tl.w[0].y);
#pragma unroll
      for (int q = 1; q < EB_TILE_F4; ++q) {
        if ((h >> 1) == q) {
          xc1 = (h & 1) ? xcB[q] : xcA[q];
          wt1 = (h & 1) ? f2(tl.w[q].z, tl.w[q].w) : f2(tl.w[q].x, tl.w[q].y);
        }
      }
      if (chain) edges_pairs<ET, 1, WEIGHTED, true>(acc, c, &xc1, &wt1);
      else edges_pairs<ET, 1, WEIGHTED, false>(acc, c, &xc1, &wt1);
    }
  }
#pragma unroll
  for (int e = 0; e < ET; ++e) {
#if EB_ACC2
    const float t = mgb_warp_sum(acc[e].T.x + acc[e].T.y);
#else
    const float t = mgb_warp_sum(acc[e].T);
#endif
    // unweighted O counts objects with z < 0: Phi = (N - O) + T, and N drops
    // out of every difference between edges
    const double o =
        WEIGHTED ? (double)mgb_warp_sum(acc[e].O)
                 : -(double)__reduce_add_sync(0xffffffffu,
                                              __float_as_int(acc[e].O));
#if EB_ACC2
    const float g = mgb_warp_sum(acc[e].G.x + acc[e].G.y);
    const float z = mgb_warp_sum(acc[e].Z.x + acc[e].Z.y);
#else
    const float g = mgb_warp_sum(acc[e].G);
    const float z = mgb_warp_sum(acc[e].Z);
#endif
    if (lane == 0) {
      s_out[4 * e + 0] = (double)t;
      s_out[4 * e + 1] = o;
      s_out[4 * e + 2] = (double)g;
      s_out[4 * e + 3] = (double)z;
    }
  }
}

// per-step shared state of a CTA
struct EbShared {
  float cs[EB_MAX_EDGES];              // zs-scaled centred edges
  double warp[eb_smem_warps()][4 * EB_ET_MAX];
  double edge[EB_MAX_EDGES][4];
  double clu[EB_MAX_CLUSTER * 3 * MGB_MAX_BINS];  // leader: pushed partials
  double tot[3 * MGB_MAX_BINS];
  double stage[MGB_MAX_WORLD * 3 * MGB_MAX_BINS];
  float pv[MGB_MAX_PARAMS];
  unsigned tok;
  int is_last;
  int go;
};

template <int ET, bool WEIGHTED>
__device__ __forceinline__ void run_group(const MgbErfBinModel& md,
                                          EbShared& sh, double ks_d,
                                          double width, bool uniform,
                                          int e_first, const EbSched& sc,
                                          int warp, int lane,
                                          const EbTile& first) {
  GroupConst<ET> c;
  c.ks = (float)ks_d;
#pragma unroll
  for (int e = 0; e < ET; ++e) c.cs[e] = sh.cs[e_first + e];
  const double d = ks_d * width;
  c.two_d = (float)(2.0 * d);
  c.nd2 = (float)(-d * d);
  c.kap = (float)exp2(-2.0 * d * d);
  c.two_jd = (float)(2.0 * EB_SKIP * d);
  c.nj2d2 = (float)(-(double)(EB_SKIP * EB_SKIP) * d * d);
  c.kapj = (float)exp2(-2.0 * EB_SKIP * d * d);
  c.za_lim = 0.f;
  if (uniform && ET > 1) {
    // largest exponent: the skip ratio, 2 J |za| d + J^2 d^2
    const double lim = (EB_CHAIN_EXP_MAX - EB_SKIP * EB_SKIP * d * d) /
                       (2.0 * EB_SKIP * d);
    c.za_lim = (float)(lim < (double)EB_ZA_MAX ? lim : (double)EB_ZA_MAX);
  }
  accumulate_group<ET, WEIGHTED>(md, c, md.x0, sc, lane, first,
                                 sh.warp[warp]);
}

template <int ET_HI, bool WEIGHTED>
__global__ void __launch_bounds__(eb_block_threads(ET_HI), eb_min_ctas(ET_HI))
erfbin_step_kernel(const __grid_constant__ MgbErfBinModel md,
                   const __grid_constant__ MgbStepArgs a,
                   const __grid_constant__ MgbPeerTable pt,
                   const __grid_constant__ EbLayout lay,
                   double* __restrict__ partials,
                   unsigned* __restrict__ counters) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int nthreads = blockDim.x;
  const int b = blockIdx.y;
  const int G = gridDim.x;
  const int K = md.nbins;
  const int E = K + 1;
  const int MP = 3 * K;  // doubles per partial
  const unsigned crank = eb_cluster_ctarank();
  const unsigned csize = eb_cluster_nctarank();
  const int NC = G / (int)csize;          // clusters = global partial rows
  const int cid = blockIdx.x / (int)csize;

  __shared__ EbShared sh;

  // this warp's edge group
  int grp = 0;
  while (grp + 1 < lay.ng && warp >= lay.w0[grp + 1]) ++grp;
  const int e_first = lay.e0[grp];
  const int et = lay.e0[grp + 1] - e_first;
  const int wpg = lay.w0[grp + 1] - lay.w0[grp];
  const int wj = warp - lay.w0[grp];
  // dynamic tail (unless MGB_STEP_STATIC_SPLIT): claim counters are words
  // EB_DYN_WORD + group*EB_NPOOL + pool of the counter block (batch == 1)
  const EbSched sched = eb_make_sched(
      md, wj, wpg, (a.flags & MGB_STEP_STATIC_SPLIT) == 0 && gridDim.y == 1,
      counters + EB_DYN_WORD + grp * EB_NPOOL);

  // Programmatic dependent launch: this grid may have been scheduled while
  // the previous step's last CTA was still in its epilogue (launch latency and
  // CTA scheduling overlap with it).  Everything it wrote -- parameters,
  // optimiser state, the ticket counter -- is visible after this wait.
  asm volatile("griddepcontrol.wait;" ::: "memory");

  // gated launch: uniform early exit on every rank (no ticket, no exchange)
  if (a.gate != nullptr && __ldg(a.gate) == 0.0f) return;

  const unsigned long long t_entry = a.trace ? mgb_globaltimer_ns() : 0ull;

  // pre-launched host-driven step: wait for the host to publish this step's
  // parameters.  Only CTA 0 talks to the host: a system-scope load of pinned
  // host memory is a PCIe round trip that is not merged with anybody else's
  // (296 CTAs polling the same word cost ~270 us per step, measured), so
  // CTA 0 polls the flag, fetches the parameters and relays both through
  // device memory (spare words of the counter block), where the other CTAs
  // wait at L2 latency.  Relay token = wait_value (go) or wait_value with
  // bit 31 set (cancelled / timed out); the host zeroes it before a run.
  bool have_params = false;
  if (a.wait_flag != nullptr) {
    unsigned* relay_tok = counters + EB_RELAY_WORD;
    float* relay_pv = reinterpret_cast<float*>(counters + EB_RELAY_WORD + 8);
    const unsigned long long t0 = mgb_globaltimer_ns();
    const unsigned long long limit =
        (unsigned long long)a.wait_timeout_ms * 1000000ull;
    if (blockIdx.x == 0) {
      if (tid == 0) {
        unsigned tok = a.wait_value | 0x80000000u;
        for (unsigned it = 0;; ++it) {
          const unsigned f = mgb_ld_sys_u32(a.wait_flag);
          if (f & 0x80000000u) break;                       // cancelled
          if ((f & 0x7fffffffu) >= a.wait_value) { tok = a.wait_value; break; }
          if ((it & 15u) == 15u && mgb_globaltimer_ns() - t0 > limit) break;
          __nanosleep(EB_HOST_POLL_NS);
        }
        __threadfence_system();   // acquire: parameters written before flag
        sh.tok = tok;
      }
      __syncthreads();
      if (!(sh.tok & 0x80000000u) && tid < md.nparams) {
        const float v = __uint_as_float(mgb_ld_sys_u32(
            reinterpret_cast<const unsigned*>(a.params) + tid));
        sh.pv[tid] = v;
        relay_pv[tid] = v;
      }
      __threadfence();
      __syncthreads();
      if (tid == 0)
        asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(relay_tok),
                     "r"(sh.tok) : "memory");
    } else {
      if (tid == 0) {
        unsigned tok = a.wait_value | 0x80000000u;
        for (unsigned it = 0;; ++it) {
          unsigned f;
          asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(f)
                       : "l"(relay_tok) : "memory");
          if ((f & 0x7fffffffu) == a.wait_value) { tok = f; break; }
          if ((it & 63u) == 63u && mgb_globaltimer_ns() - t0 > 2 * limit) break;
        }
        __threadfence();
        sh.tok = tok;
      }
      __syncthreads();
      if (!(sh.tok & 0x80000000u) && tid < md.nparams)
        sh.pv[tid] = __ldcv(relay_pv + tid);
    }
    __syncthreads();
    if (sh.tok & 0x80000000u) return;
    have_params = true;
  }

  // every CTA of the cluster is running from here on (its shared memory may be
  // written by the others once this barrier has completed; the wait sits in
  // front of the first remote store)
  eb_cluster_arrive();
  bool cluster_open = false;

  const int nsteps = (a.mode >= MGB_MODE_ADAM && a.nsteps > 1) ? a.nsteps : 1;
  const int step0 = (nsteps > 1) ? __ldcg(a.opt_step) : 0;

  for (int s = 0; s < nsteps; ++s) {
    // ---- first tile of the step: independent of the parameters, so it is in
    // flight while this CTA waits for them ------------------------------------
    EbTile first;
#pragma unroll
    for (int q = 0; q < EB_TILE_F4; ++q)
      first.w[q] = make_float4(1.f, 1.f, 1.f, 1.f);
    if (sched.first < sched.static_end)
      eb_fetch<WEIGHTED>(md, sched.first, lane, first);

    if (s > 0) {
      // persistent fit: wait until the elected CTA of the previous step has
      // published the updated parameters (it bumps *opt_step last)
      if (tid == 0) {
        int go = 1;
        const unsigned long long t0 = mgb_globaltimer_ns();
        const unsigned long long limit =
            2000000ull * (unsigned long long)(pt.world > 1
                                                  ? pt.spin_timeout_ms
                                                  : 20000u);
        for (unsigned it = 0;; ++it) {
          if (eb_ld_acquire_gpu_s32(a.opt_step) - step0 >= s) break;
          if ((it & 63u) == 63u) {
            if (__ldcv(counters + EB_ABORT_WORD) != 0u ||
                mgb_globaltimer_ns() - t0 > limit) {
              go = 0;
              break;
            }
          }
        }
        sh.go = go;
      }
      __syncthreads();
      if (!sh.go) break;
    }
    const unsigned long long t_start =
        a.trace ? (s == 0 ? t_entry : mgb_globaltimer_ns()) : 0ull;

    // ---- parameters: one coalesced read per CTA (the vector may live in
    // pinned host memory: zero-copy evaluation) --------------------------------
    if (!have_params) {
      const float* pv = (a.mode >= MGB_MODE_ADAM)
                            ? a.opt_p
                            : (a.params + (size_t)b * md.nparams);
      if (tid < md.nparams)
        sh.pv[tid] = (nsteps > 1) ? __ldcg(pv + tid) : pv[tid];
      __syncthreads();
    }
    // ---- parameters -> scaled link constants ----------------------------------
    const float p_mu = sh.pv[md.mu_idx];
    const float p_sg = sh.pv[md.sig_idx];
    const double sigma_d =
        md.sigma_link == 1 ? exp10((double)p_sg) : (double)p_sg;
    // Scaled, *centred* coordinates: with xc = x - x0 (exact in fp32 for data
    // within a factor two of x0) and ks = kappa/(sqrt2 sigma),
    //     zs_e = cs_e - ks*xc,   cs_e = (edge_e - p_mu - x0) * ks
    // cs_e is formed in fp64 so that zs never subtracts two O(10) numbers.
    const double ks_d = 1.2011224087864498 * 0.7071067811865476 / sigma_d;
    if (tid < E) {
      sh.cs[tid] =
          (float)((md.edges[tid] - ((double)p_mu + (double)md.x0)) * ks_d);
    }
    __syncthreads();

    // ---- main loop: this warp's edge group -------------------------------------
    if (et == ET_HI) {
      run_group<ET_HI, WEIGHTED>(md, sh, ks_d, lay.width, lay.uniform != 0,
                                 e_first, sched, warp, lane, first);
    } else {
      run_group<(ET_HI > 1 ? ET_HI - 1 : 1), WEIGHTED>(
          md, sh, ks_d, lay.width, lay.uniform != 0, e_first, sched, warp,
          lane, first);
    }
    // the object loop of this CTA is done: let the next launch's grid be
    // scheduled
    if (nsteps == 1)
      asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (a.trace && tid == 0) {
      unsigned long long* tr = reinterpret_cast<unsigned long long*>(a.trace) +
                               (size_t)s * (8 + 2 * (size_t)G);
      tr[8 + blockIdx.x] = t_start;
      tr[8 + G + blockIdx.x] = mgb_globaltimer_ns();
    }
    __syncthreads();

    // ---- CTA reduction in fp64: per-edge totals, then bin differences ---------
    for (int j = tid; j < 4 * E; j += nthreads) {
      const int e = j >> 2, q = j & 3;
      int g2 = 0;
      while (g2 + 1 < lay.ng && e >= lay.e0[g2 + 1]) ++g2;
      double sum = 0.0;
      for (int w = lay.w0[g2]; w < lay.w0[g2 + 1]; ++w)
        sum += sh.warp[w][4 * (e - lay.e0[g2]) + q];
      sh.edge[e][q] = sum;
    }
    __syncthreads();
    // the first remote store of this launch needs every CTA of the cluster to
    // have started
    if (!cluster_open) {
      eb_cluster_wait();
      cluster_open = true;
    }
    {
      const unsigned dst = eb_mapa(sh.clu, 0);
      for (int j = tid; j < MP; j += nthreads) {
        const int q = j / K, k = j - q * K;
        double val;
        if (q == 0) {
          val = (sh.edge[k + 1][1] - sh.edge[k][1]) +
                (sh.edge[k + 1][0] - sh.edge[k][0]);
        } else if (q == 1) {
          val = sh.edge[k + 1][2] - sh.edge[k][2];
        } else {
          val = sh.edge[k + 1][3] - sh.edge[k][3];
        }
        eb_st_cluster_f64(dst + (unsigned)((crank * MP + j) * 8), val);
      }
    }
    eb_cluster_arrive();
    eb_cluster_wait();
    bool last = false;
    if (crank == 0) {
      // leader: add the cluster's partials in rank order
      for (int j = tid; j < MP; j += nthreads) {
        double sum = 0.0;
        for (unsigned r = 0; r < csize; ++r) sum += sh.clu[r * MP + j];
        partials[((size_t)b * NC + cid) * MP + j] = sum;
      }
      // ---- last-leader election ------------------------------------------------
      __threadfence();
      __syncthreads();
      if (tid == 0) {
        const unsigned ticket = atomicAdd(&counters[b], 1u);
        sh.is_last = (ticket == (unsigned)(NC - 1));
      }
      __syncthreads();
      last = sh.is_last != 0;
    }
    if (!last) continue;   // next step (or exit): wait for the elected CTA
    __threadfence();
    unsigned long long tr1 = 0, tr2 = 0, tr3 = 0;
    if (a.trace && tid == 0) tr1 = mgb_globaltimer_ns();
    // optimiser state of parameter `lane` (warp 0): loaded now, consumed after
    // the reductions -- keeps three dependent L2 round trips out of the
    // epilogue
    int o_step = 0;
    float o_p = 0.f, o_u = 0.f, o_m = 0.f, o_v = 0.f;
    if (a.mode >= MGB_MODE_ADAM && warp == 0) {
      o_step = __ldcg(a.opt_step);
      if (lane < md.nparams) {
        o_p = __ldcg(a.opt_p + lane);
        if (a.mode == MGB_MODE_ADAM) {
          o_u = __ldcg(a.opt_u + lane);
          o_m = __ldcg(a.opt_m + lane);
          o_v = __ldcg(a.opt_v + lane);
        }
      }
    }

    // ---- fixed-order grid reduction --------------------------------------------
    mgb_grid_reduce_f64(partials + (size_t)b * NC * MP, NC, MP, sh.tot, tid,
                        nthreads);
    if (tid == 0) counters[b] = 0u;  // ready for the next step / launch
    if (sched.counter != nullptr && tid < lay.ng * EB_NPOOL)
      counters[EB_DYN_WORD + tid] = 0u;
    __syncthreads();
    if (a.trace && tid == 0) tr2 = mgb_globaltimer_ns();

    // ---- cross-GPU ordered all-reduce over NVLink ------------------------------
    if (a.reduce && pt.world > 1)
      mgb_ll_allreduce_f64(pt, sh.tot, sh.stage, 3 * K);
    if (a.trace && tid == 0) tr3 = mgb_globaltimer_ns();

    // ---- epilogue (warp 0; lane k owns bin k) ----------------------------------
    if (warp == 0) {
      const double sig = sigma_d;
      const double inv = 0.7071067811865476 / sig;           // 1/(sqrt2 sigma)
      const double rsqrtpi = 0.5641895835477563;             // 1/sqrt(pi)
      const double dsdp = md.sigma_link == 1 ? 2.302585092994046 * sig : 1.0;
      const int P = md.nparams;
      double s_k = 0.0, dmu_k = 0.0, dsg_k = 0.0;
      if (lane < K) {
        const double nrm = (double)md.norm[lane];
        s_k = nrm * sh.tot[lane];
        dmu_k = -nrm * inv * rsqrtpi * sh.tot[K + lane];
        dsg_k =
            -nrm * (rsqrtpi * EB_INV_KAPPA / sig) * dsdp * sh.tot[2 * K + lane];
        if (a.out_sumstats) a.out_sumstats[(size_t)b * K + lane] = (float)s_k;
        if (a.out_jac && a.mode >= MGB_MODE_SUMSTATS_JAC) {
          float* jr = a.out_jac + ((size_t)b * K + lane) * P;
          for (int i = 0; i < P; ++i) {
            double v = 0.0;
            if (i == md.mu_idx) v += dmu_k;
            if (i == md.sig_idx) v += dsg_k;
            jr[i] = (float)v;
          }
        }
      }
      if (a.mode >= MGB_MODE_LOSS_GRAD) {
        double l_k = 0.0, dl_k = 0.0;  // loss term and dL/ds_k
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
        const double gmu = mgb_warp_sum(dl_k * dmu_k);
        const double gsg = mgb_warp_sum(dl_k * dsg_k);
        // lane i owns parameter i from here on (the warp sums left loss / gmu
        // / gsg in every lane): the optimiser-state loads of all parameters
        // are then in flight together
        const int i = lane;
        const bool mine = i < P;
        float grad = 0.f;
        if (mine) {
          double v = 0.0;
          if (i == md.mu_idx) v += gmu;
          if (i == md.sig_idx) v += gsg;
          grad = (float)v;
        }
        if (lane == 0 && a.out_loss) a.out_loss[b] = (float)loss;
        if (mine && a.out_grad) a.out_grad[(size_t)b * P + i] = grad;
        if (a.mode >= MGB_MODE_ADAM) {
          // ---- optimiser update (replicated, deterministic on every GPU) ----
          const int step = o_step;              // updates applied so far
          const int row = step - a.step_base;   // row of this step
          if (lane == 0 && a.loss_hist && row >= 0 && row < a.traj_rows)
            a.loss_hist[row] = (float)loss;
          if (mine)
            mgb_opt_update_param_pre(a, i, P, step, row, grad, a.bound_kind[i],
                                     a.bound_lo[i], a.bound_hi[i], o_p, o_u,
                                     o_m, o_v);
          __syncwarp();                   // every lane has written its state
          if (lane == 0) {
            if (nsteps > 1) {
              // a failed exchange poisons the state: stop the fit everywhere
              if (pt.world > 1 &&
                  mgb_ld_sys_u32(mgb_ctrl(pt, pt.rank) + MGB_CTRL_ERROR) != 0u)
                counters[EB_ABORT_WORD] = 1u;
              __threadfence();
              eb_st_release_gpu_s32(a.opt_step, step + 1);
            } else {
              *a.opt_step = step + 1;
            }
          }
        }
      }
    }
    if (a.trace && tid == 0) {
      unsigned long long* tr = reinterpret_cast<unsigned long long*>(a.trace) +
                               (size_t)s * (8 + 2 * (size_t)G);
      tr[0] = t_start;
      tr[1] = tr1;
      tr[2] = tr2;
      tr[3] = tr3;
      tr[4] = mgb_globaltimer_ns();
      tr[5] = (unsigned long long)blockIdx.x;
    }
    if (nsteps > 1) __syncthreads();
  }
}

typedef void (*eb_kernel_t)(const MgbErfBinModel, const MgbStepArgs,
                            const MgbPeerTable, const EbLayout, double*,
                            unsigned*);

template <bool W>
eb_kernel_t pick_kernel(int et_hi) {
  switch (et_hi) {
    case 1: return erfbin_step_kernel<1, W>;
    case 2: return erfbin_step_kernel<2, W>;
    case 3: return erfbin_step_kernel<3, W>;
    case 4: return erfbin_step_kernel<4, W>;
#if EB_ET_MAX >= 5
    case 5: return erfbin_step_kernel<5, W>;
#endif
#if EB_ET_MAX >= 6
    case 6: return erfbin_step_kernel<6, W>;
#endif
#if EB_ET_MAX >= 7
    case 7: return erfbin_step_kernel<7, W>;
#endif
#if EB_ET_MAX >= 8
    case 8: return erfbin_step_kernel<8, W>;
#endif
#if EB_ET_MAX >= 9
    case 9: return erfbin_step_kernel<9, W>;
#endif
#if EB_ET_MAX >= 10
    case 10: return erfbin_step_kernel<10, W>;
#endif
#if EB_ET_MAX >= 11
    case 11: return erfbin_step_kernel<11, W>;
#endif
  }
  return nullptr;
}

// Edge groups and warps per group for K bins (see the comment above EbLayout).
EbLayout make_layout(const MgbErfBinModel& md, int* et_hi_out) {
  EbLayout L;
  memset(&L, 0, sizeof(L));
  const int nbins = md.nbins;
  const int E = nbins + 1;
  L.ng = (E + EB_ET_MAX - 1) / EB_ET_MAX;
  const int lo = E / L.ng, extra = E % L.ng;  // `extra` groups get lo+1 edges
  const int et_hi = lo + (extra ? 1 : 0);
  const int maxw = eb_block_threads(et_hi) / 32;
  // Warps per group: proportional to the group's edge count while one warp
  // per edge fits in a CTA (perfect balance), else the same number of warps
  // for every group; a single group takes the whole CTA.
  const int even = maxw / L.ng > 0 ? maxw / L.ng : 1;
  int e = 0, w = 0;
  for (int g = 0; g < L.ng; ++g) {
    const int et = lo + (g < extra ? 1 : 0);
    L.e0[g] = e;
    L.w0[g] = w;
    e += et;
    w += (L.ng == 1) ? maxw : ((E <= maxw) ? et : even);
  }
  L.e0[L.ng] = e;
  L.w0[L.ng] = w;
  L.nwarps = w;
  // exactly uniform spacing (to fp64 rounding), as produced by linspace
  const double width = (md.edges[nbins] - md.edges[0]) / nbins;
  bool uni = width > 0.0;
  for (int k = 0; k <= nbins && uni; ++k) {
    const double want = md.edges[0] + k * width;
    const double tol = 8.0 * 2.220446049250313e-16 *
                       (fabs(md.edges[0]) + fabs(md.edges[nbins]) + width);
    uni = fabs(md.edges[k] - want) <= tol;
  }
  const char* env = getenv("MULTIGRAD_B200_CHAIN");
  if (env && env[0] == '0') uni = false;
  L.uniform = uni ? 1 : 0;
  L.width = width;
  *et_hi_out = et_hi;
  return L;
}

int g_sm_cache[64] = {0};

int sm_count() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (dev >= 0 && dev < 64 && g_sm_cache[dev]) return g_sm_cache[dev];
  int sms = 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) !=
      cudaSuccess)
    return 0;
  if (dev >= 0 && dev < 64) g_sm_cache[dev] = sms;
  return sms;
}

// CTAs per SM so that about EB_TARGET_WARPS warps are resident (the wide
// kernels run EB_WIDE_CTAS CTA(s) per SM)
int grid_for(int nbins) {
  MgbErfBinModel tmp;
  memset(&tmp, 0, sizeof(tmp));
  tmp.nbins = nbins;
  int et_hi = 0;
  const int nwarps = make_layout(tmp, &et_hi).nwarps;
  int per_sm = EB_TARGET_WARPS / nwarps;
  const int cap = et_hi > EB_ET_NARROW ? EB_WIDE_CTAS : EB_MIN_CTAS;
  if (per_sm < 1) per_sm = 1;
  if (per_sm > cap) per_sm = cap;
  return sm_count() * per_sm;
}

// Thread-block cluster size wanted: MULTIGRAD_B200_CLUSTER (1, 2, 4 or 8;
// default 4), reduced until it divides the grid.
int cluster_for(int grid) {
  int c = 4;
  const char* env = getenv("MULTIGRAD_B200_CLUSTER");
  if (env && env[0]) c = atoi(env);
  if (c != 1 && c != 2 && c != 4 && c != 8) c = 4;
  while (c > 1 && grid % c) c >>= 1;
  return c;
}

// Largest cluster size <= the wanted one with which a whole wave of the grid
// (grid CTAs) is resident at once: a cluster is placed as a unit, so too
// large a cluster leaves SMs idle (measured: clusters of 4 with two
// 128-register CTAs per SM run the grid in 1.5 waves).  The answer is cached
// per kernel, device and launch shape.
struct EbClusterChoice {
  const void* fn;
  int dev, grid, block, want, csize;
};
EbClusterChoice g_cluster_cache[32];
int g_cluster_cached = 0;
std::mutex g_cluster_mutex;

int pick_cluster(eb_kernel_t k, int grid, int block, int* out) {
  int dev = 0;
  MGB_CUDA_CHECK(cudaGetDevice(&dev));
  const int want = cluster_for(grid);
  std::lock_guard<std::mutex> lock(g_cluster_mutex);
  for (int i = 0; i < g_cluster_cached; ++i) {
    const EbClusterChoice& c = g_cluster_cache[i];
    if (c.fn == (const void*)k && c.dev == dev && c.grid == grid &&
        c.block == block && c.want == want) {
      *out = c.csize;
      return MGB_OK;
    }
  }
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((unsigned)grid, 1, 1);
  cfg.blockDim = dim3((unsigned)block, 1, 1);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int csize = want;
  for (; csize > 1; csize >>= 1) {
    attr[0].val.clusterDim.x = (unsigned)csize;
    int max_clusters = 0;
    const cudaError_t e =
        cudaOccupancyMaxActiveClusters(&max_clusters, k, &cfg);
    if (e == cudaSuccess && max_clusters * csize >= grid) break;
    if (e != cudaSuccess) (void)cudaGetLastError();
  }
  if (g_cluster_cached < 32) {
    EbClusterChoice& c = g_cluster_cache[g_cluster_cached++];
    c.fn = (const void*)k;
    c.dev = dev;
    c.grid = grid;
    c.block = block;
    c.want = want;
    c.csize = csize;
  }
  *out = csize;
  return MGB_OK;
}

}  // namespace

// size of the counter block in front of the partial sums: one ticket word per
// batch entry, at least EB_COUNTER_WORDS words (relay, abort and claim words)
static uint64_t eb_counter_bytes(int batch) {
  const uint64_t b = ((uint64_t)batch * 4ull + 255ull) & ~255ull;
  return b < 4ull * EB_COUNTER_WORDS ? 4ull * EB_COUNTER_WORDS : b;
}

// cluster size the fused step will be launched with (diagnostics)
extern "C" int mgb_erfbin_cluster(int nbins, int weighted) {
  if (nbins < 1 || nbins > MGB_MAX_BINS) return 0;
  MgbErfBinModel tmp;
  memset(&tmp, 0, sizeof(tmp));
  tmp.nbins = nbins;
  int et_hi = 0;
  const EbLayout lay = make_layout(tmp, &et_hi);
  eb_kernel_t k =
      weighted ? pick_kernel<true>(et_hi) : pick_kernel<false>(et_hi);
  int csize = 0;
  if (!k ||
      pick_cluster(k, grid_for(nbins), lay.nwarps * 32, &csize) != MGB_OK)
    return 0;
  return csize;
}

extern "C" int mgb_erfbin_grid(int nbins) {
  if (nbins < 1 || nbins > MGB_MAX_BINS) return 0;
  return grid_for(nbins);
}

extern "C" uint64_t mgb_erfbin_scratch_bytes(int nbins, int batch) {
  if (nbins < 1 || nbins > MGB_MAX_BINS || batch < 1) return 0;
  const int grid = grid_for(nbins);
  return eb_counter_bytes(batch) + (uint64_t)batch * grid * 3ull * nbins * 8ull;
}

extern "C" int mgb_erfbin_step(const MgbErfBinModel* model,
                               const MgbStepArgs* args, const MgbPeerTable* pt,
                               void* scratch, uint64_t scratch_bytes,
                               void* stream) {
  MGB_REQUIRE(model && args && scratch, MGB_EINVAL,
              "mgb_erfbin_step: null argument");
  const MgbErfBinModel& md = *model;
  MgbStepArgs a = *args;
  {
    static const bool deterministic = [] {
      const char* e = getenv("MULTIGRAD_B200_DETERMINISTIC");
      return e && e[0] == '1';
    }();
    if (deterministic) a.flags |= MGB_STEP_STATIC_SPLIT;
    a.flags |= md.step_flags;
  }
  MGB_REQUIRE(md.nbins >= 1 && md.nbins <= MGB_MAX_BINS, MGB_ETOOBIG,
              "mgb_erfbin_step: nbins=%d outside [1,%d]", md.nbins,
              MGB_MAX_BINS);
  MGB_REQUIRE(md.nparams >= 1 && md.nparams <= MGB_MAX_PARAMS, MGB_ETOOBIG,
              "mgb_erfbin_step: nparams=%d outside [1,%d]", md.nparams,
              MGB_MAX_PARAMS);
  MGB_REQUIRE(md.mu_idx >= 0 && md.mu_idx < md.nparams && md.sig_idx >= 0 &&
                  md.sig_idx < md.nparams,
              MGB_EINVAL, "mgb_erfbin_step: link indices out of range");
  MGB_REQUIRE(md.n >= 0 && (md.n == 0 || md.x != nullptr), MGB_EINVAL,
              "mgb_erfbin_step: x is null");
  MGB_REQUIRE(((uintptr_t)md.x & 15u) == 0 && ((uintptr_t)md.w & 15u) == 0,
              MGB_EINVAL, "mgb_erfbin_step: x / w must be 16-byte aligned");
  for (int k = 0; k < md.nbins; ++k)
    MGB_REQUIRE(md.edges[k + 1] > md.edges[k], MGB_EINVAL,
                "mgb_erfbin_step: bin edges must be strictly increasing");
  MGB_REQUIRE(a.batch >= 1 && a.batch <= MGB_MAX_BATCH, MGB_ETOOBIG,
              "mgb_erfbin_step: batch=%d outside [1,%d]", a.batch,
              MGB_MAX_BATCH);
  MGB_REQUIRE(a.mode >= MGB_MODE_SUMSTATS && a.mode <= MGB_MODE_GD, MGB_EINVAL,
              "mgb_erfbin_step: bad mode %d", a.mode);
  if (a.mode >= MGB_MODE_ADAM) {
    MGB_REQUIRE(a.batch == 1, MGB_EINVAL,
                "mgb_erfbin_step: optimiser modes need batch == 1");
    MGB_REQUIRE(a.opt_p && a.opt_step, MGB_EINVAL,
                "mgb_erfbin_step: optimiser state missing");
    if (a.mode == MGB_MODE_ADAM)
      MGB_REQUIRE(a.opt_u && a.opt_m && a.opt_v, MGB_EINVAL,
                  "mgb_erfbin_step: Adam state missing");
  } else {
    MGB_REQUIRE(a.params != nullptr, MGB_EINVAL,
                "mgb_erfbin_step: params is null");
  }
  const bool persistent = a.nsteps > 1;
  MGB_REQUIRE(!persistent || (a.mode >= MGB_MODE_ADAM && !a.gate &&
                              !a.wait_flag),
              MGB_EINVAL,
              "mgb_erfbin_step: nsteps > 1 needs an optimiser mode and "
              "neither gate nor host hand-shake");
  MgbPeerTable ptv;
  if (pt) {
    ptv = *pt;
  } else {
    memset(&ptv, 0, sizeof(ptv));
    ptv.world = 1;
  }
  const bool do_reduce = a.reduce && ptv.world > 1;
  if (do_reduce) {
    MGB_REQUIRE(a.batch == 1, MGB_EINVAL,
                "mgb_erfbin_step: in-kernel GPU reduction needs batch == 1");
    MGB_REQUIRE(ptv.world <= MGB_MAX_WORLD && 6u * md.nbins <= ptv.ll_units,
                MGB_ETOOBIG, "mgb_erfbin_step: peer table too small");
  }
  int et_hi = 0;
  const EbLayout lay = make_layout(md, &et_hi);
  const int grid = grid_for(md.nbins);
  MGB_REQUIRE(grid > 0, MGB_ESTATE, "mgb_erfbin_step: no CUDA device");
  MGB_REQUIRE(a.wait_flag == nullptr ||
                  (a.batch == 1 && a.mode <= MGB_MODE_LOSS_GRAD),
              MGB_EINVAL,
              "mgb_erfbin_step: the host hand-shake needs batch == 1 and an "
              "evaluation mode");
  const uint64_t need = mgb_erfbin_scratch_bytes(md.nbins, a.batch);
  MGB_REQUIRE(scratch_bytes >= need, MGB_ETOOBIG,
              "mgb_erfbin_step: scratch %llu < %llu bytes",
              (unsigned long long)scratch_bytes, (unsigned long long)need);
  const uint64_t coff = eb_counter_bytes(a.batch);
  unsigned* counters = reinterpret_cast<unsigned*>(scratch);
  double* partials =
      reinterpret_cast<double*>(reinterpret_cast<char*>(scratch) + coff);
  eb_kernel_t k = md.w ? pick_kernel<true>(et_hi) : pick_kernel<false>(et_hi);
  MGB_REQUIRE(k != nullptr, MGB_ESTATE,
              "mgb_erfbin_step: no kernel for %d edges per group", et_hi);
  dim3 g((unsigned)grid, (unsigned)a.batch, 1), blk(lay.nwarps * 32, 1, 1);
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = g;
  cfg.blockDim = blk;
  cfg.dynamicSmemBytes = 0;
  cfg.stream = (cudaStream_t)stream;
  int csize = 1;
  {
    const int rc = pick_cluster(k, grid, (int)blk.x, &csize);
    if (rc != MGB_OK) return rc;
  }
  cudaLaunchAttribute attr[3];
  int na = 0;
  attr[na].id = cudaLaunchAttributeClusterDimension;
  attr[na].val.clusterDim.x = (unsigned)csize;
  attr[na].val.clusterDim.y = 1;
  attr[na].val.clusterDim.z = 1;
  ++na;
  if (persistent) {
    // every CTA of the grid has to be resident at once (pick_cluster checked
    // that it can be); the cooperative attribute makes the driver refuse the
    // launch instead of deadlocking if it cannot
    attr[na].id = cudaLaunchAttributeCooperative;
    attr[na].val.cooperative = 1;
    ++na;
  } else {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  MGB_CUDA_CHECK(
      cudaLaunchKernelEx(&cfg, k, md, a, ptv, lay, partials, counters));
  return MGB_OK;
}

// ---------------------------------------------------------------------------
// mgb_erfbin_step_static -- the same fused step for callers that have no way
// to obtain scratch memory.
//
// mgb_erfbin_step needs a scratch block (the per-CTA partial sums of the grid
// reduction plus the ticket counters) that the caller allocates.  Some callers
// cannot do that: a plug-in that lives inside another runtime and is only
// handed its input and output buffers (custom-call style embedding), a process
// whose whole device memory already belongs to somebody else's pool, or code
// running in a phase where calling the CUDA allocator is forbidden or would
// stall the device (cudaMalloc synchronises).  For them the scratch lives in
// the library image itself: a file-scope __device__ array of
// MGB_STATIC_SCRATCH_BYTES (1 MiB) that every GPU which loads the library
// keeps resident until the process exits.  This entry point calls no
// cudaMalloc / cudaFree and no allocating library routine.
//
// Consequences the caller has to know about:
//  * Size limit.  The request is refused with MGB_ETOOBIG when
//    mgb_erfbin_scratch_bytes(nbins, batch) exceeds MGB_STATIC_SCRATCH_BYTES
//    (batch = 1 always fits; large batched scans do not).
//  * One array per GPU, shared by every call.  It still holds whatever the
//    previous call left there; the ticket counters, the only part that must
//    start at zero, are cleared on the caller's stream before each launch.
//    Two calls that overlap on the same GPU (different streams, or different
//    host threads) use the same partial-sum slots and race: serialise them.
//  * Set-up.  The address of the array is looked up once per device, with
//    that device current, by the first call made on it
//    (cudaGetSymbolAddress: inside host code the array's name is only a
//    host-side handle, not something a kernel can dereference).  Make that
//    first call before any stream capture is started.
// ---------------------------------------------------------------------------
static __device__ __align__(256) unsigned char
    g_eb_static_scratch[MGB_STATIC_SCRATCH_BYTES];

namespace {
void* g_static_scratch_ptr[64] = {nullptr};
std::mutex g_static_scratch_mutex;

// Device address of g_eb_static_scratch on the current device (resolved once
// per device).
int static_scratch_address(void** out) {
  int dev = -1;
  MGB_CUDA_CHECK(cudaGetDevice(&dev));
  MGB_REQUIRE(dev >= 0 && dev < 64, MGB_ESTATE,
              "mgb_erfbin_step_static: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lock(g_static_scratch_mutex);
  if (g_static_scratch_ptr[dev] == nullptr) {
    void* p = nullptr;
    MGB_CUDA_CHECK(cudaGetSymbolAddress(&p, g_eb_static_scratch));
    MGB_REQUIRE(p != nullptr, MGB_ESTATE,
                "mgb_erfbin_step_static: cudaGetSymbolAddress returned null");
    g_static_scratch_ptr[dev] = p;
  }
  *out = g_static_scratch_ptr[dev];
  return MGB_OK;
}
}  // namespace

extern "C" int mgb_erfbin_step_static(const MgbErfBinModel* model,
                                      const MgbStepArgs* args,
                                      const MgbPeerTable* pt, void* stream) {
  MGB_REQUIRE(model && args, MGB_EINVAL,
              "mgb_erfbin_step_static: null argument");
  MGB_REQUIRE(model->nbins >= 1 && model->nbins <= MGB_MAX_BINS, MGB_ETOOBIG,
              "mgb_erfbin_step_static: nbins=%d outside [1,%d]", model->nbins,
              MGB_MAX_BINS);
  MGB_REQUIRE(args->batch >= 1 && args->batch <= MGB_MAX_BATCH, MGB_ETOOBIG,
              "mgb_erfbin_step_static: batch=%d outside [1,%d]", args->batch,
              MGB_MAX_BATCH);
  const uint64_t need = mgb_erfbin_scratch_bytes(model->nbins, args->batch);
  MGB_REQUIRE(need > 0, MGB_ESTATE,
              "mgb_erfbin_step_static: no CUDA device");
  MGB_REQUIRE(need <= (uint64_t)MGB_STATIC_SCRATCH_BYTES, MGB_ETOOBIG,
              "mgb_erfbin_step_static: nbins=%d, batch=%d needs %llu bytes of "
              "scratch but the built-in array holds %llu; use "
              "mgb_erfbin_step with caller-provided scratch or a smaller batch",
              model->nbins, args->batch, (unsigned long long)need,
              (unsigned long long)MGB_STATIC_SCRATCH_BYTES);
  void* scratch = nullptr;
  const int rc = static_scratch_address(&scratch);
  if (rc != MGB_OK) return rc;
  // the ticket counters must start at zero; the partial sums are overwritten
  const uint64_t counter_bytes = eb_counter_bytes(args->batch);
  MGB_CUDA_CHECK(
      cudaMemsetAsync(scratch, 0, counter_bytes, (cudaStream_t)stream));
  return mgb_erfbin_step(model, args, pt, scratch, MGB_STATIC_SCRATCH_BYTES,
                         stream);
}

// ---------------------------------------------------------------------------
// Host-driven evaluation in one native call: pinned params -> device, fused
// loss+gradient launch, loss+gradient -> pinned host buffer, stream sync.
// This is the per-evaluation body of the L-BFGS-B line search (reference
// multigrad/bfgs.py:72-77 + multigrad.py:508-538) and of any host-side
// optimiser loop; doing it in C keeps the host cost at a few microseconds.
// `params_host` (nparams floats) and `out_host` (1 + nparams floats: loss,
// gradient) should be pinned; `d_params` / `d_out` are device staging buffers
// of the same sizes.
// ---------------------------------------------------------------------------
extern "C" int mgb_erfbin_eval_host(const MgbErfBinModel* model,
                                    const MgbPeerTable* pt,
                                    const float* params_host, float* out_host,
                                    float* d_params, float* d_out,
                                    void* scratch, uint64_t scratch_bytes,
                                    void* stream) {
  MGB_REQUIRE(model && params_host && out_host && d_params && d_out,
              MGB_EINVAL, "mgb_erfbin_eval_host: null argument");
  cudaStream_t st = (cudaStream_t)stream;
  const int P = model->nparams;
  MGB_REQUIRE(P >= 1 && P <= MGB_MAX_PARAMS, MGB_ETOOBIG,
              "mgb_erfbin_eval_host: nparams=%d outside [1,%d]", P,
              MGB_MAX_PARAMS);
  MGB_CUDA_CHECK(cudaMemcpyAsync(d_params, params_host, sizeof(float) * P,
                                 cudaMemcpyHostToDevice, st));
  MgbStepArgs a;
  memset(&a, 0, sizeof(a));
  a.mode = MGB_MODE_LOSS_GRAD;
  a.batch = 1;
  a.reduce = (pt && pt->world > 1) ? 1 : 0;
  a.params = d_params;
  a.out_loss = d_out;
  a.out_grad = d_out + 1;
  const int rc = mgb_erfbin_step(model, &a, pt, scratch, scratch_bytes, stream);
  if (rc != MGB_OK) return rc;
  MGB_CUDA_CHECK(cudaMemcpyAsync(out_host, d_out, sizeof(float) * (1 + P),
                                 cudaMemcpyDeviceToHost, st));
  MGB_CUDA_CHECK(cudaStreamSynchronize(st));
  return MGB_OK;
}

// Zero-copy flavour of the above: the kernel reads the parameters straight
// from the caller's pinned host buffer and writes (loss, grad) straight into
// the pinned result buffer (both are device-accessible under unified
// addressing), so one evaluation is ONE launch and no copy-engine operation;
// the host waits by polling the result words, which it primed with a NaN
// pattern no computation produces, instead of paying a stream synchronise.
// Per evaluation P*4 bytes cross PCIe host->device and (1+P)*4 device->host.
extern "C" int mgb_erfbin_eval_host_zc(const MgbErfBinModel* model,
                                       const MgbPeerTable* pt,
                                       const float* params_host,
                                       float* out_host, void* scratch,
                                       uint64_t scratch_bytes, void* stream,
                                       uint32_t timeout_ms) {
  MGB_REQUIRE(model && params_host && out_host, MGB_EINVAL,
              "mgb_erfbin_eval_host_zc: null argument");
  const int P = model->nparams;
  MGB_REQUIRE(P >= 1 && P <= MGB_MAX_PARAMS, MGB_ETOOBIG,
              "mgb_erfbin_eval_host_zc: nparams=%d outside [1,%d]", P,
              MGB_MAX_PARAMS);
  const uint32_t kSentinel = 0x7fc0deadu;
  volatile uint32_t* o = reinterpret_cast<volatile uint32_t*>(out_host);
  for (int i = 0; i <= P; ++i) o[i] = kSentinel;
  __sync_synchronize();
  MgbStepArgs a;
  memset(&a, 0, sizeof(a));
  a.mode = MGB_MODE_LOSS_GRAD;
  a.batch = 1;
  a.reduce = (pt && pt->world > 1) ? 1 : 0;
  a.params = params_host;
  a.out_loss = out_host;
  a.out_grad = out_host + 1;
  const int rc = mgb_erfbin_step(model, &a, pt, scratch, scratch_bytes, stream);
  if (rc != MGB_OK) return rc;
  struct timespec t0;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  for (unsigned spin = 0;; ++spin) {
    bool done = true;
    for (int i = 0; i <= P; ++i) done = done && (o[i] != kSentinel);
    if (done) break;
    if ((spin & 0x3ffu) == 0x3ffu) {
      struct timespec t1;
      clock_gettime(CLOCK_MONOTONIC, &t1);
      const double ms = (t1.tv_sec - t0.tv_sec) * 1e3 +
                        (t1.tv_nsec - t0.tv_nsec) * 1e-6;
      if (ms > (double)timeout_ms) {
        // distinguish "kernel failed" from "kernel slow"
        const cudaError_t e = cudaStreamQuery((cudaStream_t)stream);
        if (e != cudaSuccess && e != cudaErrorNotReady) MGB_CUDA_CHECK(e);
        mgb_set_error("mgb_erfbin_eval_host_zc: no result after %u ms",
                      timeout_ms);
        return MGB_ESTATE;
      }
    }
#if defined(__x86_64__)
    __builtin_ia32_pause();
#endif
  }
  __sync_synchronize();
  return MGB_OK;
}

// Host-driven Adam in one native call: the optimiser state lives on the HOST
// (as in the reference, where rank 0 owns it: multigrad/adam.py:52-68) and
// every step ships the parameters to the GPU and the loss + gradient back
// (zero-copy evaluation above), but without a Python round trip per step.
// Every rank runs the same loop on bit-identical gradients, so the replicas
// stay identical.  params_host / out_host must be pinned; m, v, traj,
// loss_hist are ordinary host arrays (traj: (nsteps+1) x P with row 0 = the
// initial parameters, loss_hist: nsteps; either may be NULL).
extern "C" int mgb_erfbin_adam_host(
    const MgbErfBinModel* model, const MgbPeerTable* pt, float* params_host,
    float* out_host, float* m, float* v, int32_t step0, int32_t nsteps,
    float lr, float b1, float b2, float eps, float* traj, float* loss_hist,
    void* scratch, uint64_t scratch_bytes, void* stream,
    uint32_t timeout_ms, void* flush_buf, uint64_t flush_bytes,
    double* timed_seconds) {
  MGB_REQUIRE(model && params_host && out_host && m && v && nsteps >= 0,
              MGB_EINVAL, "mgb_erfbin_adam_host: bad arguments");
  if (timed_seconds) *timed_seconds = 0.0;
  const int P = model->nparams;
  MGB_REQUIRE(P >= 1 && P <= MGB_MAX_PARAMS, MGB_ETOOBIG,
              "mgb_erfbin_adam_host: nparams=%d outside [1,%d]", P,
              MGB_MAX_PARAMS);
  if (traj)
    for (int i = 0; i < P; ++i) traj[i] = params_host[i];
  if (nsteps == 0) return MGB_OK;
  // Pipelining: the launch for step k+1 is issued while step k computes, and
  // its grid waits (on the GPU, polling a word of out_host over PCIe) until
  // the host has published the parameters of step k+1.  The launch latency
  // therefore never sits between two steps.  Layout of out_host:
  //   [0] loss, [1..P] gradient, [MGB_HOST_FLAG_WORD] (as uint32, in its own
  //   cache line) the hand-shake word.
  const uint32_t kSentinel = 0x7fc0deadu;
  volatile uint32_t* o = reinterpret_cast<volatile uint32_t*>(out_host);
  volatile uint32_t* flag = o + MGB_HOST_FLAG_WORD;
  const char* env = getenv("MULTIGRAD_B200_HOST_PIPELINE");
  const bool pipelined = !(env && env[0] == '0');
  *flag = 0u;
  MgbStepArgs a;
  memset(&a, 0, sizeof(a));
  a.mode = MGB_MODE_LOSS_GRAD;
  a.batch = 1;
  a.reduce = (pt && pt->world > 1) ? 1 : 0;
  a.params = params_host;
  a.out_loss = out_host;
  a.out_grad = out_host + 1;
  a.wait_flag = const_cast<const uint32_t*>(
      reinterpret_cast<uint32_t*>(out_host) + MGB_HOST_FLAG_WORD);
  a.wait_timeout_ms = timeout_ms;
  int rc = MGB_OK;
  // the device-side relay of the hand-shake starts from zero
  MGB_REQUIRE(scratch != nullptr && scratch_bytes >= 4 * EB_COUNTER_WORDS, MGB_EINVAL,
              "mgb_erfbin_adam_host: scratch missing");
  MGB_CUDA_CHECK(cudaMemsetAsync(
      reinterpret_cast<char*>(scratch) + 4 * EB_RELAY_WORD, 0,
      256 - 4 * EB_RELAY_WORD, (cudaStream_t)stream));
  // Benchmarking aid: with flush_buf the L2 is overwritten (a fill larger
  // than the cache) in front of every step's grid, and the time the host
  // waits for that fill is kept out of *timed_seconds, which then sums, per
  // step, the interval [parameters published, loss+grad read back].
  cudaStream_t st = (cudaStream_t)stream;
  cudaEvent_t ev[2] = {nullptr, nullptr};
  if (flush_buf) {
    MGB_CUDA_CHECK(cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming));
    MGB_CUDA_CHECK(cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming));
  }
#define EB_FLUSH(idx)                                                         \
  do {                                                                        \
    if (flush_buf && rc == MGB_OK) {                                          \
      cudaError_t e_ = cudaMemsetAsync(flush_buf, (idx) + 1, flush_bytes, st); \
      if (e_ == cudaSuccess) e_ = cudaEventRecord(ev[(idx) & 1], st);         \
      if (e_ != cudaSuccess) {                                                \
        mgb_set_error("mgb_erfbin_adam_host: flush failed: %s",               \
                      cudaGetErrorString(e_));                                \
        rc = (int)e_;                                                         \
      }                                                                       \
    }                                                                         \
  } while (0)
  a.wait_value = 1u;
  if (pipelined) {
    EB_FLUSH(0);
    if (rc == MGB_OK)
      rc = mgb_erfbin_step(model, &a, pt, scratch, scratch_bytes, stream);
  }
  double timed = 0.0;
  for (int k = 0; k < nsteps && rc == MGB_OK; ++k) {
    if (!pipelined) EB_FLUSH(k);
    if (flush_buf && rc == MGB_OK) {
      const cudaError_t e_ = cudaEventSynchronize(ev[k & 1]);
      if (e_ != cudaSuccess) { rc = (int)e_; break; }
    }
    if (rc != MGB_OK) break;
    struct timespec t0;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    // publish step k: results primed, parameters in place, then the flag
    for (int i = 0; i <= P; ++i) o[i] = kSentinel;
    __sync_synchronize();
    *flag = (uint32_t)(k + 1);
    // queue the next step's grid behind this one (it waits for flag k+2)
    a.wait_value = (uint32_t)(pipelined ? k + 2 : k + 1);
    if (pipelined) EB_FLUSH(k + 1);
    if (rc != MGB_OK) break;
    rc = mgb_erfbin_step(model, &a, pt, scratch, scratch_bytes, stream);
    if (rc != MGB_OK) break;
    for (unsigned spin = 0;; ++spin) {
      bool done = true;
      for (int i = 0; i <= P; ++i) done = done && (o[i] != kSentinel);
      if (done) break;
      if ((spin & 0x3ffu) == 0x3ffu) {
        struct timespec t1;
        clock_gettime(CLOCK_MONOTONIC, &t1);
        const double ms = (t1.tv_sec - t0.tv_sec) * 1e3 +
                          (t1.tv_nsec - t0.tv_nsec) * 1e-6;
        if (ms > (double)timeout_ms) {
          mgb_set_error("mgb_erfbin_adam_host: no result for step %d after "
                        "%u ms", k, timeout_ms);
          rc = MGB_ESTATE;
          break;
        }
      }
#if defined(__x86_64__)
      __builtin_ia32_pause();
#endif
    }
    if (rc != MGB_OK) break;
    __sync_synchronize();
    const int step = step0 + k;
    const float c1 = 1.0f - powf(b1, (float)(step + 1));
    const float c2 = 1.0f - powf(b2, (float)(step + 1));
    if (loss_hist) loss_hist[k] = out_host[0];
    for (int i = 0; i < P; ++i) {
      const float g = out_host[1 + i];
      const float mi = (1.0f - b1) * g + b1 * m[i];
      const float vi = (1.0f - b2) * g * g + b2 * v[i];
      m[i] = mi;
      v[i] = vi;
      const float pn = params_host[i] - lr * (mi / c1) / (sqrtf(vi / c2) + eps);
      params_host[i] = pn;
      if (traj) traj[(size_t)(k + 1) * P + i] = pn;
    }
    struct timespec t2;
    clock_gettime(CLOCK_MONOTONIC, &t2);
    timed += (t2.tv_sec - t0.tv_sec) + (t2.tv_nsec - t0.tv_nsec) * 1e-9;
  }
#undef EB_FLUSH
  if (timed_seconds) *timed_seconds = timed;
  // release the one grid that is still waiting, and wait for it to leave
  __sync_synchronize();
  *flag = 0x80000000u;
  const cudaError_t e = cudaStreamSynchronize((cudaStream_t)stream);
  if (ev[0]) cudaEventDestroy(ev[0]);
  if (ev[1]) cudaEventDestroy(ev[1]);
  if (rc == MGB_OK && e != cudaSuccess) MGB_CUDA_CHECK(e);
  return rc;
}

// Host-driven evaluations with the launch latency hidden, for drivers that
// call back into Python between evaluations (scipy's L-BFGS-B): a *session*
// keeps exactly one grid pre-launched and waiting on the GPU for the next
// parameters (same hand-shake as mgb_erfbin_adam_host).
//   op 0  begin: reset the hand-shake, queue the first waiting grid
//   op 1  eval : publish params_host, queue the next waiting grid, poll the
//                result of this one into out_host[0..P]
//   op 2  end  : cancel the waiting grid and drain the stream
// *seq is the session's evaluation counter (owned by the caller, 0 at begin).
extern "C" int mgb_erfbin_host_session(
    const MgbErfBinModel* model, const MgbPeerTable* pt, float* params_host,
    float* out_host, void* scratch, uint64_t scratch_bytes, void* stream,
    uint32_t timeout_ms, int32_t op, uint32_t* seq) {
  MGB_REQUIRE(model && params_host && out_host && seq && scratch &&
                  scratch_bytes >= 4 * EB_COUNTER_WORDS,
              MGB_EINVAL, "mgb_erfbin_host_session: bad arguments");
  const int P = model->nparams;
  MGB_REQUIRE(P >= 1 && P <= MGB_MAX_PARAMS, MGB_ETOOBIG,
              "mgb_erfbin_host_session: nparams=%d outside [1,%d]", P,
              MGB_MAX_PARAMS);
  const uint32_t kSentinel = 0x7fc0deadu;
  volatile uint32_t* o = reinterpret_cast<volatile uint32_t*>(out_host);
  volatile uint32_t* flag = o + MGB_HOST_FLAG_WORD;
  cudaStream_t st = (cudaStream_t)stream;
  if (op == 2) {
    __sync_synchronize();
    *flag = 0x80000000u;
    MGB_CUDA_CHECK(cudaStreamSynchronize(st));
    return MGB_OK;
  }
  MgbStepArgs a;
  memset(&a, 0, sizeof(a));
  a.mode = MGB_MODE_LOSS_GRAD;
  a.batch = 1;
  a.reduce = (pt && pt->world > 1) ? 1 : 0;
  a.params = params_host;
  a.out_loss = out_host;
  a.out_grad = out_host + 1;
  a.wait_flag = const_cast<const uint32_t*>(
      reinterpret_cast<uint32_t*>(out_host) + MGB_HOST_FLAG_WORD);
  a.wait_timeout_ms = timeout_ms;
  if (op == 0) {
    *flag = 0u;
    *seq = 0u;
    MGB_CUDA_CHECK(cudaMemsetAsync(
        reinterpret_cast<char*>(scratch) + 4 * EB_RELAY_WORD, 0,
        256 - 4 * EB_RELAY_WORD, st));
    a.wait_value = 1u;
    return mgb_erfbin_step(model, &a, pt, scratch, scratch_bytes, stream);
  }
  MGB_REQUIRE(op == 1, MGB_EINVAL, "mgb_erfbin_host_session: bad op %d", op);
  for (int i = 0; i <= P; ++i) o[i] = kSentinel;
  __sync_synchronize();
  *seq += 1u;
  *flag = *seq;
  a.wait_value = *seq + 1u;
  int rc = mgb_erfbin_step(model, &a, pt, scratch, scratch_bytes, stream);
  if (rc != MGB_OK) return rc;
  struct timespec t0;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  for (unsigned spin = 0;; ++spin) {
    bool done = true;
    for (int i = 0; i <= P; ++i) done = done && (o[i] != kSentinel);
    if (done) break;
    if ((spin & 0x3ffu) == 0x3ffu) {
      struct timespec t1;
      clock_gettime(CLOCK_MONOTONIC, &t1);
      const double ms = (t1.tv_sec - t0.tv_sec) * 1e3 +
                        (t1.tv_nsec - t0.tv_nsec) * 1e-6;
      if (ms > (double)timeout_ms) {
        mgb_set_error("mgb_erfbin_host_session: no result after %u ms",
                      timeout_ms);
        return MGB_ESTATE;
      }
    }
#if defined(__x86_64__)
    __builtin_ia32_pause();
#endif
  }
  __sync_synchronize();
  return MGB_OK;
}
