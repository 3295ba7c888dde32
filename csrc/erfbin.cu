// K1/K2/K3: fused erf-soft-binned statistic + Jacobian + cross-CTA and
// cross-GPU reduction + loss / chain rule / optimiser epilogue.  sm_100a.
//
// What one launch does (see include/multigrad_b200.h for the contract):
//   main loop   every thread streams objects (float4 loads), and for each of
//               the E = K+1 bin edges evaluates, on packed fp32x2 lanes
//               (FFMA2/FADD2/FMUL2, two objects per instruction),
//                   zs  = kappa*(edge - mu)/(sqrt2*sigma)      (kappa^2 = log2 e)
//                   g   = 2^(-zs^2) = exp(-z^2)               (MUFU.EX2)
//                   q   = 1/2 erfc(|z|) = g * t*P(t)/2, t = 1/(1+|z|/2) (MUFU.RCP)
//               and accumulates per edge  T = sum -+q (tail part of the CDF),
//               O = sum [z>=0] (hard count), G = sum g, Z = sum zs*g.
//               Phi = O + T exactly, so bin sums never subtract numbers near 1.
//   CTA reduce  warp shuffles in fp32, cross-warp and everything after in fp64;
//               per-CTA partial (D_k, dG_k, dZ_k) goes to `scratch`.
//   grid reduce the last CTA (atomic ticket) sums the partials in a fixed
//               order -> run-to-run deterministic.
//   GPU reduce  the same CTA pushes the 3K doubles to every peer with the LL
//               protocol over NVLink and sums in rank order (mgb_common.cuh).
//   epilogue    sumstats, Jacobian, loss, dL/ds, gradient, bound-transform
//               chain rule and the Adam / GD update, trajectory + loss rows.
//
// Reference counterparts: OnePointModel._vjp (multigrad/multigrad.py:508-538),
// calc_smf_bin (tests/smf_example/smf_grad_descent.py:38-48), _adam_optimizer
// (multigrad/adam.py:52-68), transforms (adam.py:192-239), simple_grad_descent
// (multigrad/util.py:100-114).
#include <math.h>
#include "mgb_common.cuh"

#define EB_BLOCK 256
#define EB_WARPS (EB_BLOCK / 32)
#define EB_CTAS_PER_SM 2
#define EB_NBMAX 16
#define EB_PAD_EDGE 3.0e18f   // pad edge: zs^2 overflows -> g = 0, zs*g = 0
#define EB_SENTINEL_X 1.0e30f // padding object: contributes exactly zero

namespace {

struct F2 {
  float2 v;
};
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
// t = 1/(1 + 0.5 |z|) = 1/(1 + (0.5/kappa) |zs|)
#define EB_HSCALE (0.5f / EB_KAPPA)
#define EB_ZS_SKIP 12.4f  // |zs| beyond which g and q flush to zero

struct EdgeAcc {
  float2 T, O, G, Z;
};

// Contribution of a pair of objects (scaled coordinate us, weight wt) to the
// accumulators of one edge with scaled position cs.
template <bool WEIGHTED>
__device__ __forceinline__ void edge_pair(EdgeAcc& a, float cs, float2 us,
                                          float2 wt) {
  const float2 zs = vadd(f2s(cs), f2(-us.x, -us.y));
  const float2 zz = vmul(zs, zs);
  float2 g = f2(mgb_ex2(-zz.x), mgb_ex2(-zz.y));
  const float2 d = vfma(f2(fabsf(zs.x), fabsf(zs.y)), f2s(EB_HSCALE),
                        f2s(1.0f));
  const float2 t = f2(mgb_rcp(d.x), mgb_rcp(d.y));
  // Horner; coefficients pre-halved so that p = erfcx/2 / t
  float2 p = f2s(0.5f * MGB_ERFCX_C8);
  p = vfma(p, t, f2s(0.5f * MGB_ERFCX_C7));
  p = vfma(p, t, f2s(0.5f * MGB_ERFCX_C6));
  p = vfma(p, t, f2s(0.5f * MGB_ERFCX_C5));
  p = vfma(p, t, f2s(0.5f * MGB_ERFCX_C4));
  p = vfma(p, t, f2s(0.5f * MGB_ERFCX_C3));
  p = vfma(p, t, f2s(0.5f * MGB_ERFCX_C2));
  p = vfma(p, t, f2s(0.5f * MGB_ERFCX_C1));
  p = vfma(p, t, f2s(0.5f * MGB_ERFCX_C0));
  if (WEIGHTED) g = vmul(g, wt);
  const float2 q = vmul(vmul(p, t), g);  // (w) * 1/2 erfc(|z|)
  // ts = +q for z < 0, -q for z >= 0  (set the sign bit from ~sign(zs))
  float2 ts;
  ts.x = __uint_as_float(__float_as_uint(q.x) |
                         (~__float_as_uint(zs.x) & 0x80000000u));
  ts.y = __uint_as_float(__float_as_uint(q.y) |
                         (~__float_as_uint(zs.y) & 0x80000000u));
  float2 o;
  if (WEIGHTED) {
    o = f2(zs.x >= 0.0f ? wt.x : 0.0f, zs.y >= 0.0f ? wt.y : 0.0f);
  } else {
    o = f2(zs.x >= 0.0f ? 1.0f : 0.0f, zs.y >= 0.0f ? 1.0f : 0.0f);
  }
  a.T = vadd(a.T, ts);
  a.O = vadd(a.O, o);
  a.G = vadd(a.G, g);
  a.Z = vfma(zs, g, a.Z);
}

__device__ __forceinline__ float bound_inverse(int kind, float lo, float hi,
                                               float u) {
  if (kind == 1) {
    const float mid = 0.5f * (hi + lo);
    const float sc = (hi - lo) * 0.3183098861837907f;
    return mid + sc * atanf(u / sc);
  } else if (kind == 2) {
    return 0.5f * (2.0f * lo + u + sqrtf(u * u + 4.0f));
  } else if (kind == 3) {
    return 0.5f * (2.0f * hi + u - sqrtf(u * u + 4.0f));
  }
  return u;
}
// d inverse(u)/du evaluated at u = x
__device__ __forceinline__ float bound_jac(int kind, float lo, float hi,
                                           float x) {
  if (kind == 1) {
    const float sc = (hi - lo) * 0.3183098861837907f;
    const float r = x / sc;
    return 1.0f / (1.0f + r * r);
  } else if (kind == 2) {
    return 0.5f * (1.0f + x / sqrtf(x * x + 4.0f));
  } else if (kind == 3) {
    return 0.5f * (1.0f - x / sqrtf(x * x + 4.0f));
  }
  return 1.0f;
}

template <int NB, bool WEIGHTED>
__global__ void __launch_bounds__(EB_BLOCK, EB_CTAS_PER_SM)
erfbin_step_kernel(const MgbErfBinModel md, const MgbStepArgs a,
                   const MgbPeerTable pt, double* __restrict__ partials,
                   unsigned* __restrict__ counters, const int npass) {
  constexpr int E = NB + 1;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int b = blockIdx.y;
  const int G = gridDim.x;
  const int K = md.nbins;
  const int KP = npass * NB;   // padded bin count
  const int MP = 3 * KP;       // doubles per CTA partial

  __shared__ double s_warp[EB_WARPS][4 * (EB_NBMAX + 1)];
  __shared__ double s_tot[3 * MGB_MAX_BINS];
  __shared__ double s_stage[MGB_MAX_WORLD * 3 * MGB_MAX_BINS];
  __shared__ int s_is_last;

  // gated launch: uniform early exit on every rank (no ticket, no exchange)
  if (a.gate != nullptr && __ldg(a.gate) == 0.0f) return;

  // ---- parameters -> scaled link constants --------------------------------
  const float* pv =
      (a.mode >= MGB_MODE_ADAM) ? a.opt_p : (a.params + (size_t)b * md.nparams);
  const float p_mu = pv[md.mu_idx];
  const float p_sg = pv[md.sig_idx];
  const float sigma = md.sigma_link == 1 ? exp10f(p_sg) : p_sg;
  // ks = kappa / (sqrt(2) sigma):  us = (x + p_mu) * ks,  cs_e = edge_e * ks
  const float ks = EB_KAPPA * 0.7071067811865476f / sigma;
  const float us0 = p_mu * ks;

  const float4* __restrict__ x4 = reinterpret_cast<const float4*>(md.x);
  const float4* __restrict__ w4 = reinterpret_cast<const float4*>(md.w);
  const long long n4 = md.n >> 2;
  const long long gstride = (long long)G * EB_BLOCK;

  for (int pass = 0; pass < npass; ++pass) {
    float cs[E];
#pragma unroll
    for (int e = 0; e < E; ++e) {
      const int ge = pass * NB + e;
      cs[e] = (ge <= K ? md.edges[ge] : EB_PAD_EDGE) * ks;
    }
    EdgeAcc acc[E];
#pragma unroll
    for (int e = 0; e < E; ++e) {
      acc[e].T = f2s(0.f); acc[e].O = f2s(0.f);
      acc[e].G = f2s(0.f); acc[e].Z = f2s(0.f);
    }
    const float cs_lo = cs[0];
    const float cs_hi = (pass * NB + NB <= K) ? cs[NB] : md.edges[K] * ks;

    // The trip count is warp-uniform (i0 does not depend on the lane) because
    // the body votes with __all_sync; lanes past the end process padding
    // objects that contribute exactly zero.
    for (long long i0 = (long long)blockIdx.x * EB_BLOCK + warp * 32;
         i0 < n4 + 1; i0 += gstride) {
      const long long i = i0 + lane;
      float4 xv, wv = make_float4(1.f, 1.f, 1.f, 1.f);
      if (i < n4) {
        xv = __ldg(x4 + i);
        if (WEIGHTED) wv = __ldg(w4 + i);
      } else {
        xv = make_float4(EB_SENTINEL_X, EB_SENTINEL_X, EB_SENTINEL_X,
                         EB_SENTINEL_X);
        if (WEIGHTED) wv = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i == n4) {  // ragged tail: the last n % 4 objects
          const long long base = n4 << 2;
          const int rem = (int)(md.n - base);
          if (rem > 0) xv.x = md.x[base];
          if (rem > 1) xv.y = md.x[base + 1];
          if (rem > 2) xv.z = md.x[base + 2];
          if (WEIGHTED) {
            if (rem > 0) wv.x = md.w[base];
            if (rem > 1) wv.y = md.w[base + 1];
            if (rem > 2) wv.z = md.w[base + 2];
          }
        }
      }
      const float2 usA = vfma(f2(xv.x, xv.y), f2s(ks), f2s(us0));
      const float2 usB = vfma(f2(xv.z, xv.w), f2s(ks), f2s(us0));
      // objects whose every edge is saturated contribute nothing to any
      // *difference* between edges: skip them warp-uniformly
      const float umin = fminf(fminf(usA.x, usA.y), fminf(usB.x, usB.y));
      const float umax = fmaxf(fmaxf(usA.x, usA.y), fmaxf(usB.x, usB.y));
      const bool far = (cs_lo - umax > EB_ZS_SKIP) || (cs_hi - umin < -EB_ZS_SKIP);
      if (__all_sync(0xffffffffu, far)) continue;
#pragma unroll
      for (int e = 0; e < E; ++e) {
        edge_pair<WEIGHTED>(acc[e], cs[e], usA, f2(wv.x, wv.y));
        edge_pair<WEIGHTED>(acc[e], cs[e], usB, f2(wv.z, wv.w));
      }
    }

    // ---- CTA reduction: fp32 inside the warp, fp64 across warps -----------
#pragma unroll
    for (int e = 0; e < E; ++e) {
      float t = mgb_warp_sum(acc[e].T.x + acc[e].T.y);
      float o = mgb_warp_sum(acc[e].O.x + acc[e].O.y);
      float g = mgb_warp_sum(acc[e].G.x + acc[e].G.y);
      float z = mgb_warp_sum(acc[e].Z.x + acc[e].Z.y);
      if (lane == 0) {
        s_warp[warp][4 * e + 0] = (double)t;
        s_warp[warp][4 * e + 1] = (double)o;
        s_warp[warp][4 * e + 2] = (double)g;
        s_warp[warp][4 * e + 3] = (double)z;
      }
    }
    __syncthreads();
    if (tid < 4 * E) {
      double s = 0.0;
#pragma unroll
      for (int w = 0; w < EB_WARPS; ++w) s += s_warp[w][tid];
      s_warp[0][tid] = s;  // only thread `tid` touches column `tid` here
    }
    __syncthreads();
    if (tid < 3 * NB) {
      const int q = tid / NB, k = tid - q * NB;
      double val;
      if (q == 0) {
        val = (s_warp[0][4 * (k + 1) + 1] - s_warp[0][4 * k + 1]) +
              (s_warp[0][4 * (k + 1) + 0] - s_warp[0][4 * k + 0]);
      } else if (q == 1) {
        val = s_warp[0][4 * (k + 1) + 2] - s_warp[0][4 * k + 2];
      } else {
        val = s_warp[0][4 * (k + 1) + 3] - s_warp[0][4 * k + 3];
      }
      partials[((size_t)b * G + blockIdx.x) * MP + q * KP + pass * NB + k] = val;
    }
    __syncthreads();
  }

  // ---- last-CTA election ---------------------------------------------------
  __threadfence();
  if (tid == 0) {
    const unsigned ticket = atomicAdd(&counters[b], 1u);
    s_is_last = (ticket == (unsigned)(G - 1));
  }
  __syncthreads();
  if (!s_is_last) return;
  __threadfence();

  // ---- fixed-order grid reduction: 8 lanes per value -----------------------
  {
    const int sub = tid & 7, grp = tid >> 3;
    for (int j0 = 0; j0 < MP; j0 += EB_BLOCK / 8) {
      const int j = j0 + grp;
      double accd = 0.0;
      if (j < MP) {
        const double* src = partials + (size_t)b * G * MP + j;
        for (int c = sub; c < G; c += 8) accd += __ldcg(src + (size_t)c * MP);
      }
      accd += __shfl_xor_sync(0xffffffffu, accd, 4);
      accd += __shfl_xor_sync(0xffffffffu, accd, 2);
      accd += __shfl_xor_sync(0xffffffffu, accd, 1);
      if (j < MP && sub == 0) {
        // compact away the padded bins: value (q, k) -> q*K + k
        const int q = j / KP, k = j - q * KP;
        if (k < K) s_tot[q * K + k] = accd;
      }
    }
  }
  if (tid == 0) counters[b] = 0u;  // ready for the next launch
  __syncthreads();

  // ---- cross-GPU ordered all-reduce over NVLink ----------------------------
  if (a.reduce && pt.world > 1) mgb_ll_allreduce_f64(pt, s_tot, s_stage, 3 * K);

  // ---- epilogue (warp 0; lane k owns bin k) --------------------------------
  if (warp != 0) return;
  const double sig = (double)sigma;
  const double inv = 0.7071067811865476 / sig;             // 1/(sqrt2 sigma)
  const double rsqrtpi = 0.5641895835477563;               // 1/sqrt(pi)
  const double dsdp = md.sigma_link == 1 ? 2.302585092994046 * sig : 1.0;
  const int P = md.nparams;
  double s_k = 0.0, dmu_k = 0.0, dsg_k = 0.0;
  if (lane < K) {
    const double nrm = (double)md.norm[lane];
    s_k = nrm * s_tot[lane];
    dmu_k = -nrm * inv * rsqrtpi * s_tot[K + lane];
    dsg_k = -nrm * (rsqrtpi * EB_INV_KAPPA / sig) * dsdp * s_tot[2 * K + lane];
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
  if (a.mode < MGB_MODE_LOSS_GRAD) return;

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
  if (lane != 0) return;

  float grad[MGB_MAX_PARAMS];
#pragma unroll
  for (int i = 0; i < MGB_MAX_PARAMS; ++i) {
    double v = 0.0;
    if (i == md.mu_idx) v += gmu;
    if (i == md.sig_idx) v += gsg;
    grad[i] = (float)v;
  }
  if (a.out_loss) a.out_loss[b] = (float)loss;
  if (a.out_grad)
    for (int i = 0; i < P; ++i) a.out_grad[(size_t)b * P + i] = grad[i];
  if (a.mode < MGB_MODE_ADAM) return;

  // ---- optimiser update (replicated, deterministic on every GPU) ----------
  const int step = *a.opt_step;       // updates applied so far (bias corr.)
  const int row = step - a.step_base;  // row of this step in the buffers
  if (a.loss_hist && row >= 0 && row < a.traj_rows)
    a.loss_hist[row] = (float)loss;
  if (a.mode == MGB_MODE_GD) {
    for (int i = 0; i < P; ++i) {
      const float p = a.opt_p[i];
      if (a.traj && row >= 0 && row < a.traj_rows)
        a.traj[(size_t)row * P + i] = p;
      a.opt_p[i] = p - a.lr * grad[i];
    }
  } else {
    const float c1 = 1.0f - powf(a.b1, (float)(step + 1));
    const float c2 = 1.0f - powf(a.b2, (float)(step + 1));
    for (int i = 0; i < P; ++i) {
      const int kind = a.bound_kind[i];
      const float lo = a.bound_lo[i], hi = a.bound_hi[i];
      const float p = a.opt_p[i];
      float u = a.opt_u[i];
      const float gi =
          grad[i] * bound_jac(kind, lo, hi, a.jac_at_params ? p : u);
      const float m = (1.0f - a.b1) * gi + a.b1 * a.opt_m[i];
      const float v = (1.0f - a.b2) * gi * gi + a.b2 * a.opt_v[i];
      const float mhat = m / c1, vhat = v / c2;
      u = u - a.lr * mhat / (sqrtf(vhat) + a.eps);
      const float pn = bound_inverse(kind, lo, hi, u);
      a.opt_m[i] = m;
      a.opt_v[i] = v;
      a.opt_u[i] = u;
      a.opt_p[i] = pn;
      if (a.traj && row + 1 >= 0 && row + 1 < a.traj_rows)
        a.traj[(size_t)(row + 1) * P + i] = pn;
    }
  }
  *a.opt_step = step + 1;
}

typedef void (*eb_kernel_t)(const MgbErfBinModel, const MgbStepArgs,
                            const MgbPeerTable, double*, unsigned*, const int);

template <bool W>
eb_kernel_t pick_kernel(int nb) {
  switch (nb) {
#define EB_CASE(N) \
  case N:          \
    return erfbin_step_kernel<N, W>;
    EB_CASE(1) EB_CASE(2) EB_CASE(3) EB_CASE(4) EB_CASE(5) EB_CASE(6)
    EB_CASE(7) EB_CASE(8) EB_CASE(9) EB_CASE(10) EB_CASE(11) EB_CASE(12)
    EB_CASE(13) EB_CASE(14) EB_CASE(15) EB_CASE(16)
#undef EB_CASE
  }
  return nullptr;
}

int g_grid_cache[64] = {0};

}  // namespace

extern "C" int mgb_erfbin_grid(void) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (dev >= 0 && dev < 64 && g_grid_cache[dev]) return g_grid_cache[dev];
  int sms = 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) !=
      cudaSuccess)
    return 0;
  const int g = sms * EB_CTAS_PER_SM;
  if (dev >= 0 && dev < 64) g_grid_cache[dev] = g;
  return g;
}

static inline void eb_layout(int nbins, int* npass, int* nb) {
  *npass = (nbins + EB_NBMAX - 1) / EB_NBMAX;
  *nb = (nbins + *npass - 1) / *npass;
}

extern "C" uint64_t mgb_erfbin_scratch_bytes(int nbins, int batch) {
  int npass, nb;
  if (nbins < 1 || nbins > MGB_MAX_BINS || batch < 1) return 0;
  eb_layout(nbins, &npass, &nb);
  const int grid = mgb_erfbin_grid();
  const uint64_t counters = ((uint64_t)batch * 4ull + 255ull) & ~255ull;
  return counters + (uint64_t)batch * grid * 3ull * npass * nb * 8ull;
}

extern "C" int mgb_erfbin_step(const MgbErfBinModel* model,
                               const MgbStepArgs* args, const MgbPeerTable* pt,
                               void* scratch, uint64_t scratch_bytes,
                               void* stream) {
  MGB_REQUIRE(model && args && scratch, MGB_EINVAL,
              "mgb_erfbin_step: null argument");
  const MgbErfBinModel& md = *model;
  const MgbStepArgs& a = *args;
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
  int npass, nb;
  eb_layout(md.nbins, &npass, &nb);
  const int grid = mgb_erfbin_grid();
  MGB_REQUIRE(grid > 0, MGB_ESTATE, "mgb_erfbin_step: no CUDA device");
  const uint64_t need = mgb_erfbin_scratch_bytes(md.nbins, a.batch);
  MGB_REQUIRE(scratch_bytes >= need, MGB_ETOOBIG,
              "mgb_erfbin_step: scratch %llu < %llu bytes",
              (unsigned long long)scratch_bytes, (unsigned long long)need);
  const uint64_t coff = ((uint64_t)a.batch * 4ull + 255ull) & ~255ull;
  unsigned* counters = reinterpret_cast<unsigned*>(scratch);
  double* partials =
      reinterpret_cast<double*>(reinterpret_cast<char*>(scratch) + coff);
  eb_kernel_t k = md.w ? pick_kernel<true>(nb) : pick_kernel<false>(nb);
  MGB_REQUIRE(k != nullptr, MGB_ESTATE, "mgb_erfbin_step: no kernel for nb=%d",
              nb);
  dim3 g((unsigned)grid, (unsigned)a.batch, 1), blk(EB_BLOCK, 1, 1);
  k<<<g, blk, 0, (cudaStream_t)stream>>>(md, a, ptv, partials, counters,
                                         npass);
  MGB_CUDA_CHECK(cudaGetLastError());
  return MGB_OK;
}
