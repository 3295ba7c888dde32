// Shared device/host helpers for libmgb200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../include/multigrad_b200.h"

// ---------------------------------------------------------------- host side
void mgb_set_error(const char* fmt, ...);

#define MGB_CUDA_CHECK(expr)                                                  \
  do {                                                                        \
    cudaError_t _e = (expr);                                                  \
    if (_e != cudaSuccess) {                                                  \
      mgb_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr,             \
                    cudaGetErrorString(_e));                                  \
      return (int)_e;                                                         \
    }                                                                         \
  } while (0)

#define MGB_REQUIRE(cond, code, ...)                                          \
  do {                                                                        \
    if (!(cond)) {                                                            \
      mgb_set_error(__VA_ARGS__);                                             \
      return (code);                                                          \
    }                                                                         \
  } while (0)

// ---------------------------------------------------- workspace byte layout
#define MGB_CTRL_BYTES 4096ull
#define MGB_CTRL_EPOCH 0      // uint32: last completed LL exchange epoch
#define MGB_CTRL_ERROR 1      // uint32: sticky error flag (1 = spin timeout)
#define MGB_CTRL_BULK_EPOCH 2 // uint32: last completed bulk collective
#define MGB_CTRL_TICKET 3     // uint32: CTA completion ticket (one-shot push)

__host__ __device__ inline unsigned long long mgb_ll_bytes(int world,
                                                          unsigned ll_units) {
  return 2ull * (unsigned long long)world * ll_units * 8ull;
}
__host__ __device__ inline unsigned long long mgb_bulk_off(int world,
                                                           unsigned ll_units) {
  unsigned long long o = MGB_CTRL_BYTES + mgb_ll_bytes(world, ll_units);
  return (o + 1023ull) & ~1023ull;
}

// Each bulk region is split in two: [0, half) belongs to the two-shot
// all-reduce (and the K7 scatter windows), [half, bulk) to the one-shot push.
__host__ __device__ inline unsigned long long mgb_bulk_half(
    unsigned long long bulk_bytes) {
  return (bulk_bytes / 2ull) & ~1023ull;
}

// -------------------------------------------------------------- device side
#ifdef __CUDACC__

__device__ __forceinline__ unsigned long long mgb_globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// 64-bit system-scope relaxed store / load.  Naturally aligned 8-byte
// accesses are single-copy atomic, which is what the LL protocol needs: the
// flag and the payload word travel in the same store.
__device__ __forceinline__ void mgb_st_sys_u64(unsigned long long* p,
                                               unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v)
               : "memory");
}
__device__ __forceinline__ unsigned long long mgb_ld_sys_u64(
    const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void mgb_st_sys_u32(unsigned* p, unsigned v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v)
               : "memory");
}
__device__ __forceinline__ unsigned mgb_ld_sys_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p)
               : "memory");
  return v;
}

__device__ __forceinline__ unsigned* mgb_ctrl(const MgbPeerTable& pt, int r) {
  return reinterpret_cast<unsigned*>(pt.base[r]);
}
// LL slot of (parity, source rank, unit) inside rank `dst`'s workspace.
__device__ __forceinline__ unsigned long long* mgb_ll_slot(
    const MgbPeerTable& pt, int dst, unsigned parity, int src, unsigned unit) {
  char* b = reinterpret_cast<char*>(pt.base[dst]) + MGB_CTRL_BYTES;
  return reinterpret_cast<unsigned long long*>(b) +
         ((size_t)(parity * (unsigned)pt.world + (unsigned)src) * pt.ll_units +
          unit);
}

// Wait until the local slot carries `epoch`; returns the payload word.
// Bounded: after pt.spin_timeout_ms the sticky error flag is raised and 0 is
// returned so that the kernel always terminates.
__device__ __forceinline__ unsigned mgb_ll_wait(const MgbPeerTable& pt,
                                                unsigned long long* slot,
                                                unsigned epoch) {
  unsigned long long v = mgb_ld_sys_u64(slot);
  if ((unsigned)(v >> 32) == epoch) return (unsigned)v;
  const unsigned long long t0 = mgb_globaltimer_ns();
  const unsigned long long limit =
      (unsigned long long)pt.spin_timeout_ms * 1000000ull;
  unsigned it = 0;
  for (;;) {
    v = mgb_ld_sys_u64(slot);
    if ((unsigned)(v >> 32) == epoch) return (unsigned)v;
    if ((++it & 63u) == 0u) {
      if (mgb_globaltimer_ns() - t0 > limit) {
        mgb_st_sys_u32(mgb_ctrl(pt, pt.rank) + MGB_CTRL_ERROR, 1u);
        return 0u;
      }
      __nanosleep(20);
    }
  }
}

// One-shot ordered all-reduce of `m` doubles held in shared memory, executed
// by a whole CTA.  Protocol (NCCL-LL style, push model):
//   * every rank stores each 32-bit half of each double, tagged with the
//     epoch, straight into every peer's slot [parity][my rank][unit];
//   * every rank polls its *local* slots until the tag shows up, then sums
//     the W contributions in rank order -> bit-identical totals everywhere.
// Two parities suffice because a rank can only be one exchange ahead of any
// peer (it cannot finish exchange e+1 before every peer has sent e+1, which
// they do only after having consumed e).
// `sh` in: local values, out: totals.  `stage` is scratch of world*m doubles
// in shared memory.  Requires 2*m <= pt.ll_units.
// Fetch doubles that peers pushed into my slots (two tagged words per double,
// from unit `unit0` on).  Item i of count * nsrc is (si = i / count, j = i %
// count): source rank src_of(si), double j, stored to *out_of(si, j) unless
// that returns nullptr.  Four items (eight slot loads) are in flight per
// thread before any tag is looked at.
template <typename SrcOf, typename OutOf>
__device__ __forceinline__ void mgb_ll_gather(const MgbPeerTable& pt,
                                              unsigned parity, unsigned epoch,
                                              unsigned unit0, int count,
                                              int nsrc, SrcOf src_of,
                                              OutOf out_of) {
  const int tid = threadIdx.x, nt = blockDim.x;
  for (int i0 = tid; i0 < count * nsrc; i0 += 4 * nt) {
    unsigned long long lo[4], hi[4];
    double* dstp[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int i = i0 + k * nt;
      lo[k] = hi[k] = 0ull;
      dstp[k] = nullptr;
      if (i < count * nsrc) {
        const int si = i / count, j = i - si * count;
        dstp[k] = out_of(si, j);
        if (dstp[k]) {
          const int s = src_of(si);
          lo[k] = mgb_ld_sys_u64(
              mgb_ll_slot(pt, pt.rank, parity, s, unit0 + 2u * j));
          hi[k] = mgb_ld_sys_u64(
              mgb_ll_slot(pt, pt.rank, parity, s, unit0 + 2u * j + 1u));
        }
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (dstp[k]) {
        const int i = i0 + k * nt;
        const int si = i / count, j = i - si * count;
        const int s = src_of(si);
        unsigned l = (unsigned)lo[k], h = (unsigned)hi[k];
        if ((unsigned)(lo[k] >> 32) != epoch)
          l = mgb_ll_wait(
              pt, mgb_ll_slot(pt, pt.rank, parity, s, unit0 + 2u * j), epoch);
        if ((unsigned)(hi[k] >> 32) != epoch)
          h = mgb_ll_wait(
              pt, mgb_ll_slot(pt, pt.rank, parity, s, unit0 + 2u * j + 1u),
              epoch);
        *dstp[k] = __hiloint2double((int)h, (int)l);
      }
    }
  }
}

// Long vectors (a K x P Jacobian: m ~ 1300) go in two rounds -- every rank
// reduces one slice of m / W doubles (in rank order, so the totals stay
// bit-identical everywhere and equal to the one-round result) and sends the
// sums to everybody: 2 m (W-1)/W doubles leave a rank instead of m (W-1).
// All of them come from the one CTA that runs the exchange, and the remote
// stores one SM can keep in flight bound it: 8 GPUs, m = 1300 cost 114 us in
// one round.  Short vectors keep the single round (one latency less).
#ifndef MGB_LL_TWO_ROUND_MIN
#define MGB_LL_TWO_ROUND_MIN 192
#endif
__device__ __forceinline__ void mgb_ll_allreduce_two_round(
    const MgbPeerTable& pt, double* sh, double* stage, int m, unsigned parity,
    unsigned epoch) {
  const int tid = threadIdx.x, nt = blockDim.x;
  const int W = pt.world, me = pt.rank;
  const int c = (m + W - 1) / W;            // slice length
  const unsigned* words = reinterpret_cast<const unsigned*>(sh);
  // round 1: my contribution to slice d goes to rank d (units [0, 2c))
  for (int i = tid; i < 2 * c * (W - 1); i += nt) {
    const int pi = i / (2 * c), u = i - pi * (2 * c);
    int dst = me + 1 + pi;
    if (dst >= W) dst -= W;
    const int j = dst * c + (u >> 1);
    if (j < m)
      mgb_st_sys_u64(mgb_ll_slot(pt, dst, parity, me, (unsigned)u),
                     ((unsigned long long)epoch << 32) | words[2 * j + (u & 1)]);
  }
  const int lo_j = me * c;
  const int mine = lo_j < m ? (lo_j + c <= m ? c : m - lo_j) : 0;
  // stage[si * mine + j]: contribution of the si-th other rank to my slice
  if (mine > 0)
    mgb_ll_gather(pt, parity, epoch, 0u, mine, W - 1,
                  [&](int si) { return si < me ? si : si + 1; },
                  [&](int si, int j) { return stage + si * mine + j; });
  __syncthreads();
  double* red = stage + (size_t)(W - 1) * c;        // my reduced slice
  for (int j = tid; j < mine; j += nt) {
    double acc = 0.0;
    for (int s = 0; s < W; ++s)
      acc += s == me ? sh[lo_j + j] : stage[(s < me ? s : s - 1) * mine + j];
    red[j] = acc;
  }
  __syncthreads();
  // round 2: the sums of my slice go to everybody (units [2c, 4c))
  const unsigned* rw = reinterpret_cast<const unsigned*>(red);
  for (int i = tid; i < 2 * mine * (W - 1); i += nt) {
    const int pi = i / (2 * mine), u = i - pi * (2 * mine);
    int dst = me + 1 + pi;
    if (dst >= W) dst -= W;
    mgb_st_sys_u64(mgb_ll_slot(pt, dst, parity, me, (unsigned)(2 * c + u)),
                   ((unsigned long long)epoch << 32) | rw[u]);
  }
  for (int j = tid; j < mine; j += nt) sh[lo_j + j] = red[j];
  // the other slices, straight into sh (slice s ends at m)
  mgb_ll_gather(pt, parity, epoch, (unsigned)(2 * c), c, W - 1,
                [&](int si) { return si < me ? si : si + 1; },
                [&](int si, int j) -> double* {
                  const int g = (si < me ? si : si + 1) * c + j;
                  return g < m ? sh + g : nullptr;
                });
  __syncthreads();
}

__device__ __forceinline__ void mgb_ll_allreduce_f64(const MgbPeerTable& pt,
                                                     double* sh, double* stage,
                                                     int m) {
  if (pt.world <= 1) return;
  __shared__ unsigned s_epoch;
  const int tid = threadIdx.x, nt = blockDim.x;
  unsigned* ctrl = mgb_ctrl(pt, pt.rank);
  if (tid == 0) s_epoch = mgb_ld_sys_u32(ctrl + MGB_CTRL_EPOCH) + 1u;
  __syncthreads();
  const unsigned epoch = s_epoch;
  const unsigned parity = epoch & 1u;
  const int units = 2 * m;
  const int W = pt.world;
  if (m >= MGB_LL_TWO_ROUND_MIN && W > 2) {
    mgb_ll_allreduce_two_round(pt, sh, stage, m, parity, epoch);
    if (tid == 0) mgb_st_sys_u32(ctrl + MGB_CTRL_EPOCH, epoch);
    return;
  }
  const unsigned* words = reinterpret_cast<const unsigned*>(sh);
  // push to the W-1 peers, staggered so that not everyone hits rank 0 first
  for (int i = tid; i < units * (W - 1); i += nt) {
    const int pi = i / units, u = i - pi * units;
    int dst = pt.rank + 1 + pi;
    if (dst >= W) dst -= W;
    mgb_st_sys_u64(mgb_ll_slot(pt, dst, parity, pt.rank, (unsigned)u),
                   ((unsigned long long)epoch << 32) | words[u]);
  }
  // gather: double j of source s into stage[s*m+j] (own values copied)
  for (int j = tid; j < m; j += nt) stage[pt.rank * m + j] = sh[j];
  mgb_ll_gather(pt, parity, epoch, 0u, m, W - 1,
                [&](int si) { return si < pt.rank ? si : si + 1; },
                [&](int si, int j) {
                  return stage + (si < pt.rank ? si : si + 1) * m + j;
                });
  __syncthreads();
  for (int j = tid; j < m; j += nt) {
    double acc = 0.0;
    for (int s = 0; s < W; ++s) acc += stage[s * m + j];
    sh[j] = acc;
  }
  __syncthreads();
  if (tid == 0) mgb_st_sys_u32(ctrl + MGB_CTRL_EPOCH, epoch);
}

// ------------------------------------------------------------- math helpers
#include "erfcx_coeffs.inc"

__device__ __forceinline__ float mgb_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float mgb_rcp(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Deterministic column sums of the per-CTA partials [G][M] (fp64, global) into
// out[M] (shared), executed by a whole CTA after the last-CTA election.
// 8 lanes share one value; every lane keeps 8 L2 loads in flight (a naive
// dependent loop costs G/8 L2 round trips ~ 10 us for G = 296) and adds them
// in a fixed order, then the 8 lanes are combined by a fixed shuffle tree.
__device__ __forceinline__ void mgb_grid_reduce_f64(
    const double* __restrict__ partials, int G, int M, double* out, int tid,
    int nthreads) {
  const int sub = tid & 7, grp = tid >> 3, ngrp = nthreads >> 3;
  for (int j0 = 0; j0 < M; j0 += ngrp) {
    const int j = j0 + grp;
    double acc = 0.0;
    if (j < M) {
      const double* src = partials + j;
      for (int c0 = sub; c0 < G; c0 += 64) {
        double v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int c = c0 + 8 * u;
          v[u] = c < G ? __ldcg(src + (size_t)c * M) : 0.0;
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) acc += v[u];
      }
    }
    acc += __shfl_xor_sync(0xffffffffu, acc, 4);
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    if (j < M && sub == 0) out[j] = acc;
  }
}

template <typename T>
__device__ __forceinline__ T mgb_warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

#endif  // __CUDACC__
