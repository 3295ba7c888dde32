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
  const unsigned* words = reinterpret_cast<const unsigned*>(sh);
  // push to the W-1 peers, staggered so that not everyone hits rank 0 first
  for (int i = tid; i < units * (W - 1); i += nt) {
    const int pi = i / units, u = i - pi * units;
    int dst = pt.rank + 1 + pi;
    if (dst >= W) dst -= W;
    mgb_st_sys_u64(mgb_ll_slot(pt, dst, parity, pt.rank, (unsigned)u),
                   ((unsigned long long)epoch << 32) | words[u]);
  }
  // gather: thread (s, j) fetches double j of source s into stage[s*m+j]
  for (int i = tid; i < m * W; i += nt) {
    const int s = i / m, j = i - s * m;
    double val;
    if (s == pt.rank) {
      val = sh[j];
    } else {
      unsigned lo = mgb_ll_wait(
          pt, mgb_ll_slot(pt, pt.rank, parity, s, 2u * j), epoch);
      unsigned hi = mgb_ll_wait(
          pt, mgb_ll_slot(pt, pt.rank, parity, s, 2u * j + 1u), epoch);
      val = __hiloint2double((int)hi, (int)lo);
    }
    stage[i] = val;
  }
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
