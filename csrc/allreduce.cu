// K5 / K6: stand-alone collectives over the NVLink peer workspace.
//
//  * one-shot ordered LL all-reduce (latency path, bit-identical on all ranks)
//  * two-shot reduce-scatter + all-gather over peer loads / stores (bandwidth
//    path, cooperative launch; every rank reduces its slice in rank order)
//  * LL broadcast with device-side wait (BFGS parameter hand-off)
//
// Reference counterparts: reduce_sum (multigrad/multigrad.py:149-183) and the
// pickled "compute"/params broadcasts (multigrad/bfgs.py:72-77,
// multigrad/adam.py:39-49).  No MPI/NCCL call is made anywhere in this file.
#include <cooperative_groups.h>
#include "mgb_common.cuh"

namespace cg = cooperative_groups;

namespace {

#define AR_LL_BLOCK 1024
#define AR_BULK_BLOCK 512
// auto policy: one-shot push up to MGB_ONESHOT_MAX_BYTES / W bytes (8 MiB at
// 2 GPUs, 2 MiB at 8): measured crossover against the two-shot and NVLS
// kernels (benchmarks/allreduce_sweep.py --all-algos, profiles/)
#ifndef MGB_ONESHOT_MAX_BYTES
#define MGB_ONESHOT_MAX_BYTES (16ull << 20)
#endif

// ---- LL all-reduce on a global array -------------------------------------
template <typename T>
__global__ void __launch_bounds__(AR_LL_BLOCK, 1)
ll_allreduce_kernel(const MgbPeerTable pt, T* __restrict__ data, int n) {
  constexpr int WPE = sizeof(T) / 4;  // 32-bit words per element
  __shared__ unsigned s_epoch;
  const int tid = threadIdx.x, nt = blockDim.x;
  const int W = pt.world;
  unsigned* ctrl = mgb_ctrl(pt, pt.rank);
  if (tid == 0) s_epoch = mgb_ld_sys_u32(ctrl + MGB_CTRL_EPOCH) + 1u;
  __syncthreads();
  const unsigned epoch = s_epoch, parity = epoch & 1u;
  const int units = n * WPE;
  const unsigned* words = reinterpret_cast<const unsigned*>(data);
  for (int i = tid; i < units * (W - 1); i += nt) {
    const int pi = i / units, u = i - pi * units;
    int dst = pt.rank + 1 + pi;
    if (dst >= W) dst -= W;
    mgb_st_sys_u64(mgb_ll_slot(pt, dst, parity, pt.rank, (unsigned)u),
                   ((unsigned long long)epoch << 32) | words[u]);
  }
  __syncthreads();  // every local word has been read before it is replaced
  for (int j = tid; j < n; j += nt) {
    T acc = T(0);
    for (int s = 0; s < W; ++s) {
      T v;
      if (s == pt.rank) {
        v = data[j];
      } else if (WPE == 1) {
        unsigned w0 = mgb_ll_wait(
            pt, mgb_ll_slot(pt, pt.rank, parity, s, (unsigned)j), epoch);
        v = (T)__uint_as_float(w0);
      } else {
        unsigned lo = mgb_ll_wait(
            pt, mgb_ll_slot(pt, pt.rank, parity, s, 2u * j), epoch);
        unsigned hi = mgb_ll_wait(
            pt, mgb_ll_slot(pt, pt.rank, parity, s, 2u * j + 1u), epoch);
        v = (T)__hiloint2double((int)hi, (int)lo);
      }
      acc += v;
    }
    data[j] = acc;
  }
  __syncthreads();
  if (tid == 0) mgb_st_sys_u32(ctrl + MGB_CTRL_EPOCH, epoch);
}

// ---- LL broadcast -----------------------------------------------------------
__global__ void __launch_bounds__(256, 1)
ll_bcast_kernel(const MgbPeerTable pt, float* __restrict__ data, int n,
                int root) {
  __shared__ unsigned s_epoch;
  const int tid = threadIdx.x, nt = blockDim.x;
  const int W = pt.world;
  unsigned* ctrl = mgb_ctrl(pt, pt.rank);
  if (tid == 0) s_epoch = mgb_ld_sys_u32(ctrl + MGB_CTRL_EPOCH) + 1u;
  __syncthreads();
  const unsigned epoch = s_epoch, parity = epoch & 1u;
  // Ready-first: every receiver announces itself with one tagged word in a
  // unit of its own (the last one; data never uses it), and the root pushes
  // only when all have.  A broadcast is one-way, so a root that pushed at once
  // could overwrite words a slow peer has not read yet: its own all-reduce
  // contribution of two collectives ago, or -- if it only acknowledged
  // afterwards -- a pending acknowledgement.  With the announcement nobody
  // finishes collective e+1 before everybody has finished e, which is the
  // invariant the two parities rest on (tools/ll_protocol_model.py explores
  // the schedules: this form is clean, push-then-acknowledge is not).
  const unsigned ready_unit = pt.ll_units - 1u;
  if (pt.rank == root) {
    for (int s = tid; s < W; s += nt)
      if (s != root)
        (void)mgb_ll_wait(pt, mgb_ll_slot(pt, root, parity, s, ready_unit),
                          epoch);
    __syncthreads();
    for (int i = tid; i < n * (W - 1); i += nt) {
      const int pi = i / n, u = i - pi * n;
      int dst = root + 1 + pi;
      if (dst >= W) dst -= W;
      mgb_st_sys_u64(mgb_ll_slot(pt, dst, parity, root, (unsigned)u),
                     ((unsigned long long)epoch << 32) |
                         __float_as_uint(data[u]));
    }
  } else {
    if (tid == 0)
      mgb_st_sys_u64(mgb_ll_slot(pt, root, parity, pt.rank, ready_unit),
                     (unsigned long long)epoch << 32);
    for (int j = tid; j < n; j += nt)
      data[j] = __uint_as_float(mgb_ll_wait(
          pt, mgb_ll_slot(pt, pt.rank, parity, root, (unsigned)j), epoch));
  }
  __syncthreads();
  if (tid == 0) mgb_st_sys_u32(ctrl + MGB_CTRL_EPOCH, epoch);
}

// ---- cross-GPU barrier executed by one thread --------------------------------
__device__ void ll_barrier_one_thread(const MgbPeerTable& pt) {
  unsigned* ctrl = mgb_ctrl(pt, pt.rank);
  const unsigned epoch = mgb_ld_sys_u32(ctrl + MGB_CTRL_EPOCH) + 1u;
  const unsigned parity = epoch & 1u;
  __threadfence_system();
  for (int pi = 0; pi < pt.world - 1; ++pi) {
    int dst = pt.rank + 1 + pi;
    if (dst >= pt.world) dst -= pt.world;
    mgb_st_sys_u64(mgb_ll_slot(pt, dst, parity, pt.rank, 0u),
                   (unsigned long long)epoch << 32);
  }
  for (int s = 0; s < pt.world; ++s)
    if (s != pt.rank)
      (void)mgb_ll_wait(pt, mgb_ll_slot(pt, pt.rank, parity, s, 0u), epoch);
  __threadfence_system();
  mgb_st_sys_u32(ctrl + MGB_CTRL_EPOCH, epoch);
}

__device__ __forceinline__ float4 ld_peer_f4(const float4* p) {
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_peer_f4(float4* p, float4 v) {
  asm volatile("st.volatile.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p),
               "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

// ---- two-shot all-reduce, float32, n4 = number of float4 --------------------
// Phase 1: stage input into the local bulk region A.  Barrier.
// Phase 2: rank r owns float4 range [r*chunk, (r+1)*chunk): pull that range
//          from every rank's A (peer loads), sum in rank order, push the sum
//          into every rank's B (peer stores).  Barrier.
// Phase 3: copy B back to the user buffer.
__global__ void __launch_bounds__(AR_BULK_BLOCK, 1)
bulk_allreduce_f32_kernel(const MgbPeerTable pt, float* __restrict__ data,
                          long long n4, long long tail_start, int ntail) {
  cg::grid_group grid = cg::this_grid();
  const int W = pt.world;
  const unsigned long long off = mgb_bulk_off(W, pt.ll_units);
  const unsigned long long bulk = (pt.bulk_bytes + 1023ull) & ~1023ull;
  const long long gsz = (long long)gridDim.x * blockDim.x;
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  float4* locA = reinterpret_cast<float4*>((char*)pt.base[pt.rank] + off);
  float4* locB = reinterpret_cast<float4*>((char*)pt.base[pt.rank] + off + bulk);
  float4* d4 = reinterpret_cast<float4*>(data);
  // the ragged tail (< 4 floats) rides in one extra float4, zero padded
  const long long n4t = n4 + (ntail ? 1 : 0);
  for (long long i = gid; i < n4t; i += gsz) {
    float4 v;
    if (i < n4) {
      v = d4[i];
    } else {
      v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (ntail > 0) v.x = data[tail_start];
      if (ntail > 1) v.y = data[tail_start + 1];
      if (ntail > 2) v.z = data[tail_start + 2];
    }
    locA[i] = v;
  }
  __threadfence_system();
  grid.sync();
  if (gid == 0) ll_barrier_one_thread(pt);
  grid.sync();
  const long long chunk = (n4t + W - 1) / W;
  const long long lo = chunk * pt.rank;
  const long long hi = min(lo + chunk, n4t);
  for (long long i = lo + gid; i < hi; i += gsz) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    // issue all peer loads of a group of 8 ranks before the first add so
    // that they are in flight together (NVLink latency ~2 us per load)
    for (int s0 = 0; s0 < W; s0 += 8) {
      float4 v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int s = s0 + k;
        if (s < W) {
          const float4* src = reinterpret_cast<const float4*>(
                                  (const char*)pt.base[s] + off) + i;
          v[k] = (s == pt.rank) ? locA[i] : ld_peer_f4(src);
        } else {
          v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if (s0 + k < W) {
          acc.x += v[k].x; acc.y += v[k].y; acc.z += v[k].z; acc.w += v[k].w;
        }
      }
    }
    for (int s = 0; s < W; ++s) {
      float4* dst =
          reinterpret_cast<float4*>((char*)pt.base[s] + off + bulk) + i;
      if (s == pt.rank) locB[i] = acc; else st_peer_f4(dst, acc);
    }
  }
  __threadfence_system();
  grid.sync();
  if (gid == 0) ll_barrier_one_thread(pt);
  grid.sync();
  for (long long i = gid; i < n4t; i += gsz) {
    const float4 v = __ldcv(locB + i);
    if (i < n4) {
      d4[i] = v;
    } else {
      if (ntail > 0) data[tail_start] = v.x;
      if (ntail > 1) data[tail_start + 1] = v.y;
      if (ntail > 2) data[tail_start + 2] = v.z;
    }
  }
}

// ---- one-shot push all-reduce, float32 (mid-size messages) ------------------
// Every rank stores its whole input into every peer's staging area
// [parity][source rank][...] (bulk region A for even epochs, B for odd), so a
// message crosses NVLink once and there is ONE synchronisation point: CTA c
// of every rank handles the same element range and signals "my range is
// staged" with an epoch-tagged flag in LL slot unit c; after the flags of all
// peers arrived the CTA sums the W copies in rank order (bit-identical on
// every rank).  No grid-wide barrier, no second exchange.  Two parities
// suffice for the same reason as in the LL protocol: a peer can start the
// next collective (and overwrite a staging area) only one exchange ahead,
// and that one uses the other parity.  The last CTA to finish publishes the
// epoch (every CTA has read it by then).
__global__ void __launch_bounds__(AR_BULK_BLOCK, 1)
oneshot_allreduce_f32_kernel(const MgbPeerTable pt, float* __restrict__ data,
                             long long n4, long long tail_start, int ntail) {
  __shared__ unsigned s_epoch;
  const int tid = threadIdx.x;
  const int W = pt.world, me = pt.rank;
  unsigned* ctrl = mgb_ctrl(pt, me);
  if (tid == 0) s_epoch = mgb_ld_sys_u32(ctrl + MGB_CTRL_EPOCH) + 1u;
  __syncthreads();
  const unsigned epoch = s_epoch, parity = epoch & 1u;
  // upper half of region A / B: the two-shot kernel keeps the lower halves,
  // so a rank that is still copying its two-shot result out of B cannot be
  // overwritten by a peer that already started the next (one-shot) call
  const unsigned long long off =
      mgb_bulk_off(W, pt.ll_units) +
      (parity ? ((pt.bulk_bytes + 1023ull) & ~1023ull) : 0ull) +
      mgb_bulk_half(pt.bulk_bytes);
  const long long n4t = n4 + (ntail ? 1 : 0);
  const long long per = (n4t + gridDim.x - 1) / gridDim.x;
  const long long lo = per * blockIdx.x;
  const long long hi = min(lo + per, n4t);
  float4* d4 = reinterpret_cast<float4*>(data);
  // ---- push my range to every peer ----------------------------------------
  for (long long i = lo + tid; i < hi; i += AR_BULK_BLOCK) {
    float4 v;
    if (i < n4) {
      v = d4[i];
    } else {
      v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (ntail > 0) v.x = data[tail_start];
      if (ntail > 1) v.y = data[tail_start + 1];
      if (ntail > 2) v.z = data[tail_start + 2];
    }
    for (int pi = 0; pi < W - 1; ++pi) {
      int dst = me + 1 + pi;
      if (dst >= W) dst -= W;
      float4* slot = reinterpret_cast<float4*>((char*)pt.base[dst] + off) +
                     (size_t)me * n4t + i;
      st_peer_f4(slot, v);
    }
  }
  // Release/acquire by the signalling threads only: the CTA barrier makes
  // every thread's pushes visible to them, their system-scope fence is
  // cumulative.  (A system fence from all 512 threads costs microseconds.)
  __syncthreads();
  if (tid < W - 1) {
    int dst = me + 1 + tid;
    if (dst >= W) dst -= W;
    __threadfence_system();
    mgb_st_sys_u64(mgb_ll_slot(pt, dst, parity, me, blockIdx.x),
                   (unsigned long long)epoch << 32);
  }
  // ---- wait for the same range of every peer -------------------------------
  if (tid < W && tid != me) {
    (void)mgb_ll_wait(pt, mgb_ll_slot(pt, me, parity, tid, blockIdx.x), epoch);
    __threadfence_system();
  }
  __syncthreads();
  // ---- ordered sum ----------------------------------------------------------
  const float4* stage =
      reinterpret_cast<const float4*>((const char*)pt.base[me] + off);
  for (long long i = lo + tid; i < hi; i += AR_BULK_BLOCK) {
    float4 mine;
    if (i < n4) {
      mine = d4[i];
    } else {
      mine = make_float4(0.f, 0.f, 0.f, 0.f);
      if (ntail > 0) mine.x = data[tail_start];
      if (ntail > 1) mine.y = data[tail_start + 1];
      if (ntail > 2) mine.z = data[tail_start + 2];
    }
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int s0 = 0; s0 < W; s0 += 8) {
      float4 v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int s = s0 + k;
        if (s < W)
          v[k] = (s == me) ? mine : ld_peer_f4(stage + (size_t)s * n4t + i);
        else
          v[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (s0 + k < W) {
          acc.x += v[k].x; acc.y += v[k].y; acc.z += v[k].z; acc.w += v[k].w;
        }
    }
    if (i < n4) {
      d4[i] = acc;
    } else {
      if (ntail > 0) data[tail_start] = acc.x;
      if (ntail > 1) data[tail_start + 1] = acc.y;
      if (ntail > 2) data[tail_start + 2] = acc.z;
    }
  }
  // ---- the last CTA publishes the epoch ------------------------------------
  __syncthreads();
  if (tid == 0) {
    const unsigned t = atomicAdd(ctrl + MGB_CTRL_TICKET, 1u);
    if (t == gridDim.x - 1) {
      ctrl[MGB_CTRL_TICKET] = 0u;
      __threadfence();
      mgb_st_sys_u32(ctrl + MGB_CTRL_EPOCH, epoch);
    }
  }
}

int g_bulk_grid[64] = {0};

int bulk_grid() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (dev >= 0 && dev < 64 && g_bulk_grid[dev]) return g_bulk_grid[dev];
  int sms = 0, per = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(
      &per, bulk_allreduce_f32_kernel, AR_BULK_BLOCK, 0);
  if (per < 1) per = 1;
  const int g = sms;  // one CTA per SM is enough to saturate NVLink
  if (dev >= 0 && dev < 64) g_bulk_grid[dev] = g;
  return g;
}

}  // namespace

extern "C" int mgb_allreduce_sum(const MgbPeerTable* pt, void* data,
                                 uint64_t n, int dtype, int algo,
                                 void* stream) {
  MGB_REQUIRE(pt && data, MGB_EINVAL, "mgb_allreduce_sum: null argument");
  MGB_REQUIRE(dtype == 0 || dtype == 1, MGB_EINVAL,
              "mgb_allreduce_sum: dtype must be 0 (f32) or 1 (f64)");
  MGB_REQUIRE(pt->world >= 1 && pt->world <= MGB_MAX_WORLD, MGB_EINVAL,
              "mgb_allreduce_sum: bad world size %d", pt->world);
  if (pt->world == 1 || n == 0) return MGB_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const uint64_t units = n * (dtype == 1 ? 2ull : 1ull);
  bool use_ll = units <= pt->ll_units;
  if (algo == 1) {
    MGB_REQUIRE(use_ll, MGB_ETOOBIG,
                "mgb_allreduce_sum: %llu units exceed the LL capacity %u",
                (unsigned long long)units, pt->ll_units);
  } else if (algo == 2) {
    use_ll = false;
  } else if (use_ll && dtype == 0 && units > 4096 && pt->bulk_bytes > 0) {
    use_ll = false;  // past ~16 KiB the bandwidth path wins
  }
  if (use_ll) {
    if (dtype == 0)
      ll_allreduce_kernel<float><<<1, AR_LL_BLOCK, 0, st>>>(
          *pt, (float*)data, (int)n);
    else
      ll_allreduce_kernel<double><<<1, AR_LL_BLOCK, 0, st>>>(
          *pt, (double*)data, (int)n);
    MGB_CUDA_CHECK(cudaGetLastError());
    return MGB_OK;
  }
  MGB_REQUIRE(dtype == 0, MGB_EINVAL,
              "mgb_allreduce_sum: the two-shot path is float32 only "
              "(reduce float64 through the LL path in chunks)");
  MGB_REQUIRE(((uintptr_t)data & 15u) == 0, MGB_EINVAL,
              "mgb_allreduce_sum: data must be 16-byte aligned");
  const uint64_t need = ((n + 3ull) / 4ull) * 16ull;
  const uint64_t half = mgb_bulk_half(pt->bulk_bytes);
  MGB_REQUIRE(need <= half, MGB_ETOOBIG,
              "mgb_allreduce_sum: %llu bytes exceed half the bulk region "
              "(%llu)", (unsigned long long)need, (unsigned long long)half);
  MgbPeerTable ptv = *pt;
  float* d = (float*)data;
  long long n4 = (long long)(n / 4ull);
  long long tail_start = n4 * 4;
  int ntail = (int)(n - (uint64_t)tail_start);
  // One-shot push: every rank receives W-1 copies, so it pays off while the
  // message is small against the NVLink bandwidth (W-1)*n/BW, and it needs
  // W staging copies in one bulk region and one flag unit per CTA.
  const uint64_t oneshot_limit = MGB_ONESHOT_MAX_BYTES / (uint64_t)pt->world;
  const bool fits_oneshot =
      need * (uint64_t)pt->world <= pt->bulk_bytes - half;
  bool oneshot = algo == 3 || (algo == 0 && need <= oneshot_limit);
  if (algo == 3)
    MGB_REQUIRE(fits_oneshot, MGB_ETOOBIG,
                "mgb_allreduce_sum: one-shot staging needs %llu bytes, the "
                "bulk region has %llu",
                (unsigned long long)(need * pt->world),
                (unsigned long long)(pt->bulk_bytes - half));
  if (oneshot && fits_oneshot) {
    const long long n4t = n4 + (ntail ? 1 : 0);
    long long ctas = (n4t + 4 * AR_BULK_BLOCK - 1) / (4 * AR_BULK_BLOCK);
    const long long cap = pt->ll_units < 128u ? pt->ll_units : 128u;
    if (ctas > cap) ctas = cap;
    if (ctas < 1) ctas = 1;
    oneshot_allreduce_f32_kernel<<<(unsigned)ctas, AR_BULK_BLOCK, 0, st>>>(
        ptv, d, n4, tail_start, ntail);
    MGB_CUDA_CHECK(cudaGetLastError());
    return MGB_OK;
  }
  void* args[] = {&ptv, &d, &n4, &tail_start, &ntail};
  const int grid = bulk_grid();
  MGB_REQUIRE(grid > 0, MGB_ESTATE, "mgb_allreduce_sum: no CUDA device");
  MGB_CUDA_CHECK(cudaLaunchCooperativeKernel(
      (void*)bulk_allreduce_f32_kernel, dim3(grid), dim3(AR_BULK_BLOCK), args,
      0, st));
  return MGB_OK;
}

extern "C" int mgb_bcast_f32(const MgbPeerTable* pt, float* data, uint32_t n,
                             int root, void* stream) {
  MGB_REQUIRE(pt && data, MGB_EINVAL, "mgb_bcast_f32: null argument");
  MGB_REQUIRE(root >= 0 && root < pt->world, MGB_EINVAL,
              "mgb_bcast_f32: bad root %d", root);
  if (pt->world == 1 || n == 0) return MGB_OK;
  // (the last unit carries the receivers' announcements)
  MGB_REQUIRE(n + 1u <= pt->ll_units, MGB_ETOOBIG,
              "mgb_bcast_f32: %u floats exceed the LL capacity %u", n,
              pt->ll_units - 1u);
  ll_bcast_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(*pt, data, (int)n,
                                                       root);
  MGB_CUDA_CHECK(cudaGetLastError());
  return MGB_OK;
}
