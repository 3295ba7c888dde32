// N1 (multicast half) + K5 (NVLS path): NVLink-switch multicast objects and the
// in-switch all-reduce built on them.  sm_100a.
//
// A multicast object is one address range that the NVSwitch fans out to the
// same offset of a physical buffer on every GPU of the communicator:
//   multimem.ld_reduce [mc + off]  -> the switch reads offset `off` of every
//                                     GPU's buffer and returns the SUM
//   multimem.st        [mc + off]  -> the switch writes the value to offset
//                                     `off` of every GPU's buffer
// so a rank reduces its slice with ONE load and publishes it with ONE store
// per 16 bytes, independent of the number of GPUs (the peer-load two-shot
// kernel in allreduce.cu issues W loads and W stores).
//
// Host side (driver API, virtual memory management):
//   root:  mgb_mc_create -> POSIX fd (sent to the peers over a unix socket by
//          parallel/workspace.py);  peers: mgb_mc_import(fd)
//   all:   mgb_mc_add_device;  barrier;  mgb_mc_bind (physical memory of this
//          GPU bound at offset 0, unicast + multicast mappings);  barrier
//
// Reference counterpart: none -- the reference reduces on the host through
// MPI (multigrad/multigrad.py:149-183).
#include <cuda.h>
#include <stdlib.h>
#include <string.h>
#include "mgb_common.cuh"


namespace {

// The driver API is reached through cudaGetDriverEntryPoint, never through a
// link-time symbol: libmgb200.so must stay loadable on a machine without
// libcuda.so.1 (the CPU-only build check imports the library).
#define MGB_DRV_LIST(X)                                                       \
  X(cuDeviceGet)                                                 \
  X(cuDeviceGetAttribute)                                        \
  X(cuGetErrorString)                                            \
  X(cuMulticastGetGranularity)                                   \
  X(cuMemGetAllocationGranularity)                               \
  X(cuMulticastCreate)                                           \
  X(cuMemExportToShareableHandle)                                \
  X(cuMemImportFromShareableHandle)                              \
  X(cuMulticastAddDevice)                                        \
  X(cuMemCreate)                                                 \
  X(cuMulticastBindMem)                                          \
  X(cuMemAddressReserve)                                         \
  X(cuMemMap)                                                    \
  X(cuMemSetAccess)                                              \
  X(cuMemUnmap)                                                  \
  X(cuMemAddressFree)                                            \
  X(cuMulticastUnbind)                                           \
  X(cuMemRelease)

struct Drv {
#define MGB_DRV_MEMBER(name) decltype(&name) p_##name = nullptr;
  MGB_DRV_LIST(MGB_DRV_MEMBER)
#undef MGB_DRV_MEMBER
  bool ok = false;
};

Drv& drv() {
  static Drv d;
  static bool tried = false;
  if (!tried) {
    tried = true;
    bool all = true;
#define MGB_DRV_LOAD(name)                                                    \
    {                                                                         \
      void* fn = nullptr;                                                     \
      cudaDriverEntryPointQueryResult q;                                      \
      if (cudaGetDriverEntryPoint(#name, &fn, cudaEnableDefault, &q) !=       \
              cudaSuccess ||                                                  \
          fn == nullptr || q != cudaDriverEntryPointSuccess)                  \
        all = false;                                                          \
      d.p_##name = reinterpret_cast<decltype(d.p_##name)>(fn);                \
    }
    MGB_DRV_LIST(MGB_DRV_LOAD)
#undef MGB_DRV_LOAD
    d.ok = all;
  }
  return d;
}


#define MGB_CU_CHECK(expr)                                                    \
  do {                                                                        \
    CUresult _r = (expr);                                                     \
    if (_r != CUDA_SUCCESS) {                                                 \
      const char* _s = nullptr;                                               \
      drv().p_cuGetErrorString(_r, &_s);                                              \
      mgb_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr,             \
                    _s ? _s : "unknown CUresult");                            \
      return MGB_ESTATE;                                                      \
    }                                                                         \
  } while (0)

int current_device(CUdevice* dev, int* ordinal) {
  int d = 0;
  MGB_CUDA_CHECK(cudaGetDevice(&d));
  MGB_CUDA_CHECK(cudaFree(0));          // make sure the primary context exists
  MGB_REQUIRE(drv().ok, MGB_ESTATE,
              "the CUDA driver does not export the virtual-memory / multicast "
              "entry points (cudaGetDriverEntryPoint failed)");
  MGB_CU_CHECK(drv().p_cuDeviceGet(dev, d));
  if (ordinal) *ordinal = d;
  return MGB_OK;
}

CUmulticastObjectProp mc_prop(int world, uint64_t bytes) {
  CUmulticastObjectProp p;
  memset(&p, 0, sizeof(p));
  p.numDevices = (unsigned)world;
  p.size = bytes;
  p.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return p;
}

}  // namespace

extern "C" int mgb_mc_supported(int* out) {
  MGB_REQUIRE(out, MGB_EINVAL, "mgb_mc_supported: null argument");
  CUdevice dev;
  int rc = current_device(&dev, nullptr);
  if (rc != MGB_OK) return rc;
  int v = 0;
  MGB_CU_CHECK(drv().p_cuDeviceGetAttribute(
      &v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev));
  *out = v;
  return MGB_OK;
}

// Smallest size >= bytes that a multicast object of `world` devices accepts.
extern "C" int mgb_mc_round_size(int world, uint64_t bytes, uint64_t* out) {
  MGB_REQUIRE(out && world >= 2, MGB_EINVAL, "mgb_mc_round_size: bad args");
  CUmulticastObjectProp p = mc_prop(world, bytes);
  size_t gran = 0;
  MGB_CU_CHECK(drv().p_cuMulticastGetGranularity(&gran, &p,
                                         CU_MULTICAST_GRANULARITY_MINIMUM));
  // the physical allocation must satisfy its own granularity too
  CUdevice dev;
  int ord = 0;
  int rc = current_device(&dev, &ord);
  if (rc != MGB_OK) return rc;
  CUmemAllocationProp ap;
  memset(&ap, 0, sizeof(ap));
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ap.location.id = ord;
  ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t g2 = 0;
  MGB_CU_CHECK(drv().p_cuMemGetAllocationGranularity(
      &g2, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  if (g2 > gran) gran = g2;
  *out = (bytes + gran - 1) / gran * gran;
  return MGB_OK;
}

extern "C" int mgb_mc_create(int world, uint64_t bytes, uint64_t* handle_out,
                             int* fd_out) {
  MGB_REQUIRE(handle_out && fd_out && world >= 2, MGB_EINVAL,
              "mgb_mc_create: bad arguments");
  CUdevice dev;
  int rc = current_device(&dev, nullptr);
  if (rc != MGB_OK) return rc;
  CUmulticastObjectProp p = mc_prop(world, bytes);
  CUmemGenericAllocationHandle h;
  MGB_CU_CHECK(drv().p_cuMulticastCreate(&h, &p));
  int fd = -1;
  MGB_CU_CHECK(drv().p_cuMemExportToShareableHandle(
      &fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *handle_out = (uint64_t)h;
  *fd_out = fd;
  return MGB_OK;
}

extern "C" int mgb_mc_import(int fd, uint64_t* handle_out) {
  MGB_REQUIRE(handle_out && fd >= 0, MGB_EINVAL, "mgb_mc_import: bad args");
  CUdevice dev;
  int rc = current_device(&dev, nullptr);
  if (rc != MGB_OK) return rc;
  CUmemGenericAllocationHandle h;
  MGB_CU_CHECK(drv().p_cuMemImportFromShareableHandle(
      &h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  *handle_out = (uint64_t)h;
  return MGB_OK;
}

extern "C" int mgb_mc_add_device(uint64_t handle) {
  CUdevice dev;
  int rc = current_device(&dev, nullptr);
  if (rc != MGB_OK) return rc;
  MGB_CU_CHECK(drv().p_cuMulticastAddDevice((CUmemGenericAllocationHandle)handle, dev));
  return MGB_OK;
}

// After EVERY rank has added its device: allocate this GPU's physical buffer,
// bind it at offset 0 of the multicast object and map both views.
extern "C" int mgb_mc_bind(uint64_t handle, uint64_t bytes, void** uc_out,
                           void** mc_out, uint64_t* mem_out) {
  MGB_REQUIRE(uc_out && mc_out && mem_out, MGB_EINVAL, "mgb_mc_bind: bad args");
  CUdevice dev;
  int ord = 0;
  int rc = current_device(&dev, &ord);
  if (rc != MGB_OK) return rc;
  CUmemGenericAllocationHandle mc = (CUmemGenericAllocationHandle)handle;
  CUmemAllocationProp ap;
  memset(&ap, 0, sizeof(ap));
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ap.location.id = ord;
  ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle mem;
  MGB_CU_CHECK(drv().p_cuMemCreate(&mem, bytes, &ap, 0));
  MGB_CU_CHECK(drv().p_cuMulticastBindMem(mc, 0, mem, 0, bytes, 0));
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = ord;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  CUdeviceptr uc = 0, mcva = 0;
  MGB_CU_CHECK(drv().p_cuMemAddressReserve(&uc, bytes, 0, 0, 0));
  MGB_CU_CHECK(drv().p_cuMemMap(uc, bytes, 0, mem, 0));
  MGB_CU_CHECK(drv().p_cuMemSetAccess(uc, bytes, &acc, 1));
  MGB_CU_CHECK(drv().p_cuMemAddressReserve(&mcva, bytes, 0, 0, 0));
  MGB_CU_CHECK(drv().p_cuMemMap(mcva, bytes, 0, mc, 0));
  MGB_CU_CHECK(drv().p_cuMemSetAccess(mcva, bytes, &acc, 1));
  MGB_CUDA_CHECK(cudaMemset((void*)uc, 0, bytes));
  MGB_CUDA_CHECK(cudaDeviceSynchronize());
  *uc_out = (void*)uc;
  *mc_out = (void*)mcva;
  *mem_out = (uint64_t)mem;
  return MGB_OK;
}

extern "C" int mgb_mc_release(uint64_t handle, uint64_t mem, void* uc,
                              void* mc, uint64_t bytes) {
  CUdevice dev;
  int rc = current_device(&dev, nullptr);
  if (rc != MGB_OK) return rc;
  if (mc) {
    drv().p_cuMemUnmap((CUdeviceptr)mc, bytes);
    drv().p_cuMemAddressFree((CUdeviceptr)mc, bytes);
  }
  if (uc) {
    drv().p_cuMemUnmap((CUdeviceptr)uc, bytes);
    drv().p_cuMemAddressFree((CUdeviceptr)uc, bytes);
  }
  if (handle && mem)
    drv().p_cuMulticastUnbind((CUmemGenericAllocationHandle)handle, dev, 0, bytes);
  if (mem) drv().p_cuMemRelease((CUmemGenericAllocationHandle)mem);
  if (handle) drv().p_cuMemRelease((CUmemGenericAllocationHandle)handle);
  return MGB_OK;
}

// ---------------------------------------------------------------- the kernel
namespace {

#define NV_BLOCK 512
#define NV_MAX_CTAS 296
#ifndef NV_UNROLL
#define NV_UNROLL 8
#endif
#define NV_RESIDENT_CTAS 16
#ifndef NV_PIPE_DEFAULT
#define NV_PIPE_DEFAULT 0
#endif

__device__ __forceinline__ float4 mc_ld_reduce_f4(const float4* p) {
  float4 v;
  asm volatile(
      "multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
      : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
      : "l"(p)
      : "memory");
  return v;
}
__device__ __forceinline__ void mc_st_f4(float4* p, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::
                   "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

// CTA-local exchange of "range c is ready" flags between the ranks (LL slot
// unit `unit`, tagged with the epoch).  All threads of the CTA call it.
__device__ __forceinline__ void cta_rank_barrier(const MgbPeerTable& pt,
                                                 unsigned epoch,
                                                 unsigned unit) {
  const int tid = threadIdx.x, W = pt.world, me = pt.rank;
  const unsigned parity = epoch & 1u;
  // release/acquire by the signalling threads only (cumulative system
  // fences after / before the CTA barrier): a system fence from all 512
  // threads costs microseconds per barrier
  __syncthreads();
  if (tid < W - 1) {
    int dst = me + 1 + tid;
    if (dst >= W) dst -= W;
    __threadfence_system();
    mgb_st_sys_u64(mgb_ll_slot(pt, dst, parity, me, unit),
                   (unsigned long long)epoch << 32);
  }
  if (tid < W && tid != me) {
    (void)mgb_ll_wait(pt, mgb_ll_slot(pt, me, parity, tid, unit), epoch);
    __threadfence_system();
  }
  __syncthreads();
}

__device__ __forceinline__ float mc_ld_reduce_f1(const float* p) {
  float v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f32 %0, [%1];"
               : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void mc_st_f1(float* p, float v) {
  asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(p), "f"(v)
               : "memory");
}

// In-switch all-reduce, float32.  CTA c of every rank owns the same element
// range; inside it rank r reduces sub-slice r.  Three CTA-local phases, no
// grid-wide barrier:
//   1. stage my input range into my (unicast view of the) bound buffer
//   2. after every rank staged range c: multimem.ld_reduce my sub-slice and
//      multimem.st the sums to every GPU's buffer
//   3. after every rank published: copy range c of my buffer to the output
// INPLACE: the tensor already lives in the bound buffer (see
// PeerWorkspace.multicast_tensor): phases 1 and 3 disappear and the message
// crosses only the switch.
// The buffer needs no parity: a peer's phase-2 access of call k ends before it
// signals, and I stage call k+1 only after seeing every signal of call k.
template <bool INPLACE>
__global__ void __launch_bounds__(NV_BLOCK, 2)
nvls_allreduce_f32_kernel(const MgbPeerTable pt, float4* __restrict__ uc,
                          float4* __restrict__ mc, float* __restrict__ data,
                          long long n4, long long tail_start, int ntail) {
  __shared__ unsigned s_epoch;
  const int tid = threadIdx.x;
  const int W = pt.world, me = pt.rank;
  unsigned* ctrl = mgb_ctrl(pt, me);
  if (tid == 0) s_epoch = mgb_ld_sys_u32(ctrl + MGB_CTRL_EPOCH) + 1u;
  __syncthreads();
  const unsigned epoch = s_epoch;
  // INPLACE: whole float4s only, the ragged tail is reduced with scalar
  // multimem operations (the floats after it may belong to another tensor)
  const long long n4t = INPLACE ? n4 : n4 + (ntail ? 1 : 0);
  const long long per = (n4t + gridDim.x - 1) / gridDim.x;
  const long long lo = per * blockIdx.x;
  const long long hi = min(lo + per, n4t);
  float4* d4 = reinterpret_cast<float4*>(data);
  const long long hi_full = min(hi, n4);   // whole float4s of my range
  if (!INPLACE) {
    long long i = lo + tid;
    for (; i + 3 * NV_BLOCK < hi_full; i += 4 * NV_BLOCK) {
      float4 v[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) v[k] = d4[i + k * NV_BLOCK];
#pragma unroll
      for (int k = 0; k < 4; ++k) uc[i + k * NV_BLOCK] = v[k];
    }
    for (; i < hi; i += NV_BLOCK) {
      float4 v;
      if (i < n4) {
        v = d4[i];
      } else {
        v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ntail > 0) v.x = data[tail_start];
        if (ntail > 1) v.y = data[tail_start + 1];
        if (ntail > 2) v.z = data[tail_start + 2];
      }
      uc[i] = v;
    }
  }
  cta_rank_barrier(pt, epoch, blockIdx.x);
  {
    const long long len = hi > lo ? hi - lo : 0;
    const long long sub = (len + W - 1) / W;
    const long long slo = lo + sub * me;
    const long long shi = min(slo + sub, hi);
    // NV_UNROLL independent in-switch reductions in flight per thread, the
    // ragged end predicated instead of falling back to one at a time (a
    // 64 MiB message leaves each CTA and rank ~1800 float4s: less than one
    // full unrolled pass, so the old remainder loop did all the work with a
    // single outstanding request per thread)
    for (long long i = slo + tid; i < shi; i += (long long)NV_UNROLL * NV_BLOCK) {
      float4 v[NV_UNROLL];
#pragma unroll
      for (int k = 0; k < NV_UNROLL; ++k)
        if (i + k * NV_BLOCK < shi) v[k] = mc_ld_reduce_f4(mc + i + k * NV_BLOCK);
#pragma unroll
      for (int k = 0; k < NV_UNROLL; ++k)
        if (i + k * NV_BLOCK < shi) mc_st_f4(mc + i + k * NV_BLOCK, v[k]);
    }
    if (INPLACE && blockIdx.x == 0 && me == 0 && tid < ntail) {
      float* p = reinterpret_cast<float*>(mc) + tail_start + tid;
      mc_st_f1(p, mc_ld_reduce_f1(p));
    }
  }
  cta_rank_barrier(pt, epoch, NV_MAX_CTAS + blockIdx.x);
  if (!INPLACE) {
    long long i = lo + tid;
    for (; i + 3 * NV_BLOCK < hi_full; i += 4 * NV_BLOCK) {
      float4 v[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) v[k] = __ldcv(uc + i + k * NV_BLOCK);
#pragma unroll
      for (int k = 0; k < 4; ++k) d4[i + k * NV_BLOCK] = v[k];
    }
    for (; i < hi; i += NV_BLOCK) {
      const float4 v = __ldcv(uc + i);
      if (i < n4) {
        d4[i] = v;
      } else {
        if (ntail > 0) data[tail_start] = v.x;
        if (ntail > 1) data[tail_start + 1] = v.y;
        if (ntail > 2) data[tail_start + 2] = v.z;
      }
    }
  }
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


// ---- staged all-reduce as a pipeline (large ordinary tensors) --------------
// The copy phases want every SM, the in-switch phase a few CTAs with many
// requests in flight (see NV_RESIDENT_CTAS), and in the kernel above neither
// overlaps the other.  Here the roles are split inside ONE co-resident grid:
//   copier CTAs  each stage one part of one 2 MiB chunk into the bound buffer,
//                count themselves in on staged[c], and copy the same part
//                back out as soon as reduced[c] carries the epoch
//   reducer CTAs own chunks c = r, r + R, ...: wait for the local copiers of
//                the chunk, exchange "staged" flags with the peers, reduce the
//                rank's sub-slice through the switch, exchange "reduced"
//                flags, release the chunk to the copiers
// so chunk c-R is copied out while chunk c crosses the switch.  OPT-IN
// (MULTIGRAD_B200_NVLS_PIPELINE=1): measured on 2 GPUs only, 148 -> 123 us at
// 32 MiB and 253 -> 234 us at 64 MiB against the kernel above.  The grid is launched cooperatively: every CTA has to be resident
// because reducers wait for ALL copiers of a chunk.
#define NV_PIPE_CHUNK4 131072          // float4 per chunk (2 MiB)
#define NV_PIPE_MAX_CHUNKS 64
#define NV_PIPE_REDUCERS 16
#define NV_CTRL_STAGED 64              // uint32 words of the control block
#define NV_CTRL_REDUCED 128

__device__ __forceinline__ unsigned ld_acquire_gpu_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu_u32(unsigned* p, unsigned v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v)
               : "memory");
}
// thread 0 of a CTA waits until *p == want (bounded like every peer wait)
__device__ __forceinline__ void cta_wait_word(const MgbPeerTable& pt,
                                              const unsigned* p,
                                              unsigned want) {
  if (threadIdx.x == 0 && ld_acquire_gpu_u32(p) != want) {
    const unsigned long long t0 = mgb_globaltimer_ns();
    const unsigned long long limit =
        (unsigned long long)pt.spin_timeout_ms * 1000000ull;
    unsigned it = 0;
    while (ld_acquire_gpu_u32(p) != want) {
      if ((++it & 63u) == 0u) {
        if (mgb_globaltimer_ns() - t0 > limit) {
          mgb_st_sys_u32(mgb_ctrl(pt, pt.rank) + MGB_CTRL_ERROR, 1u);
          break;
        }
        __nanosleep(100);
      }
    }
  }
  __syncthreads();
}

__global__ void __launch_bounds__(NV_BLOCK, 2)
nvls_pipelined_f32_kernel(const MgbPeerTable pt, float4* __restrict__ uc,
                          float4* __restrict__ mc, float* __restrict__ data,
                          long long n4, long long tail_start, int ntail,
                          int nred) {
  __shared__ unsigned s_epoch;
  const int tid = threadIdx.x;
  const int W = pt.world, me = pt.rank;
  unsigned* ctrl = mgb_ctrl(pt, me);
  if (tid == 0) s_epoch = mgb_ld_sys_u32(ctrl + MGB_CTRL_EPOCH) + 1u;
  __syncthreads();
  const unsigned epoch = s_epoch;
  unsigned* staged = ctrl + NV_CTRL_STAGED;
  unsigned* reduced = ctrl + NV_CTRL_REDUCED;
  const long long n4t = n4 + (ntail ? 1 : 0);
  const int nchunks = (int)((n4t + NV_PIPE_CHUNK4 - 1) / NV_PIPE_CHUNK4);
  const int ncop = (int)gridDim.x - nred;
  float4* d4 = reinterpret_cast<float4*>(data);
  if ((int)blockIdx.x < nred) {
    // ------------------------------------------------------------- reducer
    for (int c = blockIdx.x; c < nchunks; c += nred) {
      cta_wait_word(pt, staged + c, (unsigned)(ncop / nchunks));
      if (tid == 0) staged[c] = 0u;          // nobody adds again in this call
      cta_rank_barrier(pt, epoch, (unsigned)c);
      const long long lo = (long long)c * NV_PIPE_CHUNK4;
      const long long hi = min(lo + (long long)NV_PIPE_CHUNK4, n4t);
      const long long sub = (hi - lo + W - 1) / W;
      const long long slo = lo + sub * me;
      const long long shi = min(slo + sub, hi);
      for (long long i = slo + tid; i < shi;
           i += (long long)NV_UNROLL * NV_BLOCK) {
        float4 v[NV_UNROLL];
#pragma unroll
        for (int k = 0; k < NV_UNROLL; ++k)
          if (i + k * NV_BLOCK < shi)
            v[k] = mc_ld_reduce_f4(mc + i + k * NV_BLOCK);
#pragma unroll
        for (int k = 0; k < NV_UNROLL; ++k)
          if (i + k * NV_BLOCK < shi) mc_st_f4(mc + i + k * NV_BLOCK, v[k]);
      }
      cta_rank_barrier(pt, epoch, (unsigned)(NV_MAX_CTAS + c));
      if (tid == 0) st_release_gpu_u32(reduced + c, epoch);
    }
  } else {
    // -------------------------------------------------------------- copier
    // copier k owns part k / nchunks of chunk k % nchunks: all chunks are
    // staged side by side (a schedule in rounds of nred chunks, staging the
    // first round first, measured slower: 249 vs 234 us at 64 MiB on 2 GPUs)
    const int k = (int)blockIdx.x - nred;
    const int parts = ncop / nchunks;
    if (k < parts * nchunks) {
      const int c = k % nchunks, part = k / nchunks;
      const long long lo = (long long)c * NV_PIPE_CHUNK4;
      const long long hi = min(lo + (long long)NV_PIPE_CHUNK4, n4t);
      const long long k0 = (long long)part * NV_BLOCK + tid;
      const long long stride = (long long)parts * NV_BLOCK;
      for (long long i = lo + k0; i < hi; i += 4 * stride) {
        float4 v[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const long long j = i + q * stride;
          if (j < hi) {
            if (j < n4) {
              v[q] = d4[j];
            } else {
              v[q] = make_float4(0.f, 0.f, 0.f, 0.f);
              if (ntail > 0) v[q].x = data[tail_start];
              if (ntail > 1) v[q].y = data[tail_start + 1];
              if (ntail > 2) v[q].z = data[tail_start + 2];
            }
          }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (i + q * stride < hi) uc[i + q * stride] = v[q];
      }
      __syncthreads();
      if (tid == 0) {
        __threadfence_system();
        atomicAdd(staged + c, 1u);
      }
      cta_wait_word(pt, reduced + c, epoch);
      for (long long i = lo + k0; i < hi; i += 4 * stride) {
        float4 v[4];
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (i + q * stride < hi) v[q] = __ldcv(uc + i + q * stride);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const long long j = i + q * stride;
          if (j < hi) {
            if (j < n4) {
              d4[j] = v[q];
            } else {
              if (ntail > 0) data[tail_start] = v[q].x;
              if (ntail > 1) data[tail_start + 1] = v[q].y;
              if (ntail > 2) data[tail_start + 2] = v[q].z;
            }
          }
        }
      }
    }
  }
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

// co-resident grid of the pipelined kernel on the current device (0: none)
int nvls_pipe_grid() {
  int dev = 0, sms = 0, per_sm = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) !=
      cudaSuccess)
    return 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(
          &per_sm, nvls_pipelined_f32_kernel, NV_BLOCK, 0) != cudaSuccess)
    return 0;
  long long g = (long long)sms * per_sm;
  if (g > NV_MAX_CTAS) g = NV_MAX_CTAS;
  return (int)g;
}

}  // namespace

extern "C" int mgb_allreduce_sum_nvls(const MgbPeerTable* pt, void* uc,
                                      void* mc, uint64_t mc_bytes, float* data,
                                      uint64_t n, void* stream) {
  MGB_REQUIRE(pt && uc && mc && data, MGB_EINVAL,
              "mgb_allreduce_sum_nvls: null argument");
  MGB_REQUIRE(pt->world >= 2 && pt->world <= MGB_MAX_WORLD, MGB_EINVAL,
              "mgb_allreduce_sum_nvls: bad world size %d", pt->world);
  if (n == 0) return MGB_OK;
  MGB_REQUIRE(((uintptr_t)data & 15u) == 0, MGB_EINVAL,
              "mgb_allreduce_sum_nvls: data must be 16-byte aligned");
  const uint64_t need = ((n + 3ull) / 4ull) * 16ull;
  MGB_REQUIRE(need <= mc_bytes, MGB_ETOOBIG,
              "mgb_allreduce_sum_nvls: %llu bytes exceed the multicast buffer "
              "(%llu)", (unsigned long long)need, (unsigned long long)mc_bytes);
  MGB_REQUIRE(pt->ll_units >= 2u * NV_MAX_CTAS, MGB_EINVAL,
              "mgb_allreduce_sum_nvls: needs %d LL units", 2 * NV_MAX_CTAS);
  long long n4 = (long long)(n / 4ull);
  long long tail_start = n4 * 4;
  int ntail = (int)(n - (uint64_t)tail_start);
  const long long n4t = n4 + (ntail ? 1 : 0);
  // a tensor that already lives in the bound buffer is reduced in place
  const uintptr_t d0 = (uintptr_t)data, u0 = (uintptr_t)uc;
  const bool inplace = d0 >= u0 && d0 + n * 4ull <= u0 + mc_bytes;
  // ~32 KiB per CTA and pass, at most NV_MAX_CTAS (one flag unit each; two
  // CTAs per SM are resident, every CTA waits for its peers)
  long long ctas = (n4t + 4 * NV_BLOCK - 1) / (4 * NV_BLOCK);
  if (ctas > NV_MAX_CTAS) ctas = NV_MAX_CTAS;
  // Resident tensors have no copy phases; their in-switch phase runs best
  // from few CTAs with many requests in flight each (8 GPUs, 64 MiB: 171 us
  // with 16 CTAs, 185 with 64, 203 with 296; tools/nvls_tune.py)
  if (inplace && ctas > NV_RESIDENT_CTAS) ctas = NV_RESIDENT_CTAS;
  if (const char* e = getenv("MULTIGRAD_B200_NVLS_CTAS")) {   // tuning aid
    const long long c = atoll(e);
    if (c >= 1 && c <= NV_MAX_CTAS && c < ctas) ctas = c;
  }
  if (ctas < 1) ctas = 1;
  // large ordinary tensors: copy phases and switch phase pipelined per chunk
  {
    static const int pipe_mode = [] {
      const char* e = getenv("MULTIGRAD_B200_NVLS_PIPELINE");
      return e ? atoi(e) : NV_PIPE_DEFAULT;
    }();
    const long long nchunks = (n4t + NV_PIPE_CHUNK4 - 1) / NV_PIPE_CHUNK4;
    if (!inplace && pipe_mode && nchunks >= NV_PIPE_REDUCERS &&
        nchunks <= NV_PIPE_MAX_CHUNKS) {
      const int grid = nvls_pipe_grid();
      if (grid >= 4 * NV_PIPE_REDUCERS &&
          grid - NV_PIPE_REDUCERS >= (int)nchunks) {   // >= 1 copier per chunk
        MgbPeerTable ptv = *pt;
        float4* ucp = (float4*)uc;
        float4* mcp = (float4*)mc;
        int nred = NV_PIPE_REDUCERS;
        void* args[] = {&ptv, &ucp, &mcp, &data, &n4, &tail_start, &ntail,
                        &nred};
        MGB_CUDA_CHECK(cudaLaunchCooperativeKernel(
            (void*)nvls_pipelined_f32_kernel, dim3(grid), dim3(NV_BLOCK), args,
            0, (cudaStream_t)stream));
        return MGB_OK;
      }
    }
  }
  if (inplace) {
    float4* mcd = (float4*)((char*)mc + (d0 - u0));
    nvls_allreduce_f32_kernel<true><<<(unsigned)ctas, NV_BLOCK, 0,
                                      (cudaStream_t)stream>>>(
        *pt, (float4*)data, mcd, data, n4, tail_start, ntail);
  } else {
    nvls_allreduce_f32_kernel<false><<<(unsigned)ctas, NV_BLOCK, 0,
                                       (cudaStream_t)stream>>>(
        *pt, (float4*)uc, (float4*)mc, data, n4, tail_start, ntail);
  }
  MGB_CUDA_CHECK(cudaGetLastError());
  return MGB_OK;
}
