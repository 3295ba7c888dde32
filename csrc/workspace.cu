// N1: peer workspace manager -- allocation, CUDA-IPC export/import, control
// words.  Host side only; the device-side protocol lives in mgb_common.cuh.
#include <stdarg.h>
#include <string.h>
#include "mgb_common.cuh"

static thread_local char g_err[512] = "";

void mgb_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* mgb_last_error(void) { return g_err; }
extern "C" int mgb_version(void) { return 100; }

// sizeof of every by-value struct of the C ABI, in declaration order, so that
// a binding can detect a stale mirror at load time instead of corrupting
// kernel arguments.
extern "C" int mgb_abi_sizes(uint32_t* out, int n) {
  const uint32_t sizes[] = {
      (uint32_t)sizeof(MgbPeerTable), (uint32_t)sizeof(MgbErfBinModel),
      (uint32_t)sizeof(MgbStepArgs),  (uint32_t)sizeof(MgbGroupArgs),
      (uint32_t)sizeof(MgbLinModel),  (uint32_t)sizeof(MgbLinArgs)};
  const int have = (int)(sizeof(sizes) / sizeof(sizes[0]));
  for (int i = 0; i < n && i < have; ++i) out[i] = sizes[i];
  return have;
}

extern "C" uint64_t mgb_ws_bytes(int world, uint32_t ll_units,
                                 uint64_t bulk_bytes) {
  // two bulk regions + a flag page per region
  uint64_t bulk = (bulk_bytes + 1023ull) & ~1023ull;
  return mgb_bulk_off(world, ll_units) + 2ull * bulk + 8192ull;
}

extern "C" int mgb_ws_alloc(uint64_t bytes, void** out_ptr) {
  MGB_REQUIRE(out_ptr != nullptr && bytes > 0, MGB_EINVAL,
              "mgb_ws_alloc: bad arguments");
  void* p = nullptr;
  MGB_CUDA_CHECK(cudaMalloc(&p, bytes));
  MGB_CUDA_CHECK(cudaMemset(p, 0, bytes));
  MGB_CUDA_CHECK(cudaDeviceSynchronize());
  *out_ptr = p;
  return MGB_OK;
}

extern "C" int mgb_ws_free(void* ptr) {
  if (ptr) MGB_CUDA_CHECK(cudaFree(ptr));
  return MGB_OK;
}

extern "C" int mgb_ws_export(void* ptr, uint8_t handle_out[64]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  cudaIpcMemHandle_t h;
  MGB_CUDA_CHECK(cudaIpcGetMemHandle(&h, ptr));
  memcpy(handle_out, &h, 64);
  return MGB_OK;
}

extern "C" int mgb_ws_import(const uint8_t handle[64], void** out_ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  void* p = nullptr;
  MGB_CUDA_CHECK(
      cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *out_ptr = p;
  return MGB_OK;
}

extern "C" int mgb_ws_unimport(void* ptr) {
  if (ptr) MGB_CUDA_CHECK(cudaIpcCloseMemHandle(ptr));
  return MGB_OK;
}

extern "C" int mgb_ws_read_ctrl(const void* local_base, int idx, uint32_t* out,
                                void* stream) {
  MGB_REQUIRE(local_base && out && idx >= 0 && idx < 1024, MGB_EINVAL,
              "mgb_ws_read_ctrl: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  MGB_CUDA_CHECK(cudaMemcpyAsync(out, (const uint32_t*)local_base + idx,
                                 sizeof(uint32_t), cudaMemcpyDeviceToHost,
                                 st));
  MGB_CUDA_CHECK(cudaStreamSynchronize(st));
  return MGB_OK;
}

// Byte offset of bulk region A inside a workspace block (region B follows at
// + round_up(bulk_bytes, 1024)).
extern "C" uint64_t mgb_ws_bulk_offset(int world, uint32_t ll_units) {
  return mgb_bulk_off(world, ll_units);
}

// K7: device-to-device copy that also accepts peer-mapped (CUDA IPC) pointers,
// i.e. a copy over NVLink when `src` lives in another rank's workspace.
// Replaces the pickled send/recv of util.scatter_nd (multigrad/util.py:65-77).
extern "C" int mgb_peer_copy(void* dst, const void* src, uint64_t bytes,
                             void* stream) {
  MGB_REQUIRE(dst && src, MGB_EINVAL, "mgb_peer_copy: null argument");
  if (bytes == 0) return MGB_OK;
  MGB_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice,
                                 (cudaStream_t)stream));
  return MGB_OK;
}
