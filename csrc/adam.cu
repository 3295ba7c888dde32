// K3 (stand-alone): Adam update on device vectors for the generic path, and
// K8: pointer jumping.  sm_100a.
//
// Reference counterparts: jax_opt.adam's opt_update as driven by
// _adam_optimizer (multigrad/adam.py:52-68); find_ultimate_top_indices
// (multigrad/diffdesi_experimental/util.py:18-28).
#include "mgb_common.cuh"

namespace {

__global__ void adam_update_kernel(float* __restrict__ params,
                                   const float* __restrict__ grad,
                                   float* __restrict__ m,
                                   float* __restrict__ v, int step, int n,
                                   float lr, float b1, float b2, float eps) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float c1 = 1.0f - powf(b1, (float)(step + 1));
  const float c2 = 1.0f - powf(b2, (float)(step + 1));
  const float g = grad[i];
  const float mi = (1.0f - b1) * g + b1 * m[i];
  const float vi = (1.0f - b2) * g * g + b2 * v[i];
  m[i] = mi;
  v[i] = vi;
  params[i] = params[i] - lr * (mi / c1) / (sqrtf(vi / c2) + eps);
}

__global__ void pointer_jump_kernel(const long long* __restrict__ in,
                                    long long* __restrict__ out, long long n,
                                    int* __restrict__ changed) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const long long a = in[i];
  const long long b = in[a];
  out[i] = b;
  if (a != b) *changed = 1;
}

}  // namespace

extern "C" int mgb_adam_update(float* params, const float* grad, float* m,
                               float* v, int32_t step, int32_t n, float lr,
                               float b1, float b2, float eps, void* stream) {
  MGB_REQUIRE(params && grad && m && v && n >= 0 && step >= 0, MGB_EINVAL,
              "mgb_adam_update: bad arguments");
  if (n == 0) return MGB_OK;
  const int blk = 128;
  adam_update_kernel<<<(n + blk - 1) / blk, blk, 0, (cudaStream_t)stream>>>(
      params, grad, m, v, step, n, lr, b1, b2, eps);
  MGB_CUDA_CHECK(cudaGetLastError());
  return MGB_OK;
}

extern "C" int mgb_pointer_jump(int64_t* indices, int64_t* tmp, int64_t n,
                                int32_t max_rounds, int32_t* rounds_out,
                                int32_t* changed_flag_dev, void* stream) {
  MGB_REQUIRE(indices && tmp && changed_flag_dev && n >= 0, MGB_EINVAL,
              "mgb_pointer_jump: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  int rounds = 0;
  if (n > 0) {
    long long* cur = (long long*)indices;
    long long* nxt = (long long*)tmp;
    const int blk = 256;
    const unsigned grid = (unsigned)((n + blk - 1) / blk);
    for (;;) {
      int changed = 0;
      MGB_CUDA_CHECK(cudaMemsetAsync(changed_flag_dev, 0, sizeof(int), st));
      pointer_jump_kernel<<<grid, blk, 0, st>>>(cur, nxt, (long long)n,
                                                changed_flag_dev);
      MGB_CUDA_CHECK(cudaGetLastError());
      MGB_CUDA_CHECK(cudaMemcpyAsync(&changed, changed_flag_dev, sizeof(int),
                                     cudaMemcpyDeviceToHost, st));
      MGB_CUDA_CHECK(cudaStreamSynchronize(st));
      long long* t = cur; cur = nxt; nxt = t;
      if (!changed) break;
      ++rounds;
      if (rounds > max_rounds) {
        if (rounds_out) *rounds_out = rounds;
        mgb_set_error("mgb_pointer_jump: not converged after %d rounds",
                      max_rounds);
        return MGB_ESTATE;
      }
    }
    if (cur != (long long*)indices)
      MGB_CUDA_CHECK(cudaMemcpyAsync(indices, cur, (size_t)n * 8,
                                     cudaMemcpyDeviceToDevice, st));
  }
  if (rounds_out) *rounds_out = rounds;
  return MGB_OK;
}
