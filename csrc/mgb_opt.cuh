// Optimiser epilogue shared by the fused step kernels (K3): bound transforms,
// Adam / fixed-step gradient descent on one parameter per thread.
//
// Reference counterparts: jax_opt.adam's update as driven by _adam_optimizer
// (multigrad/adam.py:52-68), the bound transforms (adam.py:192-239) and the
// fixed-step update of util.simple_grad_descent (util.py:100-114).
#pragma once
#include "mgb_common.cuh"

// kind: 0 = unbounded, 1 = (lo, hi), 2 = lo only, 3 = hi only
__device__ __forceinline__ float mgb_bound_inverse(int kind, float lo,
                                                   float hi, float u) {
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
__device__ __forceinline__ float mgb_bound_jac(int kind, float lo, float hi,
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

// Update parameter `i` of `P` with gradient `grad` (dL/dp_i).  `step` is the
// number of updates applied so far (the bias-correction index), `row` the row
// of this step in the trajectory buffer.  `Args` is MgbStepArgs or
// MgbLinArgs: both carry mode, opt_p/u/m/v, traj, traj_rows, lr, b1, b2, eps
// and jac_at_params under the same names.
//
// Trajectory convention: GD records the parameters *before* the update in
// row `row` (util.py:105-110); Adam records the parameters *after* it in row
// `row + 1` (row 0 is the initial guess, written by the host).
// The core takes the state of parameter `i` in registers (p, u, m0, v0) so
// that a caller can have loaded it long before the gradient is known.
template <class Args>
__device__ __forceinline__ void mgb_opt_update_param_pre(
    const Args& a, int i, int P, int step, int row, float grad, int kind,
    float lo, float hi, float p, float u, float m0, float v0) {
  if (a.mode == MGB_MODE_GD) {
    if (a.traj && row >= 0 && row < a.traj_rows)
      a.traj[(size_t)row * P + i] = p;
    a.opt_p[i] = p - a.lr * grad;
    return;
  }
  const float c1 = 1.0f - powf(a.b1, (float)(step + 1));
  const float c2 = 1.0f - powf(a.b2, (float)(step + 1));
  const float gi = grad * mgb_bound_jac(kind, lo, hi, a.jac_at_params ? p : u);
  const float m = (1.0f - a.b1) * gi + a.b1 * m0;
  const float v = (1.0f - a.b2) * gi * gi + a.b2 * v0;
  const float mhat = m / c1, vhat = v / c2;
  u = u - a.lr * mhat / (sqrtf(vhat) + a.eps);
  const float pn = mgb_bound_inverse(kind, lo, hi, u);
  a.opt_m[i] = m;
  a.opt_v[i] = v;
  a.opt_u[i] = u;
  a.opt_p[i] = pn;
  if (a.traj && row + 1 >= 0 && row + 1 < a.traj_rows)
    a.traj[(size_t)(row + 1) * P + i] = pn;
}

template <class Args>
__device__ __forceinline__ void mgb_opt_update_param(
    const Args& a, int i, int P, int step, int row, float grad, int kind,
    float lo, float hi) {
  const float p = a.opt_p[i];
  float u = 0.f, m0 = 0.f, v0 = 0.f;
  if (a.mode != MGB_MODE_GD) {
    u = a.opt_u[i];
    m0 = a.opt_m[i];
    v0 = a.opt_v[i];
  }
  mgb_opt_update_param_pre(a, i, P, step, row, grad, kind, lo, hi, p, u, m0,
                           v0);
}
