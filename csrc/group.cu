// K4 + K3: cross-group loss/gradient reduction fused with the optimiser
// update, for OnePointGroup fits whose member models evaluate on the device.
// sm_100a.
//
// Reference counterpart: OnePointGroup.calc_loss_and_grad_from_params
// (multigrad/multigrad.py:560-581) has every rank allgather pickled Python
// lists of per-model losses and gradients over the main communicator and sum
// them on the host, after which rank 0 runs the optimiser and broadcasts the
// next parameters (adam.py:39-49).  Here the member models' fused kernels
// leave (loss, grad) on the device; ONE small kernel per step then
//   * sums this rank's contributions (only the root of a model's
//     sub-communicator contributes, the others add zero),
//   * all-reduces the 1+P doubles over the main communicator with the ordered
//     LL exchange over NVLink peer memory (bit-identical totals everywhere),
//   * applies the replicated Adam / gradient-descent update and records the
//     trajectory,
// so that the next step's model kernels find their parameters already in
// place and the host never reads anything back during the fit.
#include "mgb_common.cuh"
#include "mgb_opt.cuh"

namespace {

constexpr int GP_BLOCK = 64;
constexpr int GP_M = 1 + MGB_MAX_PARAMS;

__global__ void __launch_bounds__(GP_BLOCK, 1)
group_step_kernel(const MgbGroupArgs g, const MgbStepArgs a,
                  const MgbPeerTable pt) {
  __shared__ double sh[GP_M];
  __shared__ double stage[MGB_MAX_WORLD * GP_M];
  const int tid = threadIdx.x;
  const int P = g.nparams;
  const int m = 1 + P;
  if (tid < m) {
    double v = 0.0;
    for (int k = 0; k < g.n_models; ++k) {
      if (!g.active[k]) continue;
      v += (double)(tid == 0 ? g.loss[k][0] : g.grad[k][tid - 1]);
    }
    sh[tid] = v;
  }
  __syncthreads();
  mgb_ll_allreduce_f64(pt, sh, stage, m);   // no-op for world == 1
  __syncthreads();
  const float loss = (float)sh[0];
  const bool mine = tid < P;
  const float grad = mine ? (float)sh[1 + tid] : 0.f;
  if (tid == 0 && a.out_loss) a.out_loss[0] = loss;
  if (mine && a.out_grad) a.out_grad[tid] = grad;
  if (a.mode < MGB_MODE_ADAM) return;
  const int step = *a.opt_step;
  const int row = step - a.step_base;
  if (tid == 0 && a.loss_hist && row >= 0 && row < a.traj_rows)
    a.loss_hist[row] = loss;
  if (mine)
    mgb_opt_update_param(a, tid, P, step, row, grad, a.bound_kind[tid],
                         a.bound_lo[tid], a.bound_hi[tid]);
  __syncthreads();                    // every thread has read the counter
  if (tid == 0) *a.opt_step = step + 1;
}

}  // namespace

extern "C" int mgb_group_step(const MgbGroupArgs* g, const MgbStepArgs* args,
                              const MgbPeerTable* main_pt, void* stream) {
  MGB_REQUIRE(g && args, MGB_EINVAL, "mgb_group_step: null argument");
  MGB_REQUIRE(g->nparams >= 1 && g->nparams <= MGB_MAX_PARAMS, MGB_EINVAL,
              "mgb_group_step: nparams %d outside 1..%d", g->nparams,
              MGB_MAX_PARAMS);
  MGB_REQUIRE(g->n_models >= 1 && g->n_models <= MGB_MAX_GROUP_MODELS,
              MGB_EINVAL, "mgb_group_step: n_models %d outside 1..%d",
              g->n_models, MGB_MAX_GROUP_MODELS);
  for (int k = 0; k < g->n_models; ++k)
    MGB_REQUIRE(g->loss[k] && g->grad[k], MGB_EINVAL,
                "mgb_group_step: model %d has no loss/grad buffer", k);
  MGB_REQUIRE(args->mode >= MGB_MODE_LOSS_GRAD && args->mode <= MGB_MODE_GD,
              MGB_EINVAL, "mgb_group_step: mode %d not supported", args->mode);
  if (args->mode >= MGB_MODE_ADAM) {
    MGB_REQUIRE(args->opt_p && args->opt_step, MGB_EINVAL,
                "mgb_group_step: optimiser state missing");
    if (args->mode == MGB_MODE_ADAM)
      MGB_REQUIRE(args->opt_u && args->opt_m && args->opt_v, MGB_EINVAL,
                  "mgb_group_step: Adam moments missing");
  }
  MgbPeerTable pt;
  memset(&pt, 0, sizeof(pt));
  pt.world = 1;
  if (main_pt) {
    pt = *main_pt;
    MGB_REQUIRE(pt.world >= 1 && pt.world <= MGB_MAX_WORLD, MGB_EINVAL,
                "mgb_group_step: bad world %d", pt.world);
    MGB_REQUIRE(pt.world == 1 ||
                    2u * (uint32_t)(1 + g->nparams) <= pt.ll_units,
                MGB_ETOOBIG, "mgb_group_step: payload exceeds the LL slots");
  }
  group_step_kernel<<<1, GP_BLOCK, 0, (cudaStream_t)stream>>>(*g, *args, pt);
  MGB_CUDA_CHECK(cudaGetLastError());
  return MGB_OK;
}
