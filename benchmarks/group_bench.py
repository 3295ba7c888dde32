#!/usr/bin/env python
"""W4: OnePointGroup of 3 SMF models ("three redshifts") sharing parameters.

    torchrun --nproc-per-node 8 benchmarks/group_bench.py [--halos 1e8]

The ranks are split with split_subcomms(ranks_per_group=[3, 3, 2]) (for 8
GPUs; near-even otherwise); every group holds its own catalogue shard-ed over
its GPUs and its own peer workspace.  One joint loss+gradient evaluation =
one fused launch per GPU (in-kernel reduction inside the group) + one
cross-group all-reduce of 1+P numbers over the world workspace; an Adam step
fuses that all-reduce with the replicated update (csrc/group.cu), so the fit
never reads anything back.  Reports joint loss+grad evaluations per second
and Adam steps per second (device-timed, max over ranks).  Reference counterpart: OnePointGroup
(multigrad/multigrad.py:547-607) with its two pickled allgathers.
"""
import argparse
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch
import multigrad_b200 as mg
from multigrad_b200.models import SMFModel, smf


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--halos", type=float, default=1e8,
                    help="halos per model (3 models)")
    ap.add_argument("--steps", type=int, default=100)
    args = ap.parse_args()
    comm, dev = mg.COMM, mg.parallel.local_device()
    W = comm.size
    ngroups = min(3, W)
    base, extra = divmod(W, ngroups)
    rpg = [base + (1 if g < extra else 0) for g in range(ngroups)]
    if W == 8:
        rpg = [3, 3, 2]
    sub, n, g = mg.split_subcomms(ranks_per_group=rpg)
    N = int(args.halos)
    # model g: the same power-law catalogue shifted by 0.05 dex per "redshift"
    x = smf.synthetic_log_halo_masses(N, sub.rank, sub.size, device=dev) \
        + 0.05 * g
    truth = torch.tensor([-2.0, 0.2])
    model = SMFModel(x=x, bin_edges=np.linspace(9, 10, 11),
                     volume=10.0 * N * (1.0 + 0.5 * g), comm=sub, device=dev,
                     target_sumstats=np.ones(10))
    model.set_target(model.calc_sumstats_from_params(truth))
    group = mg.OnePointGroup((model,), main_comm=comm)
    p = torch.tensor([-1.0, 0.5], device=dev)

    def timed(fn, iters):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        comm.Barrier()
        e0, e1 = (torch.cuda.Event(enable_timing=True) for _ in range(2))
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return max(comm.allgather(e0.elapsed_time(e1) / iters))

    ms_eval = timed(lambda: group.calc_loss_and_grad_from_params(p),
                    args.steps)

    def timed_fit(fit, nsteps):
        fit(5)                                       # plans, workspaces
        torch.cuda.synchronize()
        comm.Barrier()
        e0, e1 = (torch.cuda.Event(enable_timing=True) for _ in range(2))
        e0.record()
        out = fit(nsteps)
        e1.record()
        torch.cuda.synchronize()
        return out, max(comm.allgather(e0.elapsed_time(e1) / nsteps))

    # device-resident: per step one fused launch per model + one fused
    # cross-group reduce-and-update launch (csrc/group.cu)
    traj, ms_adam = timed_fit(
        lambda k: group.run_adam(p, nsteps=k, learning_rate=0.01), args.steps)
    # the generic host-driven loop on the same group, for comparison
    nh = max(args.steps // 4, 5)
    _, ms_host = timed_fit(
        lambda k: mg.OnePointModel.run_adam(group, p, nsteps=k,
                                            learning_rate=0.01, comm=comm), nh)
    loss, grad = group.calc_loss_and_grad_from_params(traj[-1])
    if comm.rank == 0:
        print(json.dumps({
            "bench": "OnePointGroup of 3 SMF models", "n_gpus": W,
            "ranks_per_group": rpg, "halos_per_model": N,
            "ms_per_joint_eval": ms_eval, "joint_evals_per_sec": 1e3 / ms_eval,
            "ms_per_adam_step": ms_adam, "adam_steps_per_sec": 1e3 / ms_adam,
            "host_driven_adam_steps_per_sec": 1e3 / ms_host,
            "final_params": traj[-1].cpu().tolist(), "final_loss": float(loss)}))
    comm.workspace().check_error()
    mg.shutdown()


if __name__ == "__main__":
    main()
