#!/usr/bin/env python
"""Joint fit of several data sets that share one parameter vector.

    python examples/group_fit.py                      # one process, two models
    torchrun --nproc-per-node 8 examples/group_fit.py # 3 groups on 3+3+2 GPUs

Every group of ranks owns one stellar-mass-function model (its own catalogue
and volume); ``OnePointGroup`` sums the losses and gradients.  With the native
models the whole fit stays on the GPUs (``csrc/group.cu``).  Reference
counterpart: ``OnePointGroup`` + ``split_subcomms``
(``multigrad/multigrad.py:88-146, 547-607``).
"""
import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))  # _bootstrap: run from the source tree
import numpy as np
import torch

import multigrad_b200 as mg
from multigrad_b200.models import SMFModel, smf


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--num-halos", type=int, default=100_000,
                    help="halos per model")
    ap.add_argument("--num-steps", type=int, default=300)
    ap.add_argument("--learning-rate", type=float, default=0.02)
    args = ap.parse_args()
    comm, dev = mg.COMM, mg.parallel.local_device()
    world = comm.size
    truth = torch.tensor([-2.0, 0.2])
    edges = np.linspace(9, 10, 11)

    def make_model(g, sub):
        """Model of data set ``g`` on communicator ``sub``."""
        rank, size = (sub.rank, sub.size) if sub is not None else (0, 1)
        x = smf.synthetic_log_halo_masses(args.num_halos, rank, size,
                                          device=dev) + 0.05 * g
        m = SMFModel(x=x, bin_edges=edges,
                     volume=10.0 * args.num_halos * (1.0 + 0.5 * g),
                     comm=sub if sub is not None else mg.Comm(None),
                     device=dev, target_sumstats=np.ones(10))
        m.set_target(m.calc_sumstats_from_params(truth))
        return m

    if world == 1:
        models = tuple(make_model(g, None) for g in range(2))
        group = mg.OnePointGroup(models, main_comm=mg.Comm(None))
        layout = "2 models in one process"
    else:
        ngroups = min(3, world)
        base, extra = divmod(world, ngroups)
        rpg = [base + (1 if g < extra else 0) for g in range(ngroups)]
        sub, ngroups, g = mg.split_subcomms(ranks_per_group=rpg)
        group = mg.OnePointGroup((make_model(g, sub),), main_comm=comm)
        layout = f"{ngroups} models on rank groups {rpg}"
    traj, losses = group.run_adam([-1.6, 0.35], nsteps=args.num_steps,
                                  learning_rate=args.learning_rate,
                                  return_loss=True)
    if comm.rank == 0:
        print(f"{layout}: loss {float(losses[0]):.4f} -> "
              f"{float(losses[-1]):.2e}, params {traj[-1].cpu().tolist()} "
              f"(truth {truth.tolist()})")
    mg.shutdown()


if __name__ == "__main__":
    main()
