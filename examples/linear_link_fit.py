#!/usr/bin/env python
"""Fit of the tensor-core model family: ``mu_h = x_h + W[h, :] . theta``.

    python examples/linear_link_fit.py
    torchrun --nproc-per-node 8 examples/linear_link_fit.py --num-halos 8000000

Every object carries a feature vector (here: a constant, which plays the role
of ``log_shmrat`` of the SMF example, and a few standardised properties) and
the mean of its log-normal is a linear function of them.  The per-object
Jacobian contraction is GEMM-shaped, so on a B200 the step runs on tcgen05
(``csrc/linlink.cu``); on a CPU the same model runs through autograd.  No
counterpart in the reference (its examples only have the two-parameter SMF).
"""
import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))  # _bootstrap: run from the source tree
import numpy as np
import torch

import multigrad_b200 as mg
from multigrad_b200.models import LinearLinkBinnedModel


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--num-halos", type=int, default=40_000)
    ap.add_argument("--num-features", type=int, default=4)
    ap.add_argument("--num-steps", type=int, default=400)
    ap.add_argument("--learning-rate", type=float, default=0.01)
    ap.add_argument("--weighted", action="store_true",
                    help="give every object a weight (e.g. completeness)")
    ap.add_argument("--polish", action="store_true",
                    help="finish with L-BFGS-B (replicated on every rank)")
    args = ap.parse_args()
    comm, dev = mg.COMM, mg.parallel.local_device()
    n = args.num_halos // comm.size
    F = args.num_features
    g = torch.Generator().manual_seed(1234 + comm.rank)
    x = (10.0 + 1.3 * torch.rand(n, generator=g)).float()
    feat = torch.cat([torch.ones(n, 1),
                      0.3 * torch.randn(n, F - 1, generator=g)], dim=1)
    truth = torch.cat([torch.tensor([-2.0]), 0.1 * torch.ones(F - 1),
                       torch.tensor([0.2])])          # theta..., sigma
    w = (0.5 + torch.rand(n, generator=g)).to(dev) if args.weighted else None
    model = LinearLinkBinnedModel(
        x=x.to(dev), features=feat.to(dev), bin_edges=np.linspace(9, 10, 11),
        volume=10.0 * args.num_halos, target_sumstats=np.ones(10), weights=w,
        loss_eps=1e-10)
    model.set_target(model.calc_sumstats_from_params(truth))
    guess = truth + torch.cat([torch.tensor([0.2]),
                               -0.1 * torch.ones(F - 1),
                               torch.tensor([0.1])])
    traj, losses = model.run_adam(guess, nsteps=args.num_steps,
                                  learning_rate=args.learning_rate,
                                  return_loss=True)
    if comm.rank == 0:
        print(f"{comm.size} rank(s), {F} features: loss "
              f"{float(losses[0]):.4f} -> {float(losses[-1]):.2e}")
        print("truth :", [round(float(v), 4) for v in truth])
        print("fitted:", [round(float(v), 4) for v in traj[-1].cpu()])
    if args.polish:
        res = model.run_bfgs(traj[-1].cpu(), maxsteps=50,
                             param_bounds=[(None, None)] * F + [(0.02, 2.0)])
        if comm.rank == 0:
            print(f"L-BFGS-B: loss {res.fun:.3e} after {res.nfev} evaluations")
            print("polished:", [round(float(v), 4) for v in res.x])
    mg.shutdown()


if __name__ == "__main__":
    main()
