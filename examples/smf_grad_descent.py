#!/usr/bin/env python
"""Fit the stellar-mass function of a synthetic halo catalogue.

    torchrun --nproc-per-node 8 examples/smf_grad_descent.py --num-halos 100_000_000
    python examples/smf_grad_descent.py --generic            # user-style torch model

Counterpart of the reference's ``tests/smf_example/smf_grad_descent.py``
(CLI flags ``--num-halos --num-steps --learning-rate`` kept; its matplotlib
plots are replaced by a printed summary and an optional ``--save`` .npz).
"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))  # _bootstrap: run from the source tree
import argparse
import time

import numpy as np
import torch

import multigrad_b200 as mg
from multigrad_b200.models import SMFModel, smf

parser = argparse.ArgumentParser(description=__doc__.split("\n")[0])
parser.add_argument("--num-halos", type=int, default=10_000)
parser.add_argument("--num-steps", type=int, default=2000)
parser.add_argument("--learning-rate", type=float, default=1e-3)
parser.add_argument("--optimizer", choices=["gd", "adam", "bfgs"], default="gd")
parser.add_argument("--generic", action="store_true",
                    help="use the two-hook torch model instead of the fused one")
parser.add_argument("--save", type=str, default=None)


def main():
    args = parser.parse_args()
    comm, dev = mg.COMM, mg.parallel.local_device()
    n = args.num_halos
    x = smf.synthetic_log_halo_masses(n, comm.rank, comm.size, device=dev)
    edges = torch.linspace(9, 10, 11)
    truth, guess = smf.ParamTuple(-2.0, 0.2), smf.ParamTuple(-1.0, 0.5)
    if args.generic:
        model = smf.MySMFModel(aux_data=dict(
            log_halo_masses=x, smf_bin_edges=edges.to(dev), volume=10.0 * n,
            target_sumstats=torch.ones(10, device=dev)))
        model.aux_data["target_sumstats"] = \
            model.calc_sumstats_from_params(truth)
    else:
        model = SMFModel(x=x, bin_edges=edges, volume=10.0 * n,
                         target_sumstats=np.ones(10))
        model.set_target(model.calc_sumstats_from_params(truth))

    t0 = time.time()
    if args.optimizer == "gd":
        res = model.run_simple_grad_descent(
            guess, nsteps=args.num_steps, learning_rate=args.learning_rate)
        params, losses = res.params, res.loss
    elif args.optimizer == "adam":
        params, losses = model.run_adam(
            guess, nsteps=args.num_steps, learning_rate=args.learning_rate,
            return_loss=True)
    else:
        out = model.run_bfgs(guess, maxsteps=args.num_steps)
        params = torch.as_tensor(np.atleast_2d(out.x))
        losses = torch.as_tensor([out.fun])
    final = params[-1].cpu().tolist()
    if dev.type == "cuda":
        torch.cuda.synchronize()
    dt = time.time() - t0
    final_smf = model.calc_sumstats_from_params(final)
    true_smf = model.calc_sumstats_from_params(truth)
    if comm.rank == 0:
        print(f"{comm.size} rank(s), {n} halos, {args.optimizer}: {dt:.3f} s")
        print(f"Initial guess : {guess}")
        print(f"Final solution: {smf.ParamTuple(*final)}")
        print(f"Truth         : {truth}")
        print(f"final loss    : {float(losses[-1]):.3e}")
        print("true SMF      :", np.array2string(true_smf.cpu().numpy(),
                                                 precision=4))
        print("fitted SMF    :", np.array2string(final_smf.cpu().numpy(),
                                                 precision=4))
        if args.save:
            np.savez(args.save, params=params.cpu().numpy(),
                     loss=losses.cpu().numpy(),
                     true_smf=true_smf.cpu().numpy(),
                     final_smf=final_smf.cpu().numpy())
    mg.shutdown()


if __name__ == "__main__":
    main()
