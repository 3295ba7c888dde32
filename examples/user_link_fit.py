#!/usr/bin/env python
"""A model of your own, fused: stellar-mass function with a MASS-DEPENDENT
scatter and (optionally) random scatter drawn inside the link.

    python examples/user_link_fit.py [--num-halos 1000000] [--stochastic]
    torchrun --nproc-per-node 8 examples/user_link_fit.py --num-halos 1e8

The reference would have you write ``calc_partial_sumstats_from_params`` in
JAX and differentiate it with ``jax.vjp`` (two passes over the objects,
``multigrad/multigrad.py:508-538``).  Here the model-specific part is the
*link* below -- eight lines of Python, no derivatives, no CUDA: it is traced,
differentiated symbolically and compiled into a fused sm_100a kernel
(``csrc/linkbin.cu``) that does statistic + Jacobian + NVLink reduction + loss
+ Adam in one launch per step.  On a machine without a GPU the very same
function runs on PyTorch tensors through the generic path.
"""
import argparse
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))

import numpy as np
import torch

import multigrad_b200 as mg
from multigrad_b200.models import LinkedBinnedModel, smf

ap = argparse.ArgumentParser()
ap.add_argument("--num-halos", type=float, default=1e6)
ap.add_argument("--num-steps", type=int, default=300)
ap.add_argument("--learning-rate", type=float, default=0.01)
ap.add_argument("--stochastic", action="store_true")
args = ap.parse_args()

comm = mg.COMM
dev = mg.parallel.local_device()
N = int(args.num_halos)
logm = smf.synthetic_log_halo_masses(N, comm.rank, comm.size, device=dev)


def link(x, feat, p, rng):
    """log stellar mass = log halo mass + p0; scatter grows with mass."""
    mu = x + p[0]
    if args.stochastic:
        mu = mu + p[3] * rng.normal(0)          # intrinsic scatter, drawn per step
    sigma = p[1] + p[2] * (x - 11.0)
    return mu, sigma


nparams = 4 if args.stochastic else 3
truth = [-2.0, 0.25, 0.05] + ([0.1] if args.stochastic else [])
guess = [-1.6, 0.35, 0.0] + ([0.05] if args.stochastic else [])
model = LinkedBinnedModel(x=logm, link=link, nparams=nparams,
                          bin_edges=np.linspace(9, 10, 11), volume=10.0 * N,
                          comm=comm, device=dev, loss_eps=1e-10, seed=1)
model.set_target(model.calc_sumstats_from_params(truth))
if comm.rank == 0:
    print(f"{N:.0e} halos on {comm.size} x {dev.type}; link compiled: "
          f"{model.fused}; stochastic: {model.spec.stochastic}")
t0 = time.perf_counter()
traj, losses = model.run_adam(guess, nsteps=args.num_steps,
                              learning_rate=args.learning_rate,
                              return_loss=True, randkey=3)
if dev.type == "cuda":
    torch.cuda.synchronize()
dt = time.perf_counter() - t0
if comm.rank == 0:
    print(f"Adam: {args.num_steps} steps in {dt:.2f} s "
          f"({args.num_steps / dt:.0f} steps/s), loss {float(losses[0]):.3e} "
          f"-> {float(losses[-1]):.3e}")
    print("final params", [round(float(v), 4) for v in traj[-1]],
          "truth", truth)
mg.shutdown()
