#!/usr/bin/env python
"""Speed test: gradient-descent iterations per second on the SMF pipeline.

    torchrun --nproc-per-node N examples/benchmark.py --num-halos 100000 --save out.txt

Counterpart of the reference's ``tests/smf_example/benchmark.py`` (same flags,
same "iterations/sec after one warm-up step" metric, same one-dict-per-line
``--save`` format).  The reference script is stale (it calls a method that no
longer exists); this one uses ``run_simple_grad_descent``.  The headline
benchmark of this repository is ``bench.py`` at the repo root.
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
parser.add_argument("--num-steps", type=int, default=100)
parser.add_argument("--learning-rate", type=float, default=1e-3)
parser.add_argument("--save", type=str, default=None)


def main():
    args = parser.parse_args()
    comm, dev = mg.COMM, mg.parallel.local_device()
    x = smf.synthetic_log_halo_masses(args.num_halos, comm.rank, comm.size,
                                      device=dev)
    model = SMFModel(x=x, bin_edges=np.linspace(9, 10, 11),
                     volume=10.0 * args.num_halos,
                     target_sumstats=np.asarray(smf.TARGET_SUMSTATS_TEST))
    guess = smf.ParamTuple(-1.0, 0.5)

    def speedtest(nsteps):
        model.run_simple_grad_descent(guess, nsteps=nsteps,
                                      learning_rate=args.learning_rate)
        if dev.type == "cuda":
            torch.cuda.synchronize()
        comm.Barrier()

    speedtest(1)                       # warm-up (workspace, module load)
    t0 = time.perf_counter()
    speedtest(args.num_steps)
    calls_per_sec = args.num_steps / (time.perf_counter() - t0)
    if comm.rank == 0:
        print(f"Benchmark with {comm.size} processes {args}")
        print("=" * 70)
        print(f"Grad descent iterations/sec = {calls_per_sec}\n")
        if args.save:
            result = dict(calls_per_sec=calls_per_sec,
                          num_processes=comm.size, num_halos=args.num_halos,
                          num_steps=args.num_steps,
                          learning_rate=args.learning_rate)
            with open(args.save, "a+") as f:
                f.write(f"{result!r}\n")
    mg.shutdown()


if __name__ == "__main__":
    main()
