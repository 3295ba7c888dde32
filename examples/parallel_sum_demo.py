#!/usr/bin/env python
"""Smallest possible multi-rank program: sum ``zeros(5) + rank`` over ranks.

    torchrun --nproc-per-node 4 examples/parallel_sum_demo.py

Counterpart of the reference's ``tests/smf_example/parallel_sum_mpi4py_demo.py``
(an ``MPI.Reduce``); here the reduction runs on the GPUs through the NVLink
kernel when CUDA is available, on gloo otherwise.
"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))  # _bootstrap: run from the source tree
import torch

import multigrad_b200 as mg

dev = mg.parallel.local_device()
mine = torch.zeros(5, device=dev) + mg.RANK
total = mg.reduce_sum(mine)
print(f"rank {mg.RANK}/{mg.N_RANKS} on {dev}: {mine.tolist()} -> {total.tolist()}")
mg.shutdown()
