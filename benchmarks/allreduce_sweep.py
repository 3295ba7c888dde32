#!/usr/bin/env python
"""W5: all-reduce latency / bandwidth sweep, 8 B - 64 MiB, ours vs NCCL.

    torchrun --nproc-per-node N benchmarks/allreduce_sweep.py [--max-mb 64]

Ours = the kernels of csrc/allreduce.cu over the NVLink peer workspace
(one-shot ordered LL for small messages, two-shot reduce-scatter/all-gather
over peer loads/stores for large ones).  Baseline = torch.distributed
all_reduce on the NCCL backend.  Device-timed with CUDA events, max over ranks;
the reference's own path (np.asarray + MPI Allreduce on host buffers,
multigrad/multigrad.py:149-183; mpi4jax in the experimental package) cannot be
installed on this image.  Prints one JSON line per size on rank 0.
"""
import argparse
import json
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
import torch.distributed as dist
import multigrad_b200 as mg
from multigrad_b200.ops import allreduce as ar


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=float, default=64)
    ap.add_argument("--iters", type=int, default=50)
    args = ap.parse_args()
    comm = mg.COMM
    dev = mg.parallel.local_device()
    world, rank = comm.size, comm.rank
    sizes = []
    n = 2
    while n * 4 <= args.max_mb * (1 << 20):
        sizes.append(n)
        n *= 4
    sizes = [2, 8, 30] + [s for s in sizes if s > 30]

    def timed(fn, iters):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        comm.Barrier()
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / iters
        return max(comm.allgather(us))

    rows = []
    for n in sizes:
        x = torch.randn(n, device=dev)
        iters = args.iters if n < (1 << 22) else max(10, args.iters // 5)
        ours = timed(lambda: ar.all_reduce_sum_(x, comm), iters)
        nccl = timed(lambda: dist.all_reduce(x, group=comm.group), iters)
        bytes_ = n * 4
        # bus bandwidth convention of nccl-tests: 2 (W-1)/W * bytes / time
        bus = lambda us: 2 * (world - 1) / world * bytes_ / (us * 1e-6) / 1e9
        row = {"bytes": bytes_, "n_gpus": world, "ours_us": round(ours, 2),
               "nccl_us": round(nccl, 2), "ours_busbw_gbs": round(bus(ours), 2),
               "nccl_busbw_gbs": round(bus(nccl), 2),
               "algo": "LL" if n <= 4096 else "two-shot"}
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
    comm.workspace().check_error()
    mg.shutdown()


if __name__ == "__main__":
    main()
