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
os.environ.setdefault("MULTIGRAD_B200_NO_TQDM", "1")
import torch
import torch.distributed as dist
import multigrad_b200 as mg
from multigrad_b200.ops import allreduce as ar


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=float, default=64)
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--all-algos", action="store_true",
                    help="also time every algorithm forced (LL / one-shot "
                         "push / two-shot) to locate the crossovers")
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
    top = int(args.max_mb * (1 << 20)) // 4
    if top not in sizes and top > sizes[-1]:
        sizes.append(top)              # the 64 MiB end point of W5
    from bench import ClockSampler

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
    mc_x = None
    if args.all_algos and comm.workspace().multicast():
        mc_x = comm.workspace().multicast_tensor(
            min(max(sizes), comm.workspace().multicast().bytes // 4))
    for n in sizes:
        x = torch.randn(n, device=dev)
        iters = args.iters if n < (1 << 22) else max(10, args.iters // 5)
        clk = ClockSampler(world) if rank == 0 else None
        if clk:
            clk.start()
        ours = timed(lambda: ar.all_reduce_sum_(x, comm), iters)
        nccl = timed(lambda: dist.all_reduce(x, group=comm.group), iters)
        forced = {}
        half = (int(comm.workspace().table.bulk_bytes) // 2) & ~1023
        if args.all_algos and n > 1024:
            if n <= comm.workspace().table.ll_units:
                forced["ll_us"] = round(timed(
                    lambda: ar.all_reduce_sum_(x, comm, algo=1), iters), 2)
            if n * 4 * world <= half and n * 4 <= (16 << 20):
                forced["oneshot_us"] = round(timed(
                    lambda: ar.all_reduce_sum_(x, comm, algo=3), iters), 2)
            if n * 4 >= (32 << 10) and comm.workspace().multicast():
                forced["nvls_us"] = round(timed(
                    lambda: ar.all_reduce_sum_(x, comm, algo=4), iters), 2)
                if mc_x is not None and n <= mc_x.numel():
                    xs = mc_x[:n]       # symmetric (multicast-resident) tensor
                    xs.copy_(x)
                    forced["nvls_inplace_us"] = round(timed(
                        lambda: ar.all_reduce_sum_(xs, comm, algo=4), iters), 2)
            if n * 4 <= half:
                forced["twoshot_us"] = round(timed(
                    lambda: ar.all_reduce_sum_(x, comm, algo=2), iters), 2)
        bytes_ = n * 4
        # bus bandwidth convention of nccl-tests: 2 (W-1)/W * bytes / time
        bus = lambda us: 2 * (world - 1) / world * bytes_ / (us * 1e-6) / 1e9
        row = {"bytes": bytes_, "n_gpus": world, "ours_us": round(ours, 2),
               "nccl_us": round(nccl, 2), "ours_busbw_gbs": round(bus(ours), 2),
               "nccl_busbw_gbs": round(bus(nccl), 2),
               **forced}
        if clk:
            c = clk.stop()
            row["clocks"] = {"sm_mhz": c.get("sm_mhz"),
                             "sm_max_mhz": c.get("sm_max_mhz"),
                             "reasons": c.get("reasons")}
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
    comm.workspace().check_error()
    mg.shutdown()


if __name__ == "__main__":
    main()
