"""Tuning aid: in-switch (NVLS) all-reduce time vs CTA count, staged and
resident, a few large sizes.  Run under torchrun, one rank per GPU."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import multigrad_b200 as mg  # noqa: E402
from multigrad_b200.ops import allreduce as ar  # noqa: E402


def main():
    comm = mg.COMM
    dev = mg.parallel.local_device()
    world = comm.size
    ws = comm.workspace()
    assert ws.multicast(), "no multicast object on this box"

    def timed(fn, iters=20):
        for _ in range(3):
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
        return max(comm.allgather(e0.elapsed_time(e1) * 1e3 / iters))

    # fabric reference: copy-engine peer copies of 64 MiB, one direction
    # (rank r -> r+1 while nobody else sends... every rank sends to its right
    # neighbour = each link direction carries one stream) and both directions
    # (every rank also sends to its left neighbour)
    if world > 1 and os.environ.get("NVT_P2P", "1") != "0":
        nb = 64 << 20
        src = torch.empty(nb, dtype=torch.uint8, device=dev)
        right = torch.device("cuda", (dev.index + 1) % world)
        left = torch.device("cuda", (dev.index - 1) % world)
        try:
            dr = torch.empty(nb, dtype=torch.uint8, device=right)
            dl = torch.empty(nb, dtype=torch.uint8, device=left)
            s2 = torch.cuda.Stream(device=dev)

            def one():
                dr.copy_(src, non_blocking=True)

            def both():
                dr.copy_(src, non_blocking=True)
                with torch.cuda.stream(s2):
                    dl.copy_(src, non_blocking=True)
                torch.cuda.current_stream().wait_stream(s2)
            t1 = timed(one, 10)
            t2 = timed(both, 10)
            if comm.rank == 0:
                print(json.dumps({
                    "p2p_copy_64MiB_us_one_neighbour": round(t1, 1),
                    "GBps_out_per_gpu": round(nb / t1 / 1e3, 1),
                    "p2p_copy_64MiB_us_two_neighbours": round(t2, 1),
                    "GBps_out_per_gpu_two": round(2 * nb / t2 / 1e3, 1)}),
                    flush=True)
            del dr, dl
        except Exception as e:  # noqa: BLE001
            if comm.rank == 0:
                print(json.dumps({"p2p_probe": f"unavailable: {e}"[:200]}))
    sizes_mb = [float(v) for v in os.environ.get("NVT_MB", "2,8,32,64").split(",")]
    ctas = [int(v) for v in os.environ.get("NVT_CTAS", "296,148,74,37").split(",")]
    mc_x = ws.multicast_tensor(int(max(sizes_mb) * (1 << 20)) // 4)
    for mb in sizes_mb:
        n = int(mb * (1 << 20)) // 4
        x = torch.randn(n, device=dev)
        xs = mc_x[:n]
        xs.copy_(x)
        # correctness of the staged and resident paths on a pattern whose sum
        # is known in closed form (ragged length on purpose)
        nn = n - 3
        idx = torch.arange(nn, device=dev, dtype=torch.float32)
        pat = (idx % 1013.0) * 0.25 + float(comm.rank + 1)
        want = (idx % 1013.0) * 0.25 * world + world * (world + 1) / 2.0
        y = pat.clone()
        ar.all_reduce_sum_(y, comm, algo=4)
        ys = mc_x[:nn]
        ys.copy_(pat)
        ar.all_reduce_sum_(ys, comm, algo=4)
        torch.cuda.synchronize()
        ok = bool(torch.equal(y, want)) and bool(torch.equal(ys, want))
        assert all(comm.allgather(ok)), f"wrong sums at {mb} MiB"
        nccl = timed(lambda: dist.all_reduce(x, group=comm.group))
        row = {"mb": mb, "n_gpus": world, "nccl_us": round(nccl, 1)}
        for c in ctas:
            os.environ["MULTIGRAD_B200_NVLS_CTAS"] = str(c)
            row[f"staged_{c}"] = round(timed(
                lambda: ar.all_reduce_sum_(x, comm, algo=4)), 1)
            row[f"resident_{c}"] = round(timed(
                lambda: ar.all_reduce_sum_(xs, comm, algo=4)), 1)
        bus = 2 * (world - 1) / world * n * 4 / 1e3
        row["best_resident_busbw"] = round(bus / min(
            v for k, v in row.items() if k.startswith("resident_")), 1)
        row["best_staged_busbw"] = round(bus / min(
            v for k, v in row.items() if k.startswith("staged_")), 1)
        row["nccl_busbw"] = round(bus / nccl, 1)
        if comm.rank == 0:
            print(json.dumps(row), flush=True)
    ws.check_error()
    mg.shutdown()


if __name__ == "__main__":
    main()
