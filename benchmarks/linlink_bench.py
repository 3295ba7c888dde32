#!/usr/bin/env python
"""Throughput of the tensor-core linear-link step (csrc/linlink.cu).

    python benchmarks/linlink_bench.py [--halos 2e7] [--features 64]
    torchrun --nproc-per-node 8 benchmarks/linlink_bench.py --halos 1.6e8

The step streams the bf16 feature matrix once (TMA), so the roofline is HBM:
bytes/step = n * (pfeat*2 + 4).  Reports ms/step, achieved GB/s per GPU and
the fraction of the measured copy bandwidth (MEASURED_PEAKS.json).
"""
import argparse
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch
import multigrad_b200 as mg
from multigrad_b200.models import LinearLinkBinnedModel
from multigrad_b200.ops._lib import MODE_ADAM
from multigrad_b200.ops.linlink import LinFitState


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--halos", type=float, default=2e7)
    ap.add_argument("--features", type=int, default=64)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--proxy", action="store_true",
                    help="also time the autograd (two-pass) formulation")
    args = ap.parse_args()
    comm, dev = mg.COMM, mg.parallel.local_device()
    n_tot, F = int(args.halos), args.features
    n = n_tot // comm.size
    g = torch.Generator(device=dev).manual_seed(comm.rank)
    x = 10.0 + 1.3 * torch.rand(n, generator=g, device=dev)
    feat = torch.empty(n, F, device=dev, dtype=torch.bfloat16)
    feat.normal_(0, 0.3, generator=g)
    feat[:, 0] = 1.0
    theta = torch.zeros(F + 1)
    theta[0], theta[-1] = -2.0, 0.2
    model = LinearLinkBinnedModel(x=x, features=feat, bin_edges=np.linspace(
        9, 10, 11), volume=10.0 * n_tot, target_sumstats=np.ones(10))
    del feat
    plan = model.plan()
    model.features = None            # the plan keeps the only copy (bf16)
    torch.cuda.empty_cache()
    model.set_target(model.calc_sumstats_from_params(theta))
    p0 = model._params(theta + 0.05)
    st = LinFitState(p0, p0, args.steps + 5, dev)
    a = plan.make_optimizer_args(MODE_ADAM, st, 0.01)
    for _ in range(5):
        plan.launch(a)
    torch.cuda.synchronize()
    comm.Barrier()
    from bench import ClockSampler
    clk = ClockSampler(comm.size) if comm.rank == 0 else None
    if clk:
        clk.start()
    e0, e1 = (torch.cuda.Event(enable_timing=True) for _ in range(2))
    e0.record()
    for _ in range(args.steps):
        plan.launch(a)
    e1.record()
    torch.cuda.synchronize()
    clocks = clk.stop() if clk else None
    ms = max(comm.allgather(e0.elapsed_time(e1) / args.steps))
    model._check()
    bytes_per_gpu = n * (plan.pfeat * 2 + 4)
    peak = 6458.7
    try:
        peak = json.load(open(Path(__file__).resolve().parents[1] /
                              "MEASURED_PEAKS.json"))["hbm_gbs"]
    except Exception:
        pass
    gbs = bytes_per_gpu / (ms * 1e-3) / 1e9
    if comm.rank == 0:
        print(json.dumps({
            "kernel": "linlink_step (tcgen05+TMA)", "n_gpus": comm.size,
            "halos_total": n_tot, "features": F, "pfeat": plan.pfeat,
            "ms_per_step": ms, "steps_per_sec": 1e3 / ms,
            "hbm_gbs_per_gpu": gbs, "frac_of_measured_copy_bw": gbs / peak,
            "final_loss": float(st.loss_hist[int(st.step.item()) - 1]),
            "l2": "inputs>L2" if bytes_per_gpu > (126 << 20) else "resident",
            "clocks": clocks}))
    mg.shutdown()


if __name__ == "__main__":
    main()
