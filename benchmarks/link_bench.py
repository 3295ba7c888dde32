#!/usr/bin/env python
"""User-defined models: the fused link kernel against the built-in kernel and
against the generic (eager PyTorch autograd) path, on one GPU.

    python benchmarks/link_bench.py [halos] [steps]

Prints one JSON line: ms per Adam step / loss+grad evaluation of
  builtin   SMFModel (mu = x + p0, sigma = p1), csrc/erfbin.cu
  linked    LinkedBinnedModel with a mass-dependent scatter
            (sigma = p1 + p2 (x - 11)), traced + compiled into csrc/linkbin.cu
  linked2   the built-in link restated as a user link (same statistic)
  generic   smf.MySMFModel through OnePointModel (torch.autograd + reduce_sum)
"""
import json
import subprocess
import sys
import time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch
import multigrad_b200 as mg
from multigrad_b200.models import smf, SMFModel, LinkedBinnedModel

N = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100_000_000
K = int(sys.argv[2]) if len(sys.argv) > 2 else 20
dev = torch.device("cuda", 0)
SELF = mg.Comm(None)
edges = np.linspace(9, 10, 11)
x = smf.synthetic_log_halo_masses(N, 0, 1, device=dev)


def timed(fn, n):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = (torch.cuda.Event(enable_timing=True) for _ in range(2))
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def clocks():
    try:
        out = subprocess.run(
            ["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm",
             "--format=csv,noheader,nounits", "-i", "0"],
            capture_output=True, text=True, timeout=10).stdout.split(",")
        return {"sm_mhz": float(out[0]), "sm_max_mhz": float(out[1])}
    except Exception:  # noqa: BLE001
        return None


res = {"halos": N, "steps": K, "clocks_before": clocks()}
b = SMFModel(x=x, bin_edges=edges, volume=10.0 * N, comm=SELF, device=dev,
             target_sumstats=np.ones(10))
b.set_target(b.calc_sumstats_from_params([-2.0, 0.2]))
fit = b.adam_fit([-1.0, 0.5], nsteps=K + 4)
res["builtin_ms_per_step"] = timed(fit.step, K)

t0 = time.perf_counter()
lk = LinkedBinnedModel(
    x=x, link=lambda x, f, p, r: (x + p[0], p[1] + p[2] * (x - 11.0)),
    nparams=3, bin_edges=edges, volume=10.0 * N, comm=SELF, device=dev,
    target_sumstats=b.target_sumstats)
lk.plan()
res["linked_build_seconds"] = time.perf_counter() - t0
pl = torch.tensor([-1.0, 0.5, 0.0], device=dev)
res["linked_ms_per_eval"] = timed(
    lambda: lk.calc_loss_and_grad_from_params(pl), K)
t0 = time.perf_counter()
lk.run_adam(pl, nsteps=K)
torch.cuda.synchronize()
res["linked_ms_per_step"] = (time.perf_counter() - t0) * 1e3 / K

lk2 = LinkedBinnedModel(x=x, link=lambda x, f, p, r: (x + p[0], p[1]),
                        nparams=2, bin_edges=edges, volume=10.0 * N, comm=SELF,
                        device=dev, target_sumstats=b.target_sumstats)
p2 = torch.tensor([-1.0, 0.5], device=dev)
res["linked2_ms_per_eval"] = timed(
    lambda: lk2.calc_loss_and_grad_from_params(p2), K)

# the generic path materialises O(10 N) floats per evaluation: 1e7 objects
ng = min(N, 10_000_000)
g = smf.MySMFModel(aux_data=dict(
    log_halo_masses=x[:ng], smf_bin_edges=torch.as_tensor(edges, device=dev),
    volume=10.0 * ng, target_sumstats=b.target_sumstats), comm=SELF,
    device=dev)
ms_g = timed(lambda: g.calc_loss_and_grad_from_params(p2), 3)
res["generic_halos"] = ng
res["generic_ms_per_eval"] = ms_g
res["generic_ms_per_eval_scaled_to_halos"] = ms_g * N / ng
res["linked_vs_builtin_objects_per_s"] = (
    res["builtin_ms_per_step"] / res["linked_ms_per_eval"])
res["clocks_after"] = clocks()
print(json.dumps(res))
