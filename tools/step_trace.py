"""Where the time of a fused step goes: %globaltimer stamps written by the
kernel itself (``MgbStepArgs.trace``), for per-step launches or one persistent
launch.

    python tools/step_trace.py [halos_per_gpu] [steps] [persistent 0|1]
    torchrun --nproc-per-node N tools/step_trace.py ...     (N GPUs)

Per step the kernel records, in ns: the start of every CTA's object loop, its
end, and in the elected CTA the times after the ticket, the grid reduction, the
NVLink exchange and the epilogue.  Printed: medians over the steps of
  loop      slowest CTA's object loop (start of the step -> last loop end)
  spread    last loop end - first loop end (imbalance between CTAs)
  ticket    last loop end -> elected CTA knows it is last (cluster + ticket)
  reduce    grid reduction of the per-cluster partials
  xchg      in-kernel LL all-reduce over NVLink (0 on one GPU)
  epi       loss / gradient / optimiser epilogue
  gap       epilogue end -> first CTA of the next step starts its loop
"""
import os
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
os.environ.setdefault("MULTIGRAD_B200_NO_TQDM", "1")
import numpy as np
import torch
import multigrad_b200 as mg
from multigrad_b200.models import smf, SMFModel
from multigrad_b200.ops.erfbin import FitState, MODE_ADAM

NPER = int(float(sys.argv[1])) if len(sys.argv) > 1 else 12_500_000
K = int(sys.argv[2]) if len(sys.argv) > 2 else 50
PERSISTENT = int(sys.argv[3]) if len(sys.argv) > 3 else 1
DET = int(sys.argv[4]) if len(sys.argv) > 4 else 0

comm = mg.COMM
rank, world = comm.rank, comm.size
dev = mg.parallel.local_device()
N = NPER * world
x = smf.synthetic_log_halo_masses(N, rank, world, device=dev)
m = SMFModel(x=x, bin_edges=np.linspace(9, 10, 11), volume=10.0 * N,
             comm=comm, device=dev, target_sumstats=np.ones(10),
             deterministic=bool(DET))
m.set_target(m.calc_sumstats_from_params([-2.0, 0.2]))
plan = m.plan()
G = plan.lib.mgb_erfbin_grid(10)
p0 = m._params([-1.0, 0.5])
st = FitState(p0, p0, K + 8, dev)
a = plan.make_optimizer_args(MODE_ADAM, st, 0.01)
for _ in range(5):
    plan.launch(a)
torch.cuda.synchronize()
comm.Barrier()
W = 8 + 2 * G
trace = torch.zeros(K * W, dtype=torch.int64, device=dev)
e0, e1 = (torch.cuda.Event(enable_timing=True) for _ in range(2))
e0.record()
if PERSISTENT:
    a.trace = trace.data_ptr()
    a.nsteps = K
    plan.launch(a)
else:
    for s in range(K):
        a.trace = trace.data_ptr() + 8 * W * s
        plan.launch(a)
e1.record()
torch.cuda.synchronize()
m._check_peers()
t = trace.cpu().numpy().reshape(K, W).astype(np.float64)
start = t[:, 8:8 + G]
end = t[:, 8 + G:8 + 2 * G]
rows = []
for s in range(1, K):
    s0 = start[s].min()
    rows.append(dict(
        loop=end[s].max() - s0, spread=end[s].max() - end[s].min(),
        ticket=t[s, 1] - end[s].max(), reduce=t[s, 2] - t[s, 1],
        xchg=t[s, 3] - t[s, 2], epi=t[s, 4] - t[s, 3],
        gap=(start[s + 1].min() - t[s, 4]) if s + 1 < K else np.nan,
        startskew=start[s].max() - s0,
        total=(start[s + 1].min() - s0) if s + 1 < K else np.nan))
med = {k: float(np.nanmedian([r[k] for r in rows])) / 1e3 for k in rows[0]}
us = e0.elapsed_time(e1) / K * 1e3
line = (f"rank {rank}/{world} halos/gpu {NPER:.2e} persistent={PERSISTENT} det={DET} "
        f"{us:7.1f} us/step | " +
        " ".join(f"{k} {v:6.2f}" for k, v in med.items()))
for r in range(world):
    if r == rank:
        print(line, flush=True)
    comm.Barrier()
mg.shutdown()
