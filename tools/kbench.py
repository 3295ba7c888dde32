"""Time the fused step kernel alone (1e8 halos, K=10) for one library build.
Usage: MULTIGRAD_B200_LIB=path python tools/kbench.py [halos] [steps]"""
import os
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch
import multigrad_b200 as mg
from multigrad_b200.models import smf, SMFModel
from multigrad_b200.ops.erfbin import FitState, MODE_ADAM, MODE_SUMSTATS

N = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100_000_000
K = int(sys.argv[2]) if len(sys.argv) > 2 else 30
dev = torch.device("cuda", 0)
SELF = mg.Comm(None)
x = smf.synthetic_log_halo_masses(N, 0, 1, device=dev)
m = SMFModel(x=x, bin_edges=np.linspace(9, 10, 11), volume=10.0 * N,
             comm=SELF, device=dev, target_sumstats=np.ones(10))
m.set_target(m.calc_sumstats_from_params([-2.0, 0.2]))
plan = m.plan()
p0 = m._params([-1.0, 0.5])
st = FitState(p0, p0, K + 5, dev)
a = plan.make_optimizer_args(MODE_ADAM, st, 0.01)
for _ in range(5):
    plan.launch(a)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
graph = None
if os.environ.get("KB_GRAPH"):
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        for _ in range(20):
            plan.launch(a)
    graph.replay()
    torch.cuda.synchronize()
e0.record()
if graph is not None:
    for _ in range(K // 20):
        graph.replay()
else:
    for _ in range(K):
        plan.launch(a)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / K
s = m.calc_sumstats_from_params([-2.0, 0.2]).cpu().numpy()
print(f"{os.environ.get('MULTIGRAD_B200_LIB', 'default'):50s} "
      f"{ms * 1e3:8.1f} us/step  {N / ms / 1e6:8.1f} Gobj/s  "
      f"p={st.p.cpu().tolist()} s0={s[0]:.6e}", flush=True)
