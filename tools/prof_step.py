"""A few per-step fused Adam launches at N halos (for ncu captures)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch
import multigrad_b200 as mg
from multigrad_b200.models import smf, SMFModel
from multigrad_b200.ops.erfbin import FitState, MODE_ADAM
N = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100_000_000
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
dev = torch.device("cuda", 0)
x = smf.synthetic_log_halo_masses(N, 0, 1, device=dev)
m = SMFModel(x=x, bin_edges=np.linspace(9, 10, 11), volume=10.0 * N,
             comm=mg.Comm(None), device=dev, target_sumstats=np.ones(10))
m.set_target(m.calc_sumstats_from_params([-2.0, 0.2]))
plan = m.plan()
p0 = m._params([-1.0, 0.5])
st = FitState(p0, p0, steps + 1, dev)
a = plan.make_optimizer_args(MODE_ADAM, st, 0.01)
for _ in range(steps):
    plan.launch(a)
torch.cuda.synchronize()
print("ok", st.p.cpu().tolist())
