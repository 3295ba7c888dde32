"""Accuracy + timing of the fused step kernel for one library build.

Prints one line: max relative error of sumstats / Jacobian against the float64
oracle for four parameter points, whether a persistent fit reproduces the
per-step launches bit for bit, and the time per Adam step for per-step
launches and for the persistent launch.

Usage: MULTIGRAD_B200_LIB=path python tools/kcheck.py [halos] [steps]
"""
import os
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch
import multigrad_b200 as mg
from multigrad_b200.models import smf, SMFModel
from multigrad_b200.ops import reference as ref
from multigrad_b200.ops.erfbin import FitState, MODE_ADAM

N = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100_000_000
K = int(sys.argv[2]) if len(sys.argv) > 2 else 30
dev = torch.device("cuda", 0)
SELF = mg.Comm(None)
tag = os.environ.get("MULTIGRAD_B200_LIB", "default").split("/")[-1]

# ---- accuracy on 1e6 objects against float64 -------------------------------
na = 1_000_000
xa = smf.synthetic_log_halo_masses(na, 0, 1, device=dev)
edges = np.linspace(9, 10, 11)
ma = SMFModel(x=xa, bin_edges=edges, volume=10.0 * na, comm=SELF, device=dev,
              target_sumstats=np.ones(10), deterministic=True)
errs = []
for p in ([-2.0, 0.2], [-1.0, 0.5], [-1.6, 0.08], [-1.2, 1.5]):
    s, j = ma.calc_sumstats_and_jac_from_params(p)
    p64 = torch.as_tensor(np.asarray(p, np.float32)).double()
    so, dmu, dsg = ref.erf_binned_partial_and_jac(
        xa.double().cpu(), torch.as_tensor(edges), torch.as_tensor(ma._norm64),
        p64[0], p64[1], "identity", None)
    es = float(((s.double().cpu() - so).abs() / so.abs().clamp_min(1e-300)).max())
    ej = max(float((j[:, 0].double().cpu() - dmu).abs().max() / dmu.abs().max()),
             float((j[:, 1].double().cpu() - dsg).abs().max() / dsg.abs().max()))
    errs.append((es, ej))
acc = " ".join(f"{a:.1e}/{b:.1e}" for a, b in errs)

# ---- persistent == per-step -------------------------------------------------
ma.set_target(ma.calc_sumstats_from_params([-2.0, 0.2]))
os.environ["MULTIGRAD_B200_PERSISTENT"] = "0"
t_step = ma.run_adam([-1.7, 0.35], nsteps=40, learning_rate=0.01)
os.environ["MULTIGRAD_B200_PERSISTENT"] = "1"
t_pers = ma.run_adam([-1.7, 0.35], nsteps=40, learning_rate=0.01)
torch.cuda.synchronize()
same = bool(torch.equal(t_step, t_pers))

# ---- timing -----------------------------------------------------------------
x = smf.synthetic_log_halo_masses(N, 0, 1, device=dev)
m = SMFModel(x=x, bin_edges=edges, volume=10.0 * N, comm=SELF, device=dev,
             target_sumstats=np.ones(10))
m.set_target(m.calc_sumstats_from_params([-2.0, 0.2]))
plan = m.plan()
p0 = m._params([-1.0, 0.5])


def timed(persistent):
    st = FitState(p0, p0, K + 5, dev)
    a = plan.make_optimizer_args(MODE_ADAM, st, 0.01)
    for _ in range(5):
        plan.launch(a)
    torch.cuda.synchronize()
    e0, e1 = (torch.cuda.Event(enable_timing=True) for _ in range(2))
    e0.record()
    if persistent:
        a.nsteps = K
        plan.launch(a)
    else:
        for _ in range(K):
            plan.launch(a)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / K * 1e3, st.p.cpu().tolist()


us_step, p_step = timed(False)
us_pers, p_pers = timed(True)
clu = plan.lib.mgb_erfbin_cluster(10, 0)
print(f"{tag:24s} clu={clu} N={N:.0e} step {us_step:7.1f} us  persistent {us_pers:7.1f} us  "
      f"same_traj={same} p_equal={p_step == p_pers} err(s/jac)=[{acc}]",
      flush=True)
