"""Per-launch fixed cost of the fused kernel by mode (graph replay, tiny N)."""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np, torch
import multigrad_b200 as mg
from multigrad_b200.models import smf, SMFModel
from multigrad_b200.ops import _lib
from multigrad_b200.ops.erfbin import FitState, MODE_ADAM, MODE_SUMSTATS, MODE_LOSS_GRAD
dev = torch.device("cuda", 0); SELF = mg.Comm(None)
N = int(float(sys.argv[1])) if len(sys.argv) > 1 else 10000
x = smf.synthetic_log_halo_masses(N, 0, 1, device=dev)
m = SMFModel(x=x, bin_edges=np.linspace(9, 10, 11), volume=10.0 * N, comm=SELF, device=dev, target_sumstats=np.ones(10))
m.set_target(m.calc_sumstats_from_params([-2.0, 0.2]))
plan = m.plan()
p = m._params([-1.0, 0.5])
def timeit(make_args, label):
    a, keep = make_args()
    plan.launch(a); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(50): plan.launch(a)
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): g.replay()
    e1.record(); torch.cuda.synchronize()
    print(f"{label:14s} {e0.elapsed_time(e1)/500*1e3:7.2f} us/launch", flush=True)
def args_mode(mode):
    def f():
        a = _lib.MgbStepArgs(); a.mode, a.batch = mode, 1
        s = torch.empty(10, device=dev); l = torch.empty(1, device=dev); g = torch.empty(2, device=dev)
        a.params = p.data_ptr(); a.out_sumstats = s.data_ptr(); a.out_loss = l.data_ptr(); a.out_grad = g.data_ptr()
        return a, (s, l, g)
    return f
def args_adam():
    st = FitState(p, p, 1000, dev)
    return plan.make_optimizer_args(MODE_ADAM, st, 0.01), st
print(os.environ.get("MULTIGRAD_B200_LIB", "default"), "N =", N)
timeit(args_mode(MODE_SUMSTATS), "sumstats")
timeit(args_mode(MODE_LOSS_GRAD), "loss+grad")
timeit(args_adam, "adam")
# empty-kernel reference: torch tiny op
z = torch.zeros(1, device=dev)
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    for _ in range(50): z.add_(1)
g.replay(); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): g.replay()
e1.record(); torch.cuda.synchronize()
print(f"{'tiny torch op':14s} {e0.elapsed_time(e1)/500*1e3:7.2f} us/launch")
