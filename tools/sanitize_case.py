"""Smallest run that touches every kernel once (for compute-sanitizer)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch
import multigrad_b200 as mg
from multigrad_b200.models import SMFModel, LinearLinkBinnedModel, smf

SELF = mg.Comm(None)
dev = torch.device("cuda", 0)
data = smf.make_smf_data(3001, comm=SELF, device=dev)
m = SMFModel(aux_data=data, comm=SELF, device=dev)
print(m.calc_loss_and_grad_from_params([-1.9, 0.3]))
print(m.run_adam([-1.9, 0.3], nsteps=3, param_bounds=[(-5, 0), (0.01, None)])[-1])
print(m.run_simple_grad_descent([-1.9, 0.3], nsteps=2).loss)
print(m.run_lhs_param_scan([-2.5, 0.15], [-1.5, 0.4], 2, 5, seed=1)[2])
w = SMFModel(x=data["log_halo_masses"], bin_edges=np.linspace(9, 10, 18),
             volume=1.0, weights=torch.rand(3001, device=dev), comm=SELF,
             device=dev)
print(w.calc_sumstats_from_params([-2.0, 0.2])[:3])
g = torch.Generator().manual_seed(0)
n, F = 700, 9
lin = LinearLinkBinnedModel(
    x=(10 + 1.3 * torch.rand(n, generator=g)).to(dev),
    features=torch.cat([torch.ones(n, 1), 0.3 * torch.randn(n, F - 1, generator=g)], 1).to(dev),
    bin_edges=np.linspace(9, 10, 11), volume=10.0 * n, comm=SELF, device=dev,
    target_sumstats=np.full(10, 1e-2))
p = torch.zeros(F + 1); p[0], p[-1] = -2.0, 0.2
print(lin.calc_loss_and_grad_from_params(p)[0])
print(lin.run_adam(p, nsteps=2)[-1][:3])
from multigrad_b200.experimental import diffdesi
print(diffdesi.find_ultimate_top_indices(torch.tensor([0, 0, 1, 2, 3], device=dev)))
torch.cuda.synchronize()
print("done")
