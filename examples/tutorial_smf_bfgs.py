#!/usr/bin/env python
"""Quick-start tutorial: fit (log f, log sigma) of a stellar-mass function
with L-BFGS-B.

    python examples/tutorial_smf_bfgs.py
    torchrun --nproc-per-node 3 examples/tutorial_smf_bfgs.py

Counterpart of the reference tutorial (``docs/source/notebooks/intro.ipynb``
and ``smf_gradient_descent.py``): 10 000 halos between 1e10 and 1e11 Msun,
``P(log M* | M_h) = N(log f + log M_h, sigma)``, ten bins on 9 < log M* < 10,
truth ``(-2, -0.5)``, initial guess ``truth + (-1.5, 0.7)``.  The reference
reports ``nit=15, nfev=31`` and a converged loss of ~5e-12.

Step 1 writes the model the way a user of the reference would (two torch
functions on ``OnePointModel``); step 2 uses the fused native model.
"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))  # _bootstrap: run from the source tree
import torch

import multigrad_b200 as mg
from multigrad_b200.models import SMFModel, smf

comm, dev = mg.COMM, mg.parallel.local_device()
log_mh = torch.log10(smf.load_halo_masses_tutorial(10_000, comm=comm)).to(dev)
edges = torch.linspace(9, 10, 11).to(dev)
true_params = torch.tensor([-2.0, -0.5])
init_params = true_params + torch.tensor([-1.5, 0.7])


class MySMFModel(mg.OnePointModel):
    """sumstats must be *additive over ranks*: return counts, not logs."""

    def calc_partial_sumstats_from_params(self, params):
        log_f, log_sigma = params[0], params[1]
        mu = log_f + self.aux_data["log_halo_masses"]
        z = (self.aux_data["smf_bin_edges"][:, None] - mu[None, :]) / (
            2 ** 0.5 * 10.0 ** log_sigma)
        cdf = 0.5 * (1 + torch.special.erf(z))
        widths = edges[1:] - edges[:-1]
        return (cdf[1:] - cdf[:-1]).sum(1) / self.aux_data["volume"] / widths

    def calc_loss_from_sumstats(self, sumstats):
        target = torch.log10(self.aux_data["target_sumstats"] + 1e-10)
        return torch.mean((torch.log10(sumstats + 1e-10) - target) ** 2)


aux = dict(log_halo_masses=log_mh, smf_bin_edges=edges, volume=1.0,
           target_sumstats=None)
model = MySMFModel(aux_data=aux)
aux["target_sumstats"] = model.calc_sumstats_from_params(true_params)

loss, grad = model.calc_loss_and_grad_from_params(true_params + 0.1)
results = model.run_bfgs(init_params)

native = SMFModel(x=log_mh, bin_edges=edges, volume=1.0, sigma_link="log10",
                  loss_eps=1e-10, target_sumstats=aux["target_sumstats"])
results_native = native.run_bfgs(init_params)

if comm.rank == 0:
    print("At true_params + [0.1, 0.1]: loss =", float(loss),
          "grad =", grad.cpu().numpy(), "(reference: 0.44032094, "
          "[2.6187496 4.2603974])")
    print("BFGS has converged:", results.success)
    print("Initial guess    =", init_params.numpy())
    print("True params      =", true_params.numpy())
    print("Converged params =", results.x)
    print("\nFull results info:\n", results, sep="")
    print("\nFused native model:", results_native.x,
          f"nit={results_native.nit} nfev={results_native.nfev}")
mg.shutdown()
