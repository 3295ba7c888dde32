"""Print kernel-vs-fp64 relative errors (run on a GPU box)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch
import multigrad_b200 as mg
from multigrad_b200.models import ErfBinnedModel
from multigrad_b200.ops import reference as ref

SELF = mg.Comm(None)
dev = torch.device("cuda", 0)
for n, k, link, p in [(20_000, 10, "identity", [-1.9, 0.3]),
                      (2_000_000, 10, "identity", [-2.0, 0.2]),
                      (2_000_000, 17, "log10", [-2.0, -0.5]),
                      (50_000_000, 10, "identity", [-1.0, 0.5])]:
    g = torch.Generator().manual_seed(0)
    x = (10.0 + 1.3 * torch.rand(n, generator=g)).float()
    edges = np.linspace(9, 10, k + 1)
    m = ErfBinnedModel(x=x.to(dev), bin_edges=edges, volume=10.0 * n,
                       sigma_link=link, comm=SELF, device=dev)
    s, jac = m.calc_sumstats_and_jac_from_params(p)
    e64 = torch.as_tensor(edges.astype(np.float32)).double()
    n64 = torch.as_tensor(m._norm64.astype(np.float32)).double()
    p64 = torch.as_tensor(np.asarray(p, np.float32)).double()
    so, dmu, dsg = ref.erf_binned_partial_and_jac(
        x.double(), e64, n64, p64[0], p64[1], link, chunk=1 << 20)
    rs = (s.double().cpu() / so - 1).abs().max().item()
    rmu = ((jac[:, 0].double().cpu() - dmu).abs().max() / dmu.abs().max()).item()
    rsg = ((jac[:, 1].double().cpu() - dsg).abs().max() / dsg.abs().max()).item()
    print(f"n={n} k={k} {link}: max rel err s={rs:.2e} dmu={rmu:.2e} "
          f"dsig={rsg:.2e}", flush=True)
