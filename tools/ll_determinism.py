"""Bitwise repeatability of the linear-link kernel (evaluate + Adam)."""
import sys
import numpy as np
import torch
import multigrad_b200 as mg
from multigrad_b200.models import LinearLinkBinnedModel

n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 2_000_000
dev = torch.device("cuda", 0)
import os
for F in [int(v) for v in os.environ.get("LLF", "64,128").split(",")]:
    g = torch.Generator().manual_seed(1)
    x = (10.0 + 1.3 * torch.rand(n, generator=g)).float().to(dev)
    feat = torch.cat([torch.ones(n, 1), 0.3 * torch.randn(n, F - 1, generator=g)], 1).to(dev)
    m = LinearLinkBinnedModel(x=x, features=feat, bin_edges=np.linspace(9, 10, 11),
                              volume=10.0 * n, comm=mg.Comm(None), device=dev)
    p = torch.cat([torch.tensor([-2.0]), 0.05 * torch.randn(F - 1, generator=g),
                   torch.tensor([0.2])])
    s0, j0 = m.calc_sumstats_and_jac_from_params(p)
    bad_s = bad_j = 0
    for i in range(200):
        s, j = m.calc_sumstats_and_jac_from_params(p)
        bad_s += int(not torch.equal(s, s0))
        bad_j += int(not torch.equal(j, j0))
        if not torch.equal(j, j0) and bad_j <= 2:
            d = (j - j0).abs()
            k = int(d.argmax())
            print("  jac differs: max", float(d.max()), "at", divmod(k, j.shape[1]),
                  "n_diff", int((d > 0).sum()))
        if not torch.equal(s, s0) and bad_s <= 2:
            print("  s differs:", (s - s0).tolist())
    m.set_target(s0 * 1.05)
    t0 = m.run_adam(p, nsteps=40, learning_rate=0.01)
    bad_t = sum(int(not torch.equal(m.run_adam(p, nsteps=40, learning_rate=0.01), t0))
                for _ in range(10))
    m._check()
    print(f"F={F}: sumstats mismatches {bad_s}/200, jac {bad_j}/200, adam {bad_t}/10")
