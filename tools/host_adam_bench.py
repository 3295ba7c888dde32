#!/usr/bin/env python
"""Host-driven Adam (run_adam_host) steps/s on one GPU, for tuning the
host<->GPU hand-shake.  usage: python tools/host_adam_bench.py [halos] [steps]
Env: MULTIGRAD_B200_HOST_PIPELINE=0 disables the pre-launched next step."""
import sys
import time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np
import torch
import multigrad_b200 as mg
from multigrad_b200.models import smf, SMFModel

N = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100_000_000
K = int(sys.argv[2]) if len(sys.argv) > 2 else 200
dev = torch.device("cuda", 0)
x = smf.synthetic_log_halo_masses(N, 0, 1, device=dev)
m = SMFModel(x=x, bin_edges=np.linspace(9, 10, 11), volume=10.0 * N,
             comm=mg.Comm(None), device=dev, target_sumstats=np.ones(10))
m.set_target(m.calc_sumstats_from_params([-2.0, 0.2]))
m.run_adam_host([-1.0, 0.5], nsteps=5)
m.run_adam([-1.0, 0.5], nsteps=5)
torch.cuda.synchronize()
t0 = time.perf_counter()
th = m.run_adam_host([-1.0, 0.5], nsteps=K)
torch.cuda.synchronize()
t_host = time.perf_counter() - t0
t0 = time.perf_counter()
td = m.run_adam([-1.0, 0.5], nsteps=K)
torch.cuda.synchronize()
t_dev = time.perf_counter() - t0
print(f"halos {N:.1e} steps {K}: host-driven {1e6 * t_host / K:8.1f} us/step, "
      f"device-resident {1e6 * t_dev / K:8.1f} us/step, "
      f"final {th[-1].tolist()} vs {td[-1].cpu().tolist()}")
