"""tcgen05 building-block check (run on a GPU box): one 128-object tile
through MMA #1 (mu = W.theta) and MMA #2 (J = G.W) vs torch."""
import ctypes as C
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from multigrad_b200.ops import _lib

lib = _lib.load()
dev = torch.device("cuda", 0)
rc_all = 0
for P in (64, 128):
    g = torch.Generator().manual_seed(P)
    W = torch.randn(128, P, generator=g).to(torch.bfloat16).to(dev)
    theta = torch.randn(P, generator=g).to(dev)
    G = torch.rand(32, 128, generator=g).to(dev)
    mu = torch.zeros(128, device=dev)
    J = torch.zeros(32, P, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    rc = lib.mgb_linlink_selftest(W.data_ptr(), theta.data_ptr(), G.data_ptr(),
                                  mu.data_ptr(), J.data_ptr(),
                                  status.data_ptr(), P, None)
    torch.cuda.synchronize()
    mu_ref = W.double() @ theta.double()
    J_ref = G.to(torch.bfloat16).double() @ W.double()
    e_mu = ((mu.double() - mu_ref).abs().max() / mu_ref.abs().max()).item()
    e_j = ((J.double() - J_ref).abs().max() / J_ref.abs().max()).item()
    print(f"P={P}: rc={rc} status={int(status)} err_mu={e_mu:.2e} "
          f"err_J={e_j:.2e}", flush=True)
    if rc or int(status) or e_mu > 1e-5 or e_j > 1e-5:
        rc_all = 1
        print("  mu", mu[:6].tolist(), "\n  ref", mu_ref[:6].tolist())
        print("  J ", J[0, :6].tolist(), "\n  ref", J_ref[0, :6].tolist())
sys.exit(rc_all)
