"""Probe the TMEM data-path layout of tcgen05.mma with M = 64."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from multigrad_b200.ops import _lib
lib = _lib.load()
dev = torch.device("cuda", 0)
P = 64
g = torch.Generator().manual_seed(1)
W = torch.randn(128, P, generator=g).to(torch.bfloat16).to(dev)
theta = torch.randn(P, generator=g).to(dev)
G = torch.rand(32, 128, generator=g).to(dev)
mu = torch.zeros(128, device=dev)
J = torch.full((128, P), float("nan"), device=dev)
status = torch.zeros(1, dtype=torch.int32, device=dev)
rc = lib.mgb_linlink_selftest(W.data_ptr(), theta.data_ptr(), G.data_ptr(),
                              mu.data_ptr(), J.data_ptr(), status.data_ptr(),
                              P + 1000, None)
torch.cuda.synchronize()
J_ref = (G.to(torch.bfloat16).double() @ W.double()).float().cpu()
Jc = J.cpu()
print("rc", rc, "status", int(status))
for r in range(32):
    d = (Jc - J_ref[r][None, :]).abs().max(dim=1).values
    lane = int(d.argmin())
    print(f"row {r:2d} -> lane {lane:3d} (err {float(d[lane]):.2e})")
