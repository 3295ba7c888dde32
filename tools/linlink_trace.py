#!/usr/bin/env python
"""Pipeline trace of the linear-link kernel (debug build with -DLL_TRACE).

    nvcc ... -DLL_TRACE -o build/variants/libmgb200_trace.so   (see below)
    python tools/linlink_trace.py [--halos 2e7] [--features 64]

Builds the traced variant if it is missing, runs a few steps, then reads the
SM-clock stamps CTA 0 recorded for 256 consecutive tiles and prints where a
tile spends its time between the TMA issue and the release of its stage:

    0 TMA issued (W, x, weights) -> 2 the tile's compute warps passed full_w
    (5 = when they started waiting: G buffer first, then the tile) -> forward
    product (ldmatrix + mma.sync), erfc math, G stores -> 3 G published ->
    4 issuer saw g_full -> Jacobian UMMAs -> 7 committed -> stage free (= event
    0 of tile + stages)
"""
import argparse
import ctypes as C
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
LIB = ROOT / "build" / "variants" / "libmgb200_trace.so"


def build():
    LIB.parent.mkdir(parents=True, exist_ok=True)
    srcs = [str(ROOT / "csrc" / f) for f in
            ("workspace.cu", "erfbin.cu", "allreduce.cu", "adam.cu",
             "linlink.cu", "group.cu", "nvls.cu")]
    subprocess.run(
        ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
         "-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-DLL_TRACE", "-I",
         str(ROOT / "include"), "-shared", "-o", str(LIB), *srcs],
        check=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--halos", type=float, default=2e7)
    ap.add_argument("--features", type=int, default=64)
    ap.add_argument("--build-only", action="store_true")
    args = ap.parse_args()
    if not LIB.exists() or args.build_only:
        build()
    if args.build_only:
        return
    os.environ["MULTIGRAD_B200_LIB"] = str(LIB)
    sys.path.insert(0, str(ROOT))
    import numpy as np
    import torch
    import multigrad_b200 as mg
    from multigrad_b200.models import LinearLinkBinnedModel
    from multigrad_b200.ops import _lib

    dev = torch.device("cuda", 0)
    n, F = int(args.halos), args.features
    g = torch.Generator(device=dev).manual_seed(0)
    x = 10.0 + 1.3 * torch.rand(n, generator=g, device=dev)
    feat = torch.empty(n, F, device=dev, dtype=torch.bfloat16)
    feat.normal_(0, 0.3, generator=g)
    feat[:, 0] = 1.0
    theta = torch.zeros(F + 1)
    theta[0], theta[-1] = -2.0, 0.2
    model = LinearLinkBinnedModel(
        x=x, features=feat, bin_edges=np.linspace(9, 10, 11),
        volume=10.0 * n, target_sumstats=np.ones(10), comm=mg.Comm(None))
    model.set_target(model.calc_sumstats_from_params(theta))
    for _ in range(3):
        model.calc_loss_and_grad_from_params(theta + 0.05)
    torch.cuda.synchronize()
    lib = _lib.load()
    NT = 256
    buf = (C.c_longlong * (8 * NT))()
    rc = lib.mgb_linlink_trace_read(buf)
    assert rc == 0, rc
    t = np.frombuffer(buf, dtype=np.int64).reshape(8, NT).astype(np.float64)
    ghz = 1.92
    stages = 8 if F <= 64 else 5
    ok = slice(8, NT - stages)

    def stat(name, d):
        d = d[ok] / ghz / 1e3
        print(f"{name:46s} mean {d.mean():6.2f} us  p10 {np.percentile(d, 10):6.2f}"
              f"  p90 {np.percentile(d, 90):6.2f}")
    period = np.diff(t[0])[ok]
    print(f"tiles traced: {NT}, stages {stages}; tile period "
          f"{period.mean() / ghz / 1e3:.3f} us")
    stat("0->2 TMA issue -> compute warps have the tile", t[2] - t[0])
    stat("5->2 compute warps waiting (G buffer + tile)", t[2] - t[5])
    stat("2->3 forward product, math, G stores", t[3] - t[2])
    stat("3->4 G published -> issuer sees g_full", t[4] - t[3])
    stat("4->7 issuer: Jacobian UMMAs issue + commit", t[7] - t[4])
    rel = np.roll(t[0], -stages)
    stat("4->0' Jacobian product -> stage free -> next TMA", rel - t[4])
    stat("0->0' stage residency", rel - t[0])
    # per compute warp: when it had its tile and when it published G,
    # relative to the earliest warp of the same tile
    sets = int(os.environ.get("LL_TRACE_SETS", "3"))
    wb = (C.c_longlong * (32 * 2 * NT))()
    assert lib.mgb_linlink_trace_read_warps(wb) == 0
    w = np.frombuffer(wb, dtype=np.int64).reshape(32, 2, NT).astype(np.float64)
    print("compute warp: set quad | busy us | start lag us | finish lag us")
    t0 = 300                                   # LL_TRACE_T0 in csrc/linlink.cu
    for st in range(sets):
        first = 8 + (st - (t0 + 8)) % sets     # first traced tile of this set
        idx = np.arange(first, NT - stages, sets)
        ws = w[st * 4:(st + 1) * 4][:, :, idx]
        first_start = ws[:, 0].min(0)
        first_done = ws[:, 1].min(0)
        for k in range(4):
            busy = (ws[k, 1] - ws[k, 0]).mean() / ghz / 1e3
            lag0 = (ws[k, 0] - first_start).mean() / ghz / 1e3
            lag1 = (ws[k, 1] - first_done).mean() / ghz / 1e3
            quad = (st * 4 + k + 2) & 3        # warp index & 3 = SM sub-partition
            print(f"   set {st} quad {quad} | {busy:5.2f} | {lag0:5.2f} | "
                  f"{lag1:5.2f}")


if __name__ == "__main__":
    main()
