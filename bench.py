#!/usr/bin/env python
"""Headline benchmark: Adam steps/s (and loss+grad evals/s) of the SMF
OnePointModel fit on a synthetic 1e8-halo catalogue (BASELINE.json config 2/3).

    python bench.py --gpus N --steps K --warmup W [--impl ours|reference|...]

For N > 1 launch under torchrun (one rank per GPU).  Rank 0 prints ONE JSON
line.  ``--impl reference`` would run the unmodified reference from
``baseline/_ref``; it cannot be installed on this image (no jax, no mpi4py, no
setuptools_scm -- see DESIGN.md), so that arm reports ``unavailable``.
``--impl proxy-nccl | proxy-host | cpu-oracle`` are the labelled stand-ins of
BASELINE.md section 3 (the reference's algorithm in PyTorch).

Timing: CUDA events on the launching stream, barrier + synchronize on both
sides, max over ranks.  The total catalogue is fixed (strong scaling).  When
a rank's shard is larger than L2 the steps run back to back ("inputs>L2");
otherwise every step is timed on its own and a 256 MiB buffer is written
between steps ("flushed").
"""
import argparse
import json
import os
import subprocess
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
os.environ.setdefault("MULTIGRAD_B200_NO_TQDM", "1")

L2_BYTES = 126 << 20
GUESS = (-1.0, 0.5)
TRUTH = (-2.0, 0.2)
LR = 0.01


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--halos", type=float, default=1e8)
    ap.add_argument("--graph", type=int, default=0)
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def reference_arm():
    why = ("reference (JAX + mpi4py) cannot run on this image: `pip install "
           "--no-index --no-build-isolation --find-links /opt/wheelhouse "
           "--target baseline/_ref /root/reference` fails with "
           "ModuleNotFoundError: setuptools_scm, and jax / jaxlib / mpi4py "
           "are neither installed nor in the wheelhouse")
    ref = ROOT / "baseline" / "_ref" / "multigrad"
    if ref.exists():
        try:
            sys.path.insert(0, str(ref.parent))
            import jax  # noqa: F401
            import mpi4py  # noqa: F401
        except Exception as e:  # noqa: BLE001
            why = f"baseline/_ref present but not importable: {e!r}"
    print(json.dumps({"impl": "reference", "unavailable": why}))
    return 0


class ClockSampler:
    """nvidia-smi clock / throttle-reason sampler running during the timed
    region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,"
         "clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpus):
        self.gpus = set(range(gpus))
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "25"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [],
                    "note": "nvidia-smi unavailable"}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
            out, _ = self.proc.communicate()
        sm, smax, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                 "sw_power_cap"]
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 8:
                continue
            try:
                if int(f[0]) not in self.gpus:
                    continue
                sm.append(float(f[1]))
                smax.append(float(f[2]))
                power.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(names, f[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None,
                "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def main():
    args = parse()
    if args.impl == "reference":
        return reference_arm()

    import numpy as np
    import torch
    import multigrad_b200 as mg
    from multigrad_b200.models import smf, SMFModel
    from multigrad_b200.ops import erfbin, allreduce

    comm = mg.COMM
    rank, world = comm.rank, comm.size
    if world != args.gpus and rank == 0:
        print(f"warning: --gpus {args.gpus} but WORLD_SIZE={world}",
              file=sys.stderr)
    if args.impl == "cpu-oracle":
        dev = torch.device("cpu")
    else:
        assert torch.cuda.is_available(), "bench.py needs a GPU"
        dev = mg.parallel.local_device()
    N = int(args.halos)
    K, W = args.steps, args.warmup
    x = smf.synthetic_log_halo_masses(N, rank, world, device=dev)
    edges = np.linspace(9, 10, 11)
    volume = 10.0 * N

    def sync_all():
        if dev.type == "cuda":
            torch.cuda.synchronize()
        comm.Barrier()

    def max_over_ranks(v):
        return max(comm.allgather(float(v)))

    shard_bytes = x.numel() * 4
    flush = dev.type == "cuda" and shard_bytes <= 2 * L2_BYTES
    flush_buf = (torch.empty(256 << 20, dtype=torch.uint8, device=dev)
                 if flush else None)

    def timed(step_fn, nsteps):
        """ms for nsteps calls of step_fn (device-timed, max over ranks)."""
        sync_all()
        if dev.type != "cuda":
            t0 = time.perf_counter()
            for _ in range(nsteps):
                step_fn()
            return max_over_ranks((time.perf_counter() - t0) * 1e3)
        if not flush:
            e0, e1 = (torch.cuda.Event(enable_timing=True) for _ in range(2))
            e0.record()
            for _ in range(nsteps):
                step_fn()
            e1.record()
            sync_all()
            return max_over_ranks(e0.elapsed_time(e1))
        evs = [(torch.cuda.Event(enable_timing=True),
                torch.cuda.Event(enable_timing=True)) for _ in range(nsteps)]
        for a, b in evs:
            flush_buf.fill_(1)
            a.record()
            step_fn()
            b.record()
        sync_all()
        return max_over_ranks(sum(a.elapsed_time(b) for a, b in evs))

    clocks = ClockSampler(world) if rank == 0 and dev.type == "cuda" else None
    extra = {}

    if args.impl == "ours":
        model = SMFModel(x=x, bin_edges=edges, volume=volume, comm=comm,
                         device=dev, target_sumstats=np.ones(10))
        target = model.calc_sumstats_from_params(TRUTH)
        model.set_target(target)
        plan = model.plan()
        # --- Adam: the public run_adam issues exactly these launches ----------
        from multigrad_b200.ops.erfbin import FitState, MODE_ADAM
        p0 = model._params(GUESS)
        st = FitState(p0, p0, K + W, dev)
        st.traj[0].copy_(p0)
        a = plan.make_optimizer_args(MODE_ADAM, st, LR)
        graph = None
        if args.graph:
            plan.launch(a)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                plan.launch(a)

        def adam_step():
            if graph is not None:
                graph.replay()
            else:
                plan.launch(a)
        for _ in range(W):
            adam_step()
        n0 = erfbin.launch_count() + allreduce.launch_count()
        if clocks:
            clocks.start()
        ms = timed(adam_step, K)
        clk = clocks.stop() if clocks else None
        launches = (K if graph is not None else
                    erfbin.launch_count() + allreduce.launch_count() - n0)
        model._check_peers()
        final = st.p.cpu().tolist()
        extra["final_params"] = final
        extra["final_loss"] = float(st.loss_hist[: int(st.step.item())][-1])
        # --- loss+grad evaluations (the BFGS line-search step) ----------------
        pe = model._params(GUESS)
        from multigrad_b200.ops.erfbin import MODE_LOSS_GRAD

        def eval_step():
            plan.evaluate(pe, MODE_LOSS_GRAD)
        for _ in range(3):
            eval_step()
        ms_eval = timed(eval_step, K)
        extra["loss_grad_evals_per_sec"] = K / (ms_eval * 1e-3)
        extra["ms_per_eval"] = ms_eval / K
        # --- run_bfgs end to end (config 3): scipy L-BFGS-B on rank 0, one
        # fused loss+grad launch per line-search evaluation, parameters handed
        # to the other GPUs by the device-side broadcast ---------------------
        if not args.no_e2e:
            # tutorial loss (log10(s + 1e-10)): finite where bins are empty,
            # so the bounded line search can recover from a long first step
            bmodel = SMFModel(x=x, bin_edges=edges, volume=volume, comm=comm,
                              device=dev, target_sumstats=target,
                              loss_eps=1e-10)
            bmodel.run_bfgs([-1.6, 0.3], maxsteps=2,        # warm the plan
                            param_bounds=[(-4.0, 0.0), (0.05, 2.0)])
            sync_all()
            t0 = time.perf_counter()
            res = bmodel.run_bfgs([-1.5, 0.35], maxsteps=40,
                                  param_bounds=[(-4.0, 0.0), (0.05, 2.0)])
            sync_all()
            dt = max_over_ranks(time.perf_counter() - t0)
            extra["run_bfgs"] = {
                "nit": int(res.nit), "nfev": int(res.nfev),
                "seconds": dt, "evals_per_sec": res.nfev / dt,
                "x": [float(v) for v in res.x], "fun": float(res.fun),
                "config": "guess (-1.5, 0.35), bounds mu in [-4, 0], sigma in "
                          "[0.05, 2], loss_eps 1e-10, maxsteps 40",
                "timing": "host wall clock, max over ranks"}
        # --- end to end through the public host API ---------------------------
        e2e = None
        if not args.no_e2e:
            h = np.array(GUESS, dtype=np.float32)
            m = np.zeros(2, np.float32)
            v = np.zeros(2, np.float32)

            def host_adam_step(i):
                nonlocal h, m, v
                loss, g = model.loss_and_grad_host(h)   # H2D 8 B, D2H 12 B
                g = g.astype(np.float32)
                m = np.float32(0.1) * g + np.float32(0.9) * m
                v = np.float32(0.001) * g * g + np.float32(0.999) * v
                mh = m / np.float32(1 - 0.9 ** (i + 1))
                vh = v / np.float32(1 - 0.999 ** (i + 1))
                h = h - np.float32(LR) * mh / (np.sqrt(vh) + np.float32(1e-8))
                return loss
            for i in range(3):
                host_adam_step(i)
            h = np.array(GUESS, dtype=np.float32)
            m[:] = 0
            v[:] = 0
            sync_all()
            e0, e1 = (torch.cuda.Event(enable_timing=True) for _ in range(2))
            e0.record()
            for i in range(K):
                host_adam_step(i)
            e1.record()
            sync_all()
            ms_py = max_over_ranks(e0.elapsed_time(e1))
            # the public host-driven fit: optimiser state on the host, every
            # step ships the parameters to the GPU and loss+grad back
            # Same L2 rule as the device-timed loop: where the shard does not
            # exceed the cache it is flushed in front of every step; the wait
            # for that fill is not part of a step (host clock per step:
            # [params published -> loss+grad read back], summed, max over
            # ranks).
            model.run_adam_host(GUESS, nsteps=5, learning_rate=LR,
                                _flush_l2=flush_buf)
            sync_all()
            traj_h, loss_h = model.run_adam_host(
                GUESS, nsteps=K, learning_rate=LR, return_loss=True,
                _flush_l2=flush_buf)
            ms_e2e = max_over_ranks(model.host_timed_seconds * 1e3)
            sync_all()
            e2e = {"value": K / (ms_e2e * 1e-3), "unit": "adam_steps/s",
                   "h2d_bytes_per_step": 8, "d2h_bytes_per_step": 12,
                   "l2": "flushed" if flush else "inputs>L2",
                   "timing": "host clock per step, params published -> "
                             "loss+grad read back, summed; max over ranks",
                   "final_params": [float(t) for t in traj_h[-1]],
                   "final_loss": float(loss_h[-1]),
                   "python_loop_steps_per_sec": K / (ms_py * 1e-3),
                   "how": "model.run_adam_host: Adam state on the host; "
                          "every step writes the params to pinned host "
                          "memory, ONE fused launch reads them over PCIe "
                          "(zero copy) and writes loss+grad into pinned host "
                          "memory, the host polls the result words and "
                          "updates (native loop); python_loop = the same "
                          "with the update written in numpy around "
                          "model.loss_and_grad_host; catalogue resident on "
                          "the GPUs.  The grid of a step is pre-launched "
                          "and already resident when its parameters are "
                          "published, so this interval holds no launch "
                          "latency (the event-bracketed device number "
                          "does)"}
        metric = "adam_steps_per_sec"
    else:
        model, step_fn, launches = build_proxy(args.impl, x, edges, volume,
                                               comm, dev, torch, np, mg, smf)
        for _ in range(W):
            step_fn()
        if clocks:
            clocks.start()
        ms = timed(step_fn, K)
        clk = clocks.stop() if clocks else None
        e2e = None
        metric = "adam_steps_per_sec"

    if rank == 0:
        val = K / (ms * 1e-3)
        out = {
            "metric": metric, "value": val, "unit": "steps/s",
            "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K,
            "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "impl": args.impl,
            "config": {
                "model": "SMF OnePointModel (erf soft-binned, K=10 bins, "
                         "P=2 params, log-MSE loss)",
                "global_batch": N, "seq_len": None,
                "parallelism": f"dp{world}",
                "halos_per_gpu": x.numel(),
                "l2": "flushed" if flush else "inputs>L2",
                "optimizer": f"adam lr={LR}", "cuda_graph": bool(args.graph),
            },
            "clocks": clk, "gpu_launches": launches,
        }
        if e2e is not None:
            out["e2e"] = e2e
        out.update(extra)
        print(json.dumps(out))
    mg.shutdown()
    return 0


def build_proxy(impl, x, edges, volume, comm, dev, torch, np, mg, smf):
    """Reference-shaped baselines: the reference's
    algorithm -- forward, all-reduce, dL/ds, VJP, all-reduce, host-driven
    Adam -- in PyTorch.  proxy-nccl: torch.distributed (NCCL) all-reduces;
    proxy-host: D2H -> gloo -> H2D staging like np.asarray + MPI;
    cpu-oracle: everything on the CPU."""
    import torch.distributed as dist
    from multigrad_b200 import adam as _adam
    from multigrad_b200.ops import reference as ref
    edges_t = torch.as_tensor(edges, dtype=torch.float32, device=dev)
    norm_t = torch.as_tensor(1.0 / (volume * np.diff(edges)),
                             dtype=torch.float32, device=dev)
    world = comm.size

    def allreduce(t):
        if world == 1:
            return t
        if impl == "proxy-host" or dev.type == "cpu":
            h = t.detach().cpu()
            dist.all_reduce(h, group=comm.group)
            return h.to(dev)
        t = t.detach().clone()
        dist.all_reduce(t, group=comm.group)
        return t

    def sumstats_fn(p):
        return ref.erf_binned_partial(x, edges_t, norm_t, p[0], p[1],
                                      chunk=1 << 24)

    with torch.no_grad():
        target = allreduce(sumstats_fn(torch.tensor(TRUTH, device=dev)))
    ltarget = torch.log10(target)

    def loss_fn(s):
        return torch.mean((torch.log10(s) - ltarget) ** 2)

    state = {"p": torch.tensor(GUESS, device=dev), "m": torch.zeros(2, device=dev),
             "v": torch.zeros(2, device=dev), "i": 0}

    def step():
        p = state["p"].clone().requires_grad_(True)
        s = sumstats_fn(p)
        tot = allreduce(s).requires_grad_(True)
        loss = loss_fn(tot)
        (dlds,) = torch.autograd.grad(loss, tot)
        (g,) = torch.autograd.grad(s, p, grad_outputs=dlds)
        g = allreduce(g)
        loss_fn(tot.detach())            # the reference evaluates it twice
        state["p"], state["m"], state["v"] = _adam.adam_update(
            state["i"], state["p"], g, state["m"], state["v"], LR)
        state["i"] += 1
    return None, step, 0


if __name__ == "__main__":
    sys.exit(main())
