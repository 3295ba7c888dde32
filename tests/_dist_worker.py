"""Multi-rank test body, run under torchrun by test_distributed_cpu.py (gloo,
CPU) and test_multigpu.py (one rank per GPU).  Every case runs on every rank;
rank 0 prints ``RESULT {case: "ok" | "<error>"}``.

Mirrors how the reference tests multi-process behaviour: the same assertions
under real multi-process launches of different sizes (reference
``tests/test_mpi.py:1-7``), never a mocked communicator.
"""
import json
import os
import sys
import traceback

import numpy as np
import torch

import multigrad_b200 as mg
from multigrad_b200 import adam
from multigrad_b200.models import smf, SMFModel
from multigrad_b200.parallel import group_assignment

COMM = mg.COMM
RANK, SIZE = COMM.rank, COMM.size
DEV = mg.parallel.local_device()
CASES = {}


def case(fn):
    CASES[fn.__name__] = fn
    return fn


@case
def reduce_sum_scalar_and_array():
    # reference tests/test_mpi.py:19-35
    expected = sum(range(SIZE))
    assert mg.reduce_sum(torch.tensor(RANK)) == expected
    assert mg.reduce_sum(float(RANK)) == float(expected)
    out = mg.reduce_sum(np.full(3, RANK, dtype=np.float64))
    assert isinstance(out, np.ndarray) and np.all(out == expected)
    t = mg.reduce_sum(torch.full((5,), float(RANK), device=DEV))
    assert t.device.type == DEV.type and torch.all(t == expected)
    gathered = COMM.allgather(float(t[0]))
    assert all(x == expected for x in gathered)
    r = mg.reduce_sum(np.ones(2) * RANK, root=SIZE - 1)
    if RANK == SIZE - 1:
        assert np.all(r == expected)
    else:
        assert np.all(r == 0)           # Q7: zeros, not uninitialised memory


@case
def comm_object_collectives():
    assert COMM.bcast({"a": RANK} if RANK == 0 else None, root=0) == {"a": 0}
    assert COMM.allgather(RANK) == list(range(SIZE))
    COMM.Barrier()
    if SIZE > 1:
        if RANK == 0:
            COMM.send(np.arange(3), dest=1)
        elif RANK == 1:
            assert np.array_equal(COMM.recv(source=0), np.arange(3))
    assert COMM.Get_rank() == RANK and COMM.Get_size() == SIZE
    assert mg.RANK == RANK and mg.N_RANKS == SIZE


@case
def scatter_nd_chunks():
    full = np.arange(7 * 3, dtype=np.float32).reshape(7, 3)
    mine = mg.util.scatter_nd(full if RANK == 0 else None, axis=0, root=0)
    want = np.array_split(full, SIZE, axis=0)[RANK]
    assert np.array_equal(np.asarray(mine), want)
    cols = mg.scatter_nd(full if RANK == 0 else None, axis=1)
    assert np.array_equal(np.asarray(cols),
                          np.array_split(full, SIZE, axis=1)[RANK])


@case
def split_subcomms_mapping():
    if SIZE < 2:
        return
    sub, n, g = mg.split_subcomms(num_groups=2)
    table, _ = group_assignment(SIZE, 2)
    assert n == 2 and g == table[RANK]
    assert sub.size == table.count(g)
    assert sub.rank == table[:RANK].count(g)
    assert mg.reduce_sum(1.0, comm=sub) == float(sub.size)
    rpg = [1, SIZE - 1]
    sub2, n2, g2 = mg.split_subcomms(ranks_per_group=rpg)
    assert n2 == 2 and g2 == (0 if RANK == 0 else 1)
    assert sub2.size == rpg[g2]
    node, n_nodes, idx = mg.split_subcomms_by_node()
    assert n_nodes == 1 and idx == 0 and node.size == SIZE
    for bad in (dict(), dict(num_groups=2, ranks_per_group=rpg),
                dict(num_groups=SIZE + 1), dict(ranks_per_group=[SIZE + 3])):
        try:
            mg.split_subcomms(**bad)
        except AssertionError:
            continue
        raise AssertionError(f"split_subcomms({bad}) should have asserted")


def _data(n):
    return smf.make_smf_data(n, comm=COMM, device=DEV)


def _pipeline(model, data):
    # reference tests/test_mpi.py:38-66 (world-size invariant constants)
    truth = smf.ParamTuple(log_shmrat=-2.0, sigma_logsm=0.2)
    s = model.calc_sumstats_from_params(truth)
    assert torch.allclose(data["target_sumstats"].cpu(), s.cpu()), s
    gd = model.run_simple_grad_descent(guess=truth, nsteps=2)
    assert torch.isclose(gd.loss[-1].cpu(), torch.tensor(0.0))
    assert torch.allclose(gd.params[-1].cpu(), torch.tensor([*truth]))
    assert torch.allclose(model.calc_dloss_dparams(truth).cpu(),
                          torch.zeros(2), atol=1e-4)
    loss, grad = model.calc_loss_and_grad_from_params(truth)
    assert torch.allclose(loss, model.calc_loss_from_params(truth))
    assert torch.allclose(grad, model.calc_dloss_dparams(truth), atol=1e-6)


@case
def pipeline_generic():
    data = _data(10_000)
    _pipeline(smf.MySMFModel(aux_data=data), data)


@case
def pipeline_native():
    data = _data(10_000)
    _pipeline(SMFModel(aux_data=data), data)


@case
def world_size_invariance():
    """loss / grad / Adam trajectory equal the single-process values."""
    data = _data(6_000)
    model = SMFModel(aux_data=data)
    full = smf.make_smf_data(6_000, comm=mg.Comm(None), device="cpu")
    ref = SMFModel(aux_data=full, comm=mg.Comm(None), device="cpu")
    p = torch.tensor([-1.9, 0.3])
    loss, grad = model.calc_loss_and_grad_from_params(p)
    rl, rg = ref.calc_loss_and_grad_from_params(p)
    assert torch.allclose(loss.cpu(), rl, rtol=2e-5), (loss, rl)
    assert torch.allclose(grad.cpu(), rg, rtol=2e-4, atol=1e-6), (grad, rg)
    traj = model.run_adam(p, nsteps=8, learning_rate=0.02)
    rt = ref.run_adam(p, nsteps=8, learning_rate=0.02)
    assert traj.shape == (9, 2)
    assert torch.allclose(traj.cpu(), rt, atol=2e-4), (traj, rt)
    # identical on every rank, bit for bit
    all_t = COMM.allgather(traj.cpu().numpy().tobytes())
    assert all(t == all_t[0] for t in all_t)
    tb = model.run_adam(p, nsteps=8, learning_rate=0.02,
                        param_bounds=[(-5.0, 0.0), (0.01, None)])
    rb = ref.run_adam(p, nsteps=8, learning_rate=0.02,
                      param_bounds=[(-5.0, 0.0), (0.01, None)])
    assert torch.allclose(tb.cpu(), rb, atol=2e-4), (tb, rb)


@case
def bfgs_all_ranks_same_result():
    x = torch.log10(smf.load_halo_masses_tutorial(10_000, comm=COMM)).to(DEV)
    edges = torch.linspace(9, 10, 11)
    true = torch.tensor([-2.0, -0.5])
    model = SMFModel(x=x, bin_edges=edges, volume=1.0, sigma_link="log10",
                     loss_eps=1e-10, target_sumstats=torch.ones(10))
    model.set_target(model.calc_sumstats_from_params(true))
    res = model.run_bfgs(true + torch.tensor([-1.5, 0.7]))
    assert res.success and np.allclose(res.x, [-2.0, -0.5], atol=2e-4), res
    xs = COMM.allgather(np.asarray(res.x).tobytes())
    assert all(v == xs[0] for v in xs)
    assert res.nit >= 5 and res.nfev >= res.nit
    gen = smf.TutorialSMFModel(aux_data=dict(
        log_halo_masses=x, smf_bin_edges=edges.to(DEV), volume=1.0,
        target_sumstats=model.target_sumstats))
    res2 = gen.run_bfgs(true + torch.tensor([-1.5, 0.7]))
    assert np.allclose(res2.x, [-2.0, -0.5], atol=2e-4), res2


@case
def gpu_bfgs_rank_skew_stress():
    """Host-driven evaluations (the body of the multi-GPU L-BFGS-B) with a
    random delay injected on every rank in front of every evaluation: 1000
    evaluations, every rank must read back the same bits each time, and a
    complete run_bfgs under the same skew must still agree on all ranks
    (stands in for the race detector that is closed on this
    pool)."""
    if DEV.type != "cuda" or SIZE < 2:
        return
    import random
    import time as _time
    from multigrad_b200.models.bfgs_device import run_bfgs_replicated
    x = torch.log10(smf.load_halo_masses_tutorial(200_000, comm=COMM)).to(DEV)
    edges = torch.linspace(9, 10, 11)
    model = SMFModel(x=x, bin_edges=edges, volume=1.0, sigma_link="log10",
                     loss_eps=1e-10, target_sumstats=torch.ones(10))
    model.set_target(model.calc_sumstats_from_params([-2.0, -0.5]))
    rnd = random.Random(1234 + RANK)
    prm = random.Random(99)                      # same points on every rank
    evaluate, end = model._host_session(np.array([-2.1, -0.4]))
    digest = []
    try:
        for i in range(1000):
            _time.sleep(rnd.random() * 3e-4 * (rnd.random() < 0.3))
            p = np.array([-2.0 + 0.5 * (prm.random() - 0.5),
                          -0.5 + 0.4 * (prm.random() - 0.5)])
            loss, grad = evaluate(p)
            digest.append(np.float64(loss).tobytes() + grad.tobytes())
    finally:
        end()
    blobs = COMM.allgather(b"".join(digest))
    assert all(b == blobs[0] for b in blobs)
    COMM.workspace().check_error()

    def skew(_n):
        _time.sleep(rnd.random() * 5e-4 * (rnd.random() < 0.5))
    res = run_bfgs_replicated(model, np.array([-3.5, 0.2]), 100, None, COMM,
                              _before_eval=skew)
    assert res.success and np.allclose(res.x, [-2.0, -0.5], atol=2e-4), res
    xs = COMM.allgather((np.asarray(res.x).tobytes(), int(res.nfev)))
    assert all(v == xs[0] for v in xs)


@case
def gpu_collective_soak():
    """Stand-in for the race detector that is closed on this pool (SURVEY.md
    5.2): MGB_SOAK_ITERS (default 2000; the committed 8-GPU log uses 10000)
    collectives of every kind in a pseudo-random order that all ranks share --
    ordered LL, one-shot push, two-shot, NVLS (staged and in place),
    broadcast, fused evaluation, fused persistent steps -- each behind a
    random per-rank delay on the host AND on the device (a spinning kernel in
    front of the collective), every result compared on the device with the
    exact expected sum; no mismatch and no sticky timeout flag allowed."""
    if DEV.type != "cuda" or SIZE < 2:
        return
    import os
    import random
    import time as _time
    from multigrad_b200.ops import allreduce as ar
    iters = int(os.environ.get("MGB_SOAK_ITERS", "2000"))
    shared = random.Random(4242)            # same choices on every rank
    mine = random.Random(777 + RANK)        # this rank's delays
    ws = COMM.workspace()
    have_mc = ws.multicast() is not None
    sym = ws.multicast_tensor(1 << 18) if have_mc else None
    sizes = {1: [1, 7, 240], 3: [1000, 65_536], 2: [50_000, 300_000],
             4: [4096, 1 << 19]}
    bad = torch.zeros(1, dtype=torch.int64, device=DEV)
    tri = SIZE * (SIZE + 1) // 2
    data = _data(20_000)
    model = SMFModel(aux_data=data, deterministic=True)   # bit-repeatable
    ref_loss, ref_grad = model.calc_loss_and_grad_from_params([-1.9, 0.3])
    ref_traj = model.run_adam([-1.9, 0.3], nsteps=3)
    kinds = ["ll", "push", "two", "bcast", "eval", "fit"] + \
        (["nvls", "nvls_inplace"] if have_mc else [])
    for i in range(iters):
        kind = shared.choice(kinds)
        mult = float(shared.randrange(1, 8))
        if mine.random() < 0.2:
            _time.sleep(mine.random() * 2e-4)
        if mine.random() < 0.5:
            torch.cuda._sleep(int(mine.random() * 200_000))    # device skew
        if kind in ("ll", "push", "two", "nvls"):
            algo = {"ll": 1, "push": 3, "two": 2, "nvls": 4}[kind]
            n = shared.choice(sizes[algo])
            t = torch.full((n,), mult * (RANK + 1), device=DEV)
            ar.all_reduce_sum_(t, COMM, algo=algo)
            bad += (t != mult * tri).any()
        elif kind == "nvls_inplace":
            sym.fill_(mult * (RANK + 1))
            ar.all_reduce_sum_(sym, COMM)
            bad += (sym != mult * tri).any()
        elif kind == "bcast":
            root = shared.randrange(SIZE)
            t = torch.full((9,), mult * (RANK + 1), device=DEV)
            ar.bcast_(t, COMM, root=root)
            bad += (t != mult * (root + 1)).any()
        elif kind == "eval":
            l, g = model.calc_loss_and_grad_from_params([-1.9, 0.3])
            bad += (l != ref_loss) | (g != ref_grad).any()
        else:
            tr = model.run_adam([-1.9, 0.3], nsteps=3)
            bad += (tr != ref_traj).any()
        if i % 500 == 499:
            torch.cuda.synchronize()
            ws.check_error()
    torch.cuda.synchronize()
    ws.check_error()
    nbad = COMM.allgather(int(bad.item()))
    assert sum(nbad) == 0, nbad
    print(f"[rank {RANK}] soak: {iters} mixed collectives, 0 mismatches",
          flush=True)


@case
def gpu_linked_model_world_invariance():
    """User-defined link on several GPUs: reduced statistic, Jacobian and a
    fused Adam fit are those of the whole catalogue on one GPU, stochastic
    draws follow the global object index."""
    if DEV.type != "cuda":
        return
    from multigrad_b200.models import LinkedBinnedModel
    n = 30_011
    g = torch.Generator().manual_seed(21)
    x_all = (10.0 + 1.3 * torch.rand(n, generator=g)).float()
    f_all = torch.randn(n, generator=g).float()
    edges = np.linspace(9, 10, 11)

    def lk(x, feat, p, rng):
        return (x + p[0] + 0.05 * feat[0] + p[3] * rng.normal(0),
                p[1] + p[2] * (x - 11.0))
    lo, hi = (n * RANK) // SIZE, (n * (RANK + 1)) // SIZE
    kw = dict(link=lk, nparams=4, bin_edges=edges, volume=10.0 * n, seed=5)
    part = LinkedBinnedModel(x=x_all[lo:hi].to(DEV),
                             features=[f_all[lo:hi].to(DEV)], **kw)
    assert part.index0 == lo
    whole = LinkedBinnedModel(x=x_all.to(DEV), features=[f_all.to(DEV)],
                              comm=mg.Comm(None), **kw)
    p = [-1.8, 0.3, 0.04, 0.1]
    s1, j1 = part.calc_sumstats_and_jac_from_params(p, step=2)
    s0, j0 = whole.calc_sumstats_and_jac_from_params(p, step=2)
    assert torch.allclose(s1, s0, rtol=2e-6)
    assert torch.allclose(j1, j0, rtol=2e-5, atol=2e-6 * float(j0.abs().max()))
    tgt = whole.calc_sumstats_from_params([-2.0, 0.25, 0.02, 0.1])
    part.set_target(tgt)
    whole.set_target(tgt)
    t1 = part.run_adam(p, nsteps=25, randkey=9)
    t0 = whole.run_adam(p, nsteps=25, randkey=9)
    assert torch.allclose(t1, t0, atol=2e-5)
    blobs = COMM.allgather(t1.cpu().numpy().tobytes())
    assert all(b == blobs[0] for b in blobs)
    COMM.workspace().check_error() if SIZE > 1 else None


@case
def group_of_models():
    """OnePointGroup over disjoint sub-communicators == sum of the models
    evaluated on all the data (config W4: multi-redshift SMF)."""
    if SIZE < 2:
        return
    vols = [1.0, 1.7]
    sub, n, g = mg.split_subcomms(num_groups=2)
    full = smf.make_smf_data(4_000, comm=mg.Comm(None), device="cpu")
    x_full = full["log_halo_masses"] + 0.05 * g     # "redshift" dependent
    x_mine = torch.tensor_split(x_full, sub.size)[sub.rank].to(DEV)
    mine = SMFModel(x=x_mine, bin_edges=full["smf_bin_edges"],
                    volume=full["volume"] * vols[g],
                    target_sumstats=full["target_sumstats"], comm=sub)
    group = mg.OnePointGroup((mine,), main_comm=COMM)
    p = torch.tensor([-1.95, 0.25])
    loss, grad = group.calc_loss_and_grad_from_params(p)
    rl, rg = 0.0, torch.zeros(2)
    for gi in range(2):
        ref = SMFModel(x=full["log_halo_masses"] + 0.05 * gi,
                       bin_edges=full["smf_bin_edges"],
                       volume=full["volume"] * vols[gi],
                       target_sumstats=full["target_sumstats"],
                       comm=mg.Comm(None), device="cpu")
        a, b = ref.calc_loss_and_grad_from_params(p)
        rl, rg = rl + a, rg + b
    assert torch.allclose(loss.cpu(), rl, rtol=2e-5), (loss, rl)
    assert torch.allclose(grad.cpu(), rg, rtol=2e-4, atol=1e-6), (grad, rg)
    traj = group.run_adam(p, nsteps=4, learning_rate=0.01)
    assert traj.shape == (5, 2)
    ts = COMM.allgather(traj.cpu().numpy().tobytes())
    assert all(t == ts[0] for t in ts)
    res = group.run_bfgs(p, maxsteps=5)
    assert np.all(np.isfinite(res.x))


@case
def experimental_scan_gd():
    from multigrad_b200.experimental import scan_gd
    data = torch.arange(10, dtype=torch.float32)
    mine = scan_gd.distribute_data(data)
    lens = COMM.allgather(len(mine))
    assert sum(lens) == 10
    tot = scan_gd.reduce_sum(torch.tensor([float(RANK)], device=DEV))
    assert float(tot) == sum(range(SIZE))

    def loss_and_grad(d, params):
        x = d["x"].to(params)
        r = params[0] * x - 2.0 * x
        return (r * r).sum(), (2 * r * x).sum().reshape(1)
    df = scan_gd.simple_grad_descent(
        {"x": mine.to(DEV)}, loss_and_grad, torch.zeros(1, device=DEV),
        learning_rate=1e-3, nsteps=50)
    assert len(df) == 50 and abs(df["params"].iloc[-1][0] - 2.0) < 0.05
    if DEV.type == "cuda":
        # the captured-graph loop (collective inside the graph) and the eager
        # loop walk the same path, without a collective launch from Python
        from multigrad_b200.ops import allreduce as ar
        n0 = ar.launch_count()
        dfe = scan_gd.simple_grad_descent(
            {"x": mine.to(DEV)}, loss_and_grad, torch.zeros(1, device=DEV),
            learning_rate=1e-3, nsteps=50, use_graph=False)
        n_eager = ar.launch_count() - n0
        assert np.allclose(np.stack(df["params"]), np.stack(dfe["params"]),
                           rtol=1e-5, atol=1e-6)
        assert np.allclose(df["loss"], dfe["loss"], rtol=1e-5)
        if SIZE > 1:
            assert n_eager == 50


@case
def gpu_scatter_nd_over_nvlink():
    """K7: scatter of a CUDA tensor from the root through peer memory."""
    if DEV.type != "cuda":
        return
    full = torch.arange(11 * 5 * 3, dtype=torch.float32).reshape(11, 5, 3)
    for axis in (0, 1):
        mine = mg.scatter_nd(full.to(DEV) if RANK == 0 else None, axis=axis)
        want = torch.tensor_split(full, SIZE, dim=axis)[RANK]
        assert mine.is_cuda and torch.equal(mine.cpu(), want), (axis, mine)
    big = torch.arange(3_000_000, dtype=torch.float64)
    mine = mg.scatter_nd(big.to(DEV) if RANK == SIZE - 1 else None,
                         root=SIZE - 1)
    assert torch.equal(mine.cpu(), torch.tensor_split(big, SIZE)[RANK])


@case
def gpu_allreduce_kernels():
    """K5: LL and two-shot all-reduce vs an exact host sum; every rank must
    hold bit-identical results."""
    if DEV.type != "cuda":
        return
    from multigrad_b200.ops import allreduce as ar
    g = torch.Generator().manual_seed(1234)
    for n, dtype, algo in [(1, torch.float32, 1), (3, torch.float32, 0),
                           (30, torch.float64, 1), (1000, torch.float32, 1),
                           (4096, torch.float32, 2), (5003, torch.float32, 0),
                           (100_001, torch.float32, 2),
                           ((1 << 20) + 3, torch.float32, 0),
                           (4097, torch.float32, 3), (70_001, torch.float32, 3),
                           (300_000, torch.float32, 3),
                           (20_000, torch.float32, 0),
                           (6000, torch.float64, 0)]:
        full = torch.randn(SIZE, n, generator=g, dtype=torch.float64)
        mine = full[RANK].to(dtype).to(DEV)
        want = full.to(dtype).double().sum(0)
        for rep in range(3):            # repeated calls reuse the LL slots
            out = ar.all_reduce_sum(mine, COMM, algo=algo)
            torch.cuda.synchronize()
            tol = 1e-6 if dtype == torch.float32 else 1e-13
            err = (out.double().cpu() - want).abs().max() / want.abs().max()
            assert err < tol * SIZE, (n, dtype, algo, rep, float(err))
            blobs = COMM.allgather(out.cpu().numpy().tobytes())
            assert all(b == blobs[0] for b in blobs), (n, dtype, algo)
        assert torch.equal(mine.cpu(), full[RANK].to(dtype))   # input intact
    # back-to-back calls of different algorithms share the workspace
    full = torch.randn(SIZE, 50_000, generator=g, dtype=torch.float64)
    want = full.float().double().sum(0)
    for rep in range(4):
        for algo in (3, 2, 3, 3, 1 if 50_000 <= COMM.workspace().table.ll_units
                     else 3, 2, 3):
            out = ar.all_reduce_sum(full[RANK].float().to(DEV), COMM, algo=algo)
            err = (out.double().cpu() - want).abs().max() / want.abs().max()
            assert err < 1e-6 * SIZE, (algo, rep, float(err))
    # in-switch reduction (NVLS) where the box offers multicast objects
    if COMM.workspace().multicast() is not None:
        for n in (5, 4096, 100_003, (1 << 22) + 1):
            full = torch.randn(SIZE, n, generator=g, dtype=torch.float64)
            want = full.float().double().sum(0)
            for rep in range(3):
                out = ar.all_reduce_sum(full[RANK].float().to(DEV), COMM,
                                        algo=4)
                err = (out.double().cpu() - want).abs().max() / \
                    want.abs().max()
                assert err < 1e-6 * SIZE, ("nvls", n, rep, float(err))
                blobs = COMM.allgather(out.cpu().numpy().tobytes())
                assert all(b == blobs[0] for b in blobs), ("nvls", n)
        out = ar.all_reduce_sum(full[RANK].float().to(DEV), COMM, algo=3 if
                                full.shape[1] * 4 * SIZE <= (16 << 20) else 2)
        assert (out.double().cpu() - want).abs().max() / want.abs().max() \
            < 1e-6 * SIZE
        # symmetric tensors (allocated in the multicast buffer): in place,
        # ragged length, neighbours untouched
        ws = COMM.workspace()
        a = ws.multicast_tensor(70_003)
        b = ws.multicast_tensor(1000)
        b.fill_(7.0)
        fa = torch.randn(SIZE, 70_003, generator=g, dtype=torch.float64)
        for rep in range(3):
            a.copy_(fa[RANK].float())
            ar.all_reduce_sum_(a, COMM)
            want = fa.float().double().sum(0)
            err = (a.double().cpu() - want).abs().max() / want.abs().max()
            assert err < 1e-6 * SIZE, ("nvls inplace", rep, float(err))
            assert bool((b == 7.0).all())
        # a staged (non-resident) reduction of a large ordinary tensor must
        # not touch the symmetric tensors: it stages through its own window
        keep_a = a.clone()
        big = torch.randn(SIZE, (5 << 20) // 4 + 3, generator=g)
        for rep in range(2):
            out = ar.all_reduce_sum(big[RANK].to(DEV), COMM, algo=4)
            want = big.double().sum(0)
            err = (out.double().cpu() - want).abs().max() / want.abs().max()
            assert err < 1e-6 * SIZE, ("nvls staged", rep, float(err))
            assert torch.equal(a, keep_a) and bool((b == 7.0).all())
        # back-to-back broadcasts from alternating roots (ack protocol)
        for rep in range(50):
            v = torch.full((9,), float(rep + RANK), device=DEV)
            ar.bcast_(v, COMM, root=rep % SIZE)
            assert torch.all(v == rep + rep % SIZE)
        print(f"[rank {RANK}] NVLS all-reduce checked", flush=True)
    # integer tensors go through float64
    t = mg.reduce_sum(torch.arange(5, device=DEV) * (RANK + 1))
    assert t.dtype == torch.int64
    assert torch.equal(t.cpu(), torch.arange(5) * sum(range(1, SIZE + 1)))
    b = torch.full((7,), float(RANK), device=DEV)
    ar.bcast_(b, COMM, root=SIZE - 1)
    assert torch.all(b == SIZE - 1)
    COMM.workspace().check_error()


@case
def gpu_fused_step_uses_one_launch():
    """The multi-GPU Adam step is ONE kernel launch per rank (no separate
    collective launch) and needs no NCCL call."""
    if DEV.type != "cuda":
        return
    from multigrad_b200.ops import erfbin, allreduce as ar
    data = _data(40_000)
    model = SMFModel(aux_data=data, deterministic=True)
    model.run_adam([-1.9, 0.3], nsteps=2)       # builds plan + workspace
    e0, a0 = erfbin.launch_count(), ar.launch_count()
    ep0 = COMM.workspace().epoch
    traj, losses = model.run_adam([-1.9, 0.3], nsteps=20, learning_rate=0.01,
                                  return_loss=True)
    torch.cuda.synchronize()
    # ONE persistent launch runs the 20 steps, exchanging once per step
    assert erfbin.launch_count() - e0 == 1 and ar.launch_count() == a0
    if SIZE > 1:
        assert COMM.workspace().epoch - ep0 == 20   # one exchange per step
    assert bool(torch.isfinite(traj).all()) and losses[-1] < losses[0]
    # per-step launches through the stepping API: same bits
    fit = model.adam_fit([-1.9, 0.3], nsteps=20, learning_rate=0.01)
    e1 = erfbin.launch_count()
    for _ in range(20):
        fit.step()
    torch.cuda.synchronize()
    assert erfbin.launch_count() - e1 == 20 and ar.launch_count() == a0
    assert torch.equal(fit.trajectory, traj)
    assert torch.equal(fit.losses, losses)
    tg = model.run_adam([-1.9, 0.3], nsteps=120, learning_rate=0.01,
                        use_graph=True)
    ts = model.run_adam([-1.9, 0.3], nsteps=120, learning_rate=0.01,
                        use_graph=False)
    assert torch.equal(tg, ts)
    # default scheduling (dynamic tail): replicas still agree bit for bit
    md = SMFModel(aux_data=data)
    td = md.run_adam([-1.9, 0.3], nsteps=60, learning_rate=0.01)
    assert torch.allclose(td, ts[:61], rtol=1e-5, atol=2e-6)
    blobs = COMM.allgather(td.cpu().numpy().tobytes())
    assert all(b == blobs[0] for b in blobs)
    # host-driven Adam (state on the host): same trajectory on every rank
    th = model.run_adam_host([-1.9, 0.3], nsteps=25, learning_rate=0.01)
    assert torch.allclose(th, ts[:26].cpu(), rtol=1e-5, atol=2e-6)
    blobs = COMM.allgather(th.numpy().tobytes())
    assert all(b == blobs[0] for b in blobs)


@case
def gpu_group_device_steps():
    """K4+K3: OnePointGroup fits run device-resident (one fused
    reduce-and-update launch per step over the main communicator) and agree
    with the generic host-driven path."""
    if DEV.type != "cuda":
        return
    from multigrad_b200.ops import group as gp
    vols = [1.0, 1.7]
    ngroups = 2 if SIZE > 1 else 1
    sub, n, g = mg.split_subcomms(num_groups=ngroups)
    full = smf.make_smf_data(30_000, comm=mg.Comm(None), device="cpu")
    x_full = full["log_halo_masses"] + 0.05 * g
    x_mine = torch.tensor_split(x_full, sub.size)[sub.rank].to(DEV)
    mine = SMFModel(x=x_mine, bin_edges=full["smf_bin_edges"],
                    volume=full["volume"] * vols[g],
                    target_sumstats=full["target_sumstats"], comm=sub)
    group = mg.OnePointGroup((mine,), main_comm=COMM)
    p = torch.tensor([-1.95, 0.25])
    bounds = [(-3.0, -1.0), (0.05, None)]
    n0 = gp.launch_count()
    traj, losses, state = group.run_adam(
        p, nsteps=12, learning_rate=0.01, param_bounds=bounds,
        return_loss=True, return_state=True)
    torch.cuda.synchronize()
    assert gp.launch_count() - n0 == 12
    want, wl = mg.OnePointModel.run_adam(
        group, p, nsteps=12, learning_rate=0.01, param_bounds=bounds,
        comm=COMM, return_loss=True)
    assert torch.allclose(traj.cpu(), want.cpu(), rtol=2e-5, atol=2e-6), \
        (traj, want)
    assert torch.allclose(losses.cpu(), wl.cpu(), rtol=2e-4, atol=1e-7)
    blobs = COMM.allgather(traj.cpu().numpy().tobytes())
    assert all(b == blobs[0] for b in blobs)
    # resuming from the returned state continues the same trajectory
    ta, sa = group.run_adam(p, nsteps=6, learning_rate=0.01,
                            param_bounds=bounds, return_state=True)
    tb = group.run_adam(ta[-1], nsteps=6, learning_rate=0.01,
                        param_bounds=bounds, init_state=sa)
    assert torch.allclose(tb[-1].cpu(), traj[-1].cpu(), rtol=1e-5, atol=1e-6)
    gd = group.run_simple_grad_descent(p, nsteps=5, learning_rate=1e-3)
    gw = mg.OnePointModel.run_simple_grad_descent(group, p, nsteps=5,
                                                  learning_rate=1e-3)
    assert torch.allclose(gd.params.cpu(), torch.as_tensor(gw.params).cpu(),
                          rtol=2e-5, atol=2e-6)
    assert torch.allclose(gd.loss.cpu(), torch.as_tensor(gw.loss).cpu(),
                          rtol=2e-4)
    COMM.workspace().check_error() if SIZE > 1 else None


@case
def gpu_linear_link_world_invariance():
    """Tensor-core linear-link model: multi-GPU loss / gradient / Adam
    trajectory equal the single-process float32 CPU values."""
    if DEV.type != "cuda":
        return
    from multigrad_b200.models import LinearLinkBinnedModel
    n, F = 24_000, 9
    g = torch.Generator().manual_seed(5)
    x = (10.0 + 1.3 * torch.rand(n, generator=g)).float()
    feat = torch.cat([torch.ones(n, 1), 0.3 * torch.randn(n, F - 1,
                                                          generator=g)], 1)
    theta = torch.cat([torch.tensor([-2.0]),
                       0.05 * torch.randn(F - 1, generator=g),
                       torch.tensor([0.2])])
    edges = np.linspace(9, 10, 11)
    ref = LinearLinkBinnedModel(x=x, features=feat, bin_edges=edges,
                                volume=10.0 * n, comm=mg.Comm(None),
                                device="cpu")
    tgt = ref.calc_sumstats_from_params(theta)
    ref.set_target(tgt)
    sl = slice(*np.linspace(0, n, SIZE + 1).astype(int)[RANK:RANK + 2])
    model = LinearLinkBinnedModel(x=x[sl].to(DEV), features=feat[sl].to(DEV),
                                  bin_edges=edges, volume=10.0 * n,
                                  target_sumstats=tgt)
    p = theta + 0.04
    loss, grad = model.calc_loss_and_grad_from_params(p)
    rl, rg = ref.calc_loss_and_grad_from_params(p)
    assert torch.allclose(loss.cpu(), rl, rtol=2e-4), (loss, rl)
    assert torch.allclose(grad.cpu(), rg, rtol=2e-3,
                          atol=2e-4 * float(rg.abs().max())), (grad, rg)
    traj = model.run_adam(p, nsteps=10, learning_rate=0.01)
    rt = ref.run_adam(p, nsteps=10, learning_rate=0.01)
    assert torch.allclose(traj.cpu(), rt, atol=5e-4), (traj[-1], rt[-1])
    blobs = COMM.allgather(traj.cpu().numpy().tobytes())
    assert all(b == blobs[0] for b in blobs)


@case
def gpu_linear_link_bfgs_and_group():
    """Tensor-core linear-link model on several GPUs: weighted objects,
    replicated L-BFGS-B (same iterate on every rank, close to the float32 CPU
    fit) and membership of a device-resident OnePointGroup."""
    if DEV.type != "cuda":
        return
    from multigrad_b200.models import LinearLinkBinnedModel
    from multigrad_b200.ops import group as gp
    n, F = 30_000, 4
    g = torch.Generator().manual_seed(11)
    x = (10.0 + 1.3 * torch.rand(n, generator=g)).float()
    w = (0.5 + torch.rand(n, generator=g)).float()
    feat = torch.cat([torch.ones(n, 1), 0.3 * torch.randn(n, F - 1,
                                                          generator=g)], 1)
    theta = torch.cat([torch.tensor([-2.0]),
                       0.05 * torch.randn(F - 1, generator=g),
                       torch.tensor([0.2])])
    edges = np.linspace(9, 10, 11)
    ref = LinearLinkBinnedModel(x=x, features=feat, bin_edges=edges,
                                volume=10.0 * n, weights=w,
                                comm=mg.Comm(None), device="cpu")
    tgt = ref.calc_sumstats_from_params(theta)
    ref.set_target(tgt)
    sl = slice(*np.linspace(0, n, SIZE + 1).astype(int)[RANK:RANK + 2])
    model = LinearLinkBinnedModel(x=x[sl].to(DEV), features=feat[sl].to(DEV),
                                  weights=w[sl].to(DEV), bin_edges=edges,
                                  volume=10.0 * n, target_sumstats=tgt)
    p = theta + 0.04
    loss, grad = model.calc_loss_and_grad_from_params(p)
    rl, rg = ref.calc_loss_and_grad_from_params(p)
    assert torch.allclose(loss.cpu(), rl, rtol=2e-4), (loss, rl)
    assert torch.allclose(grad.cpu(), rg, rtol=2e-3,
                          atol=2e-4 * float(rg.abs().max())), (grad, rg)
    bounds = [(None, None)] * F + [(0.02, 2.0)]
    res = model.run_bfgs(p, maxsteps=50, param_bounds=bounds)
    assert res.fun < 1e-3 * float(rl), res
    blobs = COMM.allgather((np.asarray(res.x).tobytes(), int(res.nfev)))
    assert all(b == blobs[0] for b in blobs)
    # group: this model (all ranks) as the only member, device-resident
    lin2 = LinearLinkBinnedModel(
        x=x[sl].to(DEV), features=torch.ones(sl.stop - sl.start, 1,
                                             device=DEV),
        bin_edges=edges, volume=10.0 * n)
    lin2.set_target(lin2.calc_sumstats_from_params([-2.0, 0.2]))
    smf_m = SMFModel(x=x[sl].to(DEV) - 0.02, bin_edges=edges, volume=10.0 * n,
                     target_sumstats=lin2.target_sumstats)
    group = mg.OnePointGroup((lin2, smf_m), main_comm=COMM)
    q = torch.tensor([-1.9, 0.3])
    assert group._device_ok(q.to(DEV))
    n0 = gp.launch_count()
    traj = group.run_adam(q, nsteps=10, learning_rate=0.01)
    torch.cuda.synchronize()
    assert gp.launch_count() - n0 == 10
    want = mg.OnePointModel.run_adam(group, q, nsteps=10, learning_rate=0.01,
                                     comm=COMM)
    assert torch.allclose(traj.cpu(), want.cpu(), rtol=5e-5, atol=5e-6), \
        (traj[-1], want[-1])
    blobs = COMM.allgather(traj.cpu().numpy().tobytes())
    assert all(b == blobs[0] for b in blobs)


def main():
    only = sys.argv[1:] or list(CASES)
    results = {}
    for name in only:
        try:
            CASES[name]()
            msg = "ok"
        except BaseException:  # noqa: BLE001
            msg = traceback.format_exc()[-1500:]
        all_msgs = COMM.allgather(msg)
        bad = [f"rank {r}: {m}" for r, m in enumerate(all_msgs) if m != "ok"]
        results[name] = "ok" if not bad else "\n".join(bad)
    if RANK == 0:
        print("RESULT " + json.dumps(results), flush=True)
    mg.shutdown()


if __name__ == "__main__":
    main()
