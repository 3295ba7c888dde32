"""GPU numerics: every CUDA kernel against a plain PyTorch reference of the
same op (float64 oracle on the same float32 inputs, plus the generic
autograd path).  All tests need one B200."""
import math

import numpy as np
import pytest
import torch

import multigrad_b200 as mg
from multigrad_b200 import adam
from multigrad_b200.models import smf, SMFModel, ErfBinnedModel
from multigrad_b200.ops import reference as ref

pytestmark = pytest.mark.gpu

SELF = mg.Comm(None)


def erfbin_launches():
    from multigrad_b200.ops import erfbin
    return erfbin.launch_count()


def dev():
    return torch.device("cuda", 0)


def oracle(x, edges, norm, p, link="identity", weights=None):
    """float64 (s, ds/dmu, ds/dsig) on the float32 inputs."""
    x64 = x.double().cpu()
    e64 = torch.as_tensor(np.asarray(edges, dtype=np.float64))
    n64 = torch.as_tensor(np.asarray(norm, dtype=np.float32)).double()
    w64 = None if weights is None else weights.double().cpu()
    p64 = torch.as_tensor(np.asarray(p, dtype=np.float32)).double()
    return ref.erf_binned_partial_and_jac(x64, e64, n64, p64[0], p64[1], link,
                                          w64)


def make(n=20_000, k=10, link="identity", weights=False, seed=0, lo=9.0,
         hi=10.0, **kw):
    g = torch.Generator().manual_seed(seed)
    x = (10.0 + 1.3 * torch.rand(n, generator=g)).float()
    edges = np.linspace(lo, hi, k + 1)
    w = torch.rand(n, generator=g).float() + 0.5 if weights else None
    m = ErfBinnedModel(x=x.to(dev()), bin_edges=edges,
                       volume=10.0 * max(n, 1),
                       weights=None if w is None else w.to(dev()),
                       sigma_link=link, comm=SELF, device=dev(), **kw)
    return m, x, w, edges


@pytest.mark.parametrize("k", [1, 2, 7, 10, 16, 17, 23, 32])
def test_sumstats_and_jacobian_vs_fp64(k):
    m, x, w, edges = make(k=k)
    p = [-1.9, 0.3]
    s, jac = m.calc_sumstats_and_jac_from_params(p)
    so, dmu, dsg = oracle(x, edges, m._norm64, p)
    assert s.shape == (k,) and jac.shape == (k, 2)
    assert torch.allclose(s.double().cpu(), so, rtol=3e-6, atol=1e-12)
    scale = max(dmu.abs().max(), dsg.abs().max())
    assert torch.allclose(jac[:, 0].double().cpu(), dmu, rtol=1e-5,
                          atol=3e-6 * float(scale))
    assert torch.allclose(jac[:, 1].double().cpu(), dsg, rtol=1e-5,
                          atol=3e-6 * float(scale))


@pytest.mark.parametrize("n", [0, 1, 3, 4, 5, 1023, 4097, 100_003])
def test_ragged_sizes(n):
    m, x, w, edges = make(n=max(n, 0), k=5)
    p = [-2.0, 0.2]
    s = m.calc_sumstats_from_params(p)
    so, _, _ = oracle(x, edges, m._norm64, p)
    assert torch.allclose(s.double().cpu(), so, rtol=3e-6, atol=1e-12)


def test_weights_and_log10_link():
    m, x, w, edges = make(k=10, link="log10", weights=True)
    p = [-2.1, -0.6]
    s, jac = m.calc_sumstats_and_jac_from_params(p)
    so, dmu, dsg = oracle(x, edges, m._norm64, p, "log10", w)
    assert torch.allclose(s.double().cpu(), so, rtol=3e-6)
    assert torch.allclose(jac[:, 0].double().cpu(), dmu, rtol=2e-5,
                          atol=3e-6 * float(dmu.abs().max()))
    assert torch.allclose(jac[:, 1].double().cpu(), dsg, rtol=2e-5,
                          atol=3e-6 * float(dsg.abs().max()))


def test_tail_accuracy_far_from_bins():
    """Objects many sigma away from every bin: the kernel keeps relative
    accuracy where erf-based differences lose it completely."""
    n = 4096
    x = torch.full((n,), 10.0)
    edges = np.linspace(11.0, 12.0, 6)          # bins 5..10 sigma above mu
    m = ErfBinnedModel(x=x.to(dev()), bin_edges=edges, volume=1.0,
                       comm=SELF, device=dev())
    p = [0.0, 0.2]
    s = m.calc_sumstats_from_params(p).double().cpu()
    so, _, _ = oracle(x, edges, m._norm64, p)
    assert so[0] > 0 and so[0] < 1e-2
    good = so > 1e-30
    assert torch.allclose(s[good], so[good], rtol=2e-5)


@pytest.mark.parametrize("weights", [False, True])
@pytest.mark.parametrize("k", [10, 20])
def test_saturated_tiles_sorted_catalogue(k, weights):
    """Narrow sigma + mass-sorted catalogue: whole warp tiles are saturated
    for some edge groups but not for others (the warp-uniform skip path must
    keep the hard counts consistent across groups)."""
    n = 60_000
    g = torch.Generator().manual_seed(3)
    x = torch.sort(10.9 + 1.3 * torch.rand(n, generator=g))[0].float()
    w = (torch.rand(n, generator=g).float() + 0.5) if weights else None
    edges = np.linspace(9.0, 10.0, k + 1)
    m = ErfBinnedModel(x=x.to(dev()), bin_edges=edges, volume=1.0 * n,
                       weights=None if w is None else w.to(dev()),
                       comm=SELF, device=dev())
    p = [-2.0, 0.02]
    s, jac = m.calc_sumstats_and_jac_from_params(p)
    so, dmu, dsg = oracle(x, edges, m._norm64, p, weights=w)
    assert torch.allclose(s.double().cpu(), so, rtol=5e-6,
                          atol=1e-7 * float(so.max()))
    assert torch.allclose(jac[:, 0].double().cpu(), dmu, rtol=2e-5,
                          atol=5e-6 * float(dmu.abs().max()))
    assert torch.allclose(jac[:, 1].double().cpu(), dsg, rtol=2e-5,
                          atol=5e-6 * float(dsg.abs().max()))


def test_parameter_layout_and_batch():
    m, x, w, edges = make(k=6, nparams=5, mu_index=3, sigma_index=1)
    P = torch.tensor([[0.0, 0.3, 7.0, -1.9, 5.0],
                      [1.0, 0.25, 0.0, -2.0, 0.0],
                      [2.0, 0.5, 0.0, -1.5, 0.0]])
    S = m.calc_sumstats_from_params(P)
    assert S.shape == (3, 6)
    for i in range(3):
        so, dmu, dsg = oracle(x, edges, m._norm64, [P[i, 3], P[i, 1]])
        assert torch.allclose(S[i].double().cpu(), so, rtol=3e-6)
    s, jac = m.calc_sumstats_and_jac_from_params(P[0])
    assert jac.shape == (6, 5)
    assert torch.all(jac[:, [0, 2, 4]] == 0)
    _, dmu0, dsg0 = oracle(x, edges, m._norm64, [P[0, 3], P[0, 1]])
    assert torch.allclose(jac[:, 3].double().cpu(), dmu0, rtol=1e-5,
                          atol=3e-6 * float(dmu0.abs().max()))
    assert torch.allclose(jac[:, 1].double().cpu(), dsg0, rtol=1e-5,
                          atol=3e-6 * float(dsg0.abs().max()))


@pytest.mark.parametrize("loss,eps", [("log_mse", 0.0), ("log_mse", 1e-10),
                                      ("mse", 0.0), ("chi2", 0.0)])
def test_loss_and_grad_vs_autograd(loss, eps):
    m, x, w, edges = make(k=10, loss=loss, loss_eps=eps,
                          inv_err=np.linspace(50, 500, 10))
    truth = [-2.0, 0.2]
    m.set_target(m.calc_sumstats_from_params(truth), m.inv_err)
    p = torch.tensor([-1.93, 0.27])
    lval, g = m.calc_loss_and_grad_from_params(p)
    # float64 autograd of the same op
    cpu = ErfBinnedModel(x=x.double(), bin_edges=edges, volume=m.volume,
                         comm=SELF, device="cpu", loss=loss, loss_eps=eps,
                         inv_err=np.linspace(50, 500, 10),
                         target_sumstats=m._target_t.cpu())
    p64 = p.double().requires_grad_(True)
    s64 = ref.erf_binned_partial(x.double(), cpu._edges_t.double(),
                                 cpu._norm_t.double(), p64[0], p64[1])
    l64 = ref.binned_loss(s64, m._target_t.cpu().double(), loss, eps,
                          cpu._inv_err_t.double())
    (g64,) = torch.autograd.grad(l64, p64)
    assert math.isclose(float(lval), float(l64), rel_tol=2e-4, abs_tol=1e-12)
    assert torch.allclose(g.double().cpu(), g64, rtol=5e-4,
                          atol=1e-5 * float(g64.abs().max()))
    assert torch.allclose(m.calc_dloss_dparams(p), g)
    assert torch.allclose(m.calc_loss_from_params(p), lval, rtol=1e-5)


def test_reference_test_constants_on_gpu():
    # reference tests/test_mpi.py:38-66 through the fused kernels
    data = smf.make_smf_data(10_000, comm=SELF, device=dev())
    model = SMFModel(aux_data=data, comm=SELF, device=dev())
    truth = smf.ParamTuple()
    s = model.calc_sumstats_from_params(truth)
    assert torch.allclose(data["target_sumstats"], s)
    gd = model.run_simple_grad_descent(guess=truth, nsteps=2)
    assert gd.loss.shape == (2,) and gd.params.shape == (2, 2)
    assert torch.isclose(gd.loss[-1].cpu(), torch.tensor(0.0))
    assert torch.allclose(gd.params[-1].cpu(), torch.tensor([*truth]))
    assert torch.allclose(gd.params[0].cpu(), torch.tensor([*truth]))
    assert torch.allclose(model.calc_dloss_dparams(truth).cpu(),
                          torch.zeros(2), atol=1e-4)


def _cpu_twin(model, x):
    return ErfBinnedModel(
        x=x, bin_edges=model._edges64, volume=model.volume,
        target_sumstats=model._target_t.cpu(), sigma_link=model.sigma_link,
        loss=model.loss, loss_eps=model.loss_eps, comm=SELF, device="cpu")


@pytest.mark.parametrize("bounds", [None, [(-5.0, 0.0), (0.01, None)],
                                    [(None, -1.0), None]])
@pytest.mark.parametrize("compat", [True, False])
def test_fused_adam_trajectory_matches_generic(bounds, compat):
    m, x, w, edges = make(n=30_000, k=10)
    m.set_target(m.calc_sumstats_from_params([-2.0, 0.2]))
    twin = _cpu_twin(m, x)
    guess = [-1.7, 0.35]
    kw = dict(nsteps=25, learning_rate=0.02, param_bounds=bounds,
              reference_compat=compat)
    traj, losses, state = m.run_adam(guess, return_loss=True,
                                     return_state=True, **kw)
    rt, rl, rstate = twin.run_adam(guess, return_loss=True,
                                   return_state=True, **kw)
    assert traj.shape == (26, 2) and losses.shape == (25,)
    assert torch.allclose(traj.cpu(), rt, atol=3e-4), (traj[-1], rt[-1])
    assert torch.allclose(losses.cpu(), rl, rtol=2e-3, atol=1e-7)
    assert state.step == 25
    assert torch.allclose(state.m.cpu(), rstate.m, rtol=1e-2, atol=1e-5)
    # resume continues the same trajectory
    t2 = m.run_adam(traj[-1], nsteps=5, learning_rate=0.02,
                    param_bounds=bounds, reference_compat=compat,
                    init_state=state)
    full = m.run_adam(guess, nsteps=30, learning_rate=0.02,
                      param_bounds=bounds, reference_compat=compat)
    assert torch.allclose(t2[-1], full[-1], atol=2e-5)


def test_fused_adam_cuda_graph_equals_stream_launches():
    m, x, w, edges = make(n=20_000, k=10, deterministic=True)
    m.set_target(m.calc_sumstats_from_params([-2.0, 0.2]))
    a = m.run_adam([-1.7, 0.35], nsteps=130, learning_rate=0.01,
                   use_graph=False)
    b = m.run_adam([-1.7, 0.35], nsteps=130, learning_rate=0.01,
                   use_graph=True)
    assert torch.equal(a, b)            # deterministic reduction order


@pytest.mark.parametrize("k", [10, 20])
def test_persistent_fit_equals_per_step_launches(k, monkeypatch):
    """One persistent launch of n steps (CTAs stay resident, parameters
    handed on through device memory) reproduces n single-step launches bit
    for bit in deterministic mode -- Adam with bounds and plain GD."""
    m, x, w, edges = make(n=300_000, k=k, deterministic=True)
    m.set_target(m.calc_sumstats_from_params([-2.0, 0.2]))
    bounds = [(-3.0, 0.0), (0.05, None)]
    out = {}
    for mode in ("0", "1"):
        monkeypatch.setenv("MULTIGRAD_B200_PERSISTENT", mode)
        n0 = erfbin_launches()
        traj, loss = m.run_adam([-1.7, 0.35], nsteps=60, learning_rate=0.01,
                                param_bounds=bounds, return_loss=True)
        gd = m.run_simple_grad_descent([-1.9, 0.25], nsteps=25,
                                       learning_rate=1e-3)
        torch.cuda.synchronize()
        out[mode] = (traj.clone(), loss.clone(), gd.params.clone(),
                     gd.loss.clone(), erfbin_launches() - n0)
    assert out["0"][4] == 85 and out["1"][4] == 2
    for a, b in zip(out["0"][:4], out["1"][:4]):
        assert torch.equal(a, b)
    assert bool(torch.isfinite(out["1"][0]).all())


def test_dynamic_tail_matches_static_split():
    """Default scheduling (last tenth of the tiles claimed through an atomic
    counter) against the bit-reproducible static split: same numbers up to
    fp32 summation order, and the static split repeats itself exactly."""
    md, x, w, edges = make(n=400_000, k=10)
    ms, _, _, _ = make(n=400_000, k=10, deterministic=True)
    p = [-1.8, 0.3]
    sd, jd = md.calc_sumstats_and_jac_from_params(p)
    ss, js = ms.calc_sumstats_and_jac_from_params(p)
    ss2, js2 = ms.calc_sumstats_and_jac_from_params(p)
    assert torch.equal(ss, ss2) and torch.equal(js, js2)
    assert torch.allclose(sd, ss, rtol=2e-6)
    assert torch.allclose(jd, js, rtol=1e-5, atol=2e-6 * float(js.abs().max()))
    so, dmu, dsg = oracle(x, edges, md._norm64, p)
    assert torch.allclose(sd.double().cpu(), so, rtol=3e-6, atol=1e-12)


def test_non_uniform_edges_take_the_per_edge_path():
    """Edges that are not uniformly spaced cannot use the Gaussian
    recurrence; same accuracy through the per-edge path."""
    n = 50_000
    g = torch.Generator().manual_seed(5)
    x = (10.0 + 1.3 * torch.rand(n, generator=g)).float()
    edges = np.array([9.0, 9.05, 9.2, 9.3, 9.55, 9.6, 9.8, 10.0, 10.3])
    m = ErfBinnedModel(x=x.to(dev()), bin_edges=edges, volume=10.0 * n,
                       comm=SELF, device=dev())
    for p in ([-1.8, 0.3], [-1.0, 0.12]):
        s, jac = m.calc_sumstats_and_jac_from_params(p)
        so, dmu, dsg = oracle(x, edges, m._norm64, p)
        assert torch.allclose(s.double().cpu(), so, rtol=3e-6, atol=1e-12)
        assert torch.allclose(jac[:, 0].double().cpu(), dmu, rtol=1e-5,
                              atol=3e-6 * float(dmu.abs().max()))
        assert torch.allclose(jac[:, 1].double().cpu(), dsg, rtol=1e-5,
                              atol=3e-6 * float(dsg.abs().max()))


@pytest.mark.parametrize("sigma", [0.03, 0.08, 0.2, 0.7, 2.5])
def test_recurrence_and_per_edge_paths_agree(sigma, monkeypatch):
    """Uniform bins over a range of widths (recurrence on, off by its own
    accuracy criterion, and forced off): all within the float64 tolerances."""
    m, x, w, edges = make(n=100_000, k=10)
    p = [-1.6, sigma]
    so, dmu, dsg = oracle(x, edges, m._norm64, p)
    for chain in ("1", "0"):
        monkeypatch.setenv("MULTIGRAD_B200_CHAIN", chain)
        s, jac = m.calc_sumstats_and_jac_from_params(p)
        assert torch.allclose(s.double().cpu(), so, rtol=3e-6, atol=1e-12)
        assert torch.allclose(jac[:, 0].double().cpu(), dmu, rtol=1e-5,
                              atol=3e-6 * float(dmu.abs().max()))
        assert torch.allclose(jac[:, 1].double().cpu(), dsg, rtol=1e-5,
                              atol=3e-6 * float(dsg.abs().max()))


def test_fused_gd_matches_generic():
    m, x, w, edges = make(n=10_000, k=10)
    m.set_target(m.calc_sumstats_from_params([-2.0, 0.2]))
    twin = _cpu_twin(m, x)
    r = m.run_simple_grad_descent([-1.9, 0.25], nsteps=10, learning_rate=0.01)
    rr = twin.run_simple_grad_descent([-1.9, 0.25], nsteps=10,
                                      learning_rate=0.01)
    assert torch.allclose(r.params.cpu(), rr.params, atol=1e-4)
    assert torch.allclose(r.loss.cpu(), rr.loss, rtol=2e-3, atol=1e-8)


@pytest.mark.parametrize("bounds", [None, [(-5.0, 0.0), (0.01, None)]])
def test_group_device_steps_match_generic(bounds):
    """K4+K3 on one GPU: two models in one OnePointGroup, device-resident
    Adam / GD against the generic host-driven loop on the same group."""
    from multigrad_b200.ops import group as gp
    m1, *_ = make(n=20_000, k=10, seed=1)
    m2, *_ = make(n=15_000, k=8, seed=2, lo=9.2, hi=10.2)
    m1.set_target(m1.calc_sumstats_from_params([-2.0, 0.2]))
    m2.set_target(m2.calc_sumstats_from_params([-2.0, 0.2]))
    group = mg.OnePointGroup((m1, m2), main_comm=SELF)
    guess = [-1.8, 0.3]
    n0 = gp.launch_count()
    traj, losses = group.run_adam(guess, nsteps=15, learning_rate=0.02,
                                  param_bounds=bounds, return_loss=True)
    assert gp.launch_count() - n0 == 15
    want, wl = mg.OnePointModel.run_adam(
        group, guess, nsteps=15, learning_rate=0.02, param_bounds=bounds,
        comm=SELF, return_loss=True)
    assert traj.shape == (16, 2)
    assert torch.allclose(traj.cpu(), want.cpu(), rtol=2e-5, atol=2e-6)
    assert torch.allclose(losses.cpu(), wl.cpu(), rtol=2e-4, atol=1e-8)
    gd = group.run_simple_grad_descent(guess, nsteps=6, learning_rate=1e-3)
    gw = mg.OnePointModel.run_simple_grad_descent(group, guess, nsteps=6,
                                                  learning_rate=1e-3)
    assert torch.allclose(gd.params.cpu(), torch.as_tensor(gw.params).cpu(),
                          rtol=2e-5, atol=2e-6)
    assert torch.allclose(gd.loss.cpu(), torch.as_tensor(gw.loss).cpu(),
                          rtol=2e-4)


def test_bfgs_tutorial_on_gpu():
    # docs/source/notebooks/intro.ipynb:258-281 through the fused kernel
    x = torch.log10(smf.load_halo_masses_tutorial(10_000, comm=SELF))
    true = torch.tensor([-2.0, -0.5])
    m = SMFModel(x=x.to(dev()), bin_edges=np.linspace(9, 10, 11), volume=1.0,
                 sigma_link="log10", loss_eps=1e-10, comm=SELF, device=dev(),
                 target_sumstats=torch.ones(10))
    m.set_target(m.calc_sumstats_from_params(true))
    loss, grad = m.calc_loss_and_grad_from_params(true + 0.1)
    assert abs(float(loss) - 0.44032094) < 3e-5
    assert np.allclose(grad.cpu().numpy(), [2.6187496, 4.2603974], rtol=5e-5)
    res = m.run_bfgs(true + torch.tensor([-1.5, 0.7]))
    assert res.success and res.fun < 1e-9
    assert np.allclose(res.x, [-2.0, -0.5], atol=2e-4)
    assert 10 <= res.nit <= 25
    l, g = m.loss_and_grad_host([-1.9, -0.4])
    assert abs(l - 0.44032094) < 3e-5 and g.shape == (2,)


def test_host_eval_zero_copy_equals_copy_path(monkeypatch):
    """loss_and_grad_host: the zero-copy flavour (kernel reads/writes pinned
    host memory, host polls) returns the same bits as the memcpy flavour."""
    pts = [[-1.9, 0.3], [-2.2, 0.15], [-1.2, 0.6]]
    out = {}
    for zc in ("1", "0"):
        monkeypatch.setenv("MULTIGRAD_B200_ZEROCOPY", zc)
        # (static tile split: the default hands the last tiles out through
        # an atomic counter, which changes the fp32 summation order)
        m, *_ = make(n=50_000, k=10, seed=3, deterministic=True)
        m.set_target(m.calc_sumstats_from_params([-2.0, 0.2]))
        out[zc] = [m.loss_and_grad_host(p) for p in pts for _ in range(3)]
        assert m._host[8] == (zc == "1")
    for (l1, g1), (l0, g0) in zip(out["1"], out["0"]):
        assert l1 == l0 and np.array_equal(g1, g0)
    ref_l, ref_g = m.calc_loss_and_grad_from_params(pts[-1])
    assert abs(out["1"][-1][0] - float(ref_l)) < 1e-6 * max(1, abs(float(ref_l)))
    assert np.allclose(out["1"][-1][1], ref_g.cpu().numpy(), rtol=1e-6)


def test_run_adam_host_matches_device_adam():
    """Host-driven Adam (state on the host, native loop) follows the same
    trajectory as the device-resident fit, and resumes."""
    m, *_ = make(n=40_000, k=10, seed=4)
    m.set_target(m.calc_sumstats_from_params([-2.0, 0.2]))
    guess = [-1.7, 0.35]
    th, lh, sh = m.run_adam_host(guess, nsteps=30, learning_rate=0.02,
                                 return_loss=True, return_state=True)
    td, ld = m.run_adam(guess, nsteps=30, learning_rate=0.02,
                        return_loss=True)
    assert th.shape == (31, 2) and not th.is_cuda
    assert torch.allclose(th, td.cpu(), rtol=1e-5, atol=2e-6)
    # (the two Adam implementations round differently: ~1e-6 in the
    # parameters, amplified near the minimum of the loss)
    assert torch.allclose(lh, ld.cpu(), rtol=5e-3, atol=1e-6)
    assert sh.step == 30
    t2 = m.run_adam_host(th[-1], nsteps=5, learning_rate=0.02, init_state=sh)
    full = m.run_adam_host(guess, nsteps=35, learning_rate=0.02)
    assert torch.allclose(t2[-1], full[-1], rtol=1e-6, atol=1e-7)


def test_custom_loss_uses_sumstats_plus_jacobian_path():
    class Custom(ErfBinnedModel):
        def calc_loss_from_sumstats(self, sumstats, sumstats_aux=None,
                                    randkey=None):
            return torch.sum(torch.sqrt(sumstats + 1e-6))

    g = torch.Generator().manual_seed(1)
    x = (10.0 + 1.3 * torch.rand(8000, generator=g)).float()
    m = Custom(x=x.to(dev()), bin_edges=np.linspace(9, 10, 8), volume=80.0,
               comm=SELF, device=dev())
    twin = Custom(x=x, bin_edges=np.linspace(9, 10, 8), volume=80.0,
                  comm=SELF, device="cpu")
    p = [-1.95, 0.22]
    loss, grad = m.calc_loss_and_grad_from_params(p)
    rl, rg = twin.calc_loss_and_grad_from_params(p)
    assert torch.allclose(loss.cpu(), rl, rtol=1e-5)
    assert torch.allclose(grad.cpu(), rg, rtol=2e-4, atol=1e-7)
    traj = m.run_adam(p, nsteps=3)
    assert traj.shape == (4, 2) and traj.is_cuda


def test_lhs_scan_batched_matches_loop():
    m, x, w, edges = make(n=5_000, k=10)
    m.set_target(m.calc_sumstats_from_params([-2.0, 0.2]))
    params, sumstats, losses = m.run_lhs_param_scan(
        [-2.5, 0.15], [-1.5, 0.4], 2, 20, seed=5)
    assert sumstats.shape == (20, 10) and losses.shape == (20,)
    for i in (0, 7, 19):
        s = m.calc_sumstats_from_params(params[i])
        assert np.allclose(sumstats[i], s.cpu().numpy(), rtol=1e-6)
        assert np.isclose(losses[i], float(m.calc_loss_from_params(params[i])),
                          rtol=1e-5)


def test_generic_model_on_gpu_uses_device_tensors():
    data = smf.make_smf_data(5_000, comm=SELF, device=dev())
    model = smf.MySMFModel(aux_data=data, comm=SELF, device=dev())
    loss, grad = model.calc_loss_and_grad_from_params([-1.9, 0.3])
    assert loss.is_cuda and grad.is_cuda
    native = SMFModel(aux_data=data, comm=SELF, device=dev())
    nl, ng = native.calc_loss_and_grad_from_params([-1.9, 0.3])
    assert torch.allclose(loss, nl, rtol=2e-4)
    assert torch.allclose(grad, ng, rtol=2e-3, atol=1e-6)


def test_adam_kernel_matches_formula():
    import ctypes as C
    from multigrad_b200.ops import _lib
    lib = _lib.load()
    n = 37
    g = torch.Generator().manual_seed(2)
    p = torch.randn(n, generator=g)
    grad = torch.randn(n, generator=g)
    m0, v0 = torch.rand(n, generator=g), torch.rand(n, generator=g)
    dp, dg, dm, dv = (t.clone().to(dev()) for t in (p, grad, m0, v0))
    rc = lib.mgb_adam_update(dp.data_ptr(), dg.data_ptr(), dm.data_ptr(),
                             dv.data_ptr(), 4, n, 0.01, 0.9, 0.999, 1e-8,
                             C.c_void_p(torch.cuda.current_stream().cuda_stream))
    _lib.check(rc)
    rp, rm, rv = adam.adam_update(4, p, grad, m0, v0, 0.01)
    assert torch.allclose(dp.cpu(), rp, rtol=1e-5, atol=1e-7)
    assert torch.allclose(dm.cpu(), rm, rtol=1e-6)
    assert torch.allclose(dv.cpu(), rv, rtol=1e-5)


def test_pointer_jump_kernel():
    from multigrad_b200.experimental import diffdesi
    rng = np.random.default_rng(0)
    n = 5000
    parent = np.arange(n)
    for i in range(1, n):                 # random forest, depth up to ~log n
        if rng.random() < 0.8:
            parent[i] = rng.integers(0, i)
    want = diffdesi.find_ultimate_top_indices(parent)
    got = diffdesi.find_ultimate_top_indices(
        torch.as_tensor(parent, device=dev()))
    assert got.is_cuda and np.array_equal(got.cpu().numpy(), want)
    chain = np.maximum(np.arange(200) - 1, 0)          # depth 199 > 2**50? no
    assert np.array_equal(
        diffdesi.find_ultimate_top_indices(
            torch.as_tensor(chain, device=dev())).cpu().numpy(),
        np.zeros(200, dtype=np.int64))
    s, r = diffdesi.sort_all_by_ultimate_top_dump(
        torch.as_tensor(parent, device=dev()),
        [torch.arange(n, device=dev())],
        [torch.as_tensor(parent, device=dev())])
    s2, r2 = diffdesi.sort_all_by_ultimate_top_dump(parent, [np.arange(n)],
                                                    [parent])
    assert np.array_equal(r[0].cpu().numpy(), r2[0])


def test_library_is_loaded_and_errors_are_loud():
    import ctypes as C
    from multigrad_b200.ops import _lib
    lib = _lib.load()
    assert any("libmgb200" in line for line in open("/proc/self/maps"))
    pt = _lib.MgbPeerTable()
    pt.world = 1
    x = torch.zeros(8, device=dev())
    rc = lib.mgb_allreduce_sum(C.byref(pt), x.data_ptr(), 8, 7, 0, None)
    assert rc != 0 and b"dtype" in lib.mgb_last_error()
    with pytest.raises(ValueError):
        ErfBinnedModel(x=x, bin_edges=np.linspace(0, 1, 40), comm=SELF,
                       device=dev()).plan()
