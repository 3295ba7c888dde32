"""Optimiser drivers, bound transforms, random keys, LHS scan, aux flags --
the surfaces the reference leaves untested."""
import math

import numpy as np
import pytest
import torch

import multigrad_b200 as mg
from multigrad_b200 import adam, util
from multigrad_b200.models import smf, SMFModel
from multigrad_b200.utils import rng

BOUNDS = [None, (0.0, 2.0), (1.0, None), (None, 3.0), (-np.inf, np.inf)]


def test_transform_roundtrip_and_ranges():
    p = torch.tensor([0.3, 0.7, 1.5, 2.5, -4.0])
    b = adam.normalize_bounds(BOUNDS, 5)
    assert b[4] is None and b[0] is None
    u = adam.apply_transforms(p, b)
    back = adam.apply_inverse_transforms(u, b)
    assert torch.allclose(back, p, atol=1e-6)
    # arbitrary unbounded values land inside the bounds
    uu = torch.tensor([5.0, -40.0, -30.0, 60.0, 7.0])
    q = adam.apply_inverse_transforms(uu, b)
    assert 0.0 < q[1] < 2.0 and q[2] > 1.0 and q[3] < 3.0
    # formulas of reference adam.py:202-239
    mid, s = 1.0, 2.0 / math.pi
    assert math.isclose(float(u[1]), s * math.tan((0.7 - mid) / s), rel_tol=1e-5)
    assert math.isclose(float(u[2]), 1.5 - 1.0 + 1.0 / (1.0 - 1.5), rel_tol=1e-6)
    assert math.isclose(float(u[3]), 2.5 - 3.0 + 1.0 / (3.0 - 2.5), rel_tol=1e-6)


def test_transform_jacobian_matches_autograd():
    b = adam.normalize_bounds(BOUNDS, 5)
    u = torch.tensor([0.3, -0.7, 1.5, -2.5, 4.0], dtype=torch.float64)
    J = torch.autograd.functional.jacobian(
        lambda x: adam.apply_inverse_transforms(x, b), u)
    diag = torch.stack([adam.inverse_transform_jacobian_diag(u[i], b[i])
                        for i in range(5)])
    assert torch.allclose(torch.diagonal(J), diag, atol=1e-10)
    assert torch.allclose(J - torch.diag(torch.diagonal(J)),
                          torch.zeros(5, 5, dtype=torch.float64))


def test_adam_update_formula():
    # jax.example_libraries.optimizers.adam
    g = torch.tensor([0.5, -2.0])
    p = torch.tensor([1.0, 1.0])
    m = v = torch.zeros(2)
    lr, b1, b2, eps = 0.01, 0.9, 0.999, 1e-8
    for i in range(3):
        p, m, v = adam.adam_update(i, p, g, m, v, lr)
    m_ref = v_ref = np.zeros(2)
    x = np.ones(2)
    gn = g.numpy().astype(np.float64)
    for i in range(3):
        m_ref = (1 - b1) * gn + b1 * m_ref
        v_ref = (1 - b2) * gn ** 2 + b2 * v_ref
        x = x - lr * (m_ref / (1 - b1 ** (i + 1))) / (
            np.sqrt(v_ref / (1 - b2 ** (i + 1))) + eps)
    assert np.allclose(p.numpy(), x, rtol=1e-5)


def _quadratic(target):
    def fn(x, _data, **kw):
        d = x - target
        return (d * d).sum(), 2 * d
    return fn


def test_run_adam_shapes_and_convergence():
    tgt = torch.tensor([1.0, -2.0])
    traj = adam.run_adam(_quadratic(tgt), torch.zeros(2), None, nsteps=400,
                         learning_rate=0.05)
    assert traj.shape == (401, 2) and torch.all(traj[0] == 0)
    assert torch.allclose(traj[-1], tgt, atol=1e-2)
    traj2, losses, state = adam.run_adam(
        _quadratic(tgt), torch.zeros(2), None, nsteps=10, learning_rate=0.05,
        return_loss=True, return_state=True)
    assert losses.shape == (10,) and state.step == 10
    # resume == uninterrupted
    traj_a, st = adam.run_adam(_quadratic(tgt), torch.zeros(2), None,
                               nsteps=5, learning_rate=0.05,
                               return_state=True)
    traj_b = adam.run_adam(_quadratic(tgt), traj_a[-1], None, nsteps=5,
                           learning_rate=0.05, init_state=st)
    assert torch.allclose(traj_b[-1], traj2[-1], atol=1e-7)


@pytest.mark.parametrize("compat", [True, False])
def test_run_adam_bounded_stays_in_bounds(compat):
    tgt = torch.tensor([5.0, -5.0])     # outside the box: pushes to the walls
    bounds = [(0.0, 2.0), (-1.0, None)]
    traj = adam.run_adam(_quadratic(tgt), torch.tensor([1.0, 0.0]), None,
                         nsteps=200, param_bounds=bounds, learning_rate=0.1,
                         reference_compat=compat)
    assert traj.shape == (201, 2)
    assert torch.all(traj[:, 0] > 0) and torch.all(traj[:, 0] < 2)
    assert torch.all(traj[:, 1] > -1)
    assert traj[-1, 0] > 1.8 and traj[-1, 1] < -0.8


def test_bounded_adam_compat_switch_changes_trajectory():
    tgt = torch.tensor([1.5])
    kw = dict(nsteps=5, param_bounds=[(0.0, 2.0)], learning_rate=0.1)
    a = adam.run_adam(_quadratic(tgt), torch.tensor([0.5]), None,
                      reference_compat=True, **kw)
    b = adam.run_adam(_quadratic(tgt), torch.tensor([0.5]), None,
                      reference_compat=False, **kw)
    assert a.shape == b.shape == (6, 1)
    # first step of Adam is sign-only, later steps differ
    assert not torch.allclose(a[-1], b[-1], atol=1e-7)


def test_randkey_helpers():
    k = rng.init_randkey(42)
    assert rng.is_key(k) and torch.equal(k, rng.init_randkey(42))
    assert not torch.equal(k, rng.init_randkey(43))
    a, b = rng.split(k)
    assert not torch.equal(a, b) and not torch.equal(a, k)
    assert torch.equal(rng.gen_new_key(k), rng.split(k, 1)[0])
    assert torch.equal(rng.init_randkey(k), k)
    with pytest.raises(AssertionError):
        rng.init_randkey(1.5)
    u1 = rng.uniform(a, (4,))
    assert torch.equal(u1, rng.uniform(a, (4,)))
    assert not torch.equal(u1, rng.uniform(b, (4,)))


class _NoisyModel(mg.OnePointModel):
    """sumstats = params + noise(randkey); loss = |s - 1|^2; with aux."""
    def calc_partial_sumstats_from_params(self, params, randkey=None):
        noise = 0.0 if randkey is None else 1e-3 * rng.normal(randkey, (2,))
        self.keys.append(None if randkey is None else randkey.clone())
        return params * 1.0 + noise, {"n": 2}

    def calc_loss_from_sumstats(self, sumstats, sumstats_aux=None,
                                randkey=None):
        assert sumstats_aux == {"n": 2}
        return ((sumstats - 1.0) ** 2).sum(), {"mean": sumstats.mean().item()}


def test_has_aux_flags_and_randkey_plumbing():
    m = _NoisyModel(sumstats_func_has_aux=True, loss_func_has_aux=True,
                    device="cpu")
    m.keys = []
    (loss, aux), grad = m.calc_loss_and_grad_from_params([0.0, 2.0])
    assert abs(float(loss) - 2.0) < 1e-6 and "mean" in aux
    assert torch.allclose(grad, torch.tensor([-2.0, 2.0]))
    s, saux = m.calc_sumstats_from_params([0.0, 2.0])
    assert saux == {"n": 2}
    g, laux = m.calc_dloss_dsumstats(s, saux)
    assert torch.allclose(g, torch.tensor([-2.0, 2.0])) and "mean" in laux
    # per-step keys differ; const_randkey freezes the key
    m.keys = []
    m.run_adam([0.0, 2.0], nsteps=3, randkey=7)
    ks = [k for k in m.keys if k is not None]
    assert len(ks) == 3 and not torch.equal(ks[0], ks[1])
    m.keys = []
    m.run_adam([0.0, 2.0], nsteps=3, randkey=7, const_randkey=True)
    ks = [k for k in m.keys if k is not None]
    assert len(ks) == 3 and torch.equal(ks[0], ks[1])
    with pytest.raises(AssertionError):
        m.run_adam([0.0, 2.0], nsteps=1, const_randkey=True)
    # BFGS holds the key constant
    m.keys = []
    res = m.run_bfgs([0.0, 2.0], randkey=3)
    ks = [k for k in m.keys if k is not None]
    assert len(ks) >= 2 and all(torch.equal(ks[0], k) for k in ks)
    assert np.allclose(res.x, [1.0, 1.0], atol=5e-3)


def test_abstract_hooks_raise():
    m = mg.OnePointModel(device="cpu")
    with pytest.raises(NotImplementedError):
        m.calc_partial_sumstats_from_params([0.0])
    with pytest.raises(NotImplementedError):
        m.calc_loss_from_sumstats(torch.zeros(1))
    assert m == m and m != mg.OnePointModel(device="cpu")
    assert isinstance(hash(m), int)


def test_simple_grad_descent_function_forms():
    f = lambda x: ((x - 3.0) ** 2).sum()          # noqa: E731
    res = util.simple_grad_descent(f, torch.zeros(2), 50, 0.1)
    assert res.loss.shape == (50,) and res.params.shape == (50, 2)
    assert torch.all(res.params[0] == 0)          # pre-update recording (Q6)
    assert res.loss[-1] < res.loss[0]
    res2 = util.simple_grad_descent(
        f, torch.zeros(2), 50, 0.1, grad_loss_func=lambda x: 2 * (x - 3.0))
    assert torch.allclose(res.params, res2.params, atol=1e-6)
    res3 = util.simple_grad_descent(
        lambda x: (f(x), x.detach().sum()), torch.zeros(2), 5, 0.1,
        has_aux=True)
    assert res3.aux.shape == (5,)


def test_lhs_sampler_and_scan():
    pts = util.latin_hypercube_sampler([-3, 0.1], [-1, 0.5], 2, 16, seed=1)
    assert pts.shape == (16, 2)
    assert np.all(pts[:, 0] >= -3) and np.all(pts[:, 0] <= -1)
    assert np.all(pts[:, 1] >= 0.1) and np.all(pts[:, 1] <= 0.5)
    # one draw per stratum in every dimension
    strata = np.floor((pts[:, 0] + 3) / 2 * 16).astype(int)
    assert sorted(strata) == list(range(16))
    assert np.array_equal(
        pts, util.latin_hypercube_sampler([-3, 0.1], [-1, 0.5], 2, 16, seed=1))
    data = smf.make_smf_data(2_000)
    for model in (smf.MySMFModel(aux_data=data),
                  SMFModel(aux_data=data, device="cpu")):
        params, sumstats, losses = model.run_lhs_param_scan(
            [-2.5, 0.15], [-1.5, 0.4], 2, 6, seed=3)
        assert params.shape == (6, 2) and sumstats.shape == (6, 10)
        assert losses.shape == (6,) and np.all(np.isfinite(losses))
        s0 = model.calc_sumstats_from_params(params[0])
        assert np.allclose(sumstats[0], s0.numpy(), rtol=1e-5)


def test_group_single_process_sums_models():
    data = smf.make_smf_data(2_000)
    a = SMFModel(aux_data=data, device="cpu")
    b = SMFModel(aux_data=dict(data, volume=2.0 * data["volume"]),
                 device="cpu")
    group = mg.OnePointGroup((a, b))
    p = torch.tensor([-1.9, 0.3])
    loss, grad = group.calc_loss_and_grad_from_params(p)
    la, ga = a.calc_loss_and_grad_from_params(p)
    lb, gb = b.calc_loss_and_grad_from_params(p)
    assert torch.allclose(loss, la + lb) and torch.allclose(grad, ga + gb)
    traj = group.run_adam(p, nsteps=3)
    assert traj.shape == (4, 2)
    gd = group.run_simple_grad_descent(p, nsteps=2, learning_rate=1e-3)
    assert gd.params.shape == (2, 2)
    res = group.run_bfgs(p, maxsteps=3)
    assert res.x.shape == (2,)
    assert isinstance(hash(group), int) and group == group
    single = mg.OnePointGroup(a)
    assert isinstance(single.models, tuple)


def test_checkpoint_resume_roundtrip(tmp_path):
    from multigrad_b200.utils import checkpoint, timing
    tgt = torch.tensor([1.0, -2.0])
    traj, losses, st = adam.run_adam(_quadratic(tgt), torch.zeros(2), None,
                                     nsteps=7, learning_rate=0.05,
                                     return_loss=True, return_state=True)
    path = str(tmp_path / "ckpt.npz")
    checkpoint.save_adam_checkpoint(path, traj, st, losses)
    t2, st2, l2 = checkpoint.load_adam_checkpoint(path)
    assert torch.equal(t2, traj) and st2.step == 7 and torch.equal(l2, losses)
    resumed = adam.run_adam(_quadratic(tgt), t2[-1], None, nsteps=5,
                            learning_rate=0.05, init_state=st2)
    full = adam.run_adam(_quadratic(tgt), torch.zeros(2), None, nsteps=12,
                         learning_rate=0.05)
    assert torch.allclose(resumed[-1], full[-1], atol=1e-7)
    timer = timing.StepTimer()
    for _ in range(3):
        with timer.step(), timing.nvtx_range("step"):
            _quadratic(tgt)(torch.zeros(2), None)
    assert len(timer.times_ms()) == 3
    assert timing.device_time_ms(lambda: None, iters=2, warmup=1) >= 0.0


def test_run_adam_host_falls_back_to_generic_on_cpu():
    """``run_adam_host`` (host-state Adam of the native family) is the plain
    generic Adam when there is no GPU: same trajectory as ``run_adam``."""
    from multigrad_b200.models import smf, SMFModel
    data = smf.make_smf_data(2_000, comm=mg.Comm(None), device="cpu")
    m = SMFModel(aux_data=data, comm=mg.Comm(None), device="cpu")
    a = m.run_adam_host([-1.8, 0.3], nsteps=4, learning_rate=0.01)
    b = m.run_adam([-1.8, 0.3], nsteps=4, learning_rate=0.01)
    assert a.shape == (5, 2) and torch.allclose(a, b)
