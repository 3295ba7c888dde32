"""CPU path of LinearLinkBinnedModel: same driver surface as on the GPU
(weights, any bin count, host evaluation, L-BFGS-B, host-state Adam), checked
against a float64 restatement of the model."""
import numpy as np
import torch

import multigrad_b200 as mg
from multigrad_b200.models import LinearLinkBinnedModel
from multigrad_b200.ops import reference as ref

SELF = mg.Comm(None)


def make(n=3000, F=4, k=14, weighted=True, seed=0, **kw):
    g = torch.Generator().manual_seed(seed)
    x = (10.0 + 1.3 * torch.rand(n, generator=g)).float()
    feat = torch.cat([torch.ones(n, 1),
                      0.3 * torch.randn(n, F - 1, generator=g)], 1)
    w = (0.3 + torch.rand(n, generator=g)).float() if weighted else None
    edges = np.linspace(9.0, 10.0, k + 1)
    model = LinearLinkBinnedModel(x=x, features=feat, bin_edges=edges,
                                  volume=10.0 * n, weights=w, comm=SELF,
                                  device="cpu", **kw)
    theta = torch.cat([torch.tensor([-2.0]),
                       0.05 * torch.randn(F - 1, generator=g),
                       torch.tensor([0.25])])
    return model, theta


def oracle(model, p):
    p = p.double()
    f = model.features.double()
    mu = model.x.double() + f @ p[:model.nfeat]
    w = None if model.weights is None else model.weights.double()
    return ref.erf_binned_partial(
        mu, torch.as_tensor(model._edges64), torch.as_tensor(model._norm64),
        torch.zeros((), dtype=torch.float64), p[model.nfeat], "identity", w)


def test_sumstats_match_the_float64_restatement():
    for weighted in (False, True):
        model, p = make(weighted=weighted)
        assert len(model._slices) == 2           # 14 bins -> two slices
        s = model.calc_sumstats_from_params(p)
        assert s.shape == (14,)
        assert torch.allclose(s.double(), oracle(model, p), rtol=2e-5,
                              atol=1e-9)
        s2, jac = model.calc_sumstats_and_jac_from_params(p)
        assert jac.shape == (14, 5)
        p64 = p.double().clone().requires_grad_(True)
        jo = torch.autograd.functional.jacobian(
            lambda q: oracle(model, q), p.double())
        assert torch.allclose(jac.double(), jo, rtol=2e-3,
                              atol=2e-5 * float(jo.abs().max()))


def test_unit_weights_are_the_unweighted_model():
    m0, p = make(weighted=False)
    m1 = LinearLinkBinnedModel(
        x=m0.x, features=m0.features, bin_edges=m0._edges64, volume=m0.volume,
        weights=torch.ones_like(m0.x), comm=SELF, device="cpu")
    assert torch.allclose(m0.calc_sumstats_from_params(p),
                          m1.calc_sumstats_from_params(p), rtol=1e-6)


def test_drivers_on_the_cpu_path():
    # (a floor inside the log keeps the first, long line-search step finite)
    model, truth = make(n=4000, F=3, k=8, loss_eps=1e-8)
    model.set_target(model.calc_sumstats_from_params(truth))
    guess = truth + 0.05
    l0, g0 = model.calc_loss_and_grad_from_params(guess)
    lh, gh = model.loss_and_grad_host(guess.numpy())
    assert abs(lh - float(l0)) <= 1e-6 * abs(float(l0))
    assert np.allclose(gh, g0.numpy(), rtol=1e-5, atol=1e-9)
    res = model.run_bfgs(guess, maxsteps=40,
                         param_bounds=[(None, None)] * 3 + [(0.02, 2.0)])
    assert res.fun < 1e-3 * float(l0)
    th, lh_hist = model.run_adam_host(guess, nsteps=8, learning_rate=0.01,
                                      return_loss=True)
    td, ld = model.run_adam(guess, nsteps=8, learning_rate=0.01,
                            return_loss=True)
    assert torch.allclose(th, td, atol=2e-5)
    assert torch.allclose(lh_hist, ld, rtol=5e-3, atol=1e-7)
    assert model.device_loss_grad_launcher(guess, torch.zeros(5)) is None
