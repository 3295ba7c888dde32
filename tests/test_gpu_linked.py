"""Fused kernels of user-defined links (csrc/linkbin.cu compiled per link)
against float64 PyTorch oracles of the same link function."""
import numpy as np
import pytest
import torch

import multigrad_b200 as mg
from multigrad_b200 import link as L
from multigrad_b200.models import ErfBinnedModel, LinkedBinnedModel

pytestmark = pytest.mark.gpu
SELF = mg.Comm(None)


def dev():
    return torch.device("cuda", 0)


def link_mass_sigma(x, feat, p, rng):          # the "user" model, 2 lines
    return x + p[0], p[1] + p[2] * (x - 11.0)


def link_rich(x, feat, p, rng):
    mu = x + p[0] + p[3] * L.tanh(feat[0])
    return mu, L.exp10(p[1] + p[2] * (x - 11.0)) + 0.02 * L.abs(feat[1])


def oracle(fn, x, feats, edges, norm, p, weights=None, seed=0, step=0,
           index0=0):
    """float64 (s[K], ds/dp[K, P]) by autograd of the torch link."""
    x64 = x.double().cpu()
    f64 = [f.double().cpu() for f in feats]
    e64 = torch.as_tensor(np.asarray(edges, dtype=np.float64))
    n64 = torch.as_tensor(np.asarray(norm, dtype=np.float64))
    w64 = None if weights is None else weights.double().cpu()
    idx = index0 + torch.arange(x64.numel())

    def s_of(q):
        rng = L.TorchRng(seed, step, idx)
        mu, sg = fn(x64, f64, q, rng)
        sg = sg + torch.zeros_like(x64)
        z = (e64[:, None] - mu[None, :]) / (np.sqrt(2.0) * sg[None, :])
        cdf = 0.5 * (1 + torch.erf(z))
        d = cdf[1:] - cdf[:-1]
        if w64 is not None:
            d = d * w64[None, :]
        return d.sum(1) * n64
    p64 = torch.as_tensor(np.asarray(p, dtype=np.float32)).double()
    return s_of(p64), torch.autograd.functional.jacobian(s_of, p64)


def data(n=60_000, seed=0, weights=False):
    g = torch.Generator().manual_seed(seed)
    x = (10.0 + 1.3 * torch.rand(n, generator=g)).float()
    f0 = torch.randn(n, generator=g).float()
    f1 = torch.randn(n, generator=g).float()
    w = (torch.rand(n, generator=g).float() + 0.5) if weights else None
    return x, [f0, f1], w


def check(s, jac, so, jo):
    assert torch.allclose(s.double().cpu(), so, rtol=5e-6, atol=1e-12)
    for k in range(jo.shape[1]):
        col = jo[:, k]
        assert torch.allclose(jac[:, k].double().cpu(), col, rtol=2e-5,
                              atol=5e-6 * float(col.abs().max())), k


@pytest.mark.parametrize("n", [60_000, 60_003, 1])
def test_mass_dependent_scatter_vs_fp64_autograd(n):
    x, _, _ = data(n)
    edges = np.linspace(9, 10, 11)
    m = LinkedBinnedModel(x=x.to(dev()), link=link_mass_sigma, nparams=3,
                          bin_edges=edges, volume=10.0 * n, comm=SELF,
                          device=dev())
    for p in ([-1.8, 0.3, 0.05], [-1.2, 0.5, -0.1]):
        s, jac = m.calc_sumstats_and_jac_from_params(p)
        so, jo = oracle(link_mass_sigma, x, [], edges, m._norm64, p)
        check(s, jac, so, jo)


def test_features_weights_and_nonlinear_link():
    x, feats, w = data(50_000, seed=3, weights=True)
    edges = np.linspace(8.8, 10.2, 8)
    m = LinkedBinnedModel(x=x.to(dev()), features=[f.to(dev()) for f in feats],
                          weights=w.to(dev()), link=link_rich, nparams=4,
                          bin_edges=edges, volume=1.0e5, comm=SELF,
                          device=dev())
    p = [-1.5, -0.55, 0.08, 0.15]
    s, jac = m.calc_sumstats_and_jac_from_params(p)
    so, jo = oracle(link_rich, x, feats, edges, m._norm64, p, weights=w)
    check(s, jac, so, jo)


def test_restating_the_builtin_link_gives_the_builtin_numbers():
    x, _, _ = data(80_000, seed=5)
    edges = np.linspace(9, 10, 11)
    ref = ErfBinnedModel(x=x.to(dev()), bin_edges=edges, volume=8.0e5,
                         comm=SELF, device=dev())
    lm = LinkedBinnedModel(x=x.to(dev()),
                           link=lambda x, f, p, r: (x + p[0], p[1]),
                           nparams=2, bin_edges=edges, volume=8.0e5,
                           comm=SELF, device=dev())
    p = [-1.8, 0.3]
    s1, j1 = ref.calc_sumstats_and_jac_from_params(p)
    s2, j2 = lm.calc_sumstats_and_jac_from_params(p)
    assert torch.allclose(s2, s1, rtol=5e-6)
    assert torch.allclose(j2, j1, rtol=2e-5, atol=5e-6 * float(j1.abs().max()))


def test_loss_grad_and_optimisers_match_the_generic_path():
    x, _, _ = data(40_000, seed=7)
    edges = np.linspace(9, 10, 11)
    kw = dict(link=link_mass_sigma, nparams=3, bin_edges=edges,
              volume=4.0e5, comm=SELF, loss_eps=1e-10)
    m = LinkedBinnedModel(x=x.to(dev()), device=dev(), **kw)
    c = LinkedBinnedModel(x=x, device="cpu", **kw)
    truth = [-2.0, 0.25, 0.04]
    tgt = m.calc_sumstats_from_params(truth)
    m.set_target(tgt)
    c.set_target(tgt.cpu())
    p = torch.tensor([-1.8, 0.3, 0.0])
    lg, gg = m.calc_loss_and_grad_from_params(p)
    lc, gc = c.calc_loss_and_grad_from_params(p.double())
    assert torch.allclose(lg.cpu().double(), lc, rtol=2e-4)
    assert torch.allclose(gg.cpu().double(), gc, rtol=2e-3, atol=1e-6)
    bounds = [(-3.0, 0.0), (0.05, None), None]
    tg, lossg = m.run_adam(p, nsteps=40, learning_rate=0.01,
                           param_bounds=bounds, return_loss=True)
    tc, lossc = c.run_adam(p, nsteps=40, learning_rate=0.01,
                           param_bounds=bounds, return_loss=True)
    assert tg.shape == (41, 3)
    assert torch.allclose(tg.cpu().double(), tc.double(), atol=5e-4)
    assert torch.allclose(lossg.cpu().double(), lossc.double(), rtol=5e-3,
                          atol=1e-7)
    gd = m.run_simple_grad_descent(p, nsteps=10, learning_rate=1e-3)
    gc2 = c.run_simple_grad_descent(p, nsteps=10, learning_rate=1e-3)
    assert torch.allclose(gd.params.cpu().double(),
                          torch.as_tensor(gc2.params).double(), atol=1e-4)
    res = m.run_bfgs(p, maxsteps=80, param_bounds=[(-3.0, -1.0), (0.1, 0.6),
                                                    (-0.15, 0.15)])
    # (p1, p2) are nearly degenerate over this mass range: judge the fit by
    # its loss, not by the parameters
    assert res.success and res.fun < 1e-4 * float(lg), res
    assert abs(res.x[0] - truth[0]) < 0.05


def test_stochastic_link_uses_the_device_key():
    """Scatter drawn inside the link: the kernel's counter-based draws are
    those of the PyTorch twin, depend on (seed, step) only, and a shard sees
    the draws of its global indices."""
    def noisy(x, feat, p, rng):
        return x + p[0] + p[2] * rng.normal(0), p[1] + 0.0 * x
    n = 30_000
    x, _, _ = data(n, seed=9)
    edges = np.linspace(9, 10, 11)
    m = LinkedBinnedModel(x=x.to(dev()), link=noisy, nparams=3,
                          bin_edges=edges, volume=3.0e5, comm=SELF,
                          device=dev(), seed=11)
    assert m.spec.stochastic
    p = [-1.8, 0.2, 0.15]
    for step in (0, 5):
        s, jac = m.calc_sumstats_and_jac_from_params(p, step=step)
        so, jo = oracle(noisy, x, [], edges, m._norm64, p, seed=11, step=step)
        # fp32 Box-Muller vs float64: the draws agree to ~1e-6, the statistic
        # to a few 1e-5
        assert torch.allclose(s.double().cpu(), so, rtol=2e-4)
        assert torch.allclose(jac.double().cpu(), jo, rtol=5e-3,
                              atol=2e-4 * float(jo.abs().max()))
    s0, _ = m.calc_sumstats_and_jac_from_params(p, step=0)
    s0b, _ = m.calc_sumstats_and_jac_from_params(p, step=0)
    s1, _ = m.calc_sumstats_and_jac_from_params(p, step=1)
    assert torch.equal(s0, s0b) and not torch.equal(s0, s1)
    s_other, _ = m.calc_sumstats_and_jac_from_params(p, randkey=12, step=0)
    assert not torch.equal(s0, s_other)
    # per-step keys inside run_adam; const_randkey freezes them
    m.set_target(m.calc_sumstats_from_params([-2.0, 0.2, 0.1], randkey=11))
    t1 = m.run_adam(p, nsteps=12, randkey=5)
    t2 = m.run_adam(p, nsteps=12, randkey=5)
    t3 = m.run_adam(p, nsteps=12, randkey=5, const_randkey=True)
    assert torch.equal(t1, t2) and not torch.equal(t1, t3)
    with pytest.raises(ValueError):
        m.run_adam(p, nsteps=2, const_randkey=True)


def test_overridden_sumstats_hook_is_not_silently_fused():
    """A subclass of the built-in family that overrides the statistic is
    evaluated by its own function (generic autograd path), and the built-in
    fused run_adam validates its key arguments like the reference."""
    x, _, _ = data(20_000, seed=2)
    edges = np.linspace(9, 10, 11)

    class Doubled(ErfBinnedModel):
        def calc_partial_sumstats_from_params(self, params, randkey=None):
            return 2.0 * ErfBinnedModel.calc_partial_sumstats_from_params(
                self, torch.as_tensor(params).to(self.device).double()
                .requires_grad_(True), randkey)

    base = ErfBinnedModel(x=x.to(dev()), bin_edges=edges, volume=2.0e5,
                          comm=SELF, device=dev())
    dbl = Doubled(x=x.to(dev()), bin_edges=edges, volume=2.0e5, comm=SELF,
                  device=dev())
    assert base.fused and not dbl.fused
    p = [-1.8, 0.3]
    assert torch.allclose(dbl.calc_sumstats_from_params(p).float(),
                          2.0 * base.calc_sumstats_from_params(p), rtol=1e-5)
    base.set_target(base.calc_sumstats_from_params([-2.0, 0.2]))
    with pytest.raises(ValueError):
        base.run_adam(p, nsteps=2, const_randkey=True)
    with pytest.raises(AssertionError):
        base.run_adam(p, nsteps=2, randkey="not a key")
    assert base.run_adam(p, nsteps=2, randkey=3).shape == (3, 2)
