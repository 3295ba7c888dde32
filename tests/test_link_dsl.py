"""The link tracer (multigrad_b200/link.py): symbolic derivatives against
autograd, the generated CUDA header (cross-compiled, no GPU needed), the
counter-based random numbers, and the CPU path of LinkedBinnedModel."""
import shutil

import numpy as np
import pytest
import torch

import multigrad_b200 as mg
from multigrad_b200 import link as L
from multigrad_b200.models import ErfBinnedModel, LinkedBinnedModel
from multigrad_b200.ops import _lib

SELF = mg.Comm(None)


def link_mass_sigma(x, feat, p, rng):
    return x + p[0], p[1] + p[2] * (x - 12.0)


def link_rich(x, feat, p, rng):
    mu = x + p[0] + p[3] * L.tanh(feat[0]) - L.sqrt(L.abs(feat[1]) + 1.0) \
        * p[4] ** 2
    sigma = L.exp10(p[1] + p[2] * (x - 11.0)) + 0.1 * L.exp(-feat[1] / 3.0) \
        + L.log(1.0 + p[4] * p[4]) / (2.0 + feat[0] * feat[0])
    return mu, sigma


def test_symbolic_derivatives_match_autograd():
    g = torch.Generator().manual_seed(0)
    x = (10 + 2 * torch.rand(50, generator=g)).double()
    feats = [torch.randn(50, generator=g).double(),
             torch.randn(50, generator=g).double()]
    p = torch.tensor([-1.3, -0.6, 0.05, 0.2, 0.7], dtype=torch.float64)
    spec = L.trace(link_rich, 5, 2)
    assert not spec.stochastic
    assert spec.mu_mask == 0b11001 and spec.sg_mask == 0b10110
    for which, nodes in (("mu", spec.dmu), ("sigma", spec.dsg)):
        def f(q):
            mu, sg = link_rich(x, feats, q, None)
            return (mu - x) if which == "mu" else sg
        jac = torch.autograd.functional.jacobian(f, p)       # [n, P]
        for k in range(5):
            got = spec.numeric(nodes[k], x, feats, p)
            assert torch.allclose(got, jac[:, k], rtol=1e-10, atol=1e-12), \
                (which, k)


def test_shift_is_extracted_symbolically():
    spec = L.trace(link_mass_sigma, 3)
    assert "shift = t0;" in spec.header and "f2s(p[0])" in spec.header
    assert spec.mu_mask == 0b001 and spec.sg_mask == 0b110
    # a mean that is not "x + ..." still works (shift = mu - x numerically)
    spec2 = L.trace(lambda x, f, p, r: (p[0] * x, p[1]), 2)
    assert "vsub" in spec2.header and spec2.mu_mask == 0b01


def test_random_numbers_are_counter_based_and_uniform():
    idx = torch.arange(200_000)
    r = L.TorchRng(seed=7, step=3, index=idx)
    u, n = r.uniform(0), r.normal(1)
    assert 0.0 < float(u.min()) and float(u.max()) < 1.0
    assert abs(float(u.mean()) - 0.5) < 5e-3
    assert abs(float(n.mean())) < 1e-2 and abs(float(n.std()) - 1.0) < 1e-2
    # pure function of (seed, step, index, draw)
    assert torch.equal(u, L.TorchRng(7, 3, idx).uniform(0))
    assert not torch.equal(u, L.TorchRng(7, 4, idx).uniform(0))
    assert not torch.equal(u, L.TorchRng(8, 3, idx).uniform(0))
    assert not torch.equal(u, r.uniform(1))
    # a shard sees the same numbers as the whole catalogue
    part = L.TorchRng(7, 3, idx[1000:2000]).uniform(0)
    assert torch.equal(part, u[1000:2000])
    spec = L.trace(lambda x, f, p, rng: (x + p[0] + 0.1 * rng.normal(0),
                                         p[1] + 0.0 * x), 2)
    assert spec.stochastic and "ul_normal(rng, 0)" in spec.header


@pytest.mark.skipif(_lib.nvcc_path() is None, reason="needs nvcc")
def test_generated_kernel_cross_compiles():
    from multigrad_b200.ops.linkbin import build_link_library
    for fn, npar, nf in ((link_mass_sigma, 3, 0), (link_rich, 5, 2)):
        path = build_link_library(L.trace(fn, npar, nf))
        assert path.exists() and path.stat().st_size > 100_000
    if shutil.which("cuobjdump"):
        import subprocess
        sass = subprocess.run(["cuobjdump", "-sass", str(path)], check=True,
                              capture_output=True, text=True).stdout
        assert "MUFU.EX2" in sass and "FFMA2" in sass


def test_linked_model_cpu_equals_builtin_family():
    """A link that restates the built-in one gives the built-in model's
    numbers (CPU path: the same PyTorch function the GPU tests use as
    oracle)."""
    g = torch.Generator().manual_seed(1)
    x = (10.0 + 1.3 * torch.rand(5000, generator=g)).float()
    edges = np.linspace(9, 10, 11)
    ref = ErfBinnedModel(x=x, bin_edges=edges, volume=5.0e4, comm=SELF,
                         device="cpu")
    lm = LinkedBinnedModel(x=x, link=lambda x, f, p, r: (x + p[0], p[1]),
                           nparams=2, bin_edges=edges, volume=5.0e4, comm=SELF,
                           device="cpu")
    p = torch.tensor([-1.8, 0.3], dtype=torch.float64)
    s_ref = ref.calc_sumstats_from_params(p)
    s_lm = lm.calc_sumstats_from_params(p)
    assert torch.allclose(s_lm, s_ref.to(s_lm), rtol=1e-6)
    ref.set_target(ref.calc_sumstats_from_params([-2.0, 0.2]))
    lm.set_target(ref.target_sumstats)
    l1, g1 = ref.calc_loss_and_grad_from_params(p)
    l2, g2 = lm.calc_loss_and_grad_from_params(p)
    assert torch.allclose(l2, l1.to(l2), rtol=1e-6)
    assert torch.allclose(g2, g1.to(g2), rtol=1e-5)
    t1 = ref.run_adam([-1.7, 0.35], nsteps=8)
    t2 = lm.run_adam([-1.7, 0.35], nsteps=8)
    assert torch.allclose(t2, t1.to(t2), atol=1e-5)
