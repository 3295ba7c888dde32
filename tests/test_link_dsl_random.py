"""Randomised check of the link tracer: expression trees drawn from every
operator of the DSL, symbolic derivatives against torch autograd (float64),
and the torch evaluation of the traced graph against the Python function."""
import random

import pytest
import torch

from multigrad_b200 import link as L

NP, NF = 4, 2


def random_expr(rng, depth):
    """A function (x, feat, p) -> value built from DSL operators; arguments
    are kept in domains where every operator is smooth."""
    if depth == 0 or rng.random() < 0.2:
        kind = rng.choice(["x", "p", "p", "feat", "const"])
        if kind == "x":
            return lambda x, f, p: (x - 11.0) * 0.5
        if kind == "p":
            i = rng.randrange(NP)
            return lambda x, f, p, i=i: p[i]
        if kind == "feat":
            i = rng.randrange(NF)
            return lambda x, f, p, i=i: f[i]
        c = rng.uniform(-1.5, 1.5)
        return lambda x, f, p, c=c: c + 0.0 * p[0]
    op = rng.choice(["add", "sub", "mul", "div", "pow", "exp", "log", "log10",
                     "exp10", "sqrt", "tanh", "abs", "neg"])
    a = random_expr(rng, depth - 1)
    if op in ("add", "sub", "mul", "div"):
        b = random_expr(rng, depth - 1)
        if op == "add":
            return lambda x, f, p: a(x, f, p) + b(x, f, p)
        if op == "sub":
            return lambda x, f, p: a(x, f, p) - b(x, f, p)
        if op == "mul":
            return lambda x, f, p: a(x, f, p) * b(x, f, p)
        return lambda x, f, p: a(x, f, p) / (2.0 + L.tanh(b(x, f, p)))
    if op == "pow":
        k = rng.choice([2, 3])
        return lambda x, f, p: L.tanh(a(x, f, p)) ** k
    if op == "exp":
        return lambda x, f, p: L.exp(L.tanh(a(x, f, p)))
    if op == "exp10":
        return lambda x, f, p: L.exp10(0.5 * L.tanh(a(x, f, p)))
    if op == "log":
        return lambda x, f, p: L.log(2.0 + L.tanh(a(x, f, p)))
    if op == "log10":
        return lambda x, f, p: L.log10(2.0 + L.tanh(a(x, f, p)))
    if op == "sqrt":
        return lambda x, f, p: L.sqrt(1.5 + L.tanh(a(x, f, p)))
    if op == "tanh":
        return lambda x, f, p: L.tanh(a(x, f, p))
    if op == "abs":
        return lambda x, f, p: L.abs(a(x, f, p) + 0.3)
    return lambda x, f, p: -a(x, f, p)


@pytest.mark.parametrize("seed", range(12))
def test_random_links_differentiate_like_autograd(seed):
    rng = random.Random(seed)
    mu_e, sg_e = random_expr(rng, 4), random_expr(rng, 4)

    def link(x, feat, p, r):
        return x + mu_e(x, feat, p), 0.3 + 0.1 * L.tanh(sg_e(x, feat, p)) \
            + 0.05 * p[1] * p[1]

    spec = L.trace(link, NP, NF)
    g = torch.Generator().manual_seed(seed)
    x = (10 + 2 * torch.rand(40, generator=g)).double()
    feats = [torch.randn(40, generator=g).double() for _ in range(NF)]
    p = (0.8 * torch.randn(NP, generator=g)).double()
    mu, sg = link(x, feats, p, None)
    tmu, tsg = spec.torch_eval(x, feats, p)
    assert torch.allclose(tmu + 0.0 * x, mu + 0.0 * x, rtol=1e-12, atol=1e-12)
    assert torch.allclose(tsg + 0.0 * x, sg + 0.0 * x, rtol=1e-12, atol=1e-12)
    for which, nodes, mask in (("mu", spec.dmu, spec.mu_mask),
                               ("sigma", spec.dsg, spec.sg_mask)):
        def f(q):
            m, s = link(x, feats, q, None)
            return ((m - x) if which == "mu" else s) + 0.0 * x
        jac = torch.autograd.functional.jacobian(f, p)
        for k in range(NP):
            got = torch.as_tensor(spec.numeric(nodes[k], x, feats, p),
                                  dtype=torch.float64) + 0.0 * x
            assert torch.allclose(got, jac[:, k], rtol=1e-9, atol=1e-11), \
                (seed, which, k)
            if not (mask >> k) & 1:            # declared independent of p[k]
                assert float(jac[:, k].abs().max()) == 0.0
    assert "ul_eval" in spec.header and len(spec.digest) >= 16


def test_every_operator_compiles():
    """One link using every operator of the DSL goes through the CUDA code
    generator and nvcc (cross-compiled: no GPU needed)."""
    from multigrad_b200.ops import _lib
    from multigrad_b200.ops.linkbin import build_link_library
    if _lib.nvcc_path() is None:
        pytest.skip("needs nvcc")

    def link(x, feat, p, rng):
        a = L.tanh(p[0] * feat[0]) + L.exp(-L.abs(feat[1])) * p[1]
        b = L.log(2.0 + a * a) / (1.5 + L.sqrt(1.0 + feat[0] * feat[0]))
        c = L.log10(3.0 + L.tanh(b)) - L.exp10(0.1 * p[2]) ** 2
        mu = x + a - b * c + 0.05 * rng.normal(0) + 0.01 * rng.uniform(1)
        sigma = 0.25 + 0.1 * L.tanh(c) + (p[3] * p[3]) / (4.0 + L.abs(x - 11.0))
        return mu, sigma

    spec = L.trace(link, NP, NF)
    assert spec.stochastic
    path = build_link_library(spec)
    assert path.exists() and path.stat().st_size > 100_000
