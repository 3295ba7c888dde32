"""Plain-PyTorch restatements of every CUDA op (component N3 of SURVEY.md):
the correctness oracle for the kernels, the CPU execution path of the native
models, and the timed CPU baseline.

Everything here is ordinary differentiable torch code in whatever dtype the
inputs carry (tests use float64 as the oracle and float32 as "the same op in
plain PyTorch").
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch

SQRT2 = math.sqrt(2.0)
LN10 = math.log(10.0)

__all__ = ["erf_binned_partial", "erf_binned_partial_and_jac",
           "binned_loss", "adam_step_reference", "LOSS_KINDS",
           "SIGMA_LINKS"]

LOSS_KINDS = {"log_mse": 0, "mse": 1, "chi2": 2}
SIGMA_LINKS = {"identity": 0, "log10": 1}


def _sigma(p_sig, sigma_link: str):
    return 10.0 ** p_sig if sigma_link == "log10" else p_sig


def erf_binned_partial(x: torch.Tensor, edges: torch.Tensor,
                       norm: torch.Tensor, p_mu, p_sig,
                       sigma_link: str = "identity",
                       weights: Optional[torch.Tensor] = None,
                       chunk: int = 1 << 20) -> torch.Tensor:
    """``s_k = norm_k * sum_h w_h [Phi((e_{k+1}-mu_h)/(sqrt2 sigma)) -
    Phi((e_k-mu_h)/(sqrt2 sigma))]`` with ``mu_h = x_h + p_mu``.

    The erf-CDF soft histogram of the reference example
    (``tests/smf_example/smf_grad_descent.py:32-76``) with each edge's CDF
    evaluated once.  Differentiable w.r.t. ``p_mu`` and ``p_sig``.
    """
    sigma = _sigma(p_sig, sigma_link)
    inv = 1.0 / (SQRT2 * sigma)
    out = None
    for lo in range(0, max(x.numel(), 1), chunk):
        xs = x[lo:lo + chunk]
        z = (edges[:, None] - (xs + p_mu)[None, :]) * inv
        # 0.5*erfc(-z) keeps relative accuracy in the lower tail
        cdf = 0.5 * torch.special.erfc(-z)
        d = cdf[1:] - cdf[:-1]
        if weights is not None:
            d = d * weights[lo:lo + chunk][None, :]
        part = d.sum(dim=1)
        out = part if out is None else out + part
    return out * norm


def erf_binned_partial_and_jac(x, edges, norm, p_mu, p_sig,
                               sigma_link="identity", weights=None,
                               chunk: int = 1 << 20
                               ) -> Tuple[torch.Tensor, torch.Tensor,
                                          torch.Tensor]:
    """Closed-form single-pass ``(s, ds/dp_mu, ds/dp_sig)`` -- the formulation
    the fused kernel uses: with ``g = exp(-z^2)/sqrt(pi)``,
    ``dPhi/dmu = -g/(sqrt2 sigma)`` and ``dPhi/dsigma = -z g / sigma``."""
    sigma = _sigma(p_sig, sigma_link)
    dsdp = LN10 * sigma if sigma_link == "log10" else 1.0
    inv = 1.0 / (SQRT2 * sigma)
    s = dmu = dsg = None
    for lo in range(0, max(x.numel(), 1), chunk):
        xs = x[lo:lo + chunk]
        z = (edges[:, None] - (xs + p_mu)[None, :]) * inv
        cdf = 0.5 * torch.special.erfc(-z)
        g = torch.exp(-z * z) / math.sqrt(math.pi)
        zg = z * g
        d, dg, dzg = cdf[1:] - cdf[:-1], g[1:] - g[:-1], zg[1:] - zg[:-1]
        if weights is not None:
            w = weights[lo:lo + chunk][None, :]
            d, dg, dzg = d * w, dg * w, dzg * w
        a, b, c = d.sum(1), dg.sum(1), dzg.sum(1)
        s = a if s is None else s + a
        dmu = b if dmu is None else dmu + b
        dsg = c if dsg is None else dsg + c
    return s * norm, -dmu * inv * norm, -dsg / sigma * dsdp * norm


def binned_loss(sumstats: torch.Tensor, target: torch.Tensor,
                kind: str = "log_mse", eps: float = 0.0,
                inv_err: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Built-in losses of the fused epilogue.

    ``log_mse``: ``mean((log10(s+eps) - log10(t+eps))**2)`` -- the loss of both
    reference examples (``smf_grad_descent.py:78-82`` with ``eps=0``,
    ``docs/source/notebooks/smf_gradient_descent.py:86-91`` with
    ``eps=1e-10``); ``mse``: ``mean((s-t)**2)``;
    ``chi2``: ``sum(((s-t)*inv_err)**2)``.
    """
    if kind == "log_mse":
        return torch.mean((torch.log10(sumstats + eps)
                           - torch.log10(target + eps)) ** 2)
    if kind == "mse":
        return torch.mean((sumstats - target) ** 2)
    if kind == "chi2":
        return torch.sum(((sumstats - target) * inv_err) ** 2)
    raise ValueError(f"unknown loss kind {kind!r}")


def adam_step_reference(step, params, grad, m, v, lr, b1=0.9, b2=0.999,
                        eps=1e-8):
    """Same formula as :func:`multigrad_b200.adam.adam_update`."""
    from ..adam import adam_update
    return adam_update(step, params, grad, m, v, lr, b1, b2, eps)
