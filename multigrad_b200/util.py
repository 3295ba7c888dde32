"""Utility functions: fixed-step gradient descent, result tuple, Latin
hypercube sampling and array scattering.

Parity targets: reference ``multigrad/util.py`` (``GradDescentResult``
``:50-53``, ``latin_hypercube_sampler`` ``:56-62``, ``scatter_nd`` ``:65-77``,
``simple_grad_descent`` ``:80-134``).
"""
from __future__ import annotations

import os
from typing import Any, NamedTuple, Union

import numpy as np
import torch
from scipy.stats import qmc

from .parallel.collectives import scatter_nd  # noqa: F401  (re-export)

__all__ = ["simple_grad_descent", "GradDescentResult",
           "latin_hypercube_sampler", "scatter_nd", "trange"]


def _rank() -> int:
    try:
        return int(os.environ.get("RANK", "0"))
    except ValueError:
        return 0


try:  # progress bar on rank 0 only (reference ``util.py:27-47``)
    if _rank() or os.environ.get("MULTIGRAD_B200_NO_TQDM"):
        raise ImportError("Only show progress bar on RANK=0 task")
    from tqdm import auto as _tqdm
except ImportError:  # pragma: no cover
    _tqdm = None


def trange(n, desc=None, **kw):
    """``tqdm.trange`` on rank 0 when tqdm is importable and stderr is a
    terminal, else a plain ``range``."""
    if _tqdm is None or not os.isatty(2):
        return range(n)
    return _tqdm.trange(n, desc=desc, **kw)


class GradDescentResult(NamedTuple):
    loss: torch.Tensor
    params: torch.Tensor
    aux: Union[torch.Tensor, list]


def latin_hypercube_sampler(xmin, xmax, n_dim, num_evaluations,
                            seed=None, optimization=None):
    """Latin-hypercube draws scaled to ``[xmin, xmax]`` (host, float64)."""
    xmin = np.zeros(n_dim) + xmin
    xmax = np.zeros(n_dim) + xmax
    sampler = qmc.LatinHypercube(n_dim, seed=seed, optimization=optimization)
    unit_hypercube = sampler.random(num_evaluations)
    return qmc.scale(unit_hypercube, xmin, xmax)


def _as_tensor(x, like=None) -> torch.Tensor:
    if isinstance(x, torch.Tensor):
        return x
    dev = like.device if isinstance(like, torch.Tensor) else None
    return torch.as_tensor(np.asarray(x, dtype=np.float32), device=dev)


def simple_grad_descent(
    loss_func,
    guess,
    nsteps,
    learning_rate,
    loss_and_grad_func=None,
    grad_loss_func=None,
    has_aux=False,
    **kwargs,
):
    """Fixed-learning-rate gradient descent.

    At every step the loss and gradient are evaluated at the current
    parameters, ``(loss, params_before_update, aux)`` is recorded and then
    ``params -= learning_rate * grad``.  The parameters after the final update
    are *not* recorded (reference ``util.py:100-114``; quirk Q6, kept).

    There is no broadcast: every rank applies the identical update to its own
    replica.  ``loss_and_grad_func`` defaults to autograd of ``loss_func``.
    """
    if loss_and_grad_func is None:
        if grad_loss_func is None:
            vg = torch.func.grad_and_value(loss_func, has_aux=has_aux,
                                           **kwargs)

            def loss_and_grad_func(params):
                g, v = vg(params)
                return v, g
        else:
            def loss_and_grad_func(params):
                return (loss_func(params), grad_loss_func(params))

    params = _as_tensor(guess)
    if not params.is_floating_point():
        params = params.to(torch.float32)
    loss_hist, param_hist, aux_hist = [], [], []
    for _ in trange(nsteps, desc="Simple Gradient Descent Progress"):
        (loss, grad), aux = loss_and_grad_func(params), None
        if has_aux:
            (loss, aux), grad = loss, grad
        loss = _as_tensor(loss, like=params)
        grad = _as_tensor(grad, like=params)
        loss_hist.append(loss.detach().reshape(()))
        param_hist.append(params.detach().to(loss.device))
        aux_hist.append(aux)
        params = params.detach().to(grad.device) - learning_rate * grad.detach()
    loss_t = torch.stack(loss_hist) if loss_hist else torch.zeros(0)
    params_t = (torch.stack(param_hist) if param_hist
                else torch.zeros((0,) + tuple(params.shape)))
    aux_out: Any = aux_hist
    if has_aux:
        try:
            aux_out = torch.stack([_as_tensor(a) for a in aux_hist])
        except (TypeError, RuntimeError, ValueError):
            pass
    return GradDescentResult(loss=loss_t, params=params_t, aux=aux_out)
