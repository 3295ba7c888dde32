"""L-BFGS-B driver (layer L3).

Parity target: reference ``multigrad/bfgs.py:32-113``.  scipy's L-BFGS-B runs
in float64 on the host of communicator rank 0 with native box bounds (no
change of variables).  Every loss+gradient evaluation is a collective over
``comm``; workers are told what to evaluate by the root.

Control plane: the generic path uses cold-path object broadcasts like the
reference (``"compute"`` + params, ``"exit"``, then the result).  Models that
own a GPU peer workspace replace this by a device-side parameter push (see
``models/erf_binned.py``: ``run_bfgs`` there launches one fused
loss+gradient kernel per line-search evaluation and workers wait *on the
GPU*).

Robustness beyond the reference: if the root raises inside
the optimiser an ``"abort"`` task is broadcast so that workers raise too
instead of blocking forever.
"""
from __future__ import annotations

import numpy as np
import scipy.optimize
import torch

from .util import trange
from .utils.rng import init_randkey

__all__ = ["run_bfgs"]


def _to_host_pair(res):
    loss, grad = res
    if isinstance(loss, tuple):
        loss = loss[0]
    loss = float(torch.as_tensor(loss).detach().cpu().double())
    grad = torch.as_tensor(grad).detach().cpu().double().numpy()
    return loss, np.ascontiguousarray(grad)


def run_bfgs(loss_and_grad_fn, params, maxsteps=100, param_bounds=None,
             randkey=None, comm=None):
    """Minimise with ``scipy.optimize.minimize(method="L-BFGS-B", jac=True)``.

    Parameters
    ----------
    loss_and_grad_fn : callable ``(params, **kw) -> (loss, grad)``; collective
    params : array-like starting point
    maxsteps : int, ``maxiter`` of L-BFGS-B
    param_bounds : sequence of ``None | (low, high)``, native box bounds
    randkey : int | key, held **constant** over the whole run
    comm : Comm or None

    Returns
    -------
    ``scipy.optimize.OptimizeResult`` (identical on every rank) with
    ``message, success, status, fun, x, nit, jac, nfev, njev, hess_inv``.
    """
    kwargs = {}
    if randkey is not None:
        kwargs["randkey"] = init_randkey(randkey)
    if param_bounds is not None and hasattr(param_bounds, "tolist"):
        param_bounds = param_bounds.tolist()
    x0 = np.asarray(torch.as_tensor(params).detach().cpu().double().numpy()
                    if isinstance(params, torch.Tensor)
                    else np.asarray(params, dtype=np.float64))
    multi = comm is not None and comm.size > 1

    if comm is None or comm.rank == 0:
        pbar = trange(maxsteps, desc="BFGS Gradient Descent Progress",
                      leave=True)

        def loss_and_grad_fn_root(p):
            if multi:
                comm.bcast("compute", root=0)
                comm.bcast(np.asarray(p), root=0)
            return _to_host_pair(loss_and_grad_fn(
                torch.as_tensor(np.asarray(p, dtype=np.float32)), **kwargs))

        def callback(*_args, **_kwargs):
            if hasattr(pbar, "update"):
                pbar.update()

        try:
            result = scipy.optimize.minimize(
                loss_and_grad_fn_root, x0=x0, method="L-BFGS-B", jac=True,
                options=dict(maxiter=maxsteps), callback=callback,
                bounds=param_bounds)
        except BaseException:
            if multi:
                comm.bcast("abort", root=0)
            raise
        finally:
            if hasattr(pbar, "close"):
                pbar.close()
        if multi:
            comm.bcast("exit", root=0)
            comm.bcast([*result.keys()], root=0)
            comm.bcast([*result.values()], root=0)
    else:
        while True:
            task = comm.bcast(None, root=0)
            if task == "compute":
                p = comm.bcast(None, root=0)
                loss_and_grad_fn(
                    torch.as_tensor(np.asarray(p, dtype=np.float32)),
                    **kwargs)
            elif task == "exit":
                break
            elif task == "abort":
                raise RuntimeError(
                    "run_bfgs: the root rank raised inside the optimiser")
            else:
                raise ValueError("task %s not recognized!" % task)
        keys = comm.bcast(None, root=0)
        vals = comm.bcast(None, root=0)
        result = scipy.optimize.OptimizeResult(dict(zip(keys, vals)))
    return result
