"""K6: multi-GPU L-BFGS-B without a control plane.

The reference runs scipy's L-BFGS-B on rank 0 and drives the other ranks with
two pickled MPI broadcasts per loss+gradient evaluation (``"compute"`` +
params) plus an ``"exit"`` and a broadcast of the result object
(``multigrad/bfgs.py:72-111``).

Here the fused evaluation leaves *bit-identical* loss and gradient on every
rank (rank-ordered NVLink sum inside the kernel), so -- exactly as for Adam --
every rank can run the same deterministic optimiser on its own copy: scipy's
L-BFGS-B is a serial float64 code, identical inputs give identical iterates on
every rank.  Each rank therefore drives its own GPU through an *evaluation
session* (``mgb_erfbin_host_session``: one grid always pre-launched and
waiting on the GPU for the next parameters in pinned host memory, results
written straight back into pinned host memory), and the only thing that
crosses GPUs is the 3K-double exchange inside the evaluation kernel.  No
parameter hand-off, no exit message, no result broadcast: one evaluation costs
the kernel plus the optimiser's own host time, at every GPU count.

The ranks verify afterwards that they really did the same thing (evaluation
count and final iterate, one cold allgather); a disagreement -- which would
mean a non-deterministic host optimiser -- raises on every rank.
"""
from __future__ import annotations

import numpy as np
import scipy.optimize
import torch

from .. import util


def _np64(v):
    if isinstance(v, torch.Tensor):
        return v.detach().cpu().double().numpy()
    return np.asarray(v, dtype=np.float64)


def run_bfgs_replicated(model, guess, maxsteps, param_bounds, comm,
                        _before_eval=None):
    """L-BFGS-B on every rank of ``comm`` over a device evaluation session of
    ``model`` (see the module docstring).  ``_before_eval`` is a test hook
    called in front of every evaluation (rank-skew injection)."""
    if param_bounds is not None and hasattr(param_bounds, "tolist"):
        param_bounds = param_bounds.tolist()
    x0 = _np64(guess).ravel()
    pbar = util.trange(maxsteps, desc="BFGS Gradient Descent Progress",
                       leave=True)

    def callback(*_a, **_k):
        if hasattr(pbar, "update"):
            pbar.update()

    evaluate, end_session = model._host_session(x0)
    nev = 0

    def fn(p):
        nonlocal nev
        if _before_eval is not None:
            _before_eval(nev)
        nev += 1
        return evaluate(p)

    err = None
    result = None
    try:
        result = scipy.optimize.minimize(
            fn, x0=x0, method="L-BFGS-B", jac=True,
            options=dict(maxiter=maxsteps), callback=callback,
            bounds=param_bounds)
    except BaseException as e:  # noqa: BLE001 - the waiting grid must go
        err = e
    finally:
        try:
            end_session()
        except Exception as e2:  # noqa: BLE001
            if err is None:
                err = e2
        if hasattr(pbar, "close"):
            pbar.close()
    # cold path: did every rank walk the same path?
    mine = (repr(err) if err is not None else None, nev,
            None if result is None else np.asarray(result.x).tobytes())
    everyone = comm.allgather(mine)
    if err is not None:
        raise err
    bad = [r for r, o in enumerate(everyone) if o[0] is not None]
    if bad:
        raise RuntimeError(f"run_bfgs: rank(s) {bad} raised "
                           f"{everyone[bad[0]][0]}")
    if any(o[1:] != everyone[0][1:] for o in everyone):
        raise RuntimeError(
            "run_bfgs: the ranks' optimisers diverged (evaluation counts "
            f"{[o[1] for o in everyone]}): host L-BFGS-B is not deterministic "
            "on this machine")
    model._check_peers()
    return result
