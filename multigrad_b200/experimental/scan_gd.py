"""Device-resident collectives inside the optimisation loop (component C14).

Parity target: the reference's experimental ``multigrad/mpi4jax/multigrad.py``
-- ``distribute_data`` (``:17-23``), ``reduce_sum`` (``:27-29``) and
``simple_grad_descent`` (``:38-61``), the only place where it issues
collectives from inside a compiled program (``jax.lax.scan`` +
``mpi4jax.allreduce/bcast``).

Here that is the normal mode of operation: the gradient is reduced on the
device by our NVLink kernel and every rank applies the identical update, so
the per-step ``bcast`` of the reference (``:52``) disappears.
"""
from __future__ import annotations

import math

import pandas as pd
import torch

from ..parallel import comm as _comm
from ..parallel.collectives import reduce_sum as _reduce_sum

__all__ = ["distribute_data", "reduce_sum", "simple_grad_descent"]


def distribute_data(data, comm=None):
    """Contiguous ``ceil(N/size)`` slice of ``data`` for this rank."""
    comm = _comm.get_world() if comm is None else comm
    chunksize = math.ceil(len(data) / comm.size)
    start = chunksize * comm.rank
    return data[start:start + chunksize]


def reduce_sum(partial_value, comm=None):
    """All-reduce (sum) that keeps tensors on their device."""
    comm = _comm.get_world() if comm is None else comm
    return _reduce_sum(partial_value, comm=comm)


def simple_grad_descent(data_dict, loss_and_grad_func, guess,
                        learning_rate=0.01, nsteps=100, comm=None):
    """Gradient descent where ``loss_and_grad_func(data_dict, params)``
    returns this rank's *partial* loss and gradient; both are summed over
    ranks on the device each step.  Returns a DataFrame with columns
    ``loss`` and ``params`` like the reference."""
    comm = _comm.get_world() if comm is None else comm
    params = torch.as_tensor(guess).detach().clone()
    losses, steps = [], []
    for _ in range(nsteps):
        loss, grad = loss_and_grad_func(data_dict, params)
        packed = torch.cat([torch.as_tensor(loss).reshape(1).to(params),
                            torch.as_tensor(grad).reshape(-1).to(params)])
        packed = _reduce_sum(packed.detach(), comm=comm)
        losses.append(packed[0])
        steps.append(params)
        params = params - learning_rate * packed[1:].reshape(params.shape)
    loss_t = torch.stack(losses).cpu().numpy()
    params_t = torch.stack(steps).cpu().numpy()
    return pd.DataFrame(dict(loss=list(loss_t), params=list(params_t)))
