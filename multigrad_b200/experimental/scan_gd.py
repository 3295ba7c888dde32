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
                        learning_rate=0.01, nsteps=100, comm=None,
                        use_graph=None):
    """Gradient descent where ``loss_and_grad_func(data_dict, params)``
    returns this rank's *partial* loss and gradient; both are summed over
    ranks on the device each step.  Returns a DataFrame with columns
    ``loss`` and ``params`` like the reference.

    The reference compiles the loop with ``jax.lax.scan`` and issues
    ``mpi4jax.allreduce`` / ``bcast`` from inside it
    (``multigrad/mpi4jax/multigrad.py:38-58``).  The equivalent here, on CUDA:
    the whole step -- the user's PyTorch function, our NVLink all-reduce
    kernel (whose epoch counter lives on the device, so it can be replayed),
    the update and the history rows -- is captured ONCE into a CUDA graph and
    replayed ``nsteps`` times; no collective is launched from Python inside
    the loop and nothing is read back until the end.  ``use_graph=False`` (or
    a CPU tensor) runs the plain eager loop.  (A fused model needs none of
    this: ``model.run_simple_grad_descent`` is one persistent kernel.)"""
    comm = _comm.get_world() if comm is None else comm
    params = torch.as_tensor(guess).detach().clone()
    if use_graph is None:
        use_graph = params.is_cuda
    if use_graph and params.is_cuda and nsteps > 0:
        return _graph_grad_descent(data_dict, loss_and_grad_func, params,
                                   learning_rate, nsteps, comm)
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


def _graph_grad_descent(data_dict, fn, params, lr, nsteps, comm):
    """One captured step (user function + in-graph all-reduce + update +
    history rows addressed by a device-side step index), replayed."""
    from ..ops.allreduce import all_reduce_sum_
    dev = params.device
    shape = params.shape
    work = params.detach().to(torch.float32).contiguous().clone()
    n = work.numel()
    packed = torch.zeros(1 + n, dtype=torch.float32, device=dev)
    traj = torch.zeros(nsteps, n, dtype=torch.float32, device=dev)
    hist = torch.zeros(nsteps, dtype=torch.float32, device=dev)
    idx = torch.zeros(1, dtype=torch.int64, device=dev)
    world = comm is not None and comm.size > 1
    if world:
        comm.workspace()                 # collective set-up outside capture

    def step():
        loss, grad = fn(data_dict, work.view(shape))
        packed[0:1].copy_(torch.as_tensor(loss).reshape(1))
        packed[1:].copy_(torch.as_tensor(grad).reshape(-1))
        if world:
            all_reduce_sum_(packed, comm, algo=1)      # ordered LL exchange
        traj.index_copy_(0, idx, work.view(1, n))
        hist.index_copy_(0, idx, packed[0:1])
        work.sub_(lr * packed[1:])
        idx.add_(1)

    start = work.clone()
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(side):        # warm-up outside capture
        step()
    torch.cuda.current_stream(dev).wait_stream(side)
    torch.cuda.synchronize(dev)
    work.copy_(start)
    idx.zero_()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        step()
    work.copy_(start)
    idx.zero_()
    for _ in range(nsteps):
        graph.replay()
    torch.cuda.synchronize(dev)
    if world:
        comm.workspace().check_error()
    loss_t = hist.cpu().numpy()
    params_t = traj.cpu().numpy().reshape((nsteps,) + tuple(shape))
    return pd.DataFrame(dict(loss=list(loss_t), params=list(params_t)))
