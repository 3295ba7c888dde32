"""Communication primitives (layer L1 of SURVEY.md).

Public surface mirrors the reference: ``reduce_sum`` (reference
``multigrad/multigrad.py:149-183``), ``split_subcomms`` (``:88-146``),
``split_subcomms_by_node`` (``:48-85``) and ``scatter_nd``
(``multigrad/util.py:65-77``).

Behavioural differences, all deliberate:

* CUDA tensors are reduced **on the device** by our own NVLink kernel
  (``ops/allreduce.py``); the reference forces a device->host copy and a host
  MPI call (Q8).
* ``root=r``: non-root ranks get zeros instead of uninitialised memory (Q7).
"""
from __future__ import annotations

import math
from typing import Optional, Sequence

import numpy as np
import torch
import torch.distributed as dist

from . import comm as _comm
from .comm import Comm

__all__ = ["reduce_sum", "split_subcomms", "split_subcomms_by_node",
           "scatter_nd", "group_assignment"]

_DEFAULT = object()


def _resolve(comm):
    if comm is _DEFAULT:
        return _comm.get_world()
    return comm


def reduce_sum(value, root: Optional[int] = None, comm=_DEFAULT):
    """Sum ``value`` over every rank of ``comm``.

    Parameters
    ----------
    value : scalar | numpy array | torch tensor (CPU or CUDA)
    root : int, optional
        If given only this rank receives the sum (others receive zeros).
    comm : Comm, optional
        Defaults to the world communicator; ``None`` means "no communicator"
        and returns ``value`` unchanged, as in the reference.

    Returns
    -------
    Same kind as the input: Python scalar for scalars, ndarray for ndarrays,
    tensor (same device) for tensors.
    """
    comm = _resolve(comm)
    if comm is None:
        return value
    return_to_scalar = not hasattr(value, "__len__")
    if isinstance(value, torch.Tensor):
        return _reduce_tensor(value, root, comm)
    arr = np.asarray(value)
    if comm.size == 1:
        total = arr.copy()
    elif root is None:
        total = comm.Allreduce(arr)
    else:
        total = comm.Reduce(arr, root=root)
    if return_to_scalar:
        total = total.tolist()
    return total


def _reduce_tensor(value: torch.Tensor, root, comm: Comm) -> torch.Tensor:
    if comm.size == 1:
        return value.clone()
    if value.is_cuda:
        from ..ops import allreduce as _ar
        total = _ar.all_reduce_sum(value, comm)
        if root is not None and comm.rank != root:
            total = torch.zeros_like(total)
        return total
    total = value.detach().clone().contiguous()
    if root is None:
        dist.all_reduce(total, group=comm.group)
    else:
        dist.reduce(total, dst=comm.global_rank(root), group=comm.group)
        if comm.rank != root:
            total = torch.zeros_like(total)
    return total


def group_assignment(size: int, num_groups: Optional[int] = None,
                     ranks_per_group: Optional[Sequence[int]] = None):
    """Rank -> group index table used by :func:`split_subcomms`.

    Reproduces the reference mapping (``multigrad.py:112-128``; the tables in
    SURVEY.md C4): with ``num_groups=G`` rank ``r`` takes the first label of
    the ``r``-th near-even chunk of ``repeat(arange(G), ceil(size/G))``.
    """
    main_msg = "Specify either num_groups OR ranks_per_group"
    if num_groups is not None:
        assert ranks_per_group is None, main_msg
        assert size >= num_groups, \
            "Cannot create more subcomms than there are ranks"
        num_groups = int(num_groups)
        # NB: *not* truncated to ``size`` labels -- the reference slices the
        # row axis, which is a no-op, and the surplus labels shape the chunks.
        labels = np.repeat(np.arange(num_groups),
                           math.ceil(size / num_groups))
    else:
        assert ranks_per_group is not None, main_msg
        assert sum(ranks_per_group) == size, \
            "The sum of ranks_per_group must equal comm.size"
        num_groups = len(ranks_per_group)
        labels = np.repeat(np.arange(num_groups), ranks_per_group)
    table = [int(chunk[0]) for chunk in np.array_split(labels, size)]
    return table, num_groups


def split_subcomms(num_groups: Optional[int] = None,
                   ranks_per_group: Optional[Sequence[int]] = None,
                   comm=_DEFAULT):
    """Split ``comm`` into sub-communicators (not grouped by node).

    Returns ``(sub_comm, num_groups, group_index)`` like the reference
    (``multigrad.py:88-146``).  Each sub-communicator later grows its own
    NVLink peer workspace, so a group's reductions only involve its GPUs.
    """
    comm = _resolve(comm)
    assert comm is not None, "Cannot split subcomms without a communicator"
    table, num_groups = group_assignment(comm.size, num_groups,
                                         ranks_per_group)
    mine = table[comm.rank]
    sub = comm.Split(color=mine, key=comm.rank)
    sub.Set_name(f"{comm.name}.{mine}".replace("WORLD.", ""))
    return sub, num_groups, mine


def split_subcomms_by_node(comm=_DEFAULT):
    """Split ``comm`` by host name (reference ``multigrad.py:48-85``).

    On the single 8xB200 NVSwitch box this returns one group containing every
    rank; the multi-node case works over gloo/NCCL but has no peer workspace.
    """
    comm = _resolve(comm)
    assert comm is not None, "Cannot split subcomms without a communicator"
    node = _comm.processor_name()
    nodes = sorted(set(comm.allgather(node)))
    idx = nodes.index(node)
    sub = comm.Split(color=idx, key=comm.rank)
    sub.Set_name(f"{comm.name}.{idx}".replace("WORLD.", ""))
    return sub, len(nodes), idx


def scatter_nd(array, axis: int = 0, comm=_DEFAULT, root: int = 0):
    """Scatter an n-d array from ``root``: rank ``i`` gets the ``i``-th of
    ``size`` near-even chunks along ``axis`` (reference ``util.py:65-77``).

    Accepts numpy arrays or torch tensors on the root.  Chunks travel over the
    cold-path (gloo) point-to-point channel; each rank then uploads its shard
    to its own GPU with one pinned H2D copy (see ``utils.to_device_shard``).
    """
    comm = _resolve(comm)
    if comm is None or comm.size == 1:
        return array
    if comm.rank == root:
        if isinstance(array, torch.Tensor):
            splits = list(torch.tensor_split(array.cpu(), comm.size, dim=axis))
        else:
            splits = np.array_split(np.asarray(array), comm.size, axis=axis)
        ans = splits[root]
        for i in range(comm.size):
            if i != root:
                comm.send(splits[i], dest=i)
        return ans
    return comm.recv(source=root)
