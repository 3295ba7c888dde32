"""Device collectives over the NVLink peer workspace (K5/K6 of SURVEY.md).

``all_reduce_sum`` is what :func:`multigrad_b200.reduce_sum` calls for CUDA
tensors: our own kernel, no NCCL/MPI (reference: host-staged MPI Allreduce,
``multigrad/multigrad.py:149-183``).
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from . import _lib

__all__ = ["all_reduce_sum", "all_reduce_sum_", "bcast_", "launch_count"]

_LAUNCHES = 0
# auto policy: messages from this size on go through the NVLink-switch
# reduction when the communicator has a multicast buffer (measured crossover,
# benchmarks/allreduce_sweep.py --all-algos)
NVLS_MIN_BYTES = int(os.environ.get("MULTIGRAD_B200_NVLS_MIN_BYTES", 2 << 20))
NVLS_MIN_WORLD = 4      # with 2 GPUs the peer-store kernels win at every size
# tensors allocated with PeerWorkspace.multicast_tensor() need no staging
NVLS_INPLACE_MIN_BYTES = int(os.environ.get(
    "MULTIGRAD_B200_NVLS_INPLACE_MIN_BYTES", 32 << 10))


def launch_count() -> int:
    return _LAUNCHES


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def all_reduce_sum_(t: torch.Tensor, comm, algo: int = 0) -> torch.Tensor:
    """In-place sum of a contiguous CUDA float32/float64 tensor over ``comm``.

    ``algo``: 0 auto, 3 one-shot push (mid-size float32),
    1 one-shot LL (ordered, bit-identical on every rank),
    2 two-shot reduce-scatter/all-gather over peer loads and stores,
    4 in-switch reduction through an NVLink multicast object (NVLS).
    """
    global _LAUNCHES
    if comm is None or comm.size == 1 or t.numel() == 0:
        return t
    assert t.is_cuda and t.is_contiguous()
    ws = comm.workspace()
    lib = ws.lib
    if t.dtype == torch.float32:
        dt = 0
    elif t.dtype == torch.float64:
        dt = 1
    else:
        raise TypeError(f"all_reduce_sum: unsupported dtype {t.dtype}")
    n = t.numel()
    with torch.cuda.device(t.device):
        if dt == 1 and algo != 1 and 2 * n > ws.table.ll_units:
            # float64 bulk: chunk through the LL path
            flat = t.view(-1)
            step = ws.table.ll_units // 2
            for lo in range(0, n, step):
                c = flat[lo:lo + step]
                _lib.check(lib.mgb_allreduce_sum(
                    C.byref(ws.table), c.data_ptr(), c.numel(), 1, 1,
                    _stream()), "mgb_allreduce_sum")
                _LAUNCHES += 1
            return t
        # large float32 messages: reduce inside the NVLink switch when the
        # communicator has a multicast buffer (collective lazy set-up)
        # A tensor from PeerWorkspace.multicast_tensor() (same offset on every
        # rank, checked when it was allocated) is reduced in place; every rank
        # has to pass the same kind of tensor to a collective call.
        resident = (dt == 0 and ws._mc and (algo == 4 or (
                    algo == 0 and comm.size >= NVLS_MIN_WORLD)) and
                    ws._mc.uc <= t.data_ptr() and
                    t.data_ptr() + n * 4 <= ws._mc.uc + ws._mc.alloc_bytes and
                    n * 4 >= NVLS_INPLACE_MIN_BYTES)
        if resident or (dt == 0 and (algo == 4 or (
                algo == 0 and n * 4 >= NVLS_MIN_BYTES
                and comm.size >= NVLS_MIN_WORLD))):
            mc = ws.multicast()
            if mc is None and algo == 4:
                raise RuntimeError("all_reduce_sum: NVLS multicast is not "
                                   "available on this communicator")
            if mc is not None:
                flat = t.view(-1)
                if resident:
                    # reduced in place by the kernel, no staging
                    _lib.check(lib.mgb_allreduce_sum_nvls(
                        C.byref(ws.table), mc.uc, mc.mc, mc.bytes,
                        flat.data_ptr(), n, _stream()),
                        "mgb_allreduce_sum_nvls")
                    _LAUNCHES += 1
                    return t
                # staged in chunks through the upper half of the bound buffer
                # (the lower half belongs to the symmetric tensors)
                step = mc.stage_bytes // 4
                for lo in range(0, n, step):
                    c = flat[lo:lo + step]
                    _lib.check(lib.mgb_allreduce_sum_nvls(
                        C.byref(ws.table), mc.uc + mc.stage_off,
                        mc.mc + mc.stage_off, mc.stage_bytes,
                        c.data_ptr(), c.numel(), _stream()),
                        "mgb_allreduce_sum_nvls")
                    _LAUNCHES += 1
                return t
        # the bandwidth kernels own the lower half of each bulk region
        half = (int(ws.table.bulk_bytes) // 2) & ~1023
        if dt == 0 and algo != 1 and n * 4 > half:
            flat = t.view(-1)
            step = half // 4
            for lo in range(0, n, step):
                c = flat[lo:lo + step]
                _lib.check(lib.mgb_allreduce_sum(
                    C.byref(ws.table), c.data_ptr(), c.numel(), 0, algo,
                    _stream()), "mgb_allreduce_sum")
                _LAUNCHES += 1
            return t
        _lib.check(lib.mgb_allreduce_sum(C.byref(ws.table), t.data_ptr(), n,
                                         dt, algo, _stream()),
                   "mgb_allreduce_sum")
        _LAUNCHES += 1
    return t


def all_reduce_sum(t: torch.Tensor, comm, algo: int = 0) -> torch.Tensor:
    """Out-of-place :func:`all_reduce_sum_`; other dtypes are reduced in
    float64 (exact for integers below 2**53) and cast back."""
    if comm is None or comm.size == 1:
        return t.clone()
    if t.dtype in (torch.float32, torch.float64):
        out = t.detach().contiguous().clone()
        if out.data_ptr() % 16:
            out = out.clone(memory_format=torch.contiguous_format)
        return all_reduce_sum_(out, comm, algo)
    work = t.detach().to(torch.float64).contiguous().clone()
    all_reduce_sum_(work, comm, algo)
    return work.to(t.dtype)


def bcast_(t: torch.Tensor, comm, root: int = 0) -> torch.Tensor:
    """Device-side broadcast of a small float32 tensor from ``root``: the
    root pushes into every peer's LL slots; non-root kernels wait on the GPU
    (K6, the BFGS parameter hand-off)."""
    global _LAUNCHES
    if comm is None or comm.size == 1 or t.numel() == 0:
        return t
    assert t.is_cuda and t.is_contiguous() and t.dtype == torch.float32
    ws = comm.workspace()
    with torch.cuda.device(t.device):
        _lib.check(ws.lib.mgb_bcast_f32(C.byref(ws.table), t.data_ptr(),
                                        t.numel(), root, _stream()),
                   "mgb_bcast_f32")
    _LAUNCHES += 1
    return t
