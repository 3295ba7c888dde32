"""NVLink peer workspace of a communicator (component N1 of SURVEY.md).

Every rank ``cudaMalloc`` s one block, exports it with CUDA IPC, and maps the
blocks of all other ranks of the (sub)communicator, so that kernels can load
and store peer memory directly over NVLink/NVSwitch.  The handles travel once
over the cold-path object channel (gloo); after that no host-side message is
needed by any collective: synchronisation is by epoch-tagged flags inside the
workspace (``csrc/mgb_common.cuh``).

This replaces the communicator plumbing of the reference
(``multigrad/multigrad.py:15-27`` MPI bootstrap, ``:88-146`` ``Split``): a
sub-communicator made by ``split_subcomms`` simply builds its own workspace
over its own GPUs.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List

import torch

from ..ops import _lib

__all__ = ["PeerWorkspace"]

DEFAULT_LL_UNITS = 8192            # 8-byte slots per (parity, source)
DEFAULT_BULK_BYTES = 64 << 20      # each of the two bulk regions
DEFAULT_TIMEOUT_MS = 20000


class PeerWorkspace:
    """Symmetric peer-mapped memory + the ``MgbPeerTable`` kernels consume."""

    def __init__(self, comm, ll_units: int = None, bulk_bytes: int = None,
                 spin_timeout_ms: int = None):
        if not torch.cuda.is_available():
            raise RuntimeError("PeerWorkspace needs CUDA")
        self.comm = comm
        self.lib = _lib.load()
        self.device = torch.device("cuda", torch.cuda.current_device())
        ll_units = int(ll_units or os.environ.get(
            "MULTIGRAD_B200_LL_UNITS", DEFAULT_LL_UNITS))
        bulk_bytes = int(bulk_bytes if bulk_bytes is not None else
                         os.environ.get("MULTIGRAD_B200_BULK_BYTES",
                                        DEFAULT_BULK_BYTES))
        timeout = int(spin_timeout_ms or os.environ.get(
            "MULTIGRAD_B200_SPIN_TIMEOUT_MS", DEFAULT_TIMEOUT_MS))
        world = comm.size
        if world > _lib.MGB_MAX_WORLD:
            raise ValueError(f"peer workspace supports <= "
                             f"{_lib.MGB_MAX_WORLD} ranks, got {world}")
        self.bytes = int(self.lib.mgb_ws_bytes(world, ll_units, bulk_bytes))
        local = C.c_void_p()
        _lib.check(self.lib.mgb_ws_alloc(self.bytes, C.byref(local)),
                   "mgb_ws_alloc")
        self._local = local
        self._imported: List[C.c_void_p] = []
        table = _lib.MgbPeerTable()
        table.world, table.rank = world, comm.rank
        table.ll_units, table.spin_timeout_ms = ll_units, timeout
        table.bulk_bytes = bulk_bytes
        if world > 1:
            handle = C.create_string_buffer(64)
            _lib.check(self.lib.mgb_ws_export(local, handle), "mgb_ws_export")
            handles = comm.allgather(bytes(handle.raw))
            for r, h in enumerate(handles):
                if r == comm.rank:
                    table.base[r] = local.value
                    continue
                p = C.c_void_p()
                _lib.check(self.lib.mgb_ws_import(h, C.byref(p)),
                           f"mgb_ws_import(rank {r})")
                self._imported.append(p)
                table.base[r] = p.value
            comm.Barrier()      # everyone mapped everyone before first use
        else:
            table.base[0] = local.value
        self.table = table
        self._closed = False

    # ------------------------------------------------------------------ state
    def read_ctrl(self, idx: int) -> int:
        out = C.c_uint32()
        _lib.check(self.lib.mgb_ws_read_ctrl(
            self._local, idx, C.byref(out),
            C.c_void_p(torch.cuda.current_stream().cuda_stream)),
            "mgb_ws_read_ctrl")
        return int(out.value)

    @property
    def epoch(self) -> int:
        return self.read_ctrl(0)

    def check_error(self) -> None:
        """Raise if any in-kernel wait timed out (sticky device flag)."""
        if self.read_ctrl(1):
            raise RuntimeError(
                f"rank {self.comm.rank}: a device-side wait on a peer GPU "
                f"timed out after {self.table.spin_timeout_ms} ms "
                "(a rank died, diverged, or never launched the collective)")

    def close(self) -> None:
        if self._closed:
            return
        self._closed = True
        try:
            torch.cuda.synchronize()
            if self.comm.size > 1:
                self.comm.Barrier()
            for p in self._imported:
                self.lib.mgb_ws_unimport(p)
            self.lib.mgb_ws_free(self._local)
        except Exception:
            pass

    def __del__(self):
        # never run collectives from a finaliser
        pass
