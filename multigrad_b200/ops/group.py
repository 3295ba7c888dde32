"""Python face of the fused group step (``csrc/group.cu``, K4 + K3).

A :class:`GroupStepper` drives an ``OnePointGroup`` fit without any host
read-back: per step every member model enqueues its fused loss+gradient
kernel (reading the parameters from the shared device-resident optimiser
state), then ONE ``mgb_group_step`` launch sums the contributions over the
main communicator's NVLink workspace and applies the replicated update.

Reference counterpart: ``OnePointGroup.calc_loss_and_grad_from_params``
(``multigrad/multigrad.py:560-581``) + the root-driven Adam loop
(``multigrad/adam.py:39-68``).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import torch

from . import _lib
from ._lib import MgbGroupArgs, MgbStepArgs, MODE_LOSS_GRAD

__all__ = ["GroupStepper", "launch_count", "fill_optimizer_args"]

_LAUNCHES = 0


def launch_count() -> int:
    """Number of ``mgb_group_step`` launches issued by this process."""
    return _LAUNCHES


def fill_optimizer_args(a: MgbStepArgs, mode: int, state, lr: float,
                        bounds: Optional[Sequence] = None,
                        jac_at_params: bool = True, b1=0.9, b2=0.999,
                        eps=1e-8) -> MgbStepArgs:
    """Point ``a`` at a :class:`~multigrad_b200.ops.erfbin.FitState`."""
    a.mode, a.batch = mode, 1
    a.jac_at_params = 1 if jac_at_params else 0
    a.opt_p, a.opt_u = state.p.data_ptr(), state.u.data_ptr()
    a.opt_m, a.opt_v = state.m.data_ptr(), state.v.data_ptr()
    a.opt_step = state.step.data_ptr()
    a.traj = state.traj.data_ptr()
    a.loss_hist = state.loss_hist.data_ptr()
    a.traj_rows = state.traj.shape[0]
    a.step_base = state.step_base
    a.lr, a.b1, a.b2, a.eps = float(lr), float(b1), float(b2), float(eps)
    a.out_loss = state.loss.data_ptr()
    a.out_grad = state.grad.data_ptr()
    if bounds is not None:
        for i, bd in enumerate(bounds):
            if bd is None:
                continue
            lo, hi = bd
            if lo is not None and hi is not None:
                a.bound_kind[i] = 1
                a.bound_lo[i], a.bound_hi[i] = lo, hi
            elif lo is not None:
                a.bound_kind[i], a.bound_lo[i] = 2, lo
            elif hi is not None:
                a.bound_kind[i], a.bound_hi[i] = 3, hi
    return a


class GroupStepper:
    """Launch plan for the device-resident steps of one ``OnePointGroup``.

    Parameters
    ----------
    models : the member models this rank evaluates; each must provide
        ``device_loss_grad_launcher(params, out)`` (see
        ``ErfBinnedModel.device_loss_grad_launcher``)
    main_comm : communicator the group sums over (``None`` = this process)
    params : CUDA float32 tensor ``[P]`` the model kernels read (for an
        optimiser run: ``FitState.p``)
    """

    def __init__(self, models, main_comm, params: torch.Tensor):
        self.lib = _lib.load()
        self.device = params.device
        P = params.numel()
        if not 1 <= P <= _lib.MGB_MAX_PARAMS:
            raise ValueError(f"device group path needs 1..{_lib.MGB_MAX_PARAMS}"
                             f" parameters, got {P}")
        if len(models) > _lib.MGB_MAX_GROUP_MODELS:
            raise ValueError("too many models on one rank for the device "
                             "group path")
        self.P = P
        self.params = params
        self.contrib = torch.zeros(len(models), 1 + P, dtype=torch.float32,
                                   device=self.device)
        self.launchers = [
            m.device_loss_grad_launcher(params, self.contrib[i])
            for i, m in enumerate(models)]
        g = MgbGroupArgs()
        g.n_models, g.nparams = len(models), P
        for i, m in enumerate(models):
            g.active[i] = 1 if (m.comm is None or m.comm.rank == 0) else 0
            g.loss[i] = self.contrib[i].data_ptr()
            g.grad[i] = self.contrib[i].data_ptr() + 4
        self.gargs = g
        self.workspace = None
        self._pt_ref = None
        if main_comm is not None and main_comm.size > 1:
            self.workspace = main_comm.workspace()
            self._pt_ref = C.byref(self.workspace.table)

    def step(self, args: MgbStepArgs) -> None:
        """Enqueue one group step: member-model kernels + the fused
        reduce/update launch.  Nothing is read back."""
        global _LAUNCHES
        for launch in self.launchers:
            launch()
        rc = self.lib.mgb_group_step(
            C.byref(self.gargs), C.byref(args), self._pt_ref,
            C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream))
        if rc:
            _lib.check(rc, "mgb_group_step")
        _LAUNCHES += 1

    def evaluate(self, out_loss: torch.Tensor, out_grad: torch.Tensor) -> None:
        """Joint loss+gradient at ``self.params`` into device tensors."""
        a = MgbStepArgs()
        a.mode, a.batch = MODE_LOSS_GRAD, 1
        a.out_loss, a.out_grad = out_loss.data_ptr(), out_grad.data_ptr()
        self.step(a)

    def check_error(self) -> None:
        if self.workspace is not None:
            self.workspace.check_error()
