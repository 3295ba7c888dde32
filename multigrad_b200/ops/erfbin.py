"""Python face of the fused erf-soft-binned kernel (``csrc/erfbin.cu``).

:class:`ErfBinPlan` owns everything one model needs to launch
``mgb_erfbin_step``: the by-value model description, the scratch buffer for
the cross-CTA reduction, the (optional) peer table of the communicator and
pre-built argument structs, so that a launch costs one ``ctypes`` call.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import numpy as np
import torch

from . import _lib
from ._lib import (MgbErfBinModel, MgbPeerTable, MgbStepArgs, MODE_ADAM,
                   MODE_GD, MODE_LOSS_GRAD, MODE_SUMSTATS, MODE_SUMSTATS_JAC)
from .reference import LOSS_KINDS, SIGMA_LINKS

__all__ = ["ErfBinPlan", "launch_count"]

_LAUNCHES = 0


def launch_count() -> int:
    """Number of ``mgb_erfbin_step`` launches issued by this process."""
    return _LAUNCHES


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


class ErfBinPlan:
    """Launch plan for one soft-binned model on one GPU.

    Parameters
    ----------
    x : CUDA float32 tensor, this rank's shard of object coordinates
    edges : K+1 bin edges;  norm : K normalisations (``1/(V*width)``)
    target, inv_err : K targets / inverse errors for the built-in losses
    nparams, mu_idx, sig_idx, sigma_link, loss, loss_eps : model description
    weights : optional CUDA float32 per-object weights
    peer_table : ``MgbPeerTable`` of the communicator (None = single GPU)
    """

    def __init__(self, x: torch.Tensor, edges, norm, target=None,
                 inv_err=None, nparams: int = 2, mu_idx: int = 0,
                 sig_idx: int = 1, sigma_link: str = "identity",
                 loss: str = "log_mse", loss_eps: float = 0.0,
                 weights: Optional[torch.Tensor] = None,
                 peer_table: Optional[MgbPeerTable] = None):
        if not x.is_cuda:
            raise ValueError("ErfBinPlan needs CUDA tensors")
        self.lib = _lib.load()
        self.device = x.device
        x = x.detach()
        if x.dtype != torch.float32 or not x.is_contiguous() \
                or x.data_ptr() % 16:
            x = x.to(torch.float32).contiguous().clone()
        self.x = x
        if weights is not None:
            weights = weights.detach().to(self.device, torch.float32)
            if not weights.is_contiguous() or weights.data_ptr() % 16:
                weights = weights.contiguous().clone()
            assert weights.numel() == x.numel()
        self.w = weights
        edges = np.asarray(edges, dtype=np.float64).ravel()
        K = edges.size - 1
        if K < 1 or K > _lib.MGB_MAX_BINS:
            raise ValueError(f"need 1..{_lib.MGB_MAX_BINS} bins, got {K}")
        if np.any(np.diff(edges) <= 0):
            raise ValueError("bin edges must be strictly increasing")
        if not (1 <= nparams <= _lib.MGB_MAX_PARAMS):
            raise ValueError(f"need 1..{_lib.MGB_MAX_PARAMS} params")
        self.K, self.P = K, int(nparams)
        md = MgbErfBinModel()
        md.x, md.w, md.n = _ptr(x), _ptr(weights), x.numel()
        md.nbins, md.nparams = K, self.P
        md.mu_idx, md.sig_idx = int(mu_idx), int(sig_idx)
        md.sigma_link = SIGMA_LINKS[sigma_link]
        md.loss_kind = LOSS_KINDS[loss]
        md.loss_eps = float(loss_eps)
        # centring constant (see csrc/erfbin.cu): mean of the shard rounded to
        # a multiple of 1/64 so that x - x0 is exact in fp32
        x0 = float(x.double().mean().item()) if x.numel() else 0.0
        md.x0 = round(x0 * 64.0) / 64.0 if np.isfinite(x0) else 0.0
        norm = np.broadcast_to(np.asarray(norm, dtype=np.float64), (K,))
        tgt = (np.ones(K) if target is None
               else np.asarray(target, dtype=np.float64).ravel())
        ierr = (np.ones(K) if inv_err is None
                else np.broadcast_to(np.asarray(inv_err, np.float64), (K,)))
        for k in range(K + 1):
            md.edges[k] = float(edges[k])
        for k in range(K):
            md.norm[k] = float(norm[k])
            md.target[k] = float(tgt[k])
            md.inv_err[k] = float(ierr[k])
        self.model = md
        self.pt = peer_table
        self._pt_ref = C.byref(peer_table) if peer_table is not None else None
        self.world = peer_table.world if peer_table is not None else 1
        self._scratch = {}
        self._scratch_for(1)

    # True: fully static tile split, bit-reproducible results (2-3% slower);
    # also forced by MULTIGRAD_B200_DETERMINISTIC=1 (read by the library).
    # Kept in the model struct (MgbErfBinModel.step_flags) so that it reaches
    # every native entry point, the host evaluations and L-BFGS-B sessions
    # included, not only launches whose arguments are built here.
    @property
    def deterministic(self) -> bool:
        return bool(self.model.step_flags & _lib.STEP_STATIC_SPLIT)

    @deterministic.setter
    def deterministic(self, value: bool) -> None:
        if value:
            self.model.step_flags |= _lib.STEP_STATIC_SPLIT
        else:
            self.model.step_flags &= ~_lib.STEP_STATIC_SPLIT

    # ------------------------------------------------------------------ utils
    def set_target(self, target, inv_err=None) -> None:
        tgt = np.asarray(target, dtype=np.float64).ravel()
        assert tgt.size == self.K
        for k in range(self.K):
            self.model.target[k] = float(tgt[k])
        if inv_err is not None:
            ie = np.broadcast_to(np.asarray(inv_err, np.float64), (self.K,))
            for k in range(self.K):
                self.model.inv_err[k] = float(ie[k])

    def _scratch_for(self, batch: int) -> torch.Tensor:
        s = self._scratch.get(batch)
        if s is None:
            with torch.cuda.device(self.device):
                nbytes = int(self.lib.mgb_erfbin_scratch_bytes(self.K, batch))
            if nbytes == 0:
                raise _lib.MgbError("mgb_erfbin_scratch_bytes returned 0")
            s = torch.zeros(nbytes, dtype=torch.uint8, device=self.device)
            self._scratch[batch] = s
        return s

    def launch(self, args: MgbStepArgs) -> None:
        global _LAUNCHES
        s = self._scratch_for(args.batch)
        if torch.cuda.current_device() != self.device.index:
            torch.cuda.set_device(self.device)     # one process per GPU
        rc = self.lib.mgb_erfbin_step(
            C.byref(self.model), C.byref(args), self._pt_ref,
            s.data_ptr(), s.numel(),
            C.c_void_p(torch.cuda.current_stream().cuda_stream))
        if rc:
            _lib.check(rc, "mgb_erfbin_step")
        _LAUNCHES += 1

    # ------------------------------------------------------- evaluation modes
    def evaluate(self, params: torch.Tensor, mode: int, reduce: bool = True):
        """Run one launch in mode 0/1/2 for ``params`` of shape ``(P,)`` or
        ``(B, P)``.  Returns a dict of freshly allocated output tensors."""
        p = params.detach().to(self.device, torch.float32)
        single = p.dim() == 1
        p = p.reshape(-1, self.P).contiguous()
        B = p.shape[0]
        f32 = dict(dtype=torch.float32, device=self.device)
        out = {"sumstats": torch.empty(B, self.K, **f32)}
        a = MgbStepArgs()
        a.mode, a.batch = mode, B
        a.reduce = 1 if (reduce and self.world > 1) else 0
        a.params = p.data_ptr()
        a.out_sumstats = out["sumstats"].data_ptr()
        if mode >= MODE_SUMSTATS_JAC:
            out["jac"] = torch.empty(B, self.K, self.P, **f32)
            a.out_jac = out["jac"].data_ptr()
        if mode >= MODE_LOSS_GRAD:
            out["loss"] = torch.empty(B, **f32)
            out["grad"] = torch.empty(B, self.P, **f32)
            a.out_loss = out["loss"].data_ptr()
            a.out_grad = out["grad"].data_ptr()
        self.launch(a)
        out["_keepalive"] = p
        if single:
            for k in ("sumstats", "jac", "loss", "grad"):
                if k in out:
                    out[k] = out[k][0]
        return out

    # ---------------------------------------------------------- optimisation
    def make_optimizer_args(self, mode: int, state: "FitState", lr: float,
                            bounds: Optional[Sequence] = None,
                            jac_at_params: bool = True, b1=0.9, b2=0.999,
                            eps=1e-8, outputs: bool = False,
                            nsteps: int = 1) -> MgbStepArgs:
        """Argument block of an optimiser launch.  ``nsteps > 1`` makes it a
        *persistent* launch: that many steps inside one kernel (see
        ``MgbStepArgs.nsteps`` in ``include/multigrad_b200.h``)."""
        a = MgbStepArgs()
        a.mode, a.batch = mode, 1
        a.nsteps = int(nsteps)
        a.reduce = 1 if self.world > 1 else 0
        a.jac_at_params = 1 if jac_at_params else 0
        a.opt_p, a.opt_u = state.p.data_ptr(), state.u.data_ptr()
        a.opt_m, a.opt_v = state.m.data_ptr(), state.v.data_ptr()
        a.opt_step = state.step.data_ptr()
        a.traj = state.traj.data_ptr()
        a.loss_hist = state.loss_hist.data_ptr()
        a.traj_rows = state.traj.shape[0]
        a.step_base = state.step_base
        a.lr, a.b1, a.b2, a.eps = float(lr), float(b1), float(b2), float(eps)
        if outputs:
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


class FitState:
    """Device-resident optimiser state + trajectory / loss-history buffers
    written by the fused kernel (no per-step host traffic)."""

    def __init__(self, params: torch.Tensor, uparams: torch.Tensor,
                 nsteps: int, device, step0: int = 0, m=None, v=None,
                 rows: Optional[int] = None):
        P = params.numel()
        f32 = dict(dtype=torch.float32, device=device)
        self.p = params.detach().to(**f32).clone().contiguous()
        self.u = uparams.detach().to(**f32).clone().contiguous()
        self.m = (torch.zeros(P, **f32) if m is None
                  else m.detach().to(**f32).clone())
        self.v = (torch.zeros(P, **f32) if v is None
                  else v.detach().to(**f32).clone())
        self.step = torch.full((1,), step0, dtype=torch.int32, device=device)
        self.step_base = int(step0)
        rows = (nsteps + 1) if rows is None else rows
        self.traj = torch.zeros(rows, P, **f32)
        self.loss_hist = torch.zeros(rows, **f32)
        self.loss = torch.zeros(1, **f32)
        self.grad = torch.zeros(P, **f32)
