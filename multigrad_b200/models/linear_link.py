"""Native model family with a per-object *linear predictor*:

    mu_h = x_h + sum_j features[h, j] * theta_j,      sigma = link(p_sigma)

with ``params = (theta_0 .. theta_{F-1}, p_sigma)``.  The features are basis
functions of object properties (a constant column reproduces the plain shift
of :class:`SMFModel`; more columns give a mass- or environment-dependent
stellar-to-halo-mass relation).  ``d sumstats / d theta`` is a tall-skinny
GEMM over the object axis, which the CUDA path runs on the tensor cores
(``csrc/linlink.cu``: TMA-streamed bf16 feature tiles, tcgen05 UMMA with TMEM
accumulators), fused with the erfc evaluation, the cross-GPU reduction and
the optimiser.  Features are stored in bf16; that quantised matrix *is* the
model (forward and Jacobian use the same values), and theta is split into
three bf16 planes so the forward product is accurate to fp32.

The reference has no such model class; it is what a user of the reference
would write as a JAX sumstats function with a ``features @ theta`` term and
differentiate with ``jax.vjp`` (``multigrad/multigrad.py:508-538``).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any

import numpy as np
import torch

from .. import adam as _adam
from ..core import OnePointModel
from ..ops import reference as _ref
from ..util import GradDescentResult

__all__ = ["LinearLinkBinnedModel"]


def _np(v):
    if isinstance(v, torch.Tensor):
        return v.detach().cpu().double().numpy()
    return np.asarray(v, dtype=np.float64)


@dataclass(eq=False)
class LinearLinkBinnedModel(OnePointModel):
    x: Any = None
    features: Any = None
    bin_edges: Any = None
    volume: float = 1.0
    target_sumstats: Any = None
    sigma_link: str = "identity"
    loss: str = "log_mse"
    loss_eps: float = 0.0
    inv_err: Any = None

    def __post_init__(self):
        super().__post_init__()
        if self.x is None or self.features is None or self.bin_edges is None:
            raise ValueError("need x, features and bin_edges")
        self.x = torch.as_tensor(self.x).detach().to(
            self.device, torch.float32).contiguous()
        # the bf16-rounded feature matrix is the model
        f = torch.as_tensor(self.features).detach().to(self.device)
        self.features = f.to(torch.bfloat16).to(torch.float32)
        self.nfeat = int(self.features.shape[1])
        self.nparams = self.nfeat + 1
        self._edges64 = _np(self.bin_edges).ravel()
        self.nbins = self._edges64.size - 1
        self._norm64 = 1.0 / (float(self.volume) * np.diff(self._edges64))
        self._edges_t = torch.as_tensor(self._edges64, dtype=torch.float32,
                                        device=self.device)
        self._norm_t = torch.as_tensor(self._norm64, dtype=torch.float32,
                                       device=self.device)
        self._target_t = None
        self._inv_err_t = None
        self._plan = None
        self._builtin_loss = (type(self).calc_loss_from_sumstats
                              is LinearLinkBinnedModel.calc_loss_from_sumstats)
        if self.target_sumstats is not None:
            self.set_target(self.target_sumstats, self.inv_err)

    @property
    def fused(self) -> bool:
        return self.device.type == "cuda"

    def set_target(self, target, inv_err=None):
        self.target_sumstats = target
        self._target_t = torch.as_tensor(_np(target), dtype=torch.float32,
                                         device=self.device)
        if inv_err is not None:
            self.inv_err = inv_err
            self._inv_err_t = torch.as_tensor(
                np.broadcast_to(_np(inv_err), (self.nbins,)).copy(),
                dtype=torch.float32, device=self.device)
        if self._plan is not None:
            self._plan.set_target(_np(target), None if inv_err is None
                                  else _np(inv_err))

    def plan(self):
        if self._plan is None:
            from ..ops.linlink import LinLinkPlan
            table = None
            if self.comm is not None and self.comm.size > 1:
                table = self.comm.workspace().table
            self._plan = LinLinkPlan(
                self.x, self.features, self._edges64, self._norm64,
                target=None if self.target_sumstats is None
                else _np(self.target_sumstats),
                inv_err=None if self.inv_err is None else _np(self.inv_err),
                sigma_link=self.sigma_link, loss=self.loss,
                loss_eps=self.loss_eps, peer_table=table)
        return self._plan

    def _check(self):
        self.plan().check_pipeline()
        if self.comm is not None and self.comm.size > 1:
            self.comm.workspace().check_error()

    # ---- the two hooks as differentiable torch (CPU path and GPU oracle) ---
    def calc_partial_sumstats_from_params(self, params, randkey=None):
        p = params if isinstance(params, torch.Tensor) else \
            self._params(params)
        if self.fused and not p.requires_grad:
            from ..ops._lib import MODE_SUMSTATS_JAC
            return self.plan().evaluate(p, MODE_SUMSTATS_JAC,
                                        reduce=False)["sumstats"]
        p = p.to(self.device)
        dt = p.dtype
        shift = self.features.to(dt) @ p[:self.nfeat]
        return _ref.erf_binned_partial(
            self.x.to(dt) + shift, self._edges_t.to(dt), self._norm_t.to(dt),
            torch.zeros((), dtype=dt, device=p.device), p[self.nfeat],
            self.sigma_link)

    def calc_loss_from_sumstats(self, sumstats, sumstats_aux=None,
                                randkey=None):
        if self._target_t is None:
            raise ValueError("no target_sumstats given")
        s = torch.as_tensor(sumstats)
        return _ref.binned_loss(
            s, self._target_t.to(s), self.loss, self.loss_eps,
            None if self._inv_err_t is None else self._inv_err_t.to(s))

    # ---- fused paths --------------------------------------------------------
    def calc_sumstats_from_params(self, params, total=True, randkey=None):
        if not self.fused:
            return super().calc_sumstats_from_params(params, total, randkey)
        from ..ops._lib import MODE_SUMSTATS_JAC
        return self.plan().evaluate(self._params(params), MODE_SUMSTATS_JAC,
                                    reduce=total)["sumstats"]

    def calc_sumstats_and_jac_from_params(self, params, total=True):
        if not self.fused:
            p = self._params(params)
            jac = torch.autograd.functional.jacobian(
                lambda q: self.calc_partial_sumstats_from_params(q), p)
            s = self.calc_partial_sumstats_from_params(p)
            if total:
                from ..parallel.collectives import reduce_sum
                s, jac = reduce_sum(s, comm=self.comm), \
                    reduce_sum(jac, comm=self.comm)
            return s, jac
        from ..ops._lib import MODE_SUMSTATS_JAC
        out = self.plan().evaluate(self._params(params), MODE_SUMSTATS_JAC,
                                   reduce=total)
        return out["sumstats"], out["jac"]

    def calc_loss_from_params(self, params, randkey=None):
        if not self.fused:
            return super().calc_loss_from_params(params, randkey)
        with torch.no_grad():
            return self.calc_loss_from_sumstats(
                self.calc_sumstats_from_params(params))

    def _vjp(self, params, randkey=None, include_loss=True):
        if not self.fused:
            return super()._vjp(params, randkey, include_loss)
        from ..ops._lib import MODE_LOSS_GRAD, MODE_SUMSTATS_JAC
        p = self._params(params)
        if self._builtin_loss:
            out = self.plan().evaluate(p, MODE_LOSS_GRAD)
            loss, grad = out["loss"], out["grad"]
        else:
            out = self.plan().evaluate(p, MODE_SUMSTATS_JAC)
            kw = {} if randkey is None else {"randkey": randkey}
            loss, dlds = self._loss_and_dloss(out["sumstats"], None, kw)
            grad = dlds.to(out["jac"]) @ out["jac"]
        return (loss, grad) if include_loss else grad

    def run_adam(self, guess, nsteps=100, param_bounds=None,
                 learning_rate=0.01, randkey=None, const_randkey=False,
                 comm=None, reference_compat=True, init_state=None,
                 return_state=False, return_loss=False):
        """Adam, one fused tensor-core launch per step."""
        if not (self.fused and self._builtin_loss):
            return super().run_adam(
                guess, nsteps, param_bounds, learning_rate, randkey,
                const_randkey, comm, reference_compat, init_state,
                return_state, return_loss)
        from ..ops._lib import MODE_ADAM
        from ..ops.linlink import LinFitState
        plan = self.plan()
        p0 = self._params(guess)
        bounds, u0 = None, p0
        if param_bounds is not None:
            bounds = _adam.normalize_bounds(param_bounds, p0.numel())
            u0 = _adam.apply_transforms(p0.cpu(), bounds).to(self.device)
        step0, m, v = 0, None, None
        if init_state is not None:
            step0, m, v = int(init_state.step), init_state.m, init_state.v
        st = LinFitState(p0, u0, nsteps, self.device, step0=step0, m=m, v=v)
        st.traj[0].copy_(p0)
        args = plan.make_optimizer_args(MODE_ADAM, st, learning_rate, bounds,
                                        jac_at_params=reference_compat)
        for _ in range(nsteps):
            plan.launch(args)
        self._check()
        out = [st.traj]
        if return_loss:
            out.append(st.loss_hist[:nsteps])
        if return_state:
            out.append(_adam.AdamState(step0 + nsteps, st.m, st.v))
        return out[0] if len(out) == 1 else tuple(out)

    def run_simple_grad_descent(self, guess, nsteps=100, learning_rate=0.01):
        if not (self.fused and self._builtin_loss):
            return super().run_simple_grad_descent(guess, nsteps,
                                                   learning_rate)
        from ..ops._lib import MODE_GD
        from ..ops.linlink import LinFitState
        plan = self.plan()
        p0 = self._params(guess)
        st = LinFitState(p0, p0, nsteps, self.device, rows=max(nsteps, 1))
        args = plan.make_optimizer_args(MODE_GD, st, learning_rate)
        for _ in range(nsteps):
            plan.launch(args)
        self._check()
        return GradDescentResult(loss=st.loss_hist[:nsteps],
                                 params=st.traj[:nsteps],
                                 aux=[None] * nsteps)
