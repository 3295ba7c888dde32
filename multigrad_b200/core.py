"""Differentiable model core: the distributed chain rule (layer L2) and the
user-facing optimiser methods (layer L4).

Parity targets: reference ``multigrad/multigrad.py`` -- ``OnePointModel``
(``:186-544``) and ``OnePointGroup`` (``:547-607``).

A user subclasses :class:`OnePointModel` and provides two differentiable
**PyTorch** functions (the reference wants JAX functions):

* ``calc_partial_sumstats_from_params(params[, randkey])`` -- this rank's
  *additive* share of the summary statistics, and
* ``calc_loss_from_sumstats(sumstats[, sumstats_aux][, randkey])`` -- total
  summary statistics -> scalar loss.

This module is the *generic* path: arbitrary callables, ``torch.autograd``
for the two vector-Jacobian products and :func:`reduce_sum` for the two
reductions (on CUDA tensors that is our NVLink kernel, not NCCL/MPI).  Model
families that have a closed-form Jacobian get a fused single-kernel path that
overrides these methods (``multigrad_b200/models/erf_binned.py``).

Quirk decisions are those of SURVEY.md section 9 (Q1 follow the code for the
``*_has_aux`` flags, Q9 group accepts ``randkey``, Q10 identity equality,
Q13 the loss is evaluated once).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Tuple, Union

import numpy as np
import torch

from . import util
from . import adam as _adam
from . import bfgs as _bfgs
from .parallel import comm as _comm
from .parallel.collectives import reduce_sum

__all__ = ["OnePointModel", "OnePointGroup"]


def _as_params(params, device=None, dtype=torch.float32) -> torch.Tensor:
    if isinstance(params, torch.Tensor):
        p = params.detach()
        if not p.is_floating_point():
            p = p.to(dtype)
    else:
        p = torch.as_tensor(np.asarray(params, dtype=np.float64)).to(dtype)
    if device is not None and p.device != torch.device(device):
        p = p.to(device)
    return p


@dataclass(eq=False)
class OnePointModel:
    """Data-parallel differentiable one-point model (abstract).

    Parameters
    ----------
    aux_data : Any
        Anything the two user functions need (typically a dict holding this
        rank's shard of objects as tensors on ``device``).
    comm : Comm, optional
        Communicator to sum over; defaults to the world communicator.
    loss_func_has_aux : bool
        ``calc_loss_from_sumstats(...) -> (loss, aux)``.
    sumstats_func_has_aux : bool
        ``calc_partial_sumstats_from_params(x) -> (y, aux)`` and the loss is
        called as ``calc_loss_from_sumstats(y, aux)``.
        (The reference's class docstring has these two swapped; this follows
        its code, ``multigrad.py:394-396,422-423,518-528``.)
    device : torch.device | str, optional
        Where parameters live.  Defaults to this process's GPU when CUDA is
        available, else CPU.
    """
    aux_data: Any = None
    comm: Any = None
    loss_func_has_aux: bool = False
    sumstats_func_has_aux: bool = False
    device: Any = None

    # ------------------------------------------------------------------ hooks
    def calc_partial_sumstats_from_params(self, params, randkey=None):
        """Custom method to map parameters to summary statistics"""
        raise NotImplementedError(
            "Subclass must implement `calc_partial_sumstats_from_params`")

    def calc_loss_from_sumstats(self, sumstats, sumstats_aux=None,
                                randkey=None):
        """Custom method to map summary statistics to loss"""
        raise NotImplementedError(
            "Subclass must implement `calc_loss_from_sumstats`")

    # ------------------------------------------------------------------ init
    def __post_init__(self):
        if self.comm is None:
            self.comm = _comm.get_world()
        if self.device is None:
            self.device = _comm.local_device()
        self.device = torch.device(self.device)

    def _params(self, params) -> torch.Tensor:
        return _as_params(params, device=self.device)

    # ------------------------------------------------------------- optimisers
    def run_simple_grad_descent(self: Any, guess, nsteps=100,
                                learning_rate=0.01):
        """Fixed-learning-rate gradient descent (reference
        ``multigrad.py:226-256``).  Returns ``GradDescentResult(loss[nsteps],
        params[nsteps, ndim], aux)``; params are recorded *before* each
        update."""
        return util.simple_grad_descent(
            None, guess=self._params(guess), nsteps=nsteps,
            learning_rate=learning_rate,
            loss_and_grad_func=self.calc_loss_and_grad_from_params,
            has_aux=False)

    def run_adam(self: Any, guess, nsteps=100, param_bounds=None,
                 learning_rate=0.01, randkey=None, const_randkey=False,
                 comm=None, reference_compat=True, init_state=None,
                 return_state=False, return_loss=False):
        """Adam (reference ``multigrad.py:259-307``).

        Returns the parameter trajectory ``(nsteps+1, ndim)`` on every rank
        (plus loss history / final optimiser state when asked for).  With
        ``randkey`` a fresh key is derived for every step unless
        ``const_randkey``.
        """
        guess = self._params(guess)
        if const_randkey:
            assert randkey is not None, "Must pass randkey if const_randkey"
            fixed = _adam.init_randkey(randkey)
            randkey = None

            def loss_and_grad_fn(x, _, **kw):
                return self.calc_loss_and_grad_from_params(
                    x, randkey=fixed, **kw)
        else:
            def loss_and_grad_fn(x, _, **kw):
                return self.calc_loss_and_grad_from_params(x, **kw)
        return _adam.run_adam(
            loss_and_grad_fn, params=guess, data=None, nsteps=nsteps,
            param_bounds=param_bounds, learning_rate=learning_rate,
            randkey=randkey, reference_compat=reference_compat,
            init_state=init_state, return_state=return_state,
            return_loss=return_loss)

    def run_bfgs(self: Any, guess, maxsteps=100, param_bounds=None,
                 randkey=None, comm=None):
        """L-BFGS-B via scipy on rank 0 (reference ``multigrad.py:310-352``).
        Returns the same ``OptimizeResult`` on every rank."""
        comm = _model_comm(self) if comm is None else comm

        def fn(p, **kw):
            return self.calc_loss_and_grad_from_params(p, **kw)
        return _bfgs.run_bfgs(fn, guess, maxsteps=maxsteps,
                              param_bounds=param_bounds, randkey=randkey,
                              comm=comm)

    def run_lhs_param_scan(self, xmins, xmaxs, n_dim, num_evaluations,
                           seed=None, randkey=None):
        """Latin-hypercube scan (reference ``multigrad.py:354-388``): returns
        ``(params, sumstats, losses)`` as numpy arrays."""
        params = util.latin_hypercube_sampler(xmins, xmaxs, n_dim,
                                              num_evaluations, seed=seed)
        rk = {} if randkey is None else {"randkey": randkey}
        sumstats = [self.calc_sumstats_from_params(x, **rk) for x in params]
        if self.sumstats_func_has_aux:
            losses = [self.calc_loss_from_sumstats(*s, **rk) for s in sumstats]
            sumstats = [s[0] for s in sumstats]
        else:
            losses = [self.calc_loss_from_sumstats(s, **rk) for s in sumstats]
        if self.loss_func_has_aux:
            losses = [x[0] for x in losses]
        to_np = lambda t: torch.as_tensor(t).detach().cpu().numpy()  # noqa
        return (params, np.array([to_np(s) for s in sumstats]),
                np.array([to_np(x) for x in losses]))

    # --------------------------------------------------------------- sumstats
    def calc_sumstats_from_params(self, params, total=True, randkey=None):
        """Summary statistics at ``params``; summed over ranks if ``total``
        (reference ``multigrad.py:400-427``)."""
        kwargs = {} if randkey is None else {"randkey": randkey}
        with torch.no_grad():
            result, aux = self.calc_partial_sumstats_from_params(
                self._params(params), **kwargs), None
        if self.sumstats_func_has_aux:
            result, aux = result
        if total:
            result = reduce_sum(result, comm=self.comm)
        return (result, aux) if self.sumstats_func_has_aux else result

    # ------------------------------------------------------------------- loss
    def _loss_and_dloss(self, sumstats, sumstats_aux, kwargs):
        """One evaluation of the user loss giving ``(loss_out, dL/ds)``."""
        s = torch.as_tensor(sumstats).detach().clone().requires_grad_(True)
        args = (s, sumstats_aux) if self.sumstats_func_has_aux else (s,)
        with torch.enable_grad():
            out = self.calc_loss_from_sumstats(*args, **kwargs)
            loss = out[0] if self.loss_func_has_aux else out
            (dlds,) = torch.autograd.grad(loss, s, allow_unused=True)
        if dlds is None:
            dlds = torch.zeros_like(s)
        if self.loss_func_has_aux:
            out = (out[0].detach(), out[1])
        else:
            out = out.detach()
        return out, dlds

    def calc_dloss_dsumstats(self, sumstats, sumstats_aux=None, randkey=None):
        """``dL/d(sumstats)`` (reference ``multigrad.py:430-436``); returns
        ``(grad, aux)`` when ``loss_func_has_aux`` like ``jax.grad`` does."""
        kwargs = {} if randkey is None else {"randkey": randkey}
        out, dlds = self._loss_and_dloss(sumstats, sumstats_aux, kwargs)
        return (dlds, out[1]) if self.loss_func_has_aux else dlds

    def calc_loss_from_params(self, params, randkey=None):
        """Loss at ``params`` (reference ``multigrad.py:439-460``)."""
        kwargs = {} if randkey is None else {"randkey": randkey}
        sumstats = self.calc_sumstats_from_params(params, **kwargs)
        if not self.sumstats_func_has_aux:
            sumstats = (sumstats,)
        with torch.no_grad():
            return self.calc_loss_from_sumstats(*sumstats, **kwargs)

    def calc_dloss_dparams(self, params, randkey=None):
        """Gradient of the loss (reference ``multigrad.py:463-479``)."""
        return self._vjp(params, randkey=randkey, include_loss=False)

    def calc_loss_and_grad_from_params(self, params, randkey=None):
        """``(loss, dloss/dparams)`` in one pass (reference
        ``multigrad.py:482-505``)."""
        return self._vjp(params, randkey=randkey, include_loss=True)

    def _vjp(self, params, randkey=None, include_loss=True):
        """Distributed chain rule ``dL/dp = sum_j dL/ds_j * ds_j/dp``
        (reference ``multigrad.py:508-538``): local forward with a saved
        graph, all-reduce of the sumstats, ``dL/ds`` from the loss, local
        vector-Jacobian product, all-reduce of the gradient."""
        kwargs = {} if randkey is None else {"randkey": randkey}
        p = self._params(params).clone().requires_grad_(True)
        with torch.enable_grad():
            res = self.calc_partial_sumstats_from_params(p, **kwargs)
        aux = None
        if self.sumstats_func_has_aux:
            res, aux = res
        total = reduce_sum(res.detach(), comm=self.comm)
        loss_out, dlds = self._loss_and_dloss(total, aux, kwargs)
        if res.requires_grad:
            (g,) = torch.autograd.grad(res, p, grad_outputs=dlds.to(res),
                                       allow_unused=True)
        else:
            g = None
        if g is None:
            g = torch.zeros_like(p)
        dloss_dparams = reduce_sum(g.detach(), comm=self.comm)
        if include_loss:
            return loss_out, dloss_dparams
        return dloss_dparams

    def __hash__(self):
        return hash((self.comm.name if self.comm is not None else None,
                     id(self)))

    def __eq__(self, other):
        return self is other


def _model_comm(obj):
    return obj.main_comm if isinstance(obj, OnePointGroup) else obj.comm


@dataclass(eq=False)
class OnePointGroup:
    """Several :class:`OnePointModel` s evaluated at the same parameters on
    disjoint sub-communicators; losses and gradients are summed (reference
    ``multigrad.py:547-607``).

    Typical use: ``sub, n, g = split_subcomms(ranks_per_group=[3, 3, 2])``,
    each rank builds the model of *its* group with ``comm=sub`` and wraps it
    in ``OnePointGroup((model,), main_comm=COMM)``.
    """
    models: Union[Tuple[OnePointModel, ...], OnePointModel]
    main_comm: Any = None

    def __post_init__(self):
        if self.main_comm is None:
            self.main_comm = _comm.get_world()
        if isinstance(self.models, OnePointModel):
            self.models = (self.models,)
        self.models = tuple(self.models)
        assert isinstance(self.models[0], OnePointModel)

    @property
    def device(self):
        return self.models[0].device

    @property
    def comm(self):
        return self.main_comm

    def _params(self, params):
        return self.models[0]._params(params)

    def calc_loss_and_grad_from_params(self, params, randkey=None):
        """Sum of the member models' losses and gradients.

        Each model reduces over its own sub-communicator; only the root of a
        sub-communicator contributes its (already complete) value to the
        cross-group sum, which is a single all-reduce of ``1 + ndim`` numbers
        over ``main_comm`` (the reference pickles Python lists through two
        ``allgather`` s, ``multigrad.py:576-580``).
        """
        kw = {} if randkey is None else {"randkey": randkey}
        packed = None
        for model in self.models:
            loss, grad = model.calc_loss_and_grad_from_params(params, **kw)
            if isinstance(loss, tuple):
                loss = loss[0]
            v = torch.cat([torch.as_tensor(loss).reshape(1).to(grad),
                           grad.reshape(-1)])
            if model.comm is not None and model.comm.rank:
                v = v * 0
            packed = v if packed is None else packed + v
        total = reduce_sum(packed, comm=self.main_comm)
        return total[0], total[1:]

    # ------------------------------------------------- device-resident path
    def _device_ok(self, params: torch.Tensor) -> bool:
        """True when *every* member model on *every* rank of ``main_comm``
        can evaluate device-resident (fused kernels with a built-in loss), so
        that the fit can run through ``ops.group.GroupStepper``.  Collective
        call: the ranks agree on the answer and all take the same path."""
        ok = (params.is_cuda and params.numel() <= 8 and all(
            getattr(m, "fused", False) and getattr(m, "_builtin_loss", False)
            and hasattr(m, "device_loss_grad_launcher")
            and getattr(m, "_one_launch", True)
            and m.device == params.device for m in self.models))
        if self.main_comm is not None and self.main_comm.size > 1:
            ok = all(self.main_comm.allgather(bool(ok)))
        return bool(ok)

    def _device_stepper(self, params: torch.Tensor):
        from .ops.group import GroupStepper
        return GroupStepper(self.models, self.main_comm, params)

    def run_simple_grad_descent(self, guess, nsteps=100, learning_rate=0.01):
        """Fixed-step gradient descent on the summed loss.  With fused member
        models: per step one launch per model plus one fused
        reduce-and-update launch (``csrc/group.cu``), nothing read back."""
        p0 = self._params(guess)
        if nsteps <= 0 or not self._device_ok(p0):
            return OnePointModel.run_simple_grad_descent(
                self, guess, nsteps, learning_rate)
        from .ops.erfbin import FitState, MODE_GD
        from .ops.group import fill_optimizer_args
        from .ops._lib import MgbStepArgs
        st = FitState(p0, p0, nsteps, p0.device, rows=max(nsteps, 1))
        stepper = self._device_stepper(st.p)
        args = fill_optimizer_args(MgbStepArgs(), MODE_GD, st, learning_rate)
        for _ in range(nsteps):
            stepper.step(args)
        stepper.check_error()
        return util.GradDescentResult(loss=st.loss_hist[:nsteps],
                                      params=st.traj[:nsteps],
                                      aux=[None] * nsteps)

    def run_bfgs(self, guess, maxsteps=100, param_bounds=None, randkey=None):
        return OnePointModel.run_bfgs(
            self, guess, maxsteps, param_bounds=param_bounds,
            randkey=randkey, comm=self.main_comm)

    def run_adam(self, guess, nsteps=100, param_bounds=None,
                 learning_rate=0.01, randkey=None, const_randkey=False,
                 reference_compat=True, init_state=None, return_state=False,
                 return_loss=False):
        """Adam on the summed loss (reference ``multigrad.py:591-599``).  With
        fused member models the whole fit stays on the device: per step one
        launch per model plus one fused reduce-and-update launch."""
        p0 = self._params(guess)
        if randkey is not None or nsteps <= 0 or not self._device_ok(p0):
            return OnePointModel.run_adam(
                self, guess, nsteps, param_bounds, learning_rate, randkey,
                const_randkey=const_randkey, comm=self.main_comm,
                reference_compat=reference_compat, init_state=init_state,
                return_state=return_state, return_loss=return_loss)
        from .ops.erfbin import FitState, MODE_ADAM
        from .ops.group import fill_optimizer_args
        from .ops._lib import MgbStepArgs
        bounds, u0 = None, p0
        if param_bounds is not None:
            bounds = _adam.normalize_bounds(param_bounds, p0.numel())
            u0 = _adam.apply_transforms(p0.cpu(), bounds).to(p0.device)
        step0, m, v = 0, None, None
        if init_state is not None:
            step0, m, v = int(init_state.step), init_state.m, init_state.v
        st = FitState(p0, u0, nsteps, p0.device, step0=step0, m=m, v=v)
        st.traj[0].copy_(p0)
        stepper = self._device_stepper(st.p)
        args = fill_optimizer_args(MgbStepArgs(), MODE_ADAM, st,
                                   learning_rate, bounds,
                                   jac_at_params=reference_compat)
        for _ in range(nsteps):
            stepper.step(args)
        stepper.check_error()
        out = [st.traj]
        if return_loss:
            out.append(st.loss_hist[:nsteps])
        if return_state:
            out.append(_adam.AdamState(step0 + nsteps, st.m, st.v))
        return out[0] if len(out) == 1 else tuple(out)

    def __hash__(self):
        return hash((self.main_comm.name, self.models[0]))

    def __eq__(self, other):
        return isinstance(other, OnePointGroup) and self is other
