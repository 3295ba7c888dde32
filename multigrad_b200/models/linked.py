"""Soft-binned one-point models with a USER-DEFINED link.

The reference's contract is "write any differentiable
``calc_partial_sumstats_from_params``" (``multigrad/multigrad.py:212-223``),
differentiated by ``jax.vjp`` in a second pass over the objects
(``:508-538``).  :class:`LinkedBinnedModel` keeps that freedom where it
matters for this family -- how an object's mean and width depend on its
coordinate, its features and the parameters -- and still runs as ONE fused
launch per evaluation / optimiser step: the link is a short Python function
(see :mod:`multigrad_b200.link`) that is traced, differentiated symbolically
and compiled into the kernel of ``csrc/linkbin.cu``.

    from multigrad_b200 import link as L

    def my_link(x, feat, p, rng):            # mass-dependent scatter
        return x + p[0], p[1] + p[2] * (x - 12.0)

    model = LinkedBinnedModel(x=log_mhalo, link=my_link, nparams=3,
                              bin_edges=edges, volume=V, target_sumstats=t)
    traj = model.run_adam(guess, nsteps=200)

On CPU (or when ``params`` requires grad) the same function runs on PyTorch
tensors through the generic :class:`~multigrad_b200.core.OnePointModel`
machinery -- that is also the oracle of the GPU tests.  Stochastic links draw
from ``rng``; the key of step ``s`` is derived on the device from ``(seed,
s)`` and is the same on every rank (the reference splits a JAX key per step on
the root, ``multigrad/adam.py:60-62``).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Callable, Optional, Sequence

import numpy as np
import torch

from .. import adam as _adam
from ..core import OnePointModel
from ..link import LinkSpec, TorchRng, trace
from ..ops import reference as _ref
from ..parallel.collectives import reduce_sum
from ..util import GradDescentResult

__all__ = ["LinkedBinnedModel"]


def _np(v):
    if isinstance(v, torch.Tensor):
        return v.detach().cpu().double().numpy()
    return np.asarray(v, dtype=np.float64)


def _seed_of(randkey) -> int:
    """An integer seed from an ``int`` or a key tensor of ``utils.rng``."""
    if randkey is None:
        return 0
    if isinstance(randkey, torch.Tensor):
        w = [int(v) & 0xFFFFFFFF for v in randkey.detach().cpu().reshape(-1)]
        s = 0
        for v in w:
            s = ((s << 32) | v) & ((1 << 64) - 1)
        return s
    return int(randkey) & ((1 << 64) - 1)


@dataclass(eq=False)
class LinkedBinnedModel(OnePointModel):
    """Soft-binned statistic ``s_k = sum_h w_h [Phi_{k+1,h} - Phi_{k,h}] /
    (V width_k)`` with ``Phi_{e,h} = Phi((edge_e - mu_h) / sigma_h)`` and a
    user link ``(mu_h, sigma_h) = link(x_h, feat_h, params, rng)``.

    Parameters (besides those of :class:`OnePointModel`)
    ----------------------------------------------------
    x : this rank's shard of object coordinates
    link : ``link(x, feat, p, rng) -> (mu, sigma)``, see the module docstring
    nparams : length of the parameter vector (<= 8, and ``(1 + nparams) *
        nbins <= 96``)
    features : sequence of per-object feature tensors (<= 4), ``feat[j]``
    bin_edges, volume, target_sumstats, weights, loss, loss_eps, inv_err :
        as for :class:`~multigrad_b200.models.ErfBinnedModel`
    seed : seed of a stochastic link (``run_adam(randkey=...)`` overrides it)
    """
    x: Any = None
    link: Optional[Callable] = None
    nparams: int = 0
    features: Sequence = ()
    bin_edges: Any = None
    volume: float = 1.0
    target_sumstats: Any = None
    weights: Any = None
    loss: str = "log_mse"
    loss_eps: float = 0.0
    inv_err: Any = None
    seed: int = 0

    def __post_init__(self):
        super().__post_init__()
        if self.x is None or self.link is None or self.bin_edges is None \
                or self.nparams < 1:
            raise ValueError(
                "LinkedBinnedModel needs x, link, nparams and bin_edges")
        if isinstance(self.x, torch.Tensor) and self.x.is_cuda \
                and self.device.type != "cuda":
            self.device = self.x.device
        self.x = torch.as_tensor(self.x).detach().to(
            self.device, torch.float32).contiguous()
        self.features = [torch.as_tensor(f).detach().to(
            self.device, torch.float32).contiguous() for f in self.features]
        if self.weights is not None:
            self.weights = torch.as_tensor(self.weights).detach().to(
                self.device, torch.float32).contiguous()
        self.spec: LinkSpec = trace(self.link, self.nparams,
                                    len(self.features))
        self._edges64 = _np(self.bin_edges).ravel()
        self.nbins = self._edges64.size - 1
        self._norm64 = 1.0 / (float(self.volume) * np.diff(self._edges64))
        self._edges_t = torch.as_tensor(self._edges64, device=self.device)
        self._norm_t = torch.as_tensor(self._norm64, device=self.device)
        self._target_t = None
        self._inv_err_t = None
        self._builtin_loss = (type(self).calc_loss_from_sumstats
                              is LinkedBinnedModel.calc_loss_from_sumstats)
        # global index of this shard's first object (random streams)
        sizes = [int(self.x.numel())] if self.comm is None or \
            self.comm.size == 1 else self.comm.allgather(int(self.x.numel()))
        rank = 0 if self.comm is None else self.comm.rank
        self.index0 = int(sum(sizes[:rank]))
        self._plan = None
        if self.target_sumstats is not None:
            self.set_target(self.target_sumstats, self.inv_err)

    # ------------------------------------------------------------------ setup
    @property
    def fused(self) -> bool:
        return self.device.type == "cuda"

    def set_target(self, target, inv_err=None) -> None:
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
        """The launch plan; the first call compiles the link (a few seconds)
        and, with several ranks, allocates the communicator's peer workspace
        (collective)."""
        if self._plan is None:
            from ..ops.linkbin import LinkBinPlan
            table = None
            if self.comm is not None and self.comm.size > 1:
                table = self.comm.workspace().table
            self._plan = LinkBinPlan(
                self.spec, self.x, self.features, self._edges64,
                self._norm64,
                target=None if self.target_sumstats is None
                else _np(self.target_sumstats),
                inv_err=None if self.inv_err is None else _np(self.inv_err),
                loss=self.loss, loss_eps=self.loss_eps, weights=self.weights,
                seed=self.seed, index0=self.index0, peer_table=table)
        return self._plan

    def _check_peers(self):
        if self.comm is not None and self.comm.size > 1:
            self.comm.workspace().check_error()

    def _set_seed(self, randkey) -> None:
        if randkey is not None:
            self.seed = _seed_of(randkey)
        if self._plan is not None:
            self._plan.model.seed = int(self.seed)

    # --------------------------------------------------------------- the hooks
    def calc_partial_sumstats_from_params(self, params, randkey=None,
                                          step: int = 0):
        """This rank's share of the statistic: the kernel for plain
        parameters on CUDA, differentiable PyTorch otherwise."""
        p = params if isinstance(params, torch.Tensor) else \
            self._params(params)
        self._set_seed(randkey)
        if self.fused and not p.requires_grad:
            from ..ops.erfbin import MODE_SUMSTATS
            return self.plan().evaluate(p, MODE_SUMSTATS, reduce=False,
                                        rng_step=step)["sumstats"]
        p = p.to(self.device)
        dt = p.dtype
        idx = self.index0 + torch.arange(self.x.numel(), device=self.device)
        mu, sigma = self.spec.torch_eval(
            self.x.to(dt), [f.to(dt) for f in self.features], p,
            seed=self.seed, step=step, index=idx)
        z = (self._edges_t.to(dt)[:, None] - mu[None, :]) / \
            (np.sqrt(2.0) * sigma[None, :])
        cdf = 0.5 * (1.0 + torch.erf(z))
        d = cdf[1:] - cdf[:-1]
        if self.weights is not None:
            d = d * self.weights.to(dt)[None, :]
        return d.sum(dim=1) * self._norm_t.to(dt)

    def calc_loss_from_sumstats(self, sumstats, sumstats_aux=None,
                                randkey=None):
        if self._target_t is None:
            raise ValueError("no target_sumstats given")
        s = torch.as_tensor(sumstats)
        return _ref.binned_loss(
            s, self._target_t.to(s), self.loss, self.loss_eps,
            None if self._inv_err_t is None else self._inv_err_t.to(s))

    # ------------------------------------------------------------ fused paths
    def calc_sumstats_from_params(self, params, total=True, randkey=None):
        if not self.fused:
            return super().calc_sumstats_from_params(params, total, randkey)
        from ..ops.erfbin import MODE_SUMSTATS
        self._set_seed(randkey)
        return self.plan().evaluate(self._params(params), MODE_SUMSTATS,
                                    reduce=total)["sumstats"]

    def calc_sumstats_and_jac_from_params(self, params, total=True,
                                          randkey=None, step: int = 0):
        """``(s[K], ds/dp[K, P])`` in one pass."""
        p = self._params(params)
        self._set_seed(randkey)
        if self.fused:
            from ..ops.erfbin import MODE_SUMSTATS_JAC
            out = self.plan().evaluate(p, MODE_SUMSTATS_JAC, reduce=total,
                                       rng_step=step)
            return out["sumstats"], out["jac"]
        q = p.detach().clone().requires_grad_(True)
        jac = torch.autograd.functional.jacobian(
            lambda v: self.calc_partial_sumstats_from_params(v, step=step), q)
        s = self.calc_partial_sumstats_from_params(p, step=step)
        if total:
            s, jac = reduce_sum(s, comm=self.comm), \
                reduce_sum(jac, comm=self.comm)
        return s, jac

    def calc_loss_from_params(self, params, randkey=None):
        if not self.fused:
            return super().calc_loss_from_params(params, randkey)
        s = self.calc_sumstats_from_params(params, randkey=randkey)
        with torch.no_grad():
            return self.calc_loss_from_sumstats(s)

    def _vjp(self, params, randkey=None, include_loss=True):
        if not self.fused:
            return super()._vjp(params, randkey, include_loss)
        from ..ops.erfbin import MODE_LOSS_GRAD, MODE_SUMSTATS_JAC
        p = self._params(params)
        self._set_seed(randkey)
        if self._builtin_loss:
            if self._target_t is None:
                raise ValueError("no target_sumstats given")
            out = self.plan().evaluate(p, MODE_LOSS_GRAD)
            loss, grad = out["loss"], out["grad"]
        else:
            out = self.plan().evaluate(p, MODE_SUMSTATS_JAC)
            kw = {} if randkey is None else {"randkey": randkey}
            loss, dlds = self._loss_and_dloss(out["sumstats"], None, kw)
            grad = dlds.to(out["jac"]) @ out["jac"]
        return (loss, grad) if include_loss else grad

    # -------------------------------------------------------------- optimisers
    def _fit(self, mode, guess, nsteps, learning_rate, bounds=None,
             reference_compat=True, init_state=None, rows=None,
             const_key=False):
        from ..ops.erfbin import FitState
        from ..ops._lib import STEP_CONST_KEY
        plan = self.plan()
        p0 = self._params(guess)
        u0 = p0
        if bounds is not None:
            bounds = _adam.normalize_bounds(bounds, p0.numel())
            u0 = _adam.apply_transforms(p0.cpu(), bounds).to(self.device)
        step0, m, v = 0, None, None
        if init_state is not None:
            step0, m, v = int(init_state.step), init_state.m, init_state.v
        st = FitState(p0, u0, nsteps, self.device, step0=step0, m=m, v=v,
                      rows=rows)
        args = plan.make_optimizer_args(mode, st, learning_rate, bounds,
                                        jac_at_params=reference_compat)
        if const_key:
            args.flags |= STEP_CONST_KEY
        for _ in range(nsteps):
            plan.launch(args)
        self._check_peers()
        return st, step0

    def run_adam(self, guess, nsteps=100, param_bounds=None,
                 learning_rate=0.01, randkey=None, const_randkey=False,
                 comm=None, reference_compat=True, init_state=None,
                 return_state=False, return_loss=False):
        """Adam with one fused launch per step (sumstats + Jacobian + NVLink
        exchange + loss + chain rule + update on the device).  ``randkey``
        seeds a stochastic link; the key of step ``s`` is derived on the
        device from ``(seed, s)`` unless ``const_randkey`` freezes it."""
        if const_randkey and randkey is None:
            raise ValueError("const_randkey=True requires a randkey")
        if not (self.fused and self._builtin_loss):
            return super().run_adam(
                guess, nsteps, param_bounds, learning_rate, randkey,
                const_randkey, comm, reference_compat, init_state,
                return_state, return_loss)
        from ..ops.erfbin import MODE_ADAM
        self.plan()
        self._set_seed(randkey)
        st, step0 = self._fit(MODE_ADAM, guess, nsteps, learning_rate,
                              param_bounds, reference_compat, init_state,
                              const_key=const_randkey)
        st.traj[0].copy_(self._params(guess))
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
        from ..ops.erfbin import MODE_GD
        st, _ = self._fit(MODE_GD, guess, nsteps, learning_rate,
                          rows=max(nsteps, 1))
        return GradDescentResult(loss=st.loss_hist[:nsteps],
                                 params=st.traj[:nsteps],
                                 aux=[None] * nsteps)
