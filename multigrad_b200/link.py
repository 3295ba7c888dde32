"""Link functions for fused user-defined models.

The reference lets a user write ``calc_partial_sumstats_from_params`` as any
JAX function and differentiates it with ``jax.vjp``
(``multigrad/multigrad.py:212-223, 508-538``).  For the soft-binned family the
user-specific part is the *link*: how an object's mean and width follow from
its coordinate, its features and the parameters.  Here the link is an ordinary
Python function

    def link(x, feat, p, rng):
        mu = x + p[0]
        sigma = p[1] + p[2] * (x - 12.0)
        return mu, sigma

written with normal arithmetic and the functions of this module (``exp``,
``log``, ``log10``, ``exp10``, ``sqrt``, ``tanh``, ``abs``, ``rng.normal(i)``,
``rng.uniform(i)``).  It is used twice:

* called with **PyTorch tensors** it is the differentiable oracle / CPU path;
* called with **symbols** (:func:`trace`) it yields an expression graph that is
  differentiated symbolically with respect to every parameter and emitted as a
  CUDA device function (:func:`generate_header`), which
  ``multigrad_b200/ops/linkbin.py`` compiles into the fused kernel of
  ``csrc/linkbin.cu``.  The user writes no derivative and no CUDA.

Random draws are counter based: draw ``i`` of object ``h`` at optimiser step
``s`` is a pure function of ``(seed, s, h, i)`` -- the same on every rank and
in the PyTorch oracle (:class:`TorchRng`).
"""
from __future__ import annotations

import hashlib
import math
from typing import Callable, List, Sequence, Tuple

import torch

__all__ = ["exp", "log", "log10", "exp10", "sqrt", "tanh", "abs", "trace",
           "generate_header", "LinkSpec", "TorchRng", "Sym"]


# --------------------------------------------------------------------- graph
class Sym:
    """Node of a traced expression (packed pair of objects on the device)."""
    __slots__ = ("op", "args", "val")
    __array_priority__ = 1000

    def __init__(self, op, args=(), val=None):
        self.op, self.args, self.val = op, tuple(args), val

    # constants fold eagerly; 0 and 1 are recognised so that derivatives of
    # parameters a quantity does not depend on vanish identically
    @staticmethod
    def wrap(v) -> "Sym":
        if isinstance(v, Sym):
            return v
        if isinstance(v, torch.Tensor):
            v = float(v)
        return Sym("const", val=float(v))

    def is_const(self, v=None) -> bool:
        return self.op == "const" and (v is None or self.val == v)

    def __add__(self, o):
        o = Sym.wrap(o)
        if self.is_const() and o.is_const():
            return Sym.wrap(self.val + o.val)
        if self.is_const(0.0):
            return o
        if o.is_const(0.0):
            return self
        return Sym("add", (self, o))

    __radd__ = __add__

    def __sub__(self, o):
        o = Sym.wrap(o)
        if self.is_const() and o.is_const():
            return Sym.wrap(self.val - o.val)
        if o.is_const(0.0):
            return self
        if self.is_const(0.0):
            return -o
        return Sym("sub", (self, o))

    def __rsub__(self, o):
        return Sym.wrap(o) - self

    def __mul__(self, o):
        o = Sym.wrap(o)
        if self.is_const() and o.is_const():
            return Sym.wrap(self.val * o.val)
        if self.is_const(0.0) or o.is_const(0.0):
            return Sym.wrap(0.0)
        if self.is_const(1.0):
            return o
        if o.is_const(1.0):
            return self
        return Sym("mul", (self, o))

    __rmul__ = __mul__

    def __truediv__(self, o):
        o = Sym.wrap(o)
        if o.is_const():
            return self * (1.0 / o.val)
        if self.is_const(0.0):
            return self
        return Sym("div", (self, o))

    def __rtruediv__(self, o):
        return Sym.wrap(o) / self

    def __neg__(self):
        if self.is_const():
            return Sym.wrap(-self.val)
        if self.op == "neg":
            return self.args[0]
        return Sym("neg", (self,))

    def __pow__(self, e):
        if isinstance(e, Sym):
            if not e.is_const():
                return exp(e * log(self))
            e = e.val
        e = float(e)
        if e == 0.0:
            return Sym.wrap(1.0)
        if e == 1.0:
            return self
        if e == 2.0:
            return self * self
        if self.is_const():
            return Sym.wrap(self.val ** e)
        return Sym("powc", (self,), e)

    def __rpow__(self, base):
        # base ** self with a constant base
        base = float(base)
        if base == 10.0:
            return exp10(self)
        return exp(self * math.log(base))


def _unary(name: str, torch_fn, const_fn):
    def f(v):
        if isinstance(v, Sym):
            if v.is_const():
                return Sym.wrap(const_fn(v.val))
            return Sym(name, (v,))
        return torch_fn(torch.as_tensor(v))
    f.__name__ = name
    f.__doc__ = f"``{name}`` of a tensor or of a traced symbol."
    return f


exp = _unary("exp", torch.exp, math.exp)
log = _unary("log", torch.log, math.log)
log10 = _unary("log10", torch.log10, math.log10)
exp10 = _unary("exp10", lambda t: torch.pow(10.0, t), lambda v: 10.0 ** v)
sqrt = _unary("sqrt", torch.sqrt, math.sqrt)
tanh = _unary("tanh", torch.tanh, math.tanh)
abs = _unary("abs", torch.abs, math.fabs)  # noqa: A001 - mirrors torch.abs


class _SymRng:
    """Random draws of a traced link (leaves of the graph)."""

    def __init__(self):
        self.used = False

    def normal(self, i: int = 0) -> Sym:
        self.used = True
        return Sym("normal", val=int(i))

    def uniform(self, i: int = 0) -> Sym:
        self.used = True
        return Sym("uniform", val=int(i))


# ------------------------------------------------------------ random numbers
_M64 = (1 << 64) - 1


def _mix_int(z: int) -> int:
    z = (z + 0x9E3779B97F4A7C15) & _M64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _M64
    return z ^ (z >> 31)


def step_key(seed: int, step: int) -> int:
    """The per-step key the kernel derives (csrc/linkbin.cu)."""
    return _mix_int((seed & _M64) ^ _mix_int(
        (0x5851F42D4C957F2D + (step & 0xFFFFFFFF)) & _M64))


def _mix_t(z: torch.Tensor) -> torch.Tensor:
    """SplitMix64 finaliser on int64 tensors (two's complement wraps like the
    unsigned arithmetic of the kernel; logical shifts are emulated)."""
    def shr(v, n):
        return (v >> n) & ((1 << (64 - n)) - 1)

    def c(v):          # python int -> wrapped int64 scalar
        v &= _M64
        return v - (1 << 64) if v >= (1 << 63) else v
    z = z + c(0x9E3779B97F4A7C15)
    z = (z ^ shr(z, 30)) * c(0xBF58476D1CE4E5B9)
    z = (z ^ shr(z, 27)) * c(0x94D049BB133111EB)
    return z ^ shr(z, 31)


class TorchRng:
    """PyTorch twin of the kernel's counter-based generator: draw ``i`` of
    the objects with global indices ``index`` (int64 tensor) at step ``step``
    for ``seed``."""

    def __init__(self, seed: int, step: int, index: torch.Tensor):
        k = step_key(seed, step)
        self._key = k - (1 << 64) if k >= (1 << 63) else k
        self._index = index.to(torch.int64)

    def uniform(self, i: int = 0) -> torch.Tensor:
        h = _mix_t(self._key ^ _mix_t(self._index * 0x100 + int(i)))
        top = (h >> 40) & ((1 << 24) - 1)
        return ((top.to(torch.float32) + 0.5) * (1.0 / 16777216.0))

    def normal(self, i: int = 0) -> torch.Tensor:
        u, v = self.uniform(2 * i), self.uniform(2 * i + 1)
        return torch.sqrt(-2.0 * torch.log(u)) * torch.cos(
            2.0 * math.pi * v)


# ----------------------------------------------------------- differentiation
def _diff(node: Sym, k: int, memo: dict) -> Sym:
    """d node / d p_k by the chain rule on the traced graph."""
    key = (id(node), k)
    if key in memo:
        return memo[key]
    op, a = node.op, node.args
    zero = Sym.wrap(0.0)
    if op == "param":
        r = Sym.wrap(1.0 if node.val == k else 0.0)
    elif op in ("const", "x", "xc", "feat", "normal", "uniform"):
        r = zero
    elif op == "add":
        r = _diff(a[0], k, memo) + _diff(a[1], k, memo)
    elif op == "sub":
        r = _diff(a[0], k, memo) - _diff(a[1], k, memo)
    elif op == "neg":
        r = -_diff(a[0], k, memo)
    elif op == "mul":
        r = _diff(a[0], k, memo) * a[1] + a[0] * _diff(a[1], k, memo)
    elif op == "div":
        da, db = _diff(a[0], k, memo), _diff(a[1], k, memo)
        r = da / a[1] - (node * db) / a[1]
    elif op == "powc":
        r = (node.val * a[0] ** (node.val - 1.0)) * _diff(a[0], k, memo)
    elif op == "exp":
        r = node * _diff(a[0], k, memo)
    elif op == "exp10":
        r = (node * math.log(10.0)) * _diff(a[0], k, memo)
    elif op == "log":
        r = _diff(a[0], k, memo) / a[0]
    elif op == "log10":
        r = _diff(a[0], k, memo) / (a[0] * math.log(10.0))
    elif op == "sqrt":
        r = _diff(a[0], k, memo) / (2.0 * node)
    elif op == "tanh":
        r = (1.0 - node * node) * _diff(a[0], k, memo)
    elif op == "abs":
        # sign(a) = a / |a|
        r = (a[0] / node) * _diff(a[0], k, memo)
    else:
        raise NotImplementedError(f"no derivative rule for {op!r}")
    memo[key] = r
    return r


def _strip_x(mu: Sym, x: Sym):
    """``mu - x`` symbolically when ``mu`` is a sum with one bare ``x`` term
    (the usual ``mu = x + ...``): the kernel then adds the shift to the
    *centred* coordinate, which keeps the cancellation ``edge - mu`` exact."""
    terms: List[Tuple[float, Sym]] = []

    def walk(n, sign):
        if n.op == "add":
            walk(n.args[0], sign)
            walk(n.args[1], sign)
        elif n.op == "sub":
            walk(n.args[0], sign)
            walk(n.args[1], -sign)
        elif n.op == "neg":
            walk(n.args[0], -sign)
        else:
            terms.append((sign, n))
    walk(mu, 1.0)
    hits = [i for i, (s, n) in enumerate(terms) if n is x and s > 0]
    if len(hits) != 1:
        return mu - x
    rest = Sym.wrap(0.0)
    for i, (s, n) in enumerate(terms):
        if i == hits[0]:
            continue
        rest = rest + n if s > 0 else rest - n
    return rest


# -------------------------------------------------------------------- codegen
_FN1 = {"exp": "vexp", "log": "vlog", "log10": "vlog10", "exp10": "vexp10",
        "sqrt": "vsqrt", "tanh": "vtanh", "abs": "vabs", "neg": "vneg"}


class _Emitter:
    def __init__(self):
        self.lines: List[str] = []
        self.names: dict = {}

    def const(self, v: float) -> str:
        if v != v or v in (float("inf"), float("-inf")):
            raise ValueError("non-finite constant in a link expression")
        return f"f2s({v!r}f)" if "e" in repr(v) or "." in repr(v) \
            else f"f2s({v!r}.0f)"

    def emit(self, n: Sym) -> str:
        if id(n) in self.names:
            return self.names[id(n)]
        op, a = n.op, n.args
        if op == "const":
            return self.const(n.val)
        if op == "x":
            return "x"
        if op == "xc":
            return "xc"
        if op == "feat":
            return f"f[{n.val}]"
        if op == "param":
            expr = f"f2s(p[{n.val}])"
        elif op == "normal":
            expr = f"ul_normal(rng, {n.val})"
        elif op == "uniform":
            expr = f"ul_uniform(rng, {n.val})"
        elif op in ("add", "sub", "mul", "div"):
            fn = {"add": "vadd", "sub": "vsub", "mul": "vmul",
                  "div": "vdiv"}[op]
            expr = f"{fn}({self.emit(a[0])}, {self.emit(a[1])})"
        elif op == "powc":
            expr = f"vpowc({self.emit(a[0])}, {n.val!r}f)"
        elif op in _FN1:
            expr = f"{_FN1[op]}({self.emit(a[0])})"
        else:
            raise NotImplementedError(op)
        name = f"t{len(self.names)}"
        self.names[id(n)] = name
        self.lines.append(f"  const float2 {name} = {expr};")
        return name


class LinkSpec:
    """A traced link: symbolic ``shift = mu - x`` and ``sigma`` with their
    non-zero parameter derivatives, plus the generated CUDA header."""

    def __init__(self, fn: Callable, nparams: int, nfeat: int):
        self.fn, self.nparams, self.nfeat = fn, int(nparams), int(nfeat)
        x = Sym("x")
        feats = tuple(Sym("feat", val=j) for j in range(self.nfeat))
        params = tuple(Sym("param", val=k) for k in range(self.nparams))
        rng = _SymRng()
        out = fn(x, feats, params, rng)
        if not (isinstance(out, (tuple, list)) and len(out) == 2):
            raise TypeError("a link returns the pair (mu, sigma)")
        mu, sigma = Sym.wrap(out[0]), Sym.wrap(out[1])
        self.stochastic = rng.used
        self.shift = _strip_x(mu, x)
        self.sigma = sigma
        memo: dict = {}
        self.dmu = [_diff(self.shift, k, memo) for k in range(self.nparams)]
        self.dsg = [_diff(sigma, k, memo) for k in range(self.nparams)]
        self.mu_mask = sum(1 << k for k, d in enumerate(self.dmu)
                           if not d.is_const(0.0))
        self.sg_mask = sum(1 << k for k, d in enumerate(self.dsg)
                           if not d.is_const(0.0))
        self.header = self._generate()
        self.digest = hashlib.sha256(self.header.encode()).hexdigest()[:16]

    def _generate(self) -> str:
        em = _Emitter()
        body = [f"  shift = {em.emit(self.shift)};",
                f"  sigma = {em.emit(self.sigma)};"]
        for k in range(self.nparams):
            if (self.mu_mask >> k) & 1:
                body.append(f"  dmu[{k}] = {em.emit(self.dmu[k])};")
            if (self.sg_mask >> k) & 1:
                body.append(f"  dsg[{k}] = {em.emit(self.dsg[k])};")
        head = [
            "// generated by multigrad_b200/link.py from a user link function",
            f"#define UL_NP {self.nparams}",
            f"#define UL_NF {self.nfeat}",
            f"#define UL_MU_MASK {self.mu_mask:#x}",
            f"#define UL_SG_MASK {self.sg_mask:#x}",
            f"#define UL_STOCHASTIC {int(self.stochastic)}",
            "__device__ __forceinline__ void ul_eval(",
            "    float2 x, float2 xc, const float2* f, const float* p,",
            "    const UlRng& rng, float2& shift, float2& sigma, float2* dmu,",
            "    float2* dsg) {",
        ]
        return "\n".join(head + em.lines + body + ["}", ""])

    def numeric(self, node: Sym, x, feats, params, rng=None):
        """Evaluate a node of the traced graph on tensors (tests: the
        symbolic derivatives against autograd of the same function)."""
        memo: dict = {}

        def ev(n):
            if id(n) in memo:
                return memo[id(n)]
            a = [ev(c) for c in n.args]
            op = n.op
            if op == "const":
                r = torch.full_like(x, n.val)
            elif op == "x":
                r = x
            elif op == "feat":
                r = feats[n.val]
            elif op == "param":
                r = params[n.val] + torch.zeros_like(x)
            elif op in ("normal", "uniform"):
                r = getattr(rng, op)(n.val)
            elif op == "add":
                r = a[0] + a[1]
            elif op == "sub":
                r = a[0] - a[1]
            elif op == "mul":
                r = a[0] * a[1]
            elif op == "div":
                r = a[0] / a[1]
            elif op == "neg":
                r = -a[0]
            elif op == "powc":
                r = a[0] ** n.val
            else:
                r = globals()[op](a[0])
            memo[id(n)] = r
            return r
        return ev(node)

    # ------------------------------------------------------- PyTorch twin
    def torch_eval(self, x, feats, params, seed=0, step=0, index=None):
        """``(mu, sigma)`` of the link on tensors (differentiable)."""
        if index is None:
            index = torch.arange(x.numel(), device=x.device)
        rng = TorchRng(seed, step, index)
        mu, sigma = self.fn(x, tuple(feats), params, rng)
        return mu, sigma + torch.zeros_like(x)


def trace(fn: Callable, nparams: int, nfeat: int = 0) -> LinkSpec:
    """Trace ``fn(x, feat, p, rng) -> (mu, sigma)`` into a :class:`LinkSpec`."""
    return LinkSpec(fn, nparams, nfeat)


def generate_header(fn: Callable, nparams: int, nfeat: int = 0) -> str:
    return trace(fn, nparams, nfeat).header
