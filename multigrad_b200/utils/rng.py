"""Counter-based random keys (replaces ``jax.random`` keys of the reference,
``multigrad/adam.py:242-257``).

A key is a CPU ``torch.int64`` tensor of shape ``(2,)`` holding two unsigned
64-bit words (stored bit-for-bit in int64).  Keys are *values*: deriving a
sub-key never mutates the parent.  Derivation is a SplitMix64 hash of
``(key, index)`` done on the host, so there is no device kernel and no
host/device sync on the optimiser hot path.

Decision Q4: every rank derives the *same* per-step key;
rank-dependent streams are obtained explicitly with ``fold_in(key, rank)``.
"""
from __future__ import annotations

from typing import Sequence, Tuple, Union

import torch

__all__ = ["init_randkey", "gen_new_key", "split", "fold_in", "key_to_seed",
           "make_generator", "uniform", "normal", "is_key"]

_MASK = (1 << 64) - 1


def _mix(x: int) -> int:
    x = (x + 0x9E3779B97F4A7C15) & _MASK
    x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & _MASK
    x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & _MASK
    return x ^ (x >> 31)


def _to_signed(x: int) -> int:
    return x - (1 << 64) if x >= (1 << 63) else x


def _words(key: torch.Tensor) -> Tuple[int, int]:
    a, b = (int(v) & _MASK for v in key.tolist())
    return a, b


def _make(a: int, b: int) -> torch.Tensor:
    return torch.tensor([_to_signed(a & _MASK), _to_signed(b & _MASK)],
                        dtype=torch.int64)


def is_key(obj) -> bool:
    return (isinstance(obj, torch.Tensor) and obj.dtype == torch.int64
            and tuple(obj.shape) == (2,))


def init_randkey(randkey: Union[int, torch.Tensor]) -> torch.Tensor:
    """Return a key: build one from an ``int`` seed or validate a key."""
    if isinstance(randkey, bool):
        raise AssertionError("Invalid randkey: Must be int or key tensor")
    if isinstance(randkey, int):
        s = randkey & _MASK
        return _make(_mix(s), _mix(s ^ 0xD6E8FEB86659FD93))
    msg = f"Invalid {type(randkey)=}: Must be int or key tensor"
    assert is_key(randkey), msg
    return randkey.detach().cpu().clone()


def fold_in(key: torch.Tensor, data: int) -> torch.Tensor:
    a, b = _words(key)
    d = int(data) & _MASK
    return _make(_mix(a ^ _mix(d)), _mix(b + 0x632BE59BD9B4E019 + d))


def split(key: torch.Tensor, num: int = 2) -> Sequence[torch.Tensor]:
    """Derive ``num`` independent keys (``jax.random.split`` analogue)."""
    return [fold_in(key, i + 1) for i in range(num)]


def gen_new_key(key: torch.Tensor) -> torch.Tensor:
    """``split(key, 1)[0]`` (reference ``adam.py:254-257``)."""
    return split(key, 1)[0]


def key_to_seed(key: torch.Tensor) -> int:
    a, b = _words(key)
    return _mix(a ^ ((b << 1) & _MASK)) & ((1 << 63) - 1)


def make_generator(key: torch.Tensor, device="cpu") -> torch.Generator:
    g = torch.Generator(device=device)
    g.manual_seed(key_to_seed(key))
    return g


def uniform(key: torch.Tensor, shape, device="cpu", dtype=torch.float32):
    return torch.rand(shape, generator=make_generator(key, device),
                      device=device, dtype=dtype)


def normal(key: torch.Tensor, shape, device="cpu", dtype=torch.float32):
    return torch.randn(shape, generator=make_generator(key, device),
                       device=device, dtype=dtype)
