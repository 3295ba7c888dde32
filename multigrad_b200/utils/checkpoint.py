"""Checkpoint / resume of optimiser runs (absent from the
reference, where results only live in memory).

``save_adam_checkpoint`` stores the trajectory so far, the loss history and
the Adam moments; ``load_adam_checkpoint`` gives back the arguments to resume
with ``run_adam(guess=traj[-1], init_state=state, ...)``.  Every rank holds
identical values, so only rank 0 needs to write.
"""
from __future__ import annotations

import os
from typing import Optional

import numpy as np
import torch

__all__ = ["save_adam_checkpoint", "load_adam_checkpoint"]


def save_adam_checkpoint(path: str, trajectory: torch.Tensor, state,
                         losses: Optional[torch.Tensor] = None) -> None:
    tmp = f"{path}.tmp{os.getpid()}.npz"
    np.savez(tmp, trajectory=trajectory.detach().cpu().numpy(),
             step=np.int64(state.step), m=state.m.detach().cpu().numpy(),
             v=state.v.detach().cpu().numpy(),
             losses=(np.zeros(0, np.float32) if losses is None
                     else losses.detach().cpu().numpy()))
    os.replace(tmp, path)


def load_adam_checkpoint(path: str, device="cpu"):
    """Returns ``(trajectory, AdamState, losses)``."""
    from ..adam import AdamState      # (adam imports utils.rng: no cycle here)
    z = np.load(path)
    state = AdamState(int(z["step"]), torch.as_tensor(z["m"], device=device),
                      torch.as_tensor(z["v"], device=device))
    return (torch.as_tensor(z["trajectory"], device=device), state,
            torch.as_tensor(z["losses"], device=device))
