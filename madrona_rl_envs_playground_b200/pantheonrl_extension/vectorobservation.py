"""Batched observation record — same fields as the reference's pantheonrl_extension/vectorobservation.py:5-32."""
from dataclasses import dataclass
from typing import Optional

import torch


@dataclass(init=False)
class VectorObservation:
    """``active`` [N] bool, ``obs`` [N, ...], ``state`` [N, ...] (defaults to ``obs``), ``action_mask`` [N, A] or None."""
    active: torch.Tensor
    obs: torch.Tensor
    state: torch.Tensor
    action_mask: Optional[torch.Tensor] = None

    def __init__(self, active, obs, state=None, action_mask=None):
        self.active = active
        self.obs = obs
        self.state = obs if state is None else state
        self.action_mask = action_mask
