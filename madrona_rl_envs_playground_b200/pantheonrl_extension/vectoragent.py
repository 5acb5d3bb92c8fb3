"""Partner-agent interface of the vector env contract (reference: pantheonrl_extension/vectoragent.py:9-42).
Only the interface and the random agent live here; the PPO learners are outside the env step path."""
from abc import ABC, abstractmethod

import torch

from .vectorobservation import VectorObservation


class VectorAgent(ABC):
    @abstractmethod
    def get_action(self, obs: VectorObservation, record: bool = True) -> torch.Tensor:
        """Action tensor [N, 1] for the batched observation."""

    @abstractmethod
    def update(self, rewards: torch.Tensor, dones: torch.Tensor) -> None:
        """Reward / done feedback for the most recent recorded action."""


class RandomVectorAgent(VectorAgent):
    def __init__(self, sampler):
        self.sampler = sampler

    def get_action(self, obs, record=True):
        return self.sampler()

    def update(self, rewards, dones):
        return None
