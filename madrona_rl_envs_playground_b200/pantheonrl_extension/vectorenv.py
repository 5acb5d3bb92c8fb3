"""VectorMultiAgentEnv contract and the generic simulator-backed env.

Interface-compatible with the reference's pantheonrl_extension/vectorenv.py: ``VectorMultiAgentEnv`` (:26-252 —
ego/partner bookkeeping, ``step`` / ``reset`` on top of abstract ``n_step`` / ``n_reset``) and ``MadronaEnv``
(:262-346).  ``MadronaEnv`` here does not gather / scatter through ID tensors: the library's tensors are already
agent-major ``[P, N, ...]``, so player i's view is simply ``tensor[i]`` — zero-copy on the simulator's device.
"""
from abc import ABC, abstractmethod
from typing import List, Optional

import numpy as np
import torch

from .vectoragent import VectorAgent
from .vectorobservation import VectorObservation


class PlayerException(Exception):
    """Raised when the players of an environment are set up inconsistently."""


class VectorMultiAgentEnv(ABC):
    def __init__(self, num_envs: int, device: torch.device, ego_ind: int = 0, n_players: int = 2,
                 resample_policy: str = "default", partners: Optional[List[List[VectorAgent]]] = None):
        self.num_envs = num_envs
        self.device = device
        self.ego_ind = ego_ind
        self.n_players = n_players
        if partners is not None:
            if len(partners) != n_players - 1:
                raise PlayerException("The number of partners needs to equal the number of non-ego players")
            for plist in partners:
                if not isinstance(plist, list) or not plist:
                    raise PlayerException("Sublist for each partner must be nonempty list")
        self.partners = partners or [[] for _ in range(n_players - 1)]
        self.partnerids = [0] * (n_players - 1)
        self._obs = tuple()
        self._actions = None
        self.set_resample_policy(resample_policy)

    # ---- partner bookkeeping ----------------------------------------------------------------------------------
    def getDummyEnv(self, player_num: int):
        return self

    def _get_partner_num(self, player_num: int) -> int:
        if player_num == self.ego_ind:
            raise PlayerException("Ego agent is not set by the environment")
        return player_num - 1 if player_num > self.ego_ind else player_num

    def add_partner_agent(self, agent: VectorAgent, player_num: int = 1) -> None:
        self.partners[self._get_partner_num(player_num)].append(agent)

    def set_partnerid(self, agent_id: int, player_num: int = 1) -> None:
        partner_num = self._get_partner_num(player_num)
        assert 0 <= agent_id < len(self.partners[partner_num])
        self.partnerids[partner_num] = agent_id

    def resample_random(self) -> None:
        self.partnerids = [np.random.randint(len(plist)) for plist in self.partners]

    def resample_round_robin(self) -> None:
        self.partnerids = [(self.partnerids[0] + 1) % len(self.partners[0])]

    def set_resample_policy(self, resample_policy: str) -> None:
        if resample_policy == "default":
            resample_policy = "robin" if self.n_players == 2 else "random"
        if resample_policy == "robin" and self.n_players != 2:
            raise PlayerException("Cannot do round robin resampling for >2 players")
        if resample_policy == "robin":
            self.resample_partner = self.resample_round_robin
        elif resample_policy == "random":
            self.resample_partner = self.resample_random
        else:
            raise PlayerException(f"Invalid resampling policy: {resample_policy}")

    # ---- ego-centric stepping -----------------------------------------------------------------------------------
    def _get_actions(self, obs, ego_act=None):
        actions = []
        for player, ob in enumerate(obs):
            if player == self.ego_ind:
                actions.append(ego_act)
            else:
                p = self._get_partner_num(player)
                actions.append(self.partners[p][self.partnerids[p]].get_action(ob))
        if self._actions is None:
            self._actions = torch.stack(actions)
        else:
            torch.stack(actions, out=self._actions)
        return self._actions

    def _update_players(self, rews, done):
        for i in range(self.n_players - 1):
            playernum = i + (0 if i < self.ego_ind else 1)
            self.partners[i][self.partnerids[i]].update(rews[playernum], done)

    def step(self, action: torch.Tensor):
        acts = self._get_actions(self._obs, action)
        self._obs, rews, done, info = self.n_step(acts)
        self._update_players(rews, done)
        return self._obs[self.ego_ind], rews[self.ego_ind], done, info

    def reset(self):
        self.resample_partner()
        self._obs = self.n_reset()
        return self._obs[self.ego_ind]

    @abstractmethod
    def n_step(self, actions: torch.Tensor):
        """actions [P, N, 1] -> (list of P VectorObservation, rewards [P, N], dones [N], infos)"""

    @abstractmethod
    def n_reset(self):
        """-> list of P VectorObservation of the current (already reset) worlds"""

    def close(self, **kwargs):
        pass


class MadronaEnv(VectorMultiAgentEnv):
    """Generic env over a simulator exporting done / active / action / observation / agent_state / action_mask /
    reward tensors (Hanabi, Balance Beam)."""

    def __init__(self, num_envs, gpu_id, sim, debug_compile=True, obs_size=None, state_size=None,
                 discrete_action_size=None, env_device=None):
        self.sim = sim
        self.static_dones = sim.done_tensor().to_torch()
        self.static_active_agents = sim.active_agent_tensor().to_torch()
        self.static_actions = sim.action_tensor().to_torch()
        self.static_observations = sim.observation_tensor().to_torch()
        self.static_agent_states = sim.agent_state_tensor().to_torch()
        self.static_action_masks = sim.action_mask_tensor().to_torch()
        self.static_rewards = sim.reward_tensor().to_torch()
        self.static_worldID = sim.world_id_tensor().to_torch().to(torch.long)
        self.static_agentID = sim.agent_id_tensor().to_torch().to(torch.long)

        self.obs_size = self.static_observations.shape[2] if obs_size is None else obs_size
        self.state_size = self.static_agent_states.shape[2] if state_size is None else state_size
        self.discrete_action_size = (self.static_action_masks.shape[2] if discrete_action_size is None
                                     else discrete_action_size)
        if env_device is None:
            env_device = self.static_observations.device
        super().__init__(num_envs, device=env_device, n_players=self.static_observations.shape[0])
        self.infos = [{}] * self.num_envs

    def to_torch(self, a):
        return a.to(self.device)

    def _observations(self):
        return [VectorObservation(self.to_torch(self.static_active_agents[i].to(torch.bool)),
                                  self.to_torch(self.static_observations[i, :, :self.obs_size]),
                                  self.to_torch(self.static_agent_states[i, :, :self.state_size]),
                                  self.to_torch(self.static_action_masks[i, :, :self.discrete_action_size].to(torch.bool)))
                for i in range(self.n_players)]

    # True: sim.step() is enqueued on torch's current stream and the host is not blocked (results are stream-ordered
    # like any torch op).  False: the reference's host-synchronous step.
    stream_ordered_step = True

    def n_step(self, actions):
        self.static_actions.copy_(actions.to(self.static_actions.device))
        if self.stream_ordered_step and hasattr(self.sim, "step_stream_ordered"):
            self.sim.step_stream_ordered()
        else:
            self.sim.step()
        return (self._observations(), self.to_torch(self.static_rewards), self.to_torch(self.static_dones),
                self.infos)

    def n_reset(self):
        return self._observations()
