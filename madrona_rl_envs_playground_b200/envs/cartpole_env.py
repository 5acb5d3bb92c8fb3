"""``CartpoleMadronaTorch`` (reference: envs/cartpole_env.py:77-128): gym-VectorEnv-style ``step(actions[N]) ->
(obs[N,4], rew[N,1], done[N], infos)`` on the B200 simulator."""
from math import pi

import numpy as np
import torch

from .. import spaces
from ..simulators import CartpoleSimulator, ExecMode

X_THRESHOLD = 2.4
THETA_THRESHOLD_RADIANS = 12 * 2 * pi / 360


class CartpoleMadronaTorch:
    def __init__(self, num_envs, gpu_id, debug_compile=True, use_cpu=False, use_env_cpu=False):
        high = np.array([X_THRESHOLD * 2, np.finfo(np.float32).max, THETA_THRESHOLD_RADIANS * 2,
                         np.finfo(np.float32).max], dtype=np.float32)
        self.device = torch.device("cpu") if use_env_cpu else torch.device("cuda", gpu_id)
        self.num_envs = num_envs
        self.action_space = self.single_action_space = spaces.Discrete(2)
        self.observation_space = self.single_observation_space = spaces.Box(-high, high, dtype=np.float32)
        self.sim = CartpoleSimulator(exec_mode=ExecMode.CPU if use_cpu else ExecMode.CUDA, gpu_id=gpu_id,
                                     num_worlds=num_envs, debug_compile=debug_compile)
        self.static_dones = self.sim.reset_tensor().to_torch()
        self.static_actions = self.sim.action_tensor().to_torch()
        self.static_observations = self.sim.observation_tensor().to_torch()
        self.static_rewards = self.sim.reward_tensor().to_torch()
        self.infos = [{}] * self.num_envs

    def to_torch(self, a):
        return a.to(self.device)

    stream_ordered_step = True   # see OvercookedMadrona.stream_ordered_step

    def step(self, actions):
        self.static_actions.copy_(actions[:, None].to(self.static_actions.device))
        if self.stream_ordered_step:
            self.sim.step_stream_ordered()
        else:
            self.sim.step()
        return (self.to_torch(self.static_observations), self.to_torch(self.static_rewards),
                self.to_torch(self.static_dones[:, 0]), self.infos)

    def reset(self):
        return self.to_torch(self.static_observations)

    def close(self, **kwargs):
        pass
