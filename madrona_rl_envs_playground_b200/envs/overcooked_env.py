"""``OvercookedMadrona`` — the reference's vectorised Overcooked env class (envs/overcooked_env.py:18-119) on top of
the B200 simulator.  Same constructor, attributes and ``n_step`` / ``n_reset`` / ``step`` / ``reset`` behaviour.

The reference's ``get_obs`` (:92-102) scatters padded 336-byte rows by location id, then reshapes and transposes to a
per-player ``(N, W, H, 5P+16)`` view; the CUDA kernel writes that array directly, so here player i's observation is
``static_observations[i]`` with no indexing kernels at all.
"""
import torch

from .. import spaces
from ..overcooked_layouts import get_base_layout_params
from ..pantheonrl_extension.vectorenv import VectorMultiAgentEnv
from ..pantheonrl_extension.vectorobservation import VectorObservation
from ..simulators import ExecMode, OvercookedSimulator

NUM_ACTIONS = 6   # overcooked_ai_py Action.NUM_ACTIONS: N, S, E, W, stay, interact


class OvercookedMadrona(VectorMultiAgentEnv):
    # what envs/overcooked2_env.py changes: the simulator class, the layout parser, the channel count
    SIMULATOR = OvercookedSimulator
    LAYOUT_PARAMS = staticmethod(get_base_layout_params)
    OBJECT_CHANNELS = 16

    def __init__(self, layout_name, num_envs, gpu_id, debug_compile=True, use_cpu=False, use_env_cpu=False,
                 ego_agent_idx=0, horizon=400, num_players=None, obs_dtype=None):
        """The reference's arguments (envs/overcooked_env.py:20) plus ``obs_dtype`` (extension): ``torch.bfloat16`` or
        ``torch.float32`` make every observation the policy-ready tensor — what the agents' ``obs.float()``
        (pantheonrl_extension/vectoragent.py:353) would compute — written by the step kernel, with no int8 tensor
        and no cast pass in between."""
        self.layout_name = layout_name
        self.base_layout_params = self.LAYOUT_PARAMS(layout_name, horizon, max_num_players=num_players)
        self.width = self.base_layout_params["width"]
        self.height = self.base_layout_params["height"]
        self.num_players = self.base_layout_params["num_players"]
        self.size = self.width * self.height
        self.horizon = horizon

        self.sim = self.SIMULATOR(exec_mode=ExecMode.CPU if use_cpu else ExecMode.CUDA, gpu_id=gpu_id,
                                  num_worlds=num_envs, debug_compile=debug_compile, obs_dtype=obs_dtype,
                                  **self.base_layout_params)

        self.static_dones = self.sim.done_tensor().to_torch()
        self.static_active_agents = self.sim.active_agent_tensor().to_torch().to(torch.bool)
        self.static_actions = self.sim.action_tensor().to_torch()
        self.static_observations = self.sim.observation_tensor().to_torch()    # [P, N, W, H, C] int8 (or obs_dtype)
        self.static_rewards = self.sim.reward_tensor().to_torch()
        self.static_worldID = self.sim.world_id_tensor().to_torch().to(torch.long)
        self.static_agentID = self.sim.agent_id_tensor().to_torch().to(torch.long)

        self.obs_size = self.size * (5 * self.num_players + self.OBJECT_CHANNELS)
        self.state_size = self.obs_size

        env_device = torch.device("cpu") if use_env_cpu else torch.device("cuda", gpu_id)
        super().__init__(num_envs, device=env_device, n_players=self.num_players)
        self.infos = [{}] * self.num_envs
        self.ego_ind = ego_agent_idx
        self._obs_views = None

        self.observation_space = self._setup_observation_space()
        self.share_observation_space = self.observation_space
        self.action_space = spaces.Discrete(NUM_ACTIONS)
        self.n_reset()

    def _setup_observation_space(self):
        return spaces.MultiBinary([self.width, self.height, 5 * self.num_players + self.OBJECT_CHANNELS])

    def to_torch(self, a):
        return a.to(self.device)

    # True: sim.step() is enqueued on torch's current stream and the host is not blocked (results are stream-ordered
    # like any torch op).  False: the reference's host-synchronous step.
    stream_ordered_step = True

    def get_obs(self):
        # the simulator rewrites its tensors in place, so on the simulator's own device the per-player views never change
        if self._obs_views is None or self.device != self.static_observations.device:
            views = [VectorObservation(self.to_torch(self.static_active_agents[i]),
                                       self.to_torch(self.static_observations[i])) for i in range(self.n_players)]
            if self.device != self.static_observations.device:
                return views
            self._obs_views = views
        return list(self._obs_views)

    def n_step(self, actions):
        self.static_actions.copy_(actions.to(self.static_actions.device))
        if self.stream_ordered_step:
            self.sim.step_stream_ordered()
        else:
            self.sim.step()
        return self.get_obs(), self.to_torch(self.static_rewards), self.to_torch(self.static_dones), self.infos

    def n_reset(self):
        return self.get_obs()

    def close(self, **kwargs):
        pass
