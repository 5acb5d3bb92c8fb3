"""``BalanceMadronaTorch`` (reference: envs/balance_beam_env.py:21-40) on the B200 simulator."""
import torch

from .. import spaces
from ..pantheonrl_extension.vectorenv import MadronaEnv
from ..simulators import BalanceBeamSimulator, ExecMode

NUM_SPACES = 5
VALID_MOVES = [-2, -1, 1, 2]
BUFFER = 2
TIME = 3
SCALE = 0.2


class BalanceMadronaTorch(MadronaEnv):
    def __init__(self, num_envs, gpu_id, debug_compile=True, use_cpu=False, use_env_cpu=False):
        sim = BalanceBeamSimulator(exec_mode=ExecMode.CPU if use_cpu else ExecMode.CUDA, gpu_id=gpu_id,
                                   num_worlds=num_envs, debug_compile=debug_compile)
        device = torch.device("cpu") if use_env_cpu else None
        super().__init__(num_envs, gpu_id, sim, env_device=device)
        self.observation_space = spaces.MultiDiscrete([NUM_SPACES + 2 * BUFFER] * 2 * TIME + [TIME])
        self.action_space = spaces.Discrete(len(VALID_MOVES))
        self.share_observation_space = self.observation_space
