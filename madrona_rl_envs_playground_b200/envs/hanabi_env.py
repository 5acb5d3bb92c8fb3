"""``HanabiMadrona`` (reference: envs/hanabi_env.py:72-104) on the B200 simulator.

The reference instantiates ``hanabi_learning_environment.rl_env.HanabiEnv`` only to read the observation length and
the move count; those are closed-form in the config and computed here (the package is not needed)."""
import torch

from .. import spaces
from ..pantheonrl_extension.vectorenv import MadronaEnv
from ..simulators import ExecMode, HanabiSimulator

DEFAULT_N = 2
FULL_CONFIG = {"colors": 5, "ranks": 5, "players": DEFAULT_N, "max_information_tokens": 8, "max_life_tokens": 3,
               "observation_type": 1}
SMALL_CONFIG = {"colors": 2, "ranks": 5, "players": DEFAULT_N, "hand_size": 2, "max_information_tokens": 3,
                "max_life_tokens": 1, "observation_type": 1}
VERY_SMALL_CONFIG = {"colors": 1, "ranks": 5, "players": DEFAULT_N, "hand_size": 5, "max_information_tokens": 3,
                     "max_life_tokens": 1, "observation_type": 1}
DEFAULT_CONFIG = VERY_SMALL_CONFIG
config_choice = {"very_small": VERY_SMALL_CONFIG, "small": SMALL_CONFIG, "full": FULL_CONFIG}


def vectorized_observation_length(config, hand_size=5):
    """Length of the canonical observation encoding the simulator produces (src/hanabi_env/sim.hpp:24-30 with the
    runtime colors / ranks / token counts; hand size is 5 for < 4 players, src/hanabi_env/sim.cpp:875)."""
    c, r, p = config["colors"], config["ranks"], config["players"]
    bits = c * r
    deck = (4 + (r - 2) * 2) * c
    hands = (p - 1) * hand_size * bits + p
    board = deck - p * hand_size + bits + config["max_information_tokens"] + config["max_life_tokens"]
    discards = deck
    last_action = p + 4 + p + c + r + hand_size + hand_size + bits + 2
    knowledge = p * hand_size * (bits + c + r)
    return hands + board + discards + last_action + knowledge


def max_moves(config, hand_size=5):
    return 2 * hand_size + (config["players"] - 1) * (config["colors"] + config["ranks"])


class HanabiMadrona(MadronaEnv):
    def __init__(self, num_envs, gpu_id, debug_compile=True, config=None, use_cpu=False, use_env_cpu=False, obs_dtype=None):
        """The reference's arguments (envs/hanabi_env.py:74) plus ``obs_dtype`` (extension): ``torch.bfloat16`` or
        ``torch.float32`` observation / state rows, written by the kernels (what the agents' ``obs.float()`` computes)."""
        self.config = config if config is not None else DEFAULT_CONFIG
        cfg = self.config
        sim = HanabiSimulator(exec_mode=ExecMode.CPU if use_cpu else ExecMode.CUDA, gpu_id=gpu_id, num_worlds=num_envs,
                              colors=cfg["colors"], ranks=cfg["ranks"], players=cfg["players"],
                              max_information_tokens=cfg["max_information_tokens"],
                              max_life_tokens=cfg["max_life_tokens"], debug_compile=debug_compile, obs_dtype=obs_dtype)
        obs_size = vectorized_observation_length(cfg)
        state_size = obs_size + cfg["ranks"] * cfg["colors"] * 5
        n_moves = max_moves(cfg)
        self.observation_space = spaces.MultiBinary(obs_size)
        self.action_space = spaces.Discrete(n_moves)
        self.share_observation_space = spaces.MultiBinary(state_size)
        device = torch.device("cpu") if use_env_cpu else None
        super().__init__(num_envs=num_envs, gpu_id=gpu_id, sim=sim, debug_compile=debug_compile, obs_size=obs_size,
                         state_size=state_size, discrete_action_size=n_moves, env_device=device)
