"""``OvercookedMadrona`` of the reference's envs/overcooked2_env.py:19-123 — the JS-compatible Overcooked ("Simplecooked",
src/overcooked2_env) that the reference's trainer (train/env_utils.py:3) and notebook import — on the B200 simulator.
Same constructor (default horizon 200), attributes and ``n_step`` / ``n_reset`` behaviour; layouts are the
oldercooked_ai ones (``simple``, ``unident_s``, ``random0``, ``random1``, ``random3``, ``simple_tomato``).

As for Overcooked, the kernel writes the per-player ``(N, W, H, 5P+10)`` array that the reference's ``get_obs``
(:91-103) assembles with a scatter by location id, a reshape and a transpose, so ``get_obs`` does no indexing here.
"""
from ..overcooked2_layouts import get_base_layout_params
from ..simulators import SimplecookedSimulator
from .overcooked_env import OvercookedMadrona as _OvercookedMadrona


class OvercookedMadrona(_OvercookedMadrona):
    SIMULATOR = SimplecookedSimulator
    LAYOUT_PARAMS = staticmethod(get_base_layout_params)
    OBJECT_CHANNELS = 10

    def __init__(self, layout_name, num_envs, gpu_id, debug_compile=True, use_cpu=False, use_env_cpu=False,
                 ego_agent_idx=0, horizon=200, num_players=None):
        super().__init__(layout_name, num_envs, gpu_id, debug_compile, use_cpu, use_env_cpu, ego_agent_idx, horizon,
                         num_players)
