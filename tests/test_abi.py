"""CPU tests of the drop-in boundary: the C-ABI library builds, loads and exports every symbol include/b200envs.h
declares; argument validation that happens before any CUDA call; host-side API surface of the Python mirror."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from madrona_rl_envs_playground_b200 import _lib, build
    build.build()
    return _lib.load()


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200envs.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200env_[a-z_0-9]+)\s*\(", text)))


def test_header_symbols_exported(lib):
    from madrona_rl_envs_playground_b200 import _lib
    names = declared_symbols()
    assert len(names) >= 19
    for name in names:
        assert hasattr(lib, name), f"{name} declared in include/b200envs.h but not exported"
    assert set(names) == set(_lib.EXPORTED_SYMBOLS), "ctypes prototypes out of sync with the header"


def test_version_and_error_string(lib):
    assert b"sm_100a" in lib.b200env_version()
    assert isinstance(lib.b200env_last_error(), bytes)


def test_invalid_arguments_do_not_abort(lib):
    from madrona_rl_envs_playground_b200 import _lib
    out = ctypes.c_void_p()
    assert lib.b200env_overcooked_create(None, 0, 4, ctypes.byref(out)) == -1
    cfg = _lib.HanabiConfig(5, 5, 3, 8, 3)   # 3 players: unsupported like the reference's tensor shapes
    assert lib.b200env_hanabi_create(ctypes.byref(cfg), 0, 4, ctypes.byref(out)) == -1
    assert b"2-player" in lib.b200env_last_error()
    cfg = _lib.HanabiConfig(2, 2, 2, 8, 3)   # a deck of 8 cards cannot deal two hands of 5 (the reference draws from an empty deck)
    assert lib.b200env_hanabi_create(ctypes.byref(cfg), 0, 4, ctypes.byref(out)) == -1
    assert b"deck" in lib.b200env_last_error()
    assert lib.b200env_balance_create(0, 0, ctypes.byref(out)) == -1
    assert lib.b200env_step(None) == -1
    assert lib.b200env_state_bytes(None) == 0


def test_overcooked_layout_validation(lib):
    from madrona_rl_envs_playground_b200 import _lib
    i64 = lambda v: (ctypes.c_int64 * len(v))(*v)
    terrain = i64([2] * 300)
    z = i64([0] * 64)
    rec = i64([0] * 16)
    out = ctypes.c_void_p()
    cfg = _lib.OvercookedConfig(terrain, 20, 15, 2, z, z, 3, 0, 5, rec, rec, 400)   # 300 cells > 255
    assert lib.b200env_overcooked_create(ctypes.byref(cfg), 0, 4, ctypes.byref(out)) == -1
    assert b"255" in lib.b200env_last_error()
    cfg = _lib.OvercookedConfig(terrain, 4, 5, 65, z, z, 3, 0, 5, rec, rec, 400)    # too many players
    assert lib.b200env_overcooked_create(ctypes.byref(cfg), 0, 4, ctypes.byref(out)) == -1


def test_cpu_exec_mode_is_refused():
    from madrona_rl_envs_playground_b200 import ExecMode, _lib, simulators
    with pytest.raises(_lib.B200EnvError):
        simulators.CartpoleSimulator(ExecMode.CPU, 0, 8)


def test_python_surface():
    import madrona_rl_envs_playground_b200 as pkg
    from madrona_rl_envs_playground_b200 import envs, pantheonrl_extension as pe
    assert pkg.madrona.ExecMode.CUDA is pkg.ExecMode.CUDA
    assert issubclass(pkg.SimplecookedSimulator, pkg.OvercookedSimulator)
    from madrona_rl_envs_playground_b200.envs.overcooked2_env import OvercookedMadrona as Overcooked2Madrona
    assert issubclass(Overcooked2Madrona, envs.OvercookedMadrona) and Overcooked2Madrona.OBJECT_CHANNELS == 10
    for cls, methods in [(pkg.OvercookedSimulator, ["step", "done_tensor", "active_agent_tensor", "action_tensor",
                                                    "observation_tensor", "action_mask_tensor", "reward_tensor",
                                                    "world_id_tensor", "agent_id_tensor"]),
                         (pkg.HanabiSimulator, ["agent_state_tensor", "action_mask_tensor"]),
                         (pkg.BalanceBeamSimulator, ["agent_state_tensor"]),
                         (pkg.CartpoleSimulator, ["reset_tensor", "observation_tensor", "reward_tensor"])]:
        for m in methods:
            assert callable(getattr(cls, m))
    for cls in (envs.OvercookedMadrona, envs.HanabiMadrona, envs.BalanceMadronaTorch):
        assert issubclass(cls, pe.VectorMultiAgentEnv)
        for m in ("step", "reset", "n_step", "n_reset", "add_partner_agent"):
            assert callable(getattr(cls, m))
    from madrona_rl_envs_playground_b200.envs.hanabi_env import FULL_CONFIG, max_moves, vectorized_observation_length
    assert vectorized_observation_length(FULL_CONFIG) == 658 and max_moves(FULL_CONFIG) == 20


def test_vector_env_contract_on_cpu():
    """ego/partner plumbing of VectorMultiAgentEnv with a trivial torch env (no simulator involved)."""
    import torch
    from madrona_rl_envs_playground_b200.pantheonrl_extension import (RandomVectorAgent, VectorMultiAgentEnv,
                                                                       VectorObservation)

    class Counter(VectorMultiAgentEnv):
        def __init__(self):
            super().__init__(4, torch.device("cpu"), n_players=2)
            self.t = torch.zeros(4)

        def _obs_list(self):
            return [VectorObservation(torch.ones(4, dtype=torch.bool), self.t.clone()) for _ in range(2)]

        def n_step(self, actions):
            assert actions.shape == (2, 4, 1)
            self.t += actions.sum(0)[:, 0]
            return self._obs_list(), torch.stack([self.t, -self.t]), self.t > 5, [{}] * 4

        def n_reset(self):
            self.t.zero_()
            return self._obs_list()

    env = Counter()
    env.add_partner_agent(RandomVectorAgent(lambda: torch.ones(4, 1)))
    ob0 = env.reset()
    assert ob0.state is ob0.obs or torch.equal(ob0.state, ob0.obs)
    obs, rew, done, info = env.step(torch.full((4, 1), 2.0))
    assert torch.equal(obs.obs, torch.full((4,), 3.0)) and torch.equal(rew, torch.full((4,), 3.0))
    obs, rew, done, info = env.step(torch.full((4, 1), 2.0))
    assert done.all()
