"""CPU check of the PRODUCT's Hanabi game logic: csrc/hanabi_game.cuh — the register-record code the sm_100a kernels
inline — compiled for the host (tests/host/hanabi_host.cpp, g++) and stepped next to the oracle and the golden
vectors.  Bit-exact on every exported tensor; no GPU involved."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from helpers import hanabi_informed_actions, legal_random_actions

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "host", "hanabi_host.cpp")
INC = os.path.join(ROOT, "madrona_rl_envs_playground_b200", "csrc")


@pytest.fixture(scope="module")
def hh(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("hanabi_host") / "libhanabi_host.so")
    subprocess.run(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-Wall", "-I", INC, SRC, "-o", out], check=True)
    lib = C.CDLL(out)
    lib.hh_create.restype = C.c_void_p
    return lib


class HostHanabi:
    def __init__(self, lib, N, colors=5, ranks=5, players=2, max_information_tokens=8, max_life_tokens=3):
        assert players == 2
        self.lib, self.N = lib, N
        self.h = C.c_void_p(lib.hh_create(C.c_uint32(N), colors, ranks, max_information_tokens, max_life_tokens))
        self.obs = np.zeros((2, N, 658), np.uint8)
        self.state = np.zeros((2, N, 783), np.uint8)
        self.mask = np.zeros((2, N, 20), np.int32)
        self.active = np.zeros((2, N), np.int32)
        self.rewards = np.zeros((2, N), np.float32)
        self.done = np.zeros((N,), np.int32)
        self.read()

    def read(self):
        p = lambda a: a.ctypes.data_as(C.c_void_p)   # noqa: E731
        self.lib.hh_read(self.h, p(self.obs), p(self.state), p(self.mask), p(self.active), p(self.rewards), p(self.done))

    def step(self, actions):
        a = np.ascontiguousarray(actions, np.int32).reshape(2, self.N)
        self.lib.hh_step(self.h, a.ctypes.data_as(C.c_void_p))
        self.read()

    def __del__(self):
        self.lib.hh_destroy(self.h)


FIELDS = ("obs", "state", "mask", "active", "rewards", "done")


def test_host_logic_vs_golden(hh):
    g = np.load(os.path.join(ROOT, "tests", "golden", "hanabi_full.npz"))
    N = g["actions"].shape[2]
    sim = HostHanabi(hh, N)
    assert np.array_equal(sim.obs, g["obs0"]) and np.array_equal(sim.state, g["state0"])
    assert np.array_equal(sim.mask, g["mask0"]) and np.array_equal(sim.active, g["active0"])
    for t in range(g["actions"].shape[0]):
        sim.step(g["actions"][t])
        for k in FIELDS:
            assert np.array_equal(getattr(sim, k), g[k][t]), (k, t)


@pytest.mark.parametrize("N,steps", [(256, 400), (33, 150)])
def test_host_logic_vs_oracle(hh, N, steps):
    from oracle.bindings import OracleHanabi
    sim, orc = HostHanabi(hh, N), OracleHanabi(N)
    rng = np.random.default_rng(N)
    resets = 0
    for t in range(steps):
        for k in FIELDS:
            assert np.array_equal(getattr(sim, k), getattr(orc, k)), (k, t)
        a = legal_random_actions(rng, orc.mask)
        sim.step(a)
        orc.step(a)
        resets += int(orc.done.sum())
    assert resets > N


@pytest.mark.parametrize("cfg", [
    dict(colors=5, ranks=2),                                                    # tokens run past their maximum
    dict(colors=3, ranks=3, max_information_tokens=4, max_life_tokens=2),
    dict(colors=4, ranks=5, max_information_tokens=3, max_life_tokens=1),
    dict(colors=2, ranks=4, max_information_tokens=2, max_life_tokens=3),
])
def test_host_logic_other_rules_vs_oracle(hh, cfg):
    from oracle.bindings import OracleHanabi
    N = 128
    sim, orc = HostHanabi(hh, N, **cfg), OracleHanabi(N, **cfg)
    rng = np.random.default_rng(11)
    c, r = cfg.get("colors", 5), cfg.get("ranks", 5)
    bits = c * r
    # nominal encoded lengths (sim.cpp:54-365); the reference never writes the bytes past them
    obs_len = (5 * bits + 2) + ((4 + (r - 2) * 2) * c - 10) + bits + cfg.get("max_information_tokens", 8) \
        + cfg.get("max_life_tokens", 3) + c * (3 + 2 * (r - 2) + 1) + (2 + 4 + 2 + c + r + 5 + 5 + bits + 2) \
        + 2 * 5 * (bits + c + r)
    state_len = obs_len + 5 * bits
    for t in range(300):
        assert np.array_equal(sim.obs[..., :obs_len], orc.obs[..., :obs_len]), (cfg, "obs", t)
        assert np.array_equal(sim.state[..., :state_len], orc.state[..., :state_len]), (cfg, "state", t)
        for k in ("mask", "active", "rewards", "done"):
            assert np.array_equal(getattr(sim, k), getattr(orc, k)), (cfg, k, t)
        a = legal_random_actions(rng, orc.mask)
        sim.step(a)
        orc.step(a)
    if cfg.get("ranks") == 2:
        assert orc.max_info_encoded() > 8


def test_host_logic_token_overflow(hh):
    """Stacks completed at 8 information tokens: rows of 659 / 784 bytes whose excess lands on the next player's row."""
    from oracle.bindings import OracleHanabi
    N = 96
    sim, orc = HostHanabi(hh, N), OracleHanabi(N)
    rng = np.random.default_rng(5)
    for t in range(260):
        a = hanabi_informed_actions(rng, orc.state, orc.mask, orc.active)
        sim.step(a)
        orc.step(a)
        for k in FIELDS:
            assert np.array_equal(getattr(sim, k), getattr(orc, k)), (k, t)
    assert orc.max_info_encoded() > 8


def test_host_logic_survives_illegal_actions_under_sanitizers(tmp_path):
    """Arbitrary action values (-8 .. 39; the reference's behaviour is undefined for moves outside the legal mask) through
    the product's game logic built with AddressSanitizer + UBSan: no out-of-bounds access (deck bytes, the 27-word bit
    row), no undefined shift, every counter within its range — 4 rule sets x 64 worlds x 3 000 steps."""
    exe = str(tmp_path / "hh_fuzz")
    subprocess.run(["g++", "-O1", "-g", "-std=c++17", "-fsanitize=address,undefined", "-fno-sanitize-recover=all",
                    "-DHH_FUZZ_MAIN", "-I", INC, SRC, "-o", exe], check=True)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stdout[-2000:] + r.stderr[-4000:]
