"""TEST INFRASTRUCTURE — ctypes bindings for the two CPU checkers.  Not a product path.

* ``Oracle*``  : oracle/liboracle.so, the plain-C restatement (oracle/oracle_*.c).
* ``Ref*``     : oracle/_ref/libref_<env>.so, the reference's own Madrona CPU exec mode (oracle/ref_harness.cpp),
                 available only where oracle/Makefile's ``ref`` target was built (needs /root/reference at build time).

Both present the canonical numpy layouts documented in oracle/oracle.h.  The ``Ref*`` classes convert the
reference's world-major exported columns (external/madrona/src/core/state.cpp:511-554) into those layouts the same
way the reference's Python wrappers do with the ID tensors (envs/overcooked_env.py:92-113,
pantheonrl_extension/vectorenv.py:306-329).
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
REFERENCE_ROOT = os.environ.get("B200ENVS_REFERENCE", "/root/reference")


def build(ref=True, quiet=True):
    """Build liboracle.so and (when the reference tree is mounted) oracle/_ref/*.so."""
    targets = ["oracle"]
    if ref and os.path.isdir(os.path.join(REFERENCE_ROOT, "src")):
        targets.append("ref")
    subprocess.run(["make", "-s", "-C", HERE, "-j8", f"REF={REFERENCE_ROOT}"] + targets, check=True,
                   stdout=subprocess.DEVNULL if quiet else None)


def _ptr(a, typ):
    return None if a is None else a.ctypes.data_as(C.POINTER(typ))


def _load_oracle():
    path = os.path.join(HERE, "liboracle.so")
    if not os.path.exists(path):
        build(ref=False)
    return C.CDLL(path)


def ref_available(env="overcooked"):
    return os.path.exists(os.path.join(REF_DIR, f"libref_{env}.so"))


class OvercookedCfg(C.Structure):
    _fields_ = [("height", C.c_int32), ("width", C.c_int32), ("num_players", C.c_int32),
                ("terrain", C.c_uint8 * 256), ("start_x", C.c_uint8 * 64), ("start_y", C.c_uint8 * 64),
                ("placement_in_pot_rew", C.c_int32), ("dish_pickup_rew", C.c_int32),
                ("soup_pickup_rew", C.c_int32),
                ("recipe_values", C.c_uint8 * 16), ("recipe_times", C.c_uint8 * 16), ("horizon", C.c_int32)]

    @classmethod
    def from_params(cls, p):
        cfg = cls()
        cfg.height, cfg.width, cfg.num_players = p["height"], p["width"], p["num_players"]
        for i, t in enumerate(p["terrain"]):
            cfg.terrain[i] = t
        for i in range(p["num_players"]):
            cfg.start_x[i] = p["start_player_x"][i]
            cfg.start_y[i] = p["start_player_y"][i]
        cfg.placement_in_pot_rew = p["placement_in_pot_rew"]
        cfg.dish_pickup_rew = p["dish_pickup_rew"]
        cfg.soup_pickup_rew = p["soup_pickup_rew"]
        for i in range(16):
            cfg.recipe_values[i] = p["recipe_values"][i] & 0xFF
            cfg.recipe_times[i] = p["recipe_times"][i] & 0xFF
        cfg.horizon = p["horizon"]
        return cfg


# --------------------------------------------------------------------------------------------------------------
# plain-C oracle
# --------------------------------------------------------------------------------------------------------------
class OracleOvercooked:
    PREFIX = "oracle_overcooked"
    CHANNELS = 16

    def _fn(self, name):
        return getattr(self.lib, f"{self.PREFIX}_{name}")

    def __init__(self, params, num_worlds):
        self.lib = _load_oracle()
        self._fn("create").restype = C.c_void_p
        self.P, self.W, self.H = params["num_players"], params["width"], params["height"]
        self.Cn = 5 * self.P + self.CHANNELS
        self.N = num_worlds
        self.cfg = OvercookedCfg.from_params(params)
        self.h = C.c_void_p(self._fn("create")(C.byref(self.cfg), C.c_uint32(num_worlds)))
        self.obs = np.zeros((self.P, self.N, self.W, self.H, self.Cn), np.uint8)
        self.rewards = np.zeros((self.P, self.N), np.int32)
        self.done = np.zeros((self.N,), np.int32)
        self._fn("observe")(self.h, _ptr(self.obs, C.c_uint8))

    def step(self, actions, want_obs=True):
        a = np.ascontiguousarray(actions, np.int32).reshape(self.P, self.N)
        self._fn("step")(self.h, _ptr(a, C.c_int32),
                                        _ptr(self.obs, C.c_uint8) if want_obs else None,
                                        _ptr(self.rewards, C.c_int32), _ptr(self.done, C.c_int32))
        return self.obs, self.rewards, self.done

    def close(self):
        if self.h:
            self._fn("destroy")(self.h)
            self.h = None

    __del__ = close


class OracleHanabi:
    OBS, STATE, MOVES = 658, 783, 20

    def __init__(self, num_worlds, colors=5, ranks=5, players=2, max_information_tokens=8, max_life_tokens=3):
        self.lib = _load_oracle()
        self.lib.oracle_hanabi_create.restype = C.c_void_p
        self.N = num_worlds
        self.h = C.c_void_p(self.lib.oracle_hanabi_create(
            C.c_uint32(num_worlds), C.c_uint32(colors), C.c_uint32(ranks), C.c_uint32(players),
            C.c_uint32(max_information_tokens), C.c_uint32(max_life_tokens)))
        N = num_worlds
        self.obs = np.zeros((2, N, self.OBS), np.uint8)
        self.state = np.zeros((2, N, self.STATE), np.uint8)
        self.mask = np.zeros((2, N, self.MOVES), np.int32)
        self.active = np.zeros((2, N), np.int32)
        self.rewards = np.zeros((2, N), np.float32)
        self.done = np.zeros((N,), np.int32)
        self.read()

    def read(self):
        self.lib.oracle_hanabi_read(self.h, _ptr(self.obs, C.c_uint8), _ptr(self.state, C.c_uint8),
                                    _ptr(self.mask, C.c_int32), _ptr(self.active, C.c_int32),
                                    _ptr(self.rewards, C.c_float), _ptr(self.done, C.c_int32))

    def step(self, actions):
        a = np.ascontiguousarray(actions, np.int32).reshape(2, self.N)
        self.lib.oracle_hanabi_step(self.h, _ptr(a, C.c_int32))
        self.read()

    def max_info_encoded(self):
        return int(self.lib.oracle_hanabi_max_info_encoded(self.h))

    def close(self):
        if self.h:
            self.lib.oracle_hanabi_destroy(self.h)
            self.h = None

    __del__ = close


class OracleBalance:
    def __init__(self, num_worlds):
        self.lib = _load_oracle()
        self.lib.oracle_balance_create.restype = C.c_void_p
        self.N = num_worlds
        self.h = C.c_void_p(self.lib.oracle_balance_create(C.c_uint32(num_worlds)))
        self.obs = np.zeros((2, num_worlds, 7), np.int32)
        self.rewards = np.zeros((2, num_worlds), np.float32)
        self.done = np.zeros((num_worlds,), np.int32)
        self.read()

    def read(self):
        self.lib.oracle_balance_read(self.h, _ptr(self.obs, C.c_int32), _ptr(self.rewards, C.c_float),
                                     _ptr(self.done, C.c_int32))

    def step(self, actions):
        a = np.ascontiguousarray(actions, np.int32).reshape(2, self.N)
        self.lib.oracle_balance_step(self.h, _ptr(a, C.c_int32))
        self.read()

    def close(self):
        if self.h:
            self.lib.oracle_balance_destroy(self.h)
            self.h = None

    __del__ = close


class OracleCartpole:
    def __init__(self, num_worlds):
        self.lib = _load_oracle()
        self.lib.oracle_cartpole_create.restype = C.c_void_p
        self.N = num_worlds
        self.h = C.c_void_p(self.lib.oracle_cartpole_create(C.c_uint32(num_worlds)))
        self.state = np.zeros((num_worlds, 4), np.float32)
        self.rewards = np.zeros((num_worlds,), np.float32)
        self.done = np.zeros((num_worlds,), np.int32)
        self.read()

    def read(self):
        self.lib.oracle_cartpole_read(self.h, _ptr(self.state, C.c_float), _ptr(self.rewards, C.c_float),
                                      _ptr(self.done, C.c_int32))

    def write_state(self, state):
        s = np.ascontiguousarray(state, np.float32).reshape(self.N, 4)
        self.lib.oracle_cartpole_write_state(self.h, _ptr(s, C.c_float))

    def step(self, actions):
        a = np.ascontiguousarray(actions, np.int32).reshape(self.N)
        self.lib.oracle_cartpole_step(self.h, _ptr(a, C.c_int32))
        self.read()

    def close(self):
        if self.h:
            self.lib.oracle_cartpole_destroy(self.h)
            self.h = None

    __del__ = close


class OracleAcrobat:
    """src/acrobat_env restated (oracle_acrobat.c); same surface as OracleCartpole."""

    def __init__(self, num_worlds):
        self.lib = _load_oracle()
        self.lib.oracle_acrobat_create.restype = C.c_void_p
        self.lib.oracle_acrobat_episode_length.restype = C.c_uint32
        self.N = num_worlds
        self.h = C.c_void_p(self.lib.oracle_acrobat_create(C.c_uint32(num_worlds)))
        self.state = np.zeros((num_worlds, 4), np.float32)
        self.rewards = np.zeros((num_worlds,), np.float32)
        self.done = np.zeros((num_worlds,), np.int32)
        self.read()

    def read(self):
        self.lib.oracle_acrobat_read(self.h, _ptr(self.state, C.c_float), _ptr(self.rewards, C.c_float),
                                     _ptr(self.done, C.c_int32))

    def write_state(self, state):
        s = np.ascontiguousarray(state, np.float32).reshape(self.N, 4)
        self.lib.oracle_acrobat_write_state(self.h, _ptr(s, C.c_float))

    def episode_length(self):
        return int(self.lib.oracle_acrobat_episode_length(self.h))

    def step(self, actions):
        a = np.ascontiguousarray(actions, np.int32).reshape(self.N)
        self.lib.oracle_acrobat_step(self.h, _ptr(a, C.c_int32))
        self.read()

    def close(self):
        if self.h:
            self.lib.oracle_acrobat_destroy(self.h)
            self.h = None

    __del__ = close


class OracleOvercooked2(OracleOvercooked):
    """src/overcooked2_env ("Simplecooked") restatement: same interface, 5P+10 channels, its own terrain enum order."""
    PREFIX = "oracle_overcooked2"
    CHANNELS = 10


# --------------------------------------------------------------------------------------------------------------
# the reference's own Madrona CPU exec mode (oracle/_ref)
# --------------------------------------------------------------------------------------------------------------
def _view(ptr, shape, dtype):
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    buf = (C.c_uint8 * n).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype).reshape(shape)


class _RefBase:
    env = None

    def _load(self):
        path = os.path.join(REF_DIR, f"libref_{self.env}.so")
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} not built (make -C oracle ref; needs {REFERENCE_ROOT})")
        lib = C.CDLL(path)
        lib.ref_create.restype = C.c_void_p
        lib.ref_exported.restype = C.c_void_p
        lib.ref_exported.argtypes = [C.c_void_p, C.c_int]
        lib.ref_step.argtypes = [C.c_void_p]
        lib.ref_destroy.argtypes = [C.c_void_p]
        return lib

    def _exp(self, slot, shape, dtype):
        return _view(self.lib.ref_exported(self.h, slot), shape, dtype)

    def raw_step(self):
        self.lib.ref_step(self.h)

    def close(self):
        if getattr(self, "h", None):
            self.lib.ref_destroy(self.h)
            self.h = None

    __del__ = close


class RefOvercooked(_RefBase):
    """src/overcooked_env via Madrona CPU.  Export slots: src/overcooked_env/sim.hpp:23-38."""
    env = "overcooked"
    ROW = 5 * 64 + 16  # sizeof(LocationXObservation), sim.hpp:125-127
    CHANNELS = 16

    def __init__(self, params, num_worlds, num_workers=1):
        self.lib = self._load()
        P, W, H = params["num_players"], params["width"], params["height"]
        self.P, self.W, self.H, self.S, self.Cn, self.N = P, W, H, W * H, 5 * P + self.CHANNELS, num_worlds
        i64 = lambda v: (C.c_int64 * len(v))(*v)
        self.h = C.c_void_p(self.lib.ref_create(
            C.c_uint32(num_worlds), C.c_uint32(num_workers), i64(params["terrain"]), C.c_int64(H), C.c_int64(W),
            C.c_int64(P), i64(params["start_player_x"]), i64(params["start_player_y"]),
            C.c_int64(params["placement_in_pot_rew"]), C.c_int64(params["dish_pickup_rew"]),
            C.c_int64(params["soup_pickup_rew"]), i64(params["recipe_values"]), i64(params["recipe_times"]),
            C.c_int64(params["horizon"])))
        N, S = self.N, self.S
        self.x_done = self._exp(0, (N,), np.int32)
        self.x_active = self._exp(1, (N, P), np.int32)
        self.x_action = self._exp(2, (N, P), np.int32)
        self.x_obs = self._exp(3, (N, S, P, self.ROW), np.uint8)     # row = cell*P + player
        self.x_mask = self._exp(4, (N, P, 6), np.int32)
        self.x_reward = self._exp(5, (N, P), np.int32)
        self.x_world_id = self._exp(6, (N, P), np.int32)
        self.x_agent_id = self._exp(7, (N, P), np.int32)
        self.x_loc_world_id = self._exp(8, (N, S * P), np.int32)
        self.x_loc_id = self._exp(9, (N, S * P), np.int32)

    @property
    def obs(self):
        """canonical [P][N][W][H][C]"""
        o = self.x_obs[..., :self.Cn].reshape(self.N, self.H, self.W, self.P, self.Cn)   # n, y, x, p, c
        return np.ascontiguousarray(o.transpose(3, 0, 2, 1, 4))

    def step(self, actions):
        self.x_action[...] = np.asarray(actions, np.int32).reshape(self.P, self.N).T
        self.raw_step()
        return self.obs, np.ascontiguousarray(self.x_reward.T), self.x_done.copy()


class RefOvercooked2(RefOvercooked):
    """src/overcooked2_env ("Simplecooked") via Madrona CPU: same export slots (src/overcooked2_env/sim.hpp:23-37)."""
    env = "overcooked2"
    ROW = 5 * 2 + 10   # sizeof(LocationXObservation) with MAX_NUM_PLAYERS 2, sim.hpp:13-14,128-130
    CHANNELS = 10


class RefHanabi(_RefBase):
    env = "hanabi"

    def __init__(self, num_worlds, colors=5, ranks=5, players=2, max_information_tokens=8, max_life_tokens=3,
                 num_workers=1):
        self.lib = self._load()
        self.N = N = num_worlds
        self.h = C.c_void_p(self.lib.ref_create(
            C.c_uint32(N), C.c_uint32(num_workers), C.c_uint32(colors), C.c_uint32(ranks), C.c_uint32(players),
            C.c_uint32(max_information_tokens), C.c_uint32(max_life_tokens)))
        self.x_done = self._exp(0, (N,), np.int32)
        self.x_active = self._exp(1, (N, 2), np.int32)
        self.x_action = self._exp(2, (N, 2), np.int32)
        self.x_obs = self._exp(3, (N, 2, 658), np.uint8)
        self.x_mask = self._exp(4, (N, 2, 20), np.int32)
        self.x_reward = self._exp(5, (N, 2), np.float32)
        self.x_state = self._exp(8, (N, 2, 783), np.uint8)

    def read(self):
        t = lambda a: np.ascontiguousarray(np.swapaxes(a, 0, 1))
        self.obs, self.state, self.mask = t(self.x_obs), t(self.x_state), t(self.x_mask)
        self.active, self.rewards, self.done = t(self.x_active), t(self.x_reward), self.x_done.copy()

    def step(self, actions):
        self.x_action[...] = np.asarray(actions, np.int32).reshape(2, self.N).T
        self.raw_step()
        self.read()


class RefBalance(_RefBase):
    env = "balance"

    def __init__(self, num_worlds, num_workers=1):
        self.lib = self._load()
        self.N = N = num_worlds
        self.h = C.c_void_p(self.lib.ref_create(C.c_uint32(N), C.c_uint32(num_workers)))
        self.x_done = self._exp(0, (N,), np.int32)
        self.x_action = self._exp(2, (N, 2), np.int32)
        self.x_obs = self._exp(3, (N, 2, 7), np.int32)
        self.x_reward = self._exp(5, (N, 2), np.float32)

    def read(self):
        t = lambda a: np.ascontiguousarray(np.swapaxes(a, 0, 1))
        self.obs, self.rewards, self.done = t(self.x_obs), t(self.x_reward), self.x_done.copy()

    def step(self, actions):
        self.x_action[...] = np.asarray(actions, np.int32).reshape(2, self.N).T
        self.raw_step()
        self.read()


class RefCartpole(_RefBase):
    env = "cartpole"

    def __init__(self, num_worlds, num_workers=1):
        self.lib = self._load()
        self.N = N = num_worlds
        self.h = C.c_void_p(self.lib.ref_create(C.c_uint32(N), C.c_uint32(num_workers)))
        self.x_done = self._exp(0, (N,), np.int32)
        self.x_action = self._exp(1, (N,), np.int32)
        self.x_state = self._exp(2, (N, 4), np.float32)
        self.x_reward = self._exp(3, (N,), np.float32)

    def read(self):
        self.state, self.rewards, self.done = self.x_state.copy(), self.x_reward.copy(), self.x_done.copy()

    def write_state(self, state):
        # exported columns are re-imported at the start of every CPU step (core/state.cpp:484-509)
        self.x_state[...] = np.asarray(state, np.float32).reshape(self.N, 4)

    def step(self, actions):
        self.x_action[...] = np.asarray(actions, np.int32).reshape(self.N)
        self.raw_step()
        self.read()


class RefAcrobat(RefCartpole):
    """src/acrobat_env: same export order as Cartpole (WorldReset, Action, State, Reward; sim.hpp:16-23)."""
    env = "acrobat"
