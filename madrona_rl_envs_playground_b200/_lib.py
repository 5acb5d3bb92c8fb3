"""ctypes binding of the C-ABI library (include/b200envs.h).

The CUDA library is the product: importing a simulator without it raises — there is no CPU or PyTorch fallback.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200ENVS_LIB") or os.path.join(HERE, "libb200envs.so")   # override: kernel experiments

OK = 0
OVERCOOKED, HANABI, BALANCE, CARTPOLE, ACROBAT = 0, 1, 2, 3, 4
I8, I32, F32, BF16 = 0, 1, 2, 3
(T_DONE, T_ACTIVE_AGENT, T_ACTION, T_OBSERVATION, T_ACTION_MASK, T_REWARD, T_WORLD_ID, T_AGENT_ID,
 T_AGENT_STATE, T_LOCATION_WORLD_ID, T_LOCATION_ID, T_OBSERVATION_RAW, T_NUM_SLOTS) = range(13)


class TensorDesc(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("dtype", C.c_int32), ("ndim", C.c_int32), ("dims", C.c_int64 * 6),
                ("device", C.c_int32)]


class OvercookedConfig(C.Structure):
    _fields_ = [("terrain", C.POINTER(C.c_int64)), ("height", C.c_int64), ("width", C.c_int64),
                ("num_players", C.c_int64), ("start_player_x", C.POINTER(C.c_int64)),
                ("start_player_y", C.POINTER(C.c_int64)), ("placement_in_pot_rew", C.c_int64),
                ("dish_pickup_rew", C.c_int64), ("soup_pickup_rew", C.c_int64),
                ("recipe_values", C.POINTER(C.c_int64)), ("recipe_times", C.POINTER(C.c_int64)),
                ("horizon", C.c_int64)]


class HanabiConfig(C.Structure):
    _fields_ = [("colors", C.c_uint32), ("ranks", C.c_uint32), ("players", C.c_uint32),
                ("max_information_tokens", C.c_uint32), ("max_life_tokens", C.c_uint32)]


_PROTOTYPES = {
    "b200env_overcooked_create": (C.c_int, [C.POINTER(OvercookedConfig), C.c_int, C.c_uint32, C.POINTER(C.c_void_p)]),
    "b200env_overcooked2_create": (C.c_int, [C.POINTER(OvercookedConfig), C.c_int, C.c_uint32, C.POINTER(C.c_void_p)]),
    "b200env_overcooked_create_typed": (C.c_int, [C.POINTER(OvercookedConfig), C.c_int, C.c_int, C.c_int, C.c_uint32,
                                                  C.POINTER(C.c_void_p)]),
    "b200env_hanabi_create": (C.c_int, [C.POINTER(HanabiConfig), C.c_int, C.c_uint32, C.POINTER(C.c_void_p)]),
    "b200env_hanabi_create_typed": (C.c_int, [C.POINTER(HanabiConfig), C.c_int, C.c_int, C.c_uint32, C.POINTER(C.c_void_p)]),
    "b200env_balance_create": (C.c_int, [C.c_int, C.c_uint32, C.POINTER(C.c_void_p)]),
    "b200env_cartpole_create": (C.c_int, [C.c_int, C.c_uint32, C.POINTER(C.c_void_p)]),
    "b200env_acrobat_create": (C.c_int, [C.c_int, C.c_uint32, C.POINTER(C.c_void_p)]),
    "b200env_destroy": (None, [C.c_void_p]),
    "b200env_kind": (C.c_int, [C.c_void_p]),
    "b200env_num_worlds": (C.c_uint32, [C.c_void_p]),
    "b200env_step": (C.c_int, [C.c_void_p]),
    "b200env_step_async": (C.c_int, [C.c_void_p, C.c_void_p]),
    "b200env_tensor": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(TensorDesc)]),
    "b200env_step_host": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_void_p), C.c_int]),
    "b200env_step_host_async": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_void_p), C.c_int]),
    "b200env_step_host_wait": (C.c_int, [C.c_void_p]),
    "b200env_overcooked_export_raw": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200env_debug_stream_gate": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32]),
    "b200env_overcooked_step_into": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200env_overcooked_rollout": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                             C.c_void_p, C.c_int]),
    "b200env_random_rollout": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_uint32, C.c_uint32]),
    "b200env_random_action_hash": (C.c_uint32, [C.c_uint32, C.c_uint32, C.c_uint32]),
    "b200env_overcooked_set_rollout_fused": (C.c_int, [C.c_void_p, C.c_int]),
    "b200env_hanabi_set_rollout_fused": (C.c_int, [C.c_void_p, C.c_int]),
    "b200env_set_rollout_fused": (C.c_int, [C.c_void_p, C.c_int]),
    "b200env_overcooked_force_kernel": (C.c_int, [C.c_void_p, C.c_int]),
    "b200env_overcooked_set_lanes": (C.c_int, [C.c_void_p, C.c_int]),
    "b200env_debug_trace": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]),
    "b200env_state_bytes": (C.c_size_t, [C.c_void_p]),
    "b200env_get_state": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
    "b200env_set_state": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
    "b200env_episode_stats": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64)]),
    "b200env_cartpole_step_noalloc": (C.c_int, [C.c_int, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p,
                                                C.c_void_p, C.c_void_p]),
    "b200env_cartpole_step_noalloc_max_worlds": (C.c_uint32, []),
    "b200env_last_error": (C.c_char_p, []),
    "b200env_version": (C.c_char_p, []),
}
EXPORTED_SYMBOLS = tuple(_PROTOTYPES)

_lib = None


class B200EnvError(RuntimeError):
    pass


def load():
    """Load libb200envs.so (once).  Raises if it has not been built — no fallback path exists."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B200EnvError(
                f"{LIB_PATH} is missing: build it with `python -m madrona_rl_envs_playground_b200.build` "
                "(needs nvcc). The CUDA library is the only implementation of the step path.")
        lib = C.CDLL(LIB_PATH)
        for name, (restype, argtypes) in _PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.restype = restype
            fn.argtypes = argtypes
        _lib = lib
    return _lib


def check(rc):
    if rc != OK:
        raise B200EnvError(f"b200envs error {rc}: {load().b200env_last_error().decode()}")
