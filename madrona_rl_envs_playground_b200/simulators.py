"""Simulator objects with the reference's nanobind surface, backed by the C-ABI CUDA library.

Same constructor keywords, ``step()`` and ``<name>_tensor().to_torch()`` accessors as
``madrona_<env>_example_python.<Env>Simulator`` (src/overcooked_env/bindings.cpp:11-84, src/hanabi_env/bindings.cpp,
src/balance_beam_env/bindings.cpp, src/cartpole_env/bindings.cpp) and the ``madrona.ExecMode`` attribute of
madrona/src/python/bindings.cpp:70-79.  ``to_torch()`` is a zero-copy view of library-owned device memory
(the reference's DLPack path, madrona/src/python/bindings.cpp:109-124); the tensors stay valid and are updated in
place by every ``step()``.

Differences, all on the side of doing less work:
  * tensors are agent-major and dense, so the ID tensors are identities (``world_id[p, n] == n``);
  * Overcooked's ``observation_tensor()`` is int8 ``[P, N, W, H, 5P+16]`` (policy layout), not padded 336-byte rows;
  * ``ExecMode.CPU`` raises: the CPU implementation is the reference itself, this package is the CUDA path.
"""
import ctypes as C
import enum

from . import _lib as L


class ExecMode(enum.Enum):
    CPU = 0
    CUDA = 1


class madrona:   # namespace stand-in: `<module>.madrona.ExecMode.CUDA` as in the reference bindings
    ExecMode = ExecMode


_TYPESTR = {L.I8: "|i1", L.I32: "<i4", L.F32: "<f4", L.BF16: "<i2"}   # bfloat16 travels as int16 bits, see to_torch


class Tensor:
    """Stand-in for madrona.Tensor: ``to_torch()`` returns a zero-copy torch view of device memory."""

    def __init__(self, owner, desc):
        self._owner = owner   # keeps the simulator (and its device memory) alive while views exist
        self.ptr = desc.ptr
        self.dtype = desc.dtype
        self.shape = tuple(int(desc.dims[i]) for i in range(desc.ndim))
        self.device = int(desc.device)
        self.__cuda_array_interface__ = {
            "shape": self.shape, "typestr": _TYPESTR[self.dtype], "data": (int(self.ptr), False), "version": 2,
            "strides": None,
        }

    def to_torch(self):
        import torch
        t = torch.as_tensor(self, device=torch.device("cuda", self.device))
        # the CUDA array interface has no bfloat16 typestr: reinterpret the 16-bit words (still zero-copy)
        return t.view(torch.bfloat16) if self.dtype == L.BF16 else t


class _Simulator:
    def __init__(self, exec_mode, gpu_id, num_worlds):
        if isinstance(exec_mode, ExecMode) and exec_mode is ExecMode.CPU:
            raise L.B200EnvError("ExecMode.CPU is not provided by the B200 build: the CUDA kernels are the only "
                                 "implementation (use the reference itself for a CPU run)")
        self._lib = L.load()
        self._h = C.c_void_p()
        self.gpu_id = int(gpu_id)
        self.num_worlds = int(num_worlds)

    def _tensor(self, slot):
        d = L.TensorDesc()
        L.check(self._lib.b200env_tensor(self._h, slot, C.byref(d)))
        return Tensor(self, d)

    # --- reference surface -------------------------------------------------------------------------------------
    def step(self):
        L.check(self._lib.b200env_step(self._h))

    def done_tensor(self): return self._tensor(L.T_DONE)
    def active_agent_tensor(self): return self._tensor(L.T_ACTIVE_AGENT)
    def action_tensor(self): return self._tensor(L.T_ACTION)
    def observation_tensor(self): return self._tensor(L.T_OBSERVATION)
    def action_mask_tensor(self): return self._tensor(L.T_ACTION_MASK)
    def reward_tensor(self): return self._tensor(L.T_REWARD)
    def world_id_tensor(self): return self._tensor(L.T_WORLD_ID)
    def agent_id_tensor(self): return self._tensor(L.T_AGENT_ID)

    # --- extensions --------------------------------------------------------------------------------------------
    def step_async(self, stream=None):
        """Enqueue one step on ``stream`` (a torch.cuda.Stream, a raw handle or None) without synchronising."""
        handle = getattr(stream, "cuda_stream", stream)
        L.check(self._lib.b200env_step_async(self._h, C.c_void_p(handle or 0)))

    def step_stream_ordered(self):
        """One step enqueued on torch's CURRENT stream of the simulator's device, without blocking the host: torch
        work issued afterwards on that stream (or anything that synchronises with it, e.g. ``.cpu()``) sees the
        results.  This is what the packaged env classes call; ``step()`` keeps the reference's host-synchronous
        behaviour (cudaStreamSynchronize, madrona/src/mw/cuda_exec.cpp:1785)."""
        import torch
        self.step_async(torch.cuda.current_stream(self.gpu_id))

    def step_host(self, actions, outputs):
        """End-to-end host-buffer step: ``actions`` int32 numpy / pinned torch CPU tensor; ``outputs`` maps tensor
        slot -> preallocated host array that receives the tensor after the step."""
        slots = (C.c_int * len(outputs))(*outputs.keys())
        ptrs = (C.c_void_p * len(outputs))(*[_host_ptr(v) for v in outputs.values()])
        L.check(self._lib.b200env_step_host(self._h, C.c_void_p(_host_ptr(actions)), slots, ptrs, len(outputs)))

    def step_host_async(self, actions, outputs):
        """Pipelined ``step_host``: returns at once; the outputs are complete after the matching ``step_host_wait()``
        (at most two steps in flight; ``actions`` and the output buffers must be pinned and stay untouched until then)."""
        slots = (C.c_int * len(outputs))(*outputs.keys())
        ptrs = (C.c_void_p * len(outputs))(*[_host_ptr(v) for v in outputs.values()])
        L.check(self._lib.b200env_step_host_async(self._h, C.c_void_p(_host_ptr(actions)), slots, ptrs, len(outputs)))

    def step_host_wait(self):
        """Blocks until the oldest ``step_host_async`` step has landed in its host buffers."""
        L.check(self._lib.b200env_step_host_wait(self._h))

    def random_rollout(self, num_steps, seed=0, first_step=0, stream=None):
        """``num_steps`` x {uniform random legal actions -> step} enqueued on ``stream`` without host round trips."""
        handle = getattr(stream, "cuda_stream", stream)
        L.check(self._lib.b200env_random_rollout(self._h, C.c_void_p(handle or 0), int(num_steps), int(seed),
                                                 int(first_step)))

    def set_rollout_fused(self, fused):
        """Fused multi-step launches of ``random_rollout`` (default where available) or one launch per step; the
        tensors after the call are the same."""
        L.check(self._lib.b200env_set_rollout_fused(self._h, int(bool(fused))))

    def random_action_hash(self, seed, step, index):
        return int(self._lib.b200env_random_action_hash(int(seed), int(step), int(index)))

    def get_state(self):
        n = self._lib.b200env_state_bytes(self._h)
        buf = C.create_string_buffer(n)
        L.check(self._lib.b200env_get_state(self._h, buf, n))
        return buf.raw

    def set_state(self, blob):
        L.check(self._lib.b200env_set_state(self._h, blob, len(blob)))

    def episode_stats(self):
        out = (C.c_int64 * 2)()
        L.check(self._lib.b200env_episode_stats(self._h, out))
        return int(out[0]), int(out[1])

    def close(self):
        if self._h:
            self._lib.b200env_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _host_ptr(a):
    if hasattr(a, "data_ptr"):
        return a.data_ptr()
    return a.ctypes.data


def _i64(values):
    return (C.c_int64 * len(values))(*[int(v) for v in values])


def _obs_dtype_code(obs_dtype):
    """None / torch.int8 -> the reference's int8 export; torch.bfloat16 / torch.float32 (or their names) -> the
    policy-ready tensor written by the step kernel (b200env_overcooked_create_typed)."""
    name = str(obs_dtype).replace("torch.", "") if obs_dtype is not None else "int8"
    try:
        return {"int8": L.I8, "i8": L.I8, "bfloat16": L.BF16, "bf16": L.BF16, "float32": L.F32, "float": L.F32,
                "f32": L.F32}[name]
    except KeyError:
        raise L.B200EnvError(f"obs_dtype must be int8, bfloat16 or float32, not {obs_dtype!r}") from None


class OvercookedSimulator(_Simulator):
    _CREATE = "b200env_overcooked_create"
    _RULES = 0

    def __init__(self, exec_mode, gpu_id, num_worlds, terrain, height, width, num_players, start_player_x,
                 start_player_y, placement_in_pot_rew, dish_pickup_rew, soup_pickup_rew, recipe_values, recipe_times,
                 horizon, debug_compile=False, obs_dtype=None):
        """Reference keywords (src/overcooked_env/bindings.cpp:15-53) plus ``obs_dtype`` (extension):
        ``torch.bfloat16`` / ``torch.float32`` make ``observation_tensor()`` hold what ``obs.float()`` of the int8
        tensor would (pantheonrl_extension/vectoragent.py:353), written by the step kernel itself."""
        super().__init__(exec_mode, gpu_id, num_worlds)
        if len(terrain) < height * width or len(start_player_x) < num_players or len(start_player_y) < num_players \
                or len(recipe_values) < 16 or len(recipe_times) < 16:
            raise L.B200EnvError("layout vectors shorter than height*width / num_players / 16 recipes")
        self._keep = [_i64(terrain), _i64(start_player_x), _i64(start_player_y), _i64(recipe_values), _i64(recipe_times)]
        cfg = L.OvercookedConfig(self._keep[0], height, width, num_players, self._keep[1], self._keep[2],
                                 placement_in_pot_rew, dish_pickup_rew, soup_pickup_rew, self._keep[3], self._keep[4],
                                 horizon)
        code = _obs_dtype_code(obs_dtype)
        if code == L.I8:
            L.check(getattr(self._lib, self._CREATE)(C.byref(cfg), self.gpu_id, self.num_worlds, C.byref(self._h)))
        else:
            L.check(self._lib.b200env_overcooked_create_typed(C.byref(cfg), self._RULES, code, self.gpu_id,
                                                              self.num_worlds, C.byref(self._h)))
        self.height, self.width, self.num_players = int(height), int(width), int(num_players)

    def step_into(self, stream=None, actions=None, obs=None, rewards=None, done=None):
        """Asynchronous step reading / writing caller-supplied device tensors (rollout buffers)."""
        handle = getattr(stream, "cuda_stream", stream)
        p = lambda t: C.c_void_p(t.data_ptr() if t is not None else 0)
        L.check(self._lib.b200env_overcooked_step_into(self._h, C.c_void_p(handle or 0), p(actions), p(obs), p(rewards),
                                                       p(done)))

    def rollout(self, stream, num_steps, actions, obs=None, rewards=None, done=None):
        """``num_steps`` asynchronous steps on ``stream``: step t reads ``actions[t % len(actions)]`` (int32
        ``[A, P, N]`` device tensor) and writes slice ``t % len(obs)`` of the ``[T, ...]`` output tensors (None = the
        simulator's own tensor)."""
        handle = getattr(stream, "cuda_stream", stream)
        outs = [t for t in (obs, rewards, done) if t is not None]
        out_slices = outs[0].shape[0] if outs else 1
        if any(t.shape[0] != out_slices for t in outs):
            raise L.B200EnvError("rollout output tensors must have the same number of slices")
        p = lambda t: C.c_void_p(t.data_ptr() if t is not None else 0)
        L.check(self._lib.b200env_overcooked_rollout(self._h, C.c_void_p(handle or 0), int(num_steps), p(actions),
                                                     int(actions.shape[0]), p(obs), p(rewards), p(done), out_slices))

    def set_rollout_fused(self, fused):
        """``rollout`` as one fused launch (default where available) or one launch per step; same results."""
        L.check(self._lib.b200env_overcooked_set_rollout_fused(self._h, int(bool(fused))))

    def force_kernel(self, which):
        L.check(self._lib.b200env_overcooked_force_kernel(self._h, which))

    def set_lanes(self, lanes):
        """Lanes per world of the warp-tile kernel (0 = auto, 1, 2, 4); results do not depend on it."""
        L.check(self._lib.b200env_overcooked_set_lanes(self._h, lanes))


    # --- reference-raw export (compat) ---------------------------------------------------------
    def location_world_id_tensor(self): return self._tensor(L.T_LOCATION_WORLD_ID)
    def location_id_tensor(self): return self._tensor(L.T_LOCATION_ID)
    def raw_observation_tensor(self): return self._tensor(L.T_OBSERVATION_RAW)

    def export_raw(self, stream=None, dst=None):
        """Writes the reference's padded ``[P*S, N, 336]`` observation rows (world-major memory, as the reference's
        export) from the native tensor into ``dst`` (None = ``raw_observation_tensor()``)."""
        handle = getattr(stream, "cuda_stream", stream)
        L.check(self._lib.b200env_overcooked_export_raw(self._h, C.c_void_p(handle or 0),
                                                        C.c_void_p(dst.data_ptr() if dst is not None else 0)))


class ReferenceLayoutOvercookedSimulator(OvercookedSimulator):
    """``OvercookedSimulator`` with EXACTLY the reference module's tensor table (src/overcooked_env/bindings.cpp:72-83), for
    running the reference's own, unmodified ``envs/overcooked_env.py`` wrapper on this library: ``observation_tensor()``
    is the padded ``[P*S, N, 336]`` export, refreshed after every ``step()``, and ``location_world_id_tensor()`` /
    ``location_id_tensor()`` drive the wrapper's scatter (envs/overcooked_env.py:92-102).  It pays the extra pass over
    13x the bytes that the native layout exists to avoid; new code should use ``OvercookedSimulator``."""

    def __init__(self, *args, **kw):
        super().__init__(*args, **kw)
        self.export_raw()
        self._sync()

    def _sync(self):
        import torch
        torch.cuda.synchronize(self.gpu_id)

    def step(self):
        super().step()
        self.export_raw()
        self._sync()

    def observation_tensor(self): return self._tensor(L.T_OBSERVATION_RAW)
    def native_observation_tensor(self): return self._tensor(L.T_OBSERVATION)


class SimplecookedSimulator(OvercookedSimulator):
    """``madrona_simplecooked_example_python.SimplecookedSimulator`` (src/overcooked2_env/bindings.cpp:10-59): the
    JS-compatible Overcooked rules.  Same constructor keywords; ``terrain`` uses that module's enum order (AIR, POT,
    COUNTER, ONION_SOURCE, DISH_SOURCE, SERVING, TOMATO_SOURCE); the observation has 5P + 10 channels."""
    _CREATE = "b200env_overcooked2_create"
    _RULES = 1


class HanabiSimulator(_Simulator):
    def __init__(self, exec_mode, gpu_id, num_worlds, colors, ranks, players, max_information_tokens, max_life_tokens,
                 debug_compile=False, obs_dtype=None):
        """Reference keywords (src/hanabi_env/bindings.cpp) plus ``obs_dtype`` (extension):
        ``torch.bfloat16`` / ``torch.float32`` make ``observation_tensor()`` and ``agent_state_tensor()`` hold what
        ``.float()`` of the int8 rows would (pantheonrl_extension/hanabi_agent.py:257,292), written by the kernels."""
        super().__init__(exec_mode, gpu_id, num_worlds)
        cfg = L.HanabiConfig(colors, ranks, players, max_information_tokens, max_life_tokens)
        code = _obs_dtype_code(obs_dtype)
        if code == L.I8:
            L.check(self._lib.b200env_hanabi_create(C.byref(cfg), self.gpu_id, self.num_worlds, C.byref(self._h)))
        else:
            L.check(self._lib.b200env_hanabi_create_typed(C.byref(cfg), code, self.gpu_id, self.num_worlds, C.byref(self._h)))

    def set_rollout_fused(self, fused):
        """``random_rollout`` as one cooperative launch (default where all tiles fit on the GPU) or one per step."""
        L.check(self._lib.b200env_hanabi_set_rollout_fused(self._h, int(bool(fused))))

    def agent_state_tensor(self): return self._tensor(L.T_AGENT_STATE)


class BalanceBeamSimulator(_Simulator):
    def __init__(self, exec_mode, gpu_id, num_worlds, debug_compile=False):
        super().__init__(exec_mode, gpu_id, num_worlds)
        L.check(self._lib.b200env_balance_create(self.gpu_id, self.num_worlds, C.byref(self._h)))

    def agent_state_tensor(self):
        # the reference exports no separate state for this env; MadronaEnv reads the observation for both
        # (src/balance_beam_env/bindings.cpp:29 binds agent_state_tensor to observationTensor)
        return self._tensor(L.T_OBSERVATION)


class CartpoleSimulator(_Simulator):
    def __init__(self, exec_mode, gpu_id, num_worlds, debug_compile=False):
        super().__init__(exec_mode, gpu_id, num_worlds)
        L.check(self._lib.b200env_cartpole_create(self.gpu_id, self.num_worlds, C.byref(self._h)))

    def reset_tensor(self): return self._tensor(L.T_DONE)
    def observation_tensor(self): return self._tensor(L.T_AGENT_STATE)
    def state_tensor(self): return self._tensor(L.T_AGENT_STATE)


class AcrobatSimulator(_Simulator):
    """``madrona_acrobat_example_python.AcrobatSimulator`` (src/acrobat_env/bindings.cpp:10-33): ``reset_tensor``,
    ``action_tensor`` (0..2 = torque -1 / 0 / +1), ``observation_tensor`` (the f32 ``[N, 4]`` state), ``reward_tensor``,
    ``world_id_tensor``."""

    def __init__(self, exec_mode, gpu_id, num_worlds, debug_compile=True):
        super().__init__(exec_mode, gpu_id, num_worlds)
        L.check(self._lib.b200env_acrobat_create(self.gpu_id, self.num_worlds, C.byref(self._h)))

    def reset_tensor(self): return self._tensor(L.T_DONE)
    def observation_tensor(self): return self._tensor(L.T_AGENT_STATE)
    def state_tensor(self): return self._tensor(L.T_AGENT_STATE)


def stream_gate(gpu_id, stream, flag, timeout_ms=2000):
    """Benchmark helper: holds ``stream`` until ``flag`` (a pinned int32 torch tensor of one element) is set non-zero
    by the host (or ``timeout_ms`` pass), so that everything enqueued behind it starts without host launch latency."""
    handle = getattr(stream, "cuda_stream", stream)
    L.check(L.load().b200env_debug_stream_gate(int(gpu_id), C.c_void_p(handle or 0), C.c_void_p(flag.data_ptr()),
                                               int(timeout_ms)))
