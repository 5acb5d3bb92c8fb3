"""GPU tests of the drop-in boundary beyond the step itself (run with -m gpu): the reference-raw Overcooked export that
lets the reference's UNMODIFIED wrapper bind, the pipelined host-buffer steps, the stream gate."""
import time

import numpy as np
import pytest

from helpers import GpuSim, biased_overcooked_actions

pytestmark = pytest.mark.gpu


def _params(layout, horizon, npl=None):
    from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
    return get_base_layout_params(layout, horizon, npl)


@pytest.mark.parametrize("layout,npl,N", [("cramped_room", None, 300), ("many_player_layout", None, 5),
                                          ("multiplayer_schelling", None, 33)])
def test_reference_raw_export_rebuilds_the_native_tensor(layout, npl, N):
    """What /root/reference/envs/overcooked_env.py:92-102 does with the reference module's tensors, statement for
    statement, on ReferenceLayoutOvercookedSimulator: scatter the padded rows by (locationID, locationWorldID), reshape
    (P, H, W, N, C), transpose(1, 3) — must equal the native tensor, and both must equal the oracle."""
    import torch
    from madrona_rl_envs_playground_b200 import ExecMode, ReferenceLayoutOvercookedSimulator
    from oracle.bindings import OracleOvercooked, RefOvercooked, ref_available
    params = _params(layout, 30, npl)
    P, W, H = params["num_players"], params["width"], params["height"]
    C, S = 5 * P + 16, W * H
    sim = ReferenceLayoutOvercookedSimulator(ExecMode.CUDA, 0, N, **params)
    orc = OracleOvercooked(params, N)
    static_observations = sim.observation_tensor().to_torch()
    static_locationWorldID = sim.location_world_id_tensor().to_torch().to(torch.long)
    static_locationID = sim.location_id_tensor().to_torch().to(torch.long)
    assert tuple(static_observations.shape) == (P * S, N, 336)
    assert tuple(static_locationID.shape) == (P * S, N) == tuple(static_locationWorldID.shape)
    if ref_available("overcooked"):   # the ID tables are the reference's own, element for element
        ref = RefOvercooked(params, N)
        assert np.array_equal(static_locationID.cpu().numpy().reshape(-1), ref.x_loc_id.reshape(-1))
        assert np.array_equal(static_locationWorldID.cpu().numpy().reshape(-1), ref.x_loc_world_id.reshape(-1))
        ref.close()
    scattered = torch.empty((H * W * P, N, C), dtype=torch.int8, device="cuda")
    native = sim.native_observation_tensor().to_torch()
    action = sim.action_tensor().to_torch()
    rng = np.random.default_rng(N)

    def get_obs():
        scattered[static_locationID, static_locationWorldID, :] = static_observations[:, :, :C]
        return scattered.reshape((P, H, W, N, -1)).transpose(1, 3)

    for t in range(35):
        obs0 = get_obs()
        assert torch.equal(obs0, native), t
        assert np.array_equal(obs0.cpu().numpy().view(np.uint8), orc.obs), t
        assert int(static_observations[:, :, C:].abs().sum()) == 0   # padding bytes: zero here, uninitialised there
        a = biased_overcooked_actions(rng, (P, N))
        action.copy_(torch.as_tensor(a).reshape(action.shape).cuda())
        sim.step()
        orc.step(a)
    sim.close()


def test_reference_raw_export_is_lazy_and_int8_only():
    import torch
    from madrona_rl_envs_playground_b200 import ExecMode, OvercookedSimulator, _lib
    params = _params("cramped_room", 30)
    free0 = torch.cuda.mem_get_info()[0]
    sim = OvercookedSimulator(ExecMode.CUDA, 0, 20000, **params)
    used_native = free0 - torch.cuda.mem_get_info()[0]
    sim.raw_observation_tensor()
    used_raw = free0 - torch.cuda.mem_get_info()[0] - used_native
    assert used_raw >= 20000 * 40 * 336 > used_native   # the padded export alone is bigger than everything native
    sim.close()
    typed = OvercookedSimulator(ExecMode.CUDA, 0, 64, obs_dtype=torch.bfloat16, **params)
    with pytest.raises(_lib.B200EnvError):
        typed.export_raw()
    typed.close()


@pytest.mark.parametrize("env", ["overcooked", "hanabi"])
def test_pipelined_host_steps_match_synchronous_ones(env):
    """b200env_step_host_async / _wait: same bytes in the host buffers as b200env_step_host, step for step, with two
    steps in flight; misuse is refused, not undefined."""
    import torch
    from madrona_rl_envs_playground_b200 import ExecMode, HanabiSimulator, OvercookedSimulator, _lib
    N, T = 3000, 24
    if env == "overcooked":
        params = _params("cramped_room", 10)
        make = lambda: OvercookedSimulator(ExecMode.CUDA, 0, N, **params)
        slots = [_lib.T_OBSERVATION, _lib.T_REWARD, _lib.T_DONE]
        acts = torch.randint(0, 6, (T, 2, N), dtype=torch.int32).pin_memory()
    else:
        make = lambda: HanabiSimulator(ExecMode.CUDA, 0, N, 5, 5, 2, 8, 3)
        slots = [_lib.T_OBSERVATION, _lib.T_AGENT_STATE, _lib.T_ACTION_MASK, _lib.T_REWARD, _lib.T_DONE]
        acts = None
    a, b = make(), make()

    def host_like(sim, slot):
        t = sim._tensor(slot).to_torch()
        return torch.empty(t.shape, dtype=t.dtype).pin_memory()

    out_a = {s: host_like(a, s) for s in slots}
    out_b = [{s: host_like(b, s) for s in slots} for _ in range(2)]
    with pytest.raises(_lib.B200EnvError):
        b.step_host_wait()                      # nothing in flight
    rng = np.random.default_rng(0)
    mask = a.action_mask_tensor().to_torch().cpu().numpy()
    pending = []
    for t in range(T):
        if env == "overcooked":
            act = acts[t]
        else:   # legal moves from the synchronous twin's mask (the pipelined one lags a step behind)
            logits = rng.random(mask.shape)
            logits[mask == 0] = -1
            act = torch.as_tensor(logits.argmax(-1).astype(np.int32)).pin_memory()
        a.step_host(act, out_a)
        if env == "hanabi":
            mask = out_a[_lib.T_ACTION_MASK].numpy().copy()
        expect = {s: v.clone() for s, v in out_a.items()}
        b.step_host_async(act, out_b[t & 1])
        pending.append((t, act, expect))        # (keeps `act` alive until its wait)
        if len(pending) == 2:
            with pytest.raises(_lib.B200EnvError):
                b.step_host_async(act, out_b[t & 1])   # a third step in flight
            with pytest.raises(_lib.B200EnvError):
                b.step()                               # synchronous entry points refuse while steps are in flight
            t0, _, exp0 = pending.pop(0)
            b.step_host_wait()
            for s in slots:
                assert torch.equal(out_b[t0 & 1][s], exp0[s]), (s, t0)
    t0, _, exp0 = pending.pop(0)
    b.step_host_wait()
    for s in slots:
        assert torch.equal(out_b[t0 & 1][s], exp0[s]), (s, t0)
    # everything has landed: the synchronous calls work again and the device tensors agree
    b.step(); a.step()
    for s in slots:
        assert torch.equal(a._tensor(s).to_torch(), b._tensor(s).to_torch()), s
    a.close(); b.close()


def test_stream_gate_holds_the_stream_until_opened():
    import torch
    from madrona_rl_envs_playground_b200 import stream_gate
    torch.cuda.synchronize()
    stream = torch.cuda.Stream()
    flag = torch.zeros(1, dtype=torch.int32).pin_memory()
    x = torch.zeros(1, device="cuda")
    # Everything enqueued behind a closed gate must have been launched once before: the FIRST launch of a kernel loads
    # its module (CUDA lazy loading), which waits for running kernels — i.e. for the gate — and blocks the host until the
    # gate times out.  (bench.py's gated regions only hold kernels that already ran in its warm-up.)
    with torch.cuda.stream(stream):
        x += 1
        x -= 1
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    stream_gate(0, stream, flag, 3000)
    with torch.cuda.stream(stream):
        x += 1
    e1.record(stream)
    time.sleep(0.05)
    held = not e1.query()
    flag.fill_(1)            # open
    stream.synchronize()
    assert float(x.item()) == 1.0
    # the work behind the gate started only after the host opened it, >= 50 ms after it was enqueued
    assert held and 45.0 < e0.elapsed_time(e1) < 2500.0, (held, e0.elapsed_time(e1))
    # a gate nobody opens gives up by itself (no hung GPU)
    flag.zero_()
    t0 = time.perf_counter()
    stream_gate(0, stream, flag, 100)
    stream.synchronize()
    assert 0.05 < time.perf_counter() - t0 < 2.0


@pytest.mark.parametrize("name", ["cramped_room", "coordination_ring", "multiplayer_schelling",
                                  "asymmetric_advantages_tomato"])
@pytest.mark.parametrize("kernel", [-1, 0])
def test_overcooked_vs_dummy_mdp_golden(name, kernel):
    """The CUDA kernels against the reference's numpy DummyMDP (its own validate_step oracle,
    /root/reference/envs/overcooked_env.py:478-527): exact obs / reward / done, both step kernels."""
    from helpers import load_golden, overcooked_params_from_golden
    from madrona_rl_envs_playground_b200 import ExecMode, OvercookedSimulator
    g = load_golden("dummy_mdp_" + name)
    params = overcooked_params_from_golden(g)
    N = g["actions"].shape[2]
    s = OvercookedSimulator(ExecMode.CUDA, 0, N, **params)
    if kernel != -1:
        s.force_kernel(kernel)
    sim = GpuSim(s)
    assert np.array_equal(sim.np("observation").view(np.uint8), g["obs0"])
    for t in range(g["actions"].shape[0]):
        sim.step(g["actions"][t])
        assert np.array_equal(sim.np("observation").view(np.uint8), g["obs"][t]), (name, t)
        assert np.array_equal(sim.np("reward"), g["rewards"][t]), (name, t)
        assert np.array_equal(sim.np("done"), g["done"][t]), (name, t)
    s.close()


def test_env_classes_vs_oracle_through_the_class():
    """Row O10 of SURVEY 8(a): the reference-facing classes themselves (OvercookedMadrona.n_step / get_obs, the
    VectorMultiAgentEnv ego / partner plumbing, HanabiMadrona.n_step) against the oracle — exact obs / reward / done, the
    protocol of the reference's validate_step (/root/reference/envs/overcooked_env.py:478-527)."""
    import torch
    from madrona_rl_envs_playground_b200.envs import HanabiMadrona, OvercookedMadrona
    from madrona_rl_envs_playground_b200.envs.hanabi_env import FULL_CONFIG
    from madrona_rl_envs_playground_b200.pantheonrl_extension import RandomVectorAgent
    from oracle.bindings import OracleHanabi, OracleOvercooked
    N, horizon = 200, 25
    params = _params("cramped_room", horizon)
    rng = np.random.default_rng(10)

    # n_step: per-player VectorObservation list, rewards [P, N], dones [N]
    env, orc = OvercookedMadrona("cramped_room", N, 0, horizon=horizon), OracleOvercooked(params, N)
    obs = env.n_reset()
    for p in range(2):
        assert np.array_equal(obs[p].obs.cpu().numpy().view(np.uint8), orc.obs[p]) and bool(obs[p].active.all())
    for t in range(horizon + 6):
        a = biased_overcooked_actions(rng, (2, N))
        obs, rew, done, _ = env.n_step(torch.as_tensor(a).reshape(2, N, 1).cuda())
        o, r, d = orc.step(a)
        for p in range(2):
            assert np.array_equal(obs[p].obs.cpu().numpy().view(np.uint8), o[p]), (p, t)
        assert np.array_equal(rew.cpu().numpy(), r) and np.array_equal(done.cpu().numpy(), d), t

    # step(ego_action): the ego agent's view, with a scripted partner whose moves the test knows
    env, orc = OvercookedMadrona("cramped_room", N, 0, horizon=horizon, ego_agent_idx=1), OracleOvercooked(params, N)
    partner_moves = []
    def partner():
        partner_moves.append(rng.integers(0, 6, (N, 1)).astype(np.int32))
        return torch.as_tensor(partner_moves[-1]).cuda()
    env.add_partner_agent(RandomVectorAgent(partner), player_num=0)   # (the default player_num = 1 is the ego here: refused, as in the reference)
    ob = env.reset()
    assert np.array_equal(ob.obs.cpu().numpy().view(np.uint8), orc.obs[1])
    for t in range(horizon + 3):
        ego = rng.integers(0, 6, (N, 1)).astype(np.int32)
        ob, rew, done, _ = env.step(torch.as_tensor(ego).cuda())
        o, r, d = orc.step(np.stack([partner_moves[-1][:, 0], ego[:, 0]]))   # player 0 = partner, player 1 = ego
        assert np.array_equal(ob.obs.cpu().numpy().view(np.uint8), o[1]), t
        assert np.array_equal(rew.cpu().numpy(), r[1]) and np.array_equal(done.cpu().numpy(), d), t

    # Hanabi through MadronaEnv.n_step: obs / state / mask / active per player
    henv, horc = HanabiMadrona(N, 0, config=FULL_CONFIG), OracleHanabi(N)
    obs = henv.n_reset()
    for t in range(60):
        for p in range(2):
            assert np.array_equal(obs[p].obs.cpu().numpy().view(np.uint8), horc.obs[p]), (p, t)
            assert np.array_equal(obs[p].state.cpu().numpy().view(np.uint8), horc.state[p]), (p, t)
            assert np.array_equal(obs[p].action_mask.cpu().numpy().astype(np.int32), horc.mask[p]), (p, t)
            assert np.array_equal(obs[p].active.cpu().numpy().astype(np.int32), horc.active[p]), (p, t)
        logits = rng.random(horc.mask.shape)
        logits[horc.mask == 0] = -1
        a = logits.argmax(-1).astype(np.int32)
        obs, rew, done, _ = henv.n_step(torch.as_tensor(a).reshape(2, N, 1).cuda())
        horc.step(a)
        assert np.array_equal(rew.cpu().numpy(), horc.rewards) and np.array_equal(done.cpu().numpy(), horc.done), t
