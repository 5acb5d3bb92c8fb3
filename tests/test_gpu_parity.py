"""GPU parity tests (run with -m gpu on the B200 box): the CUDA kernels, called through the C ABI, against
  * the committed golden vectors generated from the reference's Madrona CPU exec mode, and
  * the CPU oracle on fresh seeded action streams,
bit-exact for Overcooked / Hanabi / Balance Beam, 1e-5 relative for Cartpole's float state (exact done flags away from
the thresholds), plus size-independent properties at the BASELINE sizes."""
import numpy as np
import pytest

from helpers import (OVERCOOKED2_GOLDENS, OVERCOOKED_GOLDENS, GpuSim, biased_overcooked_actions, legal_random_actions, load_golden,
                     overcooked_params_from_golden, random_policy_actions, random_policy_hash)

pytestmark = pytest.mark.gpu


def make_overcooked(params, N, kernel=-1, lanes=0):
    from madrona_rl_envs_playground_b200 import ExecMode, OvercookedSimulator
    sim = OvercookedSimulator(ExecMode.CUDA, 0, N, **params)
    if lanes:
        sim.set_lanes(lanes)
    if kernel != -1:
        sim.force_kernel(kernel)
    return GpuSim(sim)


def obs_u8(g):
    return g.np("observation").view(np.uint8)


# ------------------------------------------------------------------------------------------------------------------
# Overcooked
# ------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", OVERCOOKED_GOLDENS)
@pytest.mark.parametrize("kernel", [-1, 0])
def test_overcooked_vs_golden(name, kernel):
    g = load_golden("overcooked_" + name)
    params = overcooked_params_from_golden(g)
    N = g["actions"].shape[2]
    sim = make_overcooked(params, N, kernel)
    assert np.array_equal(obs_u8(sim), g["obs0"])
    for t in range(g["actions"].shape[0]):
        sim.step(g["actions"][t])
        assert np.array_equal(obs_u8(sim), g["obs"][t]), (name, t)
        assert np.array_equal(sim.np("reward"), g["rewards"][t]), (name, t)
        assert np.array_equal(sim.np("done"), g["done"][t]), (name, t)


@pytest.mark.parametrize("lanes", [1, 2, 4, 8])
@pytest.mark.parametrize("layout,npl,N,horizon", [
    ("cramped_room", 1, 100, 45),                    # single agent: 1 lane per world is allowed
    ("cramped_room", None, 531, 45),                 # ragged last tile for every lanes-per-world variant
    ("cramped_room", None, 77, 400),                 # odd world count: observer slabs are not 16-byte congruent
    ("asymmetric_advantages_tomato", None, 200, 45),
    ("multiplayer_schelling", None, 96, 45),         # 4 players
    ("many_player_layout", 7, 40, 45),               # 7 players on the 15x17 grid (PMAX = 8)
])
def test_overcooked_lanes_per_world(lanes, layout, npl, N, horizon):
    """The warp-tile kernel gives each world 1, 2, 4 or 8 lanes (>= players); every variant must match the oracle."""
    from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
    from oracle.bindings import OracleOvercooked
    params = get_base_layout_params(layout, horizon, npl)
    P = params["num_players"]
    from madrona_rl_envs_playground_b200._lib import B200EnvError
    try:
        sim = make_overcooked(params, N, kernel=1, lanes=lanes)
    except B200EnvError as e:
        assert "lanes per world" in str(e) or "warp-tile" in str(e)   # fewer lanes than players / tile too big
        pytest.skip(str(e))
    orc = OracleOvercooked(params, N)
    assert np.array_equal(obs_u8(sim), orc.obs)
    rng = np.random.default_rng(lanes * 1000 + N)
    for t in range(70):
        a = biased_overcooked_actions(rng, (P, N))
        sim.step(a)
        obs, rew, done = orc.step(a)
        assert np.array_equal(sim.np("reward"), rew), (layout, lanes, t)
        assert np.array_equal(sim.np("done"), done), (layout, lanes, t)
        assert np.array_equal(obs_u8(sim), obs), (layout, lanes, t)


@pytest.mark.parametrize("layout,npl,N,kernel,steps", [
    ("cramped_room", None, 1024, -1, 120),
    ("cramped_room", None, 1000, -1, 60),      # ragged last tile
    ("cramped_room", None, 37, -1, 60),        # odd world count: unaligned slabs take the byte path
    ("cramped_room", None, 64, 0, 60),         # CTA-per-world kernel on a small layout
    ("cramped_room", None, 64, 1, 60),         # warp-tile kernel forced
    ("coordination_ring", None, 256, -1, 80),
    ("asymmetric_advantages_tomato", None, 256, -1, 80),
    ("multiplayer_schelling", None, 128, -1, 80),
    ("multiplayer_schelling", None, 64, 0, 60),  # CTA-per-world kernel with 4 players
    ("many_player_layout", 8, 16, -1, 50),
    ("many_player_layout", None, 6, -1, 50),   # 40 players, 15x17
    ("many_player_layout", None, 3, -1, 30),   # odd world count with 55080-byte slabs
])
def test_overcooked_vs_oracle(layout, npl, N, kernel, steps):
    from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
    from oracle.bindings import OracleOvercooked
    horizon = 45   # crosses the auto-reset (and the urgency window, horizon - t < 40) inside the test
    params = get_base_layout_params(layout, horizon, npl)
    P = params["num_players"]
    sim = make_overcooked(params, N, kernel)
    orc = OracleOvercooked(params, N)
    assert np.array_equal(obs_u8(sim), orc.obs)
    rng = np.random.default_rng(N + steps)
    for t in range(steps):
        a = biased_overcooked_actions(rng, (P, N))
        sim.step(a)
        obs, rew, done = orc.step(a)
        assert np.array_equal(sim.np("reward"), rew), (layout, t)
        assert np.array_equal(sim.np("done"), done), (layout, t)
        assert np.array_equal(obs_u8(sim), obs), (layout, t)


def test_overcooked_full_horizon_400():
    """BASELINE config 3 semantics at a size the oracle finishes in seconds: horizon 400, 2 full episodes + change."""
    from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
    from oracle.bindings import OracleOvercooked
    params = get_base_layout_params("cramped_room", 400)
    N = 256
    sim, orc = make_overcooked(params, N), OracleOvercooked(params, N)
    rng = np.random.default_rng(400)
    total = finished = 0
    for t in range(820):
        a = rng.integers(0, 6, (2, N)).astype(np.int32)
        sim.step(a)
        obs, rew, done = orc.step(a, want_obs=(t % 7 == 0 or t in (399, 400, 799, 800)))
        assert np.array_equal(sim.np("reward"), rew), t
        assert np.array_equal(sim.np("done"), done), t
        if t % 7 == 0 or t in (399, 400, 799, 800):
            assert np.array_equal(obs_u8(sim), obs), t
        total += int(rew[0].sum())
        if t == 799:
            finished = total
    assert total > 0
    episodes, ret = sim.sim.episode_stats()
    assert episodes == 2 * N and ret == finished   # device-side statistics: two full episodes per world


def test_overcooked_properties_at_baseline_size():
    """32K worlds (BASELINE config 3): identical action streams give identical worlds; obs structure invariants."""
    import torch
    from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
    params = get_base_layout_params("cramped_room", 400)
    N, P, C = 32768, 2, 26
    sim = make_overcooked(params, N)
    gen = torch.Generator(device="cuda").manual_seed(0)
    obs = sim.t("observation")
    assert tuple(obs.shape) == (P, N, 5, 4, C)
    for t in range(60):
        a = torch.randint(0, 6, (P, 64, 1), generator=gen, device="cuda", dtype=torch.int32).repeat(1, N // 64, 1)
        sim.action.copy_(a)
        sim.sim.step()
        # worlds n and n + 64 received the same actions from the same start: their rows must be identical
        o = obs.view(P, N // 64, 64, -1)
        assert bool((o == o[:, :1]).all())
        r = sim.t("reward").view(P, N // 64, 64)
        assert bool((r == r[:, :1]).all())
    o = obs.to(torch.int32)
    # every observer sees itself exactly once (channel 0) and each other player exactly once
    assert bool((o[..., :P].sum(dim=(2, 3)) == 1).all())
    # exactly one orientation bit per player
    assert bool((o[..., P:5 * P].sum(dim=(2, 3, 4)) == P).all())
    # static terrain one-hot is identical in every world and for both observers
    assert bool((o[..., 5 * P:5 * P + 6] == o[0, 0, :, :, 5 * P:5 * P + 6]).all())
    # player 1's view of channel 1 (the other player) equals player 0's view of channel 0 (itself)
    assert bool((o[1, ..., 1] == o[0, ..., 0]).all())


@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("layout,npl,N,lanes", [
    ("cramped_room", None, 203, 2),
    ("cramped_room", None, 64, 4),
    ("cramped_room", None, 45, 8),             # odd world count: byte path of the slab store
    ("coordination_ring", None, 100, 0),
    ("multiplayer_schelling", None, 72, 4),
    ("many_player_layout", None, 3, 0),        # CTA-per-world layout: rollout falls back to one launch per step
])
def test_overcooked_rollout_vs_oracle(fused, layout, npl, N, lanes):
    """b200env_overcooked_rollout (fused T-step kernel / per-step launches) against the oracle stepped T times, then a
    plain step() to prove the on-chip state went back to HBM, then the device-side episode statistics."""
    import torch
    from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
    from oracle.bindings import OracleOvercooked
    horizon, T, A = 45, 100, 7            # two resets inside one rollout; action ring shorter than T
    params = get_base_layout_params(layout, horizon, npl)
    P, W, H = params["num_players"], params["width"], params["height"]
    C = 5 * P + 16
    g = make_overcooked(params, N, lanes=lanes)
    g.sim.set_rollout_fused(fused)
    orc = OracleOvercooked(params, N)
    rng = np.random.default_rng(N + T)
    acts = biased_overcooked_actions(rng, (A, P, N))
    d_act = torch.as_tensor(acts).cuda()
    obs = torch.zeros((T, P, N, W, H, C), dtype=torch.int8, device="cuda")
    rew = torch.zeros((T, P, N), dtype=torch.int32, device="cuda")
    done = torch.zeros((T, N), dtype=torch.int32, device="cuda")
    g.sim.rollout(None, T, d_act, obs, rew, done)
    torch.cuda.synchronize()
    obs, rew, done = obs.cpu().numpy().view(np.uint8), rew.cpu().numpy(), done.cpu().numpy()
    total = 0
    for t in range(T):
        o, r, d = orc.step(acts[t % A])
        assert np.array_equal(rew[t], r), (layout, t)
        assert np.array_equal(done[t], d), (layout, t)
        assert np.array_equal(obs[t], o), (layout, t)
        total += int(r[0].sum())
    # outputs into the simulator's own tensors (single slice), starting from the written-back state
    g.sim.rollout(None, 3, d_act[:3])
    torch.cuda.synchronize()
    for t in range(3):
        o, r, d = orc.step(acts[t])
    assert np.array_equal(obs_u8(g), o) and np.array_equal(g.np("reward"), r) and np.array_equal(g.np("done"), d)
    g.step(acts[5])
    o, r, d = orc.step(acts[5])
    assert np.array_equal(obs_u8(g), o) and np.array_equal(g.np("reward"), r) and np.array_equal(g.np("done"), d)
    episodes, ret = g.sim.episode_stats()
    assert episodes == 2 * N and ret <= total


def test_overcooked_step_into_and_checkpoint():
    import torch
    from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
    params = get_base_layout_params("cramped_room", 30)
    N = 256
    a_sim, b_sim = make_overcooked(params, N), make_overcooked(params, N)
    gen = torch.Generator(device="cuda").manual_seed(3)
    T = 8
    acts = torch.randint(0, 6, (T, 2, N), generator=gen, device="cuda", dtype=torch.int32)
    ring = torch.zeros((T,) + tuple(a_sim.t("observation").shape), dtype=torch.int8, device="cuda")
    rews = torch.zeros((T, 2, N), dtype=torch.int32, device="cuda")
    dones = torch.zeros((T, N), dtype=torch.int32, device="cuda")
    for t in range(T):   # rollout-buffer path, asynchronous
        a_sim.sim.step_into(None, acts[t], ring[t], rews[t], dones[t])
    torch.cuda.synchronize()
    blob = None
    for t in range(T):   # exported-tensor path
        b_sim.step(acts[t].cpu().numpy())
        assert torch.equal(b_sim.t("observation"), ring[t])
        assert torch.equal(b_sim.t("reward"), rews[t])
        if t == 3:
            blob = b_sim.sim.get_state()
    # restore the checkpoint taken after step 3 and replay steps 4.. : same trajectory
    b_sim.sim.set_state(blob)
    for t in range(4, T):
        b_sim.step(acts[t].cpu().numpy())
        assert torch.equal(b_sim.t("observation"), ring[t])


# ------------------------------------------------------------------------------------------------------------------
# Policy-ready observations: the bfloat16 / float tensor equals `.float()` of the int8 one
# ------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("dtype_name", ["bfloat16", "float32"])
@pytest.mark.parametrize("layout,npl,N,lanes,rules", [
    ("cramped_room", None, 203, 0, 0),
    ("cramped_room", None, 45, 8, 0),              # odd world count: byte path of the slab store
    ("coordination_ring", None, 64, 2, 0),
    ("multiplayer_schelling", None, 40, 4, 0),
    ("simple", None, 77, 0, 1),                    # Simplecooked rules
])
def test_overcooked_typed_observations_vs_oracle(dtype_name, layout, npl, N, lanes, rules):
    """Step kernel and fused rollout with a bfloat16 / float32 observation tensor against the oracle's int8 rows: every
    element must be exactly the float value of the reference's int8 element (tolerance 0: bfloat16 holds all of int8)."""
    import torch
    from madrona_rl_envs_playground_b200 import ExecMode, OvercookedSimulator, SimplecookedSimulator
    if rules:
        from madrona_rl_envs_playground_b200.overcooked2_layouts import get_base_layout_params
        from oracle.bindings import OracleOvercooked2 as Oracle
        cls, extra = SimplecookedSimulator, 10
    else:
        from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
        from oracle.bindings import OracleOvercooked as Oracle
        cls, extra = OvercookedSimulator, 16
    dtype = getattr(torch, dtype_name)
    horizon, T, A = 45, 100, 7
    params = get_base_layout_params(layout, horizon, npl) if not rules else get_base_layout_params(layout, horizon)
    P, W, H = params["num_players"], params["width"], params["height"]
    C = 5 * P + extra
    sim = cls(ExecMode.CUDA, 0, N, obs_dtype=dtype, **params)
    if lanes:
        sim.set_lanes(lanes)
    g, orc = GpuSim(sim), Oracle(params, N)
    assert g.t("observation").dtype == dtype and tuple(g.t("observation").shape) == (P, N, W, H, C)

    def as_i8(t):   # exact: compare in float, then narrow
        f = t.float().cpu().numpy()
        assert np.array_equal(f, np.round(f)) and np.abs(f).max() <= 127
        return f.astype(np.int8).view(np.uint8)

    rng = np.random.default_rng(N)
    for t in range(60):   # one launch per step, across a reset
        a = biased_overcooked_actions(rng, (P, N))
        g.step(a)
        o, r, d = orc.step(a)
        assert np.array_equal(as_i8(g.t("observation")), o), (layout, t)
        assert np.array_equal(g.np("reward"), r) and np.array_equal(g.np("done"), d)
    for fused in (True, False):   # rollout into a [T, ...] buffer of the same dtype
        sim.set_rollout_fused(fused)
        acts = biased_overcooked_actions(rng, (A, P, N))
        d_act = torch.as_tensor(acts).cuda()
        obs = torch.zeros((T, P, N, W, H, C), dtype=dtype, device="cuda")
        rew = torch.zeros((T, P, N), dtype=torch.int32, device="cuda")
        sim.rollout(None, T, d_act, obs, rew, None)
        torch.cuda.synchronize()
        for t in range(T):
            o, r, d = orc.step(acts[t % A])
            assert np.array_equal(as_i8(obs[t]), o), (layout, fused, t)
            assert np.array_equal(rew[t].cpu().numpy(), r)
        assert np.array_equal(g.np("done"), d)
    # checkpoint round trip with the wider tensor
    blob = sim.get_state()
    a = biased_overcooked_actions(rng, (P, N))
    g.step(a)
    want = g.t("observation").clone()
    sim.set_state(blob)
    g.step(a)
    assert torch.equal(g.t("observation"), want)
    sim.close()


@pytest.mark.parametrize("dtype_name", ["bfloat16", "float32"])
def test_many_player_typed_observations_vs_oracle(dtype_name):
    """The CTA-per-world kernel (40 players, 15 x 17 cells) with a bfloat16 / float32 observation tensor: two slab buffers
    of 110 KB for bfloat16, one of 220 KB for float.  Every element is exactly the float value of the oracle's int8."""
    import torch
    from madrona_rl_envs_playground_b200 import ExecMode, OvercookedSimulator
    from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
    from oracle.bindings import OracleOvercooked
    dtype = getattr(torch, dtype_name)
    N, horizon = 5, 9
    params = get_base_layout_params("many_player_layout", horizon)
    P, W, H = params["num_players"], params["width"], params["height"]
    sim = OvercookedSimulator(ExecMode.CUDA, 0, N, obs_dtype=dtype, **params)
    g, orc = GpuSim(sim), OracleOvercooked(params, N)
    assert g.t("observation").dtype == dtype and tuple(g.t("observation").shape) == (P, N, W, H, 5 * P + 16)
    rng = np.random.default_rng(3)
    for t in range(2 * horizon + 3):   # across two resets
        a = biased_overcooked_actions(rng, (P, N))
        g.step(a)
        o, r, d = orc.step(a)
        f = g.t("observation").float().cpu().numpy()
        assert np.array_equal(f, np.round(f)) and np.abs(f).max() <= 127
        assert np.array_equal(f.astype(np.int8).view(np.uint8), o), t
        assert np.array_equal(g.np("reward"), r) and np.array_equal(g.np("done"), d)
    sim.close()


def test_typed_observations_refused_where_unsupported():
    import torch
    from madrona_rl_envs_playground_b200 import ExecMode, OvercookedSimulator
    from madrona_rl_envs_playground_b200._lib import B200EnvError
    from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
    with pytest.raises(B200EnvError):
        OvercookedSimulator(ExecMode.CUDA, 0, 4, obs_dtype=torch.float16, **get_base_layout_params("cramped_room", 400))


# ------------------------------------------------------------------------------------------------------------------
# Hanabi
# ------------------------------------------------------------------------------------------------------------------
def make_hanabi(N, **kw):
    from madrona_rl_envs_playground_b200 import ExecMode, HanabiSimulator
    cfg = dict(colors=5, ranks=5, players=2, max_information_tokens=8, max_life_tokens=3)
    cfg.update(kw)
    return GpuSim(HanabiSimulator(ExecMode.CUDA, 0, N, **cfg))


def hanabi_fields(sim):
    return dict(obs=sim.np("observation").view(np.uint8), state=sim.np("agent_state").view(np.uint8),
                mask=sim.np("action_mask"), active=sim.np("active_agent"), rewards=sim.np("reward"), done=sim.np("done"))


def test_hanabi_vs_golden():
    g = load_golden("hanabi_full")
    N = g["actions"].shape[2]
    sim = make_hanabi(N)
    f = hanabi_fields(sim)
    assert np.array_equal(f["obs"], g["obs0"]) and np.array_equal(f["state"], g["state0"])
    assert np.array_equal(f["mask"], g["mask0"]) and np.array_equal(f["active"], g["active0"])
    for t in range(g["actions"].shape[0]):
        sim.step(g["actions"][t])
        f = hanabi_fields(sim)
        for k in ("obs", "state", "mask", "active", "rewards", "done"):
            assert np.array_equal(f[k], g[k][t]), (k, t)


@pytest.mark.parametrize("N,steps", [(2048, 200), (1000, 120), (33, 150)])
def test_hanabi_vs_oracle(N, steps):
    from oracle.bindings import OracleHanabi
    sim, orc = make_hanabi(N), OracleHanabi(N)
    rng = np.random.default_rng(N)
    resets = 0
    for t in range(steps):
        f = hanabi_fields(sim)
        for k in ("obs", "state", "mask", "active", "rewards", "done"):
            assert np.array_equal(f[k], getattr(orc, k)), (k, t)
        a = legal_random_actions(rng, orc.mask)
        sim.step(a)
        orc.step(a)
        resets += int(orc.done.sum())
    assert resets > 0
    assert sim.sim.episode_stats() == (N + resets, 0)   # episodes started: the initial deal of every world + the resets


@pytest.mark.parametrize("cfg", [
    dict(colors=5, ranks=2),                                                    # tokens run past their maximum
    dict(colors=3, ranks=3, max_information_tokens=4, max_life_tokens=2),
    dict(colors=4, ranks=5, max_information_tokens=3, max_life_tokens=1),
])
def test_hanabi_other_rules_vs_oracle(cfg):
    """Non-standard games take the run-time-rules kernel; short stacks make information tokens exceed their maximum,
    which lengthens the board section and shifts everything after it (variable-length row path)."""
    from oracle.bindings import OracleHanabi
    N = 512
    sim, orc = make_hanabi(N, **cfg), OracleHanabi(N, **cfg)
    rng = np.random.default_rng(11)
    # nominal encoded lengths (sim.cpp:54-365); bytes past them are never written by the reference (uninitialised in
    # its tables and in the oracle's scratch rows), so only the encoded prefix is compared
    c, r = cfg.get("colors", 5), cfg.get("ranks", 5)
    bits = c * r
    obs_len = (5 * bits + 2) + ((4 + (r - 2) * 2) * c - 10) + bits + cfg.get("max_information_tokens", 8) \
        + cfg.get("max_life_tokens", 3) + c * (3 + 2 * (r - 2) + 1) + (2 + 4 + 2 + c + r + 5 + 5 + bits + 2) \
        + 2 * 5 * (bits + c + r)
    state_len = obs_len + 5 * bits
    for t in range(300):
        f = hanabi_fields(sim)
        assert np.array_equal(f["obs"][..., :obs_len], orc.obs[..., :obs_len]), (cfg, "obs", t)
        assert np.array_equal(f["state"][..., :state_len], orc.state[..., :state_len]), (cfg, "state", t)
        for k in ("mask", "active", "rewards", "done"):
            assert np.array_equal(f[k], getattr(orc, k)), (cfg, k, t)
        a = legal_random_actions(rng, orc.mask)
        sim.step(a)
        orc.step(a)
    if cfg.get("ranks") == 2:
        assert orc.max_info_encoded() > 8


def test_hanabi_token_overflow_spills_like_the_reference():
    """Standard game, stacks completed at 8 information tokens: the 9th token makes the rows 659 / 784 bytes long and
    the excess byte lands on the next player's row (reference: out-of-row write), the rest of the row shifts by one."""
    from oracle.bindings import OracleHanabi
    from helpers import hanabi_informed_actions
    N = 96
    sim, orc = make_hanabi(N), OracleHanabi(N)
    rng = np.random.default_rng(5)
    for t in range(260):
        a = hanabi_informed_actions(rng, orc.state, orc.mask, orc.active)
        sim.step(a)
        orc.step(a)
        f = hanabi_fields(sim)
        for k in ("obs", "state", "mask", "active", "rewards", "done"):
            assert np.array_equal(f[k], getattr(orc, k)), (k, t)
    assert orc.max_info_encoded() > 8, "the policy never pushed the tokens past their maximum"


def test_hanabi_structure_at_baseline_size():
    """64K worlds (BASELINE config 4): one active agent per world, masks legal, obs is the state prefix."""
    import torch
    N = 65536
    sim = make_hanabi(N)
    gen = torch.Generator(device="cuda").manual_seed(1)
    for t in range(40):
        mask = sim.t("action_mask")
        active = sim.t("active_agent")
        assert bool((active.sum(0) == 1).all())
        logits = torch.rand(mask.shape, generator=gen, device="cuda")
        logits[mask == 0] = -1.0
        a = logits.argmax(-1, keepdim=True).to(torch.int32)
        assert bool((mask.gather(2, a.long()) == 1)[active == 1].all())
        sim.action.copy_(a)
        sim.sim.step()
        obs, state = sim.t("observation"), sim.t("agent_state")
        assert bool((obs == state[..., :658]).all())
        assert bool((sim.t("reward")[0] == sim.t("reward")[1]).all())


# ------------------------------------------------------------------------------------------------------------------
# Balance Beam / Cartpole
# ------------------------------------------------------------------------------------------------------------------
def test_balance_vs_golden_and_oracle():
    from madrona_rl_envs_playground_b200 import BalanceBeamSimulator, ExecMode
    from oracle.bindings import OracleBalance
    g = load_golden("balance")
    sim = GpuSim(BalanceBeamSimulator(ExecMode.CUDA, 0, g["actions"].shape[2]))
    assert np.array_equal(sim.np("observation"), g["obs0"])
    for t in range(g["actions"].shape[0]):
        sim.step(g["actions"][t])
        assert np.array_equal(sim.np("observation"), g["obs"][t]), t
        assert np.array_equal(sim.np("reward"), g["rewards"][t]), t
        assert np.array_equal(sim.np("done"), g["done"][t]), t
    for N in (16384, 1000):   # BASELINE config 2 size, and a ragged size; multi-tile episode numbering
        sim, orc = GpuSim(BalanceBeamSimulator(ExecMode.CUDA, 0, N)), OracleBalance(N)
        rng = np.random.default_rng(N)
        for t in range(40):
            assert np.array_equal(sim.np("observation"), orc.obs), (N, t)
            assert np.array_equal(sim.np("reward"), orc.rewards), (N, t)
            assert np.array_equal(sim.np("done"), orc.done), (N, t)
            a = rng.integers(0, 4, (2, N))
            sim.step(a)
            orc.step(a)


def test_cartpole_vs_golden_and_oracle():
    """Float parity: each step the oracle is re-seeded with the GPU state (the reference's own validate_step protocol,
    envs/cartpole_env.py:246-288), so the comparison is single-step: rtol 1e-5, done flags exact unless the CPU value
    sits within 1e-5 of a threshold."""
    from madrona_rl_envs_playground_b200 import CartpoleSimulator, ExecMode
    from oracle.bindings import OracleCartpole
    expect = np.load(__import__("os").path.join(__import__("helpers").GOLDEN, "rng_cartpole_reset_states.npy"))
    sim = GpuSim(CartpoleSimulator(ExecMode.CUDA, 0, expect.shape[0]))
    assert np.array_equal(sim.np("observation"), expect)   # RNG known answers: bit-exact

    g = load_golden("cartpole")
    N = g["actions"].shape[1]
    sim = GpuSim(CartpoleSimulator(ExecMode.CUDA, 0, N))
    assert np.array_equal(sim.np("observation"), g["state0"])
    for t in range(g["actions"].shape[0]):   # free-running against the golden trajectory
        sim.step(g["actions"][t])
        assert np.array_equal(sim.np("reset")[:, 0], g["done"][t]), t
        np.testing.assert_allclose(sim.np("observation"), g["state"][t], rtol=1e-5, atol=1e-6)

    for N in (1024, 5000):   # BASELINE config 1 size and a multi-tile size
        sim, orc = GpuSim(CartpoleSimulator(ExecMode.CUDA, 0, N)), OracleCartpole(N)
        rng = np.random.default_rng(N)
        assert np.array_equal(sim.np("observation"), orc.state)
        for t in range(300):
            a = rng.integers(0, 2, (N,))
            orc.write_state(sim.np("observation"))
            sim.step(a)
            orc.step(a)
            gs, gd = sim.np("observation"), sim.np("reset")[:, 0]
            same = gd == orc.done
            assert same.all(), (N, t, np.flatnonzero(~same))
            alive = orc.done == 0
            np.testing.assert_allclose(gs[alive], orc.state[alive], rtol=1e-5, atol=1e-7)
            # worlds that reset drew their new state from the same episode index -> bit-exact
            assert np.array_equal(gs[~alive], orc.state[~alive]), (N, t)
            assert (sim.np("reward") == 1.0).all()


def test_acrobat_vs_golden_and_oracle():
    """src/acrobat_env.  Float parity like Cartpole's: single-step comparisons from identical states
    (rtol 1e-5; the double pendulum is chaotic, free-running trajectories diverge), done flags exact unless the CPU
    value sits within 1e-5 of the termination threshold; reset states, the shared episode-length counter's truncations
    and the episode numbering bit-exact."""
    import torch
    from madrona_rl_envs_playground_b200 import AcrobatSimulator, ExecMode
    from oracle.bindings import OracleAcrobat

    def height(s):
        return -np.cos(s[:, 0]) - np.cos(s[:, 1] + s[:, 0])

    g = load_golden("acrobat")
    N = g["actions"].shape[1]
    sim = GpuSim(AcrobatSimulator(ExecMode.CUDA, 0, N))
    state = sim.t("observation")
    assert np.array_equal(sim.np("observation"), g["state0"])
    for t in range(g["actions"].shape[0]):
        state.copy_(torch.as_tensor(g["state_in"][t]))
        sim.step(g["actions"][t])
        assert np.array_equal(sim.np("reset")[:, 0], g["done"][t]), t
        np.testing.assert_allclose(sim.np("observation"), g["state"][t], rtol=1e-5, atol=1e-6)

    for N in (1, 257, 1024, 5000):   # single world, two tiles, the example size, many tiles
        sim, orc = GpuSim(AcrobatSimulator(ExecMode.CUDA, 0, N)), OracleAcrobat(N)
        state = sim.t("observation")
        rng = np.random.default_rng(N)
        assert np.array_equal(sim.np("observation"), orc.state)
        resets = terms = 0
        for t in range(150 if N > 1 else 1100):
            if t % 7 == 3:   # put ~1 % of the worlds near the upright position
                s = sim.np("observation")
                up = rng.random(N) < 0.01
                s[up] = np.float32([2.9, 0.1, 0.5, -0.5]) + rng.uniform(-0.2, 0.2, (int(up.sum()), 4)).astype(np.float32)
                state.copy_(torch.as_tensor(s))
            a = rng.integers(0, 3, (N,))
            before = sim.np("observation")
            orc.write_state(before)
            sim.step(a)
            orc.step(a)
            gs, gd = sim.np("observation"), sim.np("reset")[:, 0]
            assert (gd == orc.done).all(), (N, t, np.flatnonzero(gd != orc.done))
            alive = orc.done == 0
            np.testing.assert_allclose(gs[alive], orc.state[alive], rtol=1e-5, atol=2e-6)
            assert np.array_equal(gs[~alive], orc.state[~alive]), (N, t)   # same episode index -> same draws
            assert (sim.np("reward") == -1.0).all()
            resets += int(gd.sum())
        assert resets >= (150 * N) // 501 and resets > 0, (N, resets)


# ------------------------------------------------------------------------------------------------------------------
# Python env classes (the drop-in surface)
# ------------------------------------------------------------------------------------------------------------------
def test_env_classes_smoke():
    import torch
    from madrona_rl_envs_playground_b200.envs import (BalanceMadronaTorch, CartpoleMadronaTorch, HanabiMadrona,
                                                      OvercookedMadrona)
    from madrona_rl_envs_playground_b200.envs.hanabi_env import FULL_CONFIG
    from madrona_rl_envs_playground_b200.pantheonrl_extension import RandomVectorAgent
    N = 128
    env = OvercookedMadrona("cramped_room", N, 0, horizon=20)
    env.add_partner_agent(RandomVectorAgent(lambda: torch.randint(0, 6, (N, 1), device="cuda", dtype=torch.int32)))
    ob = env.reset()
    assert tuple(ob.obs.shape) == (N, 5, 4, 26) and ob.obs.dtype == torch.int8 and bool(ob.active.all())
    for t in range(25):
        ob, rew, done, info = env.step(torch.randint(0, 6, (N, 1), device="cuda", dtype=torch.int32))
        assert tuple(rew.shape) == (N,) and tuple(done.shape) == (N,)
        assert bool(done.all()) == (t == 19)

    fenv = OvercookedMadrona("cramped_room", N, 0, horizon=20, obs_dtype=torch.bfloat16)   # policy-ready observations
    ref_env = OvercookedMadrona("cramped_room", N, 0, horizon=20)
    fob = fenv.n_reset()
    assert fob[0].obs.dtype == torch.bfloat16 and torch.equal(fob[1].obs.float(), ref_env.n_reset()[1].obs.float())
    acts = torch.randint(0, 6, (2, N, 1), device="cuda", dtype=torch.int32)
    fob, frew, _, _ = fenv.n_step(acts)
    iob, irew, _, _ = ref_env.n_step(acts)
    assert all(torch.equal(f.obs.float(), i.obs.float()) for f, i in zip(fob, iob)) and torch.equal(frew, irew)

    from madrona_rl_envs_playground_b200.envs.overcooked2_env import OvercookedMadrona as Overcooked2Madrona
    env2 = Overcooked2Madrona("simple", N, 0, horizon=15)
    assert env2.horizon == 15 and tuple(env2.observation_space.shape) == (5, 4, 20)
    obs2 = env2.n_reset()
    assert tuple(obs2[1].obs.shape) == (N, 5, 4, 20) and obs2[0].obs.dtype == torch.int8
    for t in range(16):
        obs2, rew, done, _ = env2.n_step(torch.randint(0, 6, (2, N, 1), device="cuda", dtype=torch.int32))
        assert tuple(rew.shape) == (2, N) and bool(done.all()) == (t == 14)

    henv = HanabiMadrona(N, 0, config=FULL_CONFIG)
    obs = henv.n_reset()
    assert tuple(obs[0].obs.shape) == (N, 658) and tuple(obs[0].state.shape) == (N, 783)
    assert tuple(obs[0].action_mask.shape) == (N, 20) and bool(obs[0].active.all()) and not bool(obs[1].active.any())
    logits = torch.rand(2, N, 20, device="cuda")
    logits[torch.stack([o.action_mask for o in obs]) == 0] = -1
    obs, rew, done, _ = henv.n_step(logits.argmax(-1, keepdim=True).to(torch.int32))
    assert bool(obs[1].active.all()) and tuple(rew.shape) == (2, N)

    benv = BalanceMadronaTorch(N, 0)
    obs = benv.n_reset()
    assert tuple(obs[0].obs.shape) == (N, 7)
    obs, rew, done, _ = benv.n_step(torch.randint(0, 4, (2, N, 1), device="cuda", dtype=torch.int32))
    assert rew.dtype == torch.float32

    cenv = CartpoleMadronaTorch(N, 0)
    s0 = cenv.reset()
    s, r, d, _ = cenv.step(torch.randint(0, 2, (N,), device="cuda", dtype=torch.int32))
    assert tuple(s.shape) == (N, 4) and tuple(r.shape) == (N, 1) and tuple(d.shape) == (N,)


# ------------------------------------------------------------------------------------------------------------------
# Overcooked2 / "Simplecooked" (src/overcooked2_env): the same kernels under the JS-compatible rules
# ------------------------------------------------------------------------------------------------------------------
def make_overcooked2(params, N, lanes=0):
    from madrona_rl_envs_playground_b200 import ExecMode, SimplecookedSimulator
    sim = SimplecookedSimulator(ExecMode.CUDA, 0, N, **params)
    if lanes:
        sim.set_lanes(lanes)
    return GpuSim(sim)


@pytest.mark.parametrize("name", OVERCOOKED2_GOLDENS)
def test_overcooked2_vs_golden(name):
    from madrona_rl_envs_playground_b200.overcooked2_layouts import get_base_layout_params
    g = load_golden("overcooked2_" + name)
    params = get_base_layout_params(name, int(g["horizon"]))
    sim = make_overcooked2(params, g["actions"].shape[2])
    assert np.array_equal(obs_u8(sim), g["obs0"])
    for t in range(g["actions"].shape[0]):
        sim.step(g["actions"][t])
        assert np.array_equal(obs_u8(sim), g["obs"][t]), (name, t)
        assert np.array_equal(sim.np("reward"), g["rewards"][t]), (name, t)
        assert np.array_equal(sim.np("done"), g["done"][t]), (name, t)


@pytest.mark.parametrize("lanes", [1, 2, 4, 8])
@pytest.mark.parametrize("layout,npl,N,horizon", [
    ("simple", None, 333, 120),            # ragged last tile; soups get cooked, plated and served
    ("simple", 1, 64, 90),                 # one player: 1 lane per world allowed
    ("simple_tomato", None, 77, 150),      # tomato sources, whose terrain bit never shows; odd world count
    ("unident_s", None, 128, 200),         # two pots: dish-pickup shaping depends on both
])
def test_overcooked2_lanes_per_world_vs_oracle(lanes, layout, npl, N, horizon):
    from madrona_rl_envs_playground_b200 import simulators
    from madrona_rl_envs_playground_b200.overcooked2_layouts import get_base_layout_params
    from oracle.bindings import OracleOvercooked2
    params = get_base_layout_params(layout, horizon, npl)
    P = params["num_players"]
    if lanes < P:
        pytest.skip("fewer lanes than players")
    sim, orc = make_overcooked2(params, N, lanes), OracleOvercooked2(params, N)
    assert tuple(sim.sim.observation_tensor().to_torch().shape) == (P, N, params["width"], params["height"], 5 * P + 10)
    assert np.array_equal(obs_u8(sim), orc.obs)
    rng = np.random.default_rng(lanes + N)
    total = 0
    for t in range(420):
        a = biased_overcooked_actions(rng, (P, N))
        sim.step(a)
        obs, rew, done = orc.step(a)
        assert np.array_equal(sim.np("reward"), rew), (layout, t)
        assert np.array_equal(sim.np("done"), done), (layout, t)
        assert np.array_equal(obs_u8(sim), obs), (layout, t)
        total += int(rew.sum())
    assert total > 0
    with pytest.raises(simulators.L.B200EnvError):
        sim.sim.force_kernel(0)            # the CTA-per-world kernel does not implement these rules


@pytest.mark.parametrize("fused", [True, False])
def test_overcooked2_rollout_vs_oracle(fused):
    """the fused T-step kernel keeps the world's dishes-on-counters count on chip too"""
    import torch
    from madrona_rl_envs_playground_b200.overcooked2_layouts import get_base_layout_params
    from oracle.bindings import OracleOvercooked2
    horizon, T, A, N = 130, 300, 11, 200
    params = get_base_layout_params("simple", horizon)
    P, W, H = 2, params["width"], params["height"]
    C = 5 * P + 10
    g, orc = make_overcooked2(params, N), OracleOvercooked2(params, N)
    g.sim.set_rollout_fused(fused)
    rng = np.random.default_rng(3)
    acts = biased_overcooked_actions(rng, (A, P, N))
    d_act = torch.as_tensor(acts).cuda()
    obs = torch.zeros((T, P, N, W, H, C), dtype=torch.int8, device="cuda")
    rew = torch.zeros((T, P, N), dtype=torch.int32, device="cuda")
    done = torch.zeros((T, N), dtype=torch.int32, device="cuda")
    g.sim.rollout(None, T, d_act, obs, rew, done)
    torch.cuda.synchronize()
    obs, rew, done = obs.cpu().numpy().view(np.uint8), rew.cpu().numpy(), done.cpu().numpy()
    for t in range(T):
        o, r, d = orc.step(acts[t % A])
        assert np.array_equal(rew[t], r) and np.array_equal(done[t], d), t
        assert np.array_equal(obs[t], o), t
    assert rew.sum() > 0
    g.step(acts[5])                        # the on-chip state went back to HBM
    o, r, d = orc.step(acts[5])
    assert np.array_equal(obs_u8(g), o) and np.array_equal(g.np("reward"), r) and np.array_equal(g.np("done"), d)


# ------------------------------------------------------------------------------------------------------------------
# device-only random rollouts (b200env_random_rollout): replayed on the host against the oracle
# ------------------------------------------------------------------------------------------------------------------
def test_random_rollout_matches_host_replay():
    import torch
    from madrona_rl_envs_playground_b200 import (BalanceBeamSimulator, CartpoleSimulator, ExecMode, HanabiSimulator,
                                                 OvercookedSimulator)
    from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
    from oracle import bindings as ob
    seed = 77
    # the numpy hash equals the library's
    sim = CartpoleSimulator(ExecMode.CUDA, 0, 64)
    for step, idx in ((0, 0), (3, 17), (400, 65535), (2 ** 31, 2 ** 20 + 3)):
        assert int(random_policy_hash(seed, step, idx)) == sim.random_action_hash(seed, step, idx)
    sim.random_rollout(1, seed, first_step=9)
    torch.cuda.synchronize()
    assert np.array_equal(sim.action_tensor().to_torch().cpu().numpy().reshape(-1),
                          random_policy_actions(seed, 9, (64,), num_moves=2))
    sim.close()
    from madrona_rl_envs_playground_b200 import AcrobatSimulator
    sim = AcrobatSimulator(ExecMode.CUDA, 0, 300)   # the step kernel draws the action itself (torque -1 / 0 / +1)
    sim.random_rollout(3, seed, first_step=4)
    torch.cuda.synchronize()
    assert np.array_equal(sim.action_tensor().to_torch().cpu().numpy().reshape(-1),
                          random_policy_actions(seed, 6, (300,), num_moves=3))
    sim.close()

    N, T = 300, 150
    hb, orc = HanabiSimulator(ExecMode.CUDA, 0, N, 5, 5, 2, 8, 3), ob.OracleHanabi(N)
    hb.random_rollout(T, seed)
    torch.cuda.synchronize()
    for t in range(T):
        orc.step(random_policy_actions(seed, t, (2, N), mask=orc.mask))
    g = GpuSim(hb)
    assert np.array_equal(g.np("observation").view(np.uint8), orc.obs)
    assert np.array_equal(g.np("agent_state").view(np.uint8), orc.state)
    assert np.array_equal(g.np("action_mask"), orc.mask)
    assert np.array_equal(g.np("reward"), orc.rewards) and np.array_equal(g.np("done"), orc.done)
    hb.close()

    params = get_base_layout_params("cramped_room", 40)
    oc, orc = OvercookedSimulator(ExecMode.CUDA, 0, N, **params), ob.OracleOvercooked(params, N)
    oc.random_rollout(T, seed, first_step=5)
    torch.cuda.synchronize()
    for t in range(T):
        obs, rew, done = orc.step(random_policy_actions(seed, 5 + t, (2, N), num_moves=6))
    g = GpuSim(oc)
    assert np.array_equal(obs_u8(g), obs) and np.array_equal(g.np("reward"), rew) and np.array_equal(g.np("done"), done)
    oc.close()

    bb, orc = BalanceBeamSimulator(ExecMode.CUDA, 0, N), ob.OracleBalance(N)
    bb.random_rollout(T, seed)
    torch.cuda.synchronize()
    for t in range(T):
        orc.step(random_policy_actions(seed, t, (2, N), num_moves=4))
    g = GpuSim(bb)
    assert np.array_equal(g.np("observation"), orc.obs)
    assert np.array_equal(g.np("reward"), orc.rewards) and np.array_equal(g.np("done"), orc.done)
    bb.close()


@pytest.mark.parametrize("N", [300, 4500])
def test_hanabi_fused_random_rollout(N):
    """The cooperative T-step kernel (game records on chip, game warp running ahead of the writing warp, per-step
    episode numbering across tiles) against the oracle and against one launch per step: two rollouts back to back
    (step counter and episode counter carry over), then ordinary steps from the written-back state."""
    import torch
    from madrona_rl_envs_playground_b200 import ExecMode, HanabiSimulator
    from oracle import bindings as ob
    seed, T1, T2 = 5, 70, 45
    fused, plain, orc = (HanabiSimulator(ExecMode.CUDA, 0, N, 5, 5, 2, 8, 3) for _ in range(2)), None, ob.OracleHanabi(N)
    fused, plain = tuple(fused)
    plain.set_rollout_fused(False)
    for sim in (fused, plain):
        sim.random_rollout(T1, seed)
        sim.random_rollout(T2, seed, first_step=T1)
    torch.cuda.synchronize()
    for t in range(T1 + T2):
        orc.step(random_policy_actions(seed, t, (2, N), mask=orc.mask))
    gf, gp = GpuSim(fused), GpuSim(plain)
    for name, attr in (("observation", "obs"), ("agent_state", "state"), ("action_mask", "mask"),
                       ("active_agent", "active"), ("reward", "rewards"), ("done", "done")):
        a, b = gf.np(name), gp.np(name)
        assert np.array_equal(a, b), name
        ref = getattr(orc, attr)
        assert np.array_equal(a.view(np.uint8) if ref.dtype == np.uint8 else a, ref), name
    # the next move of the player to move is in the action tensor, like after per-step launches
    cur = orc.active.argmax(0)
    af, ap = gf.np("action").reshape(2, N), gp.np("action").reshape(2, N)
    assert np.array_equal(af[cur, np.arange(N)], ap[cur, np.arange(N)])
    # ordinary steps continue from the state the rollout wrote back (records, episode counter)
    rng = np.random.default_rng(2)
    for t in range(40):
        a = legal_random_actions(rng, orc.mask)
        gf.step(a)
        orc.step(a)
        assert np.array_equal(gf.np("observation").view(np.uint8), orc.obs), t
        assert np.array_equal(gf.np("done"), orc.done), t
    fused.close(); plain.close()


@pytest.mark.parametrize("rules,layout,N", [("overcooked", "cramped_room", 300), ("overcooked", "coordination_ring", 1000),
                                            ("simplecooked", "simple", 300)])
def test_overcooked_fused_random_rollout(rules, layout, N):
    """b200env_random_rollout on the fused T-step kernel (the kernel draws the moves itself: no sampler launches, no
    action ring): against the oracle replaying the same draws, and against one sampler + one step launch per step;
    two rollouts back to back across an episode boundary, the last draw left in the action tensor."""
    import torch
    from madrona_rl_envs_playground_b200 import ExecMode, OvercookedSimulator, SimplecookedSimulator
    from oracle import bindings as ob
    if rules == "overcooked":
        from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
        params = get_base_layout_params(layout, 60)
        make, orc = (lambda: OvercookedSimulator(ExecMode.CUDA, 0, N, **params)), ob.OracleOvercooked(params, N)
    else:
        from madrona_rl_envs_playground_b200.overcooked2_layouts import get_base_layout_params
        params = get_base_layout_params(layout, 60)
        make, orc = (lambda: SimplecookedSimulator(ExecMode.CUDA, 0, N, **params)), ob.OracleOvercooked2(params, N)
    seed, T1, T2 = 21, 50, 45
    fused, plain = make(), make()
    plain.set_rollout_fused(False)
    for sim in (fused, plain):
        sim.random_rollout(T1, seed, first_step=3)
        sim.random_rollout(T2, seed, first_step=3 + T1)
    torch.cuda.synchronize()
    P = params["num_players"]
    for t in range(T1 + T2):
        obs, rew, done = orc.step(random_policy_actions(seed, 3 + t, (P, N), num_moves=6))
    gf, gp = GpuSim(fused), GpuSim(plain)
    assert np.array_equal(obs_u8(gf), obs) and np.array_equal(gf.np("reward"), rew) and np.array_equal(gf.np("done"), done)
    for name in ("observation", "reward", "done", "action"):
        assert np.array_equal(gf.np(name), gp.np(name)), name
    assert np.array_equal(gf.np("action").reshape(P, N), random_policy_actions(seed, 3 + T1 + T2 - 1, (P, N), num_moves=6))
    rng = np.random.default_rng(8)
    for t in range(30):   # ordinary steps continue from the state the rollout wrote back (crosses the horizon)
        a = rng.integers(0, 6, size=(P, N)).astype(np.int32)
        gf.step(a)
        obs, rew, done = orc.step(a)
        assert np.array_equal(obs_u8(gf), obs) and np.array_equal(gf.np("done"), done), t
    fused.close(); plain.close()


@pytest.mark.parametrize("env", ["cartpole", "balance"])
@pytest.mark.parametrize("N", [64, 300, 40000, 310000])   # 310 000 worlds: too many tiles to be co-resident -> per-step fallback
def test_simple_env_fused_random_rollout(env, N):
    """Cartpole / Balance Beam: the cooperative T-step kernel (world in registers, per-step episode numbering across
    tiles) leaves bit-for-bit what one launch per step leaves — two rollouts back to back, then ordinary steps from the
    written-back state (hidden state, episode counter); Balance Beam also against the oracle."""
    import torch
    from madrona_rl_envs_playground_b200 import BalanceBeamSimulator, CartpoleSimulator, ExecMode
    from oracle import bindings as ob
    cls, moves, players = (CartpoleSimulator, 2, 1) if env == "cartpole" else (BalanceBeamSimulator, 4, 2)
    seed, T1, T2 = 11, 90, 37
    fused, plain = cls(ExecMode.CUDA, 0, N), cls(ExecMode.CUDA, 0, N)
    plain.set_rollout_fused(False)
    for sim in (fused, plain):
        sim.random_rollout(T1, seed)
        sim.random_rollout(T2, seed, first_step=T1)
    torch.cuda.synchronize()
    gf, gp = GpuSim(fused), GpuSim(plain)
    names = ("observation", "reward", "done", "action")
    for name in names:
        assert np.array_equal(gf.np(name), gp.np(name)), name
    assert np.array_equal(gf.np("action").reshape(-1),
                          random_policy_actions(seed, T1 + T2 - 1, (players * N,), num_moves=moves))
    if env == "balance" and N <= 300:
        orc = ob.OracleBalance(N)
        for t in range(T1 + T2):
            orc.step(random_policy_actions(seed, t, (2, N), num_moves=4))
        assert np.array_equal(gf.np("observation"), orc.obs)
        assert np.array_equal(gf.np("reward"), orc.rewards) and np.array_equal(gf.np("done"), orc.done)
    rng = np.random.default_rng(4)
    for t in range(60):   # enough steps for every Balance Beam world to reset (3-step episodes) and many poles to fall
        a = rng.integers(0, moves, size=gf.np("action").shape).astype(np.int32)
        gf.step(a)
        gp.step(a)
        for name in names[:3]:
            assert np.array_equal(gf.np(name), gp.np(name)), (name, t)
    fused.close(); plain.close()


def test_hanabi_fused_rollout_at_baseline_size():
    """BASELINE config 4 (65 536 worlds: every SM holds its 14 tiles, the memory system is saturated): the fused
    rollout must leave exactly the tensors, game records and episode counter that one launch per step leaves."""
    import torch
    from madrona_rl_envs_playground_b200 import ExecMode, HanabiSimulator
    N, seed = 65536, 11
    fused, plain = (HanabiSimulator(ExecMode.CUDA, 0, N, 5, 5, 2, 8, 3) for _ in range(2))
    plain.set_rollout_fused(False)
    first = 0
    for T in (150, 1, 77):
        for sim in (fused, plain):
            sim.random_rollout(T, seed, first_step=first)
        first += T
        torch.cuda.synchronize()
        gf, gp = GpuSim(fused), GpuSim(plain)
        for name in ("observation", "agent_state", "action_mask", "active_agent", "reward", "done"):
            assert torch.equal(gf.t(name), gp.t(name)), (name, first)
        # checkpoint blob = game records | action tensor | exported tensors | episode counter: identical except for the
        # action slot of the player who is NOT to move (ignored by the step; each path leaves its own stale value)
        bf, bp = fused.get_state(), plain.get_state()
        G, A = 25 * 4 * N, 2 * 4 * N   # GAME_WORDS = 25 (13 words of deck, 12 of record: csrc/hanabi_game.cuh)
        assert bf[:G] == bp[:G] and bf[G + A:] == bp[G + A:], first
        cur = gf.t("active_agent").argmax(0)
        af, ap = gf.t("action").reshape(2, N), gp.t("action").reshape(2, N)
        assert torch.equal(af.gather(0, cur[None]), ap.gather(0, cur[None])), first
    assert int(gf.t("done").sum()) > 0
    fused.close(); plain.close()


def test_simulators_on_two_devices_in_one_process():
    """One process may drive simulators on several GPUs (the reference passes gpu_id per Manager): kernels needing
    opted-in shared memory must work on every device.  Skipped on single-GPU boxes."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from madrona_rl_envs_playground_b200 import ExecMode, HanabiSimulator, OvercookedSimulator
    from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
    from oracle.bindings import OracleOvercooked
    params = get_base_layout_params("many_player_layout", 30)      # CTA-per-world kernel: 110 KB of dynamic smem
    P = params["num_players"]
    sims = [OvercookedSimulator(ExecMode.CUDA, d, 3, **params) for d in (0, 1)]
    orc = OracleOvercooked(params, 3)
    rng = np.random.default_rng(3)
    for t in range(8):
        a = biased_overcooked_actions(rng, (P, 3))
        obs, rew, done = orc.step(a)
        for d, sim in enumerate(sims):
            act = sim.action_tensor().to_torch()
            assert act.device.index == d
            act.copy_(torch.as_tensor(a).reshape(act.shape).to(act.device))
            sim.step()
            assert np.array_equal(sim.observation_tensor().to_torch().cpu().numpy().view(np.uint8), obs), (d, t)
    hb = HanabiSimulator(ExecMode.CUDA, 1, 64, 5, 5, 2, 8, 3)
    hb.random_rollout(20, 1)
    torch.cuda.synchronize(1)
    assert hb.observation_tensor().to_torch().device.index == 1


# ------------------------------------------------------------------------------------------------------------------
# edge cases: single world, single agent, horizon 1 (reset on every step), tiny RNG-env batches
# ------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("layout,npl,N,horizon", [
    ("cramped_room", None, 1, 7),
    ("cramped_room", 1, 1, 3),
    ("cramped_room", None, 5, 1),          # done on every step: obs is always the reset state, rewards still count
    ("coordination_ring", None, 2, 2),
    ("many_player_layout", None, 1, 2),
])
def test_overcooked_edge_sizes(layout, npl, N, horizon):
    import torch
    from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
    from oracle.bindings import OracleOvercooked
    params = get_base_layout_params(layout, horizon, npl)
    P = params["num_players"]
    sim, orc = make_overcooked(params, N), OracleOvercooked(params, N)
    assert np.array_equal(obs_u8(sim), orc.obs)
    rng = np.random.default_rng(horizon)
    acts = biased_overcooked_actions(rng, (12, P, N))
    for t in range(12):
        sim.step(acts[t])
        obs, rew, done = orc.step(acts[t])
        assert np.array_equal(obs_u8(sim), obs) and np.array_equal(sim.np("reward"), rew), (layout, t)
        assert np.array_equal(sim.np("done"), done), (layout, t)
    # and the same 12 steps again through the rollout entry point
    sim.sim.rollout(None, 12, torch.as_tensor(acts).cuda())
    torch.cuda.synchronize()
    for t in range(12):
        obs, rew, done = orc.step(acts[t])
    assert np.array_equal(obs_u8(sim), obs) and np.array_equal(sim.np("reward"), rew)
    assert np.array_equal(sim.np("done"), done)


@pytest.mark.parametrize("N", [1, 31, 257])
def test_rng_envs_small_batches(N):
    """Episode numbering across ragged tiles: Hanabi (32-world tiles) and Balance Beam (256-world tiles)."""
    from madrona_rl_envs_playground_b200 import BalanceBeamSimulator, ExecMode
    from oracle.bindings import OracleBalance, OracleHanabi
    sim, orc = make_hanabi(N), OracleHanabi(N)
    rng = np.random.default_rng(N)
    for t in range(150):
        a = legal_random_actions(rng, orc.mask)
        sim.step(a)
        orc.step(a)
        f = hanabi_fields(sim)
        for k in ("obs", "state", "mask", "active", "rewards", "done"):
            assert np.array_equal(f[k], getattr(orc, k)), (k, t)
    bb, orb = GpuSim(BalanceBeamSimulator(ExecMode.CUDA, 0, N)), OracleBalance(N)
    for t in range(60):
        a = rng.integers(0, 4, (2, N))
        bb.step(a)
        orb.step(a)
        assert np.array_equal(bb.np("observation"), orb.obs), t
        assert np.array_equal(bb.np("reward"), orb.rewards) and np.array_equal(bb.np("done"), orb.done), t


@pytest.mark.parametrize("dtype_name", ["bfloat16", "float32"])
@pytest.mark.parametrize("N,cfg", [(300, {}), (70, dict(colors=3, ranks=4, max_information_tokens=5, max_life_tokens=2))])
def test_hanabi_typed_rows_vs_oracle(dtype_name, N, cfg):
    """Policy-ready Hanabi rows (b200env_hanabi_create_typed): the bfloat16 / float tensors hold exactly the oracle's
    bytes as numbers (tolerance 0), per-step launches with resets, then the fused rollout against the int8 simulator."""
    import torch
    from madrona_rl_envs_playground_b200 import ExecMode, HanabiSimulator
    from oracle.bindings import OracleHanabi
    dtype = getattr(torch, dtype_name)
    full = dict(colors=5, ranks=5, players=2, max_information_tokens=8, max_life_tokens=3)
    full.update(cfg)
    typed = HanabiSimulator(ExecMode.CUDA, 0, N, obs_dtype=dtype, **full)
    orc = OracleHanabi(N, **full)
    g = GpuSim(typed)
    obs_t, state_t = typed.observation_tensor().to_torch(), typed.agent_state_tensor().to_torch()
    assert obs_t.dtype == dtype and tuple(obs_t.shape) == (2, N, 658) and tuple(state_t.shape) == (2, N, 783)
    rng = np.random.default_rng(N)

    from helpers import hanabi_row_lengths
    obs_len, state_len = hanabi_row_lengths(full)   # (658, 783) for the standard game: the whole row

    def check(t):
        assert np.array_equal(obs_t.float().cpu().numpy()[..., :obs_len], orc.obs[..., :obs_len].astype(np.float32)), ("obs", t)
        assert np.array_equal(state_t.float().cpu().numpy()[..., :state_len],
                              orc.state[..., :state_len].astype(np.float32)), ("state", t)
        assert np.array_equal(g.np("action_mask"), orc.mask) and np.array_equal(g.np("done"), orc.done), t

    resets = 0
    for t in range(90):
        check(t)
        a = legal_random_actions(rng, orc.mask)
        g.step(a)
        orc.step(a)
        resets += int(orc.done.sum())
    check(90)
    assert resets > 0
    typed.close()
    # fused rollout (standard game only goes through the cooperative kernel's <true, E> instantiation; others <false, E>)
    a8 = HanabiSimulator(ExecMode.CUDA, 0, N, **full)
    a16 = HanabiSimulator(ExecMode.CUDA, 0, N, obs_dtype=dtype, **full)
    for sim in (a8, a16):
        sim.random_rollout(60, 3)
        sim.random_rollout(25, 3, first_step=60)
    torch.cuda.synchronize()
    for name, n in (("observation", obs_len), ("agent_state", state_len)):
        ref, got = getattr(a8, name + "_tensor")().to_torch(), getattr(a16, name + "_tensor")().to_torch()
        assert torch.equal(got.float()[..., :n], ref.float()[..., :n]), name
    for name in ("action_mask", "reward", "done", "active_agent"):
        assert torch.equal(getattr(a8, name + "_tensor")().to_torch(), getattr(a16, name + "_tensor")().to_torch()), name
    a8.close(); a16.close()
