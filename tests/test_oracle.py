"""CPU tests: the plain-C oracle against the golden vectors generated from the reference's Madrona CPU exec mode
(tests/golden/generate.py) and, where oracle/_ref is built, against the reference live on fresh random streams."""
import json
import os

import numpy as np
import pytest

from helpers import (GOLDEN, OVERCOOKED2_GOLDENS, OVERCOOKED_GOLDENS, biased_overcooked_actions, legal_random_actions, load_golden,
                     overcooked_params_from_golden)
from oracle import bindings as ob


def test_layout_parser_matches_reference_golden():
    from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
    golden = json.load(open(os.path.join(GOLDEN, "layout_params.json")))
    assert len(golden) >= 15
    for key, expect in golden.items():
        name, npl = key.split("|")
        got = get_base_layout_params(name, 400, None if npl == "None" else int(npl))
        assert got == expect, key


def test_layout_dimensions():
    from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
    p = get_base_layout_params("cramped_room", 400)
    assert (p["width"], p["height"], p["num_players"]) == (5, 4, 2)
    assert p["recipe_values"][12] == 20 and sum(p["recipe_values"]) == 20   # only 3 onions pays
    p = get_base_layout_params("many_player_layout", 400)
    assert (p["width"], p["height"], p["num_players"]) == (15, 17, 40)
    assert get_base_layout_params("many_player_layout", 400, 8)["num_players"] == 8


@pytest.mark.parametrize("name", OVERCOOKED_GOLDENS)
def test_overcooked_oracle_vs_golden(name):
    g = load_golden("overcooked_" + name)
    params = overcooked_params_from_golden(g)
    N = g["actions"].shape[2]
    o = ob.OracleOvercooked(params, N)
    assert np.array_equal(o.obs, g["obs0"])
    for t in range(g["actions"].shape[0]):
        obs, rew, done = o.step(g["actions"][t])
        assert np.array_equal(obs, g["obs"][t]), (name, t)
        assert np.array_equal(rew, g["rewards"][t]), (name, t)
        assert np.array_equal(done, g["done"][t]), (name, t)


def test_overcooked2_layout_parser_matches_reference_golden():
    from madrona_rl_envs_playground_b200.overcooked2_layouts import get_base_layout_params
    golden = json.load(open(os.path.join(GOLDEN, "layout2_params.json")))
    assert len(golden) >= 12
    for key, expect in golden.items():
        name, npl = key.split("|")
        assert get_base_layout_params(name, 200, None if npl == "None" else int(npl)) == expect, key
    p = get_base_layout_params("simple_tomato", 200)
    assert p["terrain"][5] == 6 and p["dish_pickup_rew"] == 3      # 'T' is 6 in Simplecooked's enum; dish shaping on


@pytest.mark.parametrize("name", OVERCOOKED2_GOLDENS)
def test_overcooked2_oracle_vs_golden(name):
    from madrona_rl_envs_playground_b200.overcooked2_layouts import get_base_layout_params
    g = load_golden("overcooked2_" + name)
    params = get_base_layout_params(name, int(g["horizon"]))
    o = ob.OracleOvercooked2(params, g["actions"].shape[2])
    assert np.array_equal(o.obs, g["obs0"])
    for t in range(g["actions"].shape[0]):
        obs, rew, done = o.step(g["actions"][t])
        assert np.array_equal(obs, g["obs"][t]), (name, t)
        assert np.array_equal(rew, g["rewards"][t]), (name, t)
        assert np.array_equal(done, g["done"][t]), (name, t)


def test_hanabi_oracle_vs_golden():
    g = load_golden("hanabi_full")
    N = g["actions"].shape[2]
    o = ob.OracleHanabi(N)
    assert np.array_equal(o.obs, g["obs0"]) and np.array_equal(o.state, g["state0"])
    assert np.array_equal(o.mask, g["mask0"]) and np.array_equal(o.active, g["active0"])
    for t in range(g["actions"].shape[0]):
        o.step(g["actions"][t])
        for k in ("obs", "state", "mask", "active", "rewards", "done"):
            assert np.array_equal(getattr(o, k), g[k][t]), (k, t)


def test_balance_oracle_vs_golden():
    g = load_golden("balance")
    o = ob.OracleBalance(g["actions"].shape[2])
    assert np.array_equal(o.obs, g["obs0"])
    for t in range(g["actions"].shape[0]):
        o.step(g["actions"][t])
        assert np.array_equal(o.obs, g["obs"][t]), t
        assert np.array_equal(o.rewards, g["rewards"][t]), t
        assert np.array_equal(o.done, g["done"][t]), t


def test_cartpole_oracle_vs_golden():
    g = load_golden("cartpole")
    o = ob.OracleCartpole(g["actions"].shape[1])
    assert np.array_equal(o.state, g["state0"])
    for t in range(g["actions"].shape[0]):
        o.step(g["actions"][t])
        # same compiler flags (-ffp-contract=off) and libm as the reference build: bit-exact on this host
        assert np.array_equal(o.done, g["done"][t]), t
        np.testing.assert_allclose(o.state, g["state"][t], rtol=1e-6, atol=1e-7)


def test_acrobat_oracle_vs_golden():
    """src/acrobat_env: trajectories from the reference build with states injected near the upright position
    (tests/golden/generate.py gen_acrobat) — terminations, truncations of the Manager-wide counter, ordered resets."""
    g = load_golden("acrobat")
    o = ob.OracleAcrobat(g["actions"].shape[1])
    assert np.array_equal(o.state, g["state0"])
    for t in range(g["actions"].shape[0]):
        o.write_state(g["state_in"][t])
        o.step(g["actions"][t])
        assert np.array_equal(o.done, g["done"][t]), t
        np.testing.assert_allclose(o.state, g["state"][t], rtol=1e-6, atol=1e-7)
    assert g["done"].sum() > 50


def test_rng_known_answers():
    """episode index i -> seed -> first four draws, through the reference Cartpole reset (rng.hpp:7-36)."""
    expect = np.load(os.path.join(GOLDEN, "rng_cartpole_reset_states.npy"))
    o = ob.OracleCartpole(expect.shape[0])
    assert np.array_equal(o.state, expect)


needs_ref = pytest.mark.skipif(not ob.ref_available("overcooked"), reason="oracle/_ref not built (no /root/reference)")


@needs_ref
@pytest.mark.parametrize("layout,npl,N", [("cramped_room", None, 48), ("many_player_layout", None, 2),
                                          ("many_player_layout", 5, 8)])
def test_overcooked_oracle_vs_reference_live(layout, npl, N):
    from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
    params = get_base_layout_params(layout, 35, npl)
    P = params["num_players"]
    r, o = ob.RefOvercooked(params, N), ob.OracleOvercooked(params, N)
    rng = np.random.default_rng(7)
    assert np.array_equal(r.obs, o.obs)
    for t in range(80):
        a = biased_overcooked_actions(rng, (P, N))
        ro, rr, rd = r.step(a)
        oo, orr, od = o.step(a)
        assert np.array_equal(ro, oo) and np.array_equal(rr, orr) and np.array_equal(rd, od), t
    r.close()


@pytest.mark.skipif(not ob.ref_available("overcooked2"), reason="oracle/_ref not built (no /root/reference)")
@pytest.mark.parametrize("layout", ["simple", "simple_tomato", "unident_s"])
def test_overcooked2_oracle_vs_reference_live(layout):
    """long enough for soups to be cooked, plated and served, and for the dish-pickup shaping reward to fire"""
    from madrona_rl_envs_playground_b200.overcooked2_layouts import get_base_layout_params
    params = get_base_layout_params(layout, 300)
    N = 48
    r, o = ob.RefOvercooked2(params, N), ob.OracleOvercooked2(params, N)
    rng = np.random.default_rng(17)
    assert np.array_equal(r.obs, o.obs)
    total = 0
    for t in range(450):
        a = biased_overcooked_actions(rng, (2, N))
        ro, rr, rd = r.step(a)
        oo, orr, od = o.step(a)
        assert np.array_equal(ro, oo) and np.array_equal(rr, orr) and np.array_equal(rd, od), t
        total += int(rr.sum())
    assert total > 500
    r.close()


@needs_ref
def test_rng_envs_oracle_vs_reference_live():
    N = 96
    rng = np.random.default_rng(8)
    r, o = ob.RefHanabi(N), ob.OracleHanabi(N)
    r.read()
    for t in range(150):
        for k in ("obs", "state", "mask", "active", "rewards", "done"):
            assert np.array_equal(getattr(r, k), getattr(o, k)), (k, t)
        a = legal_random_actions(rng, r.mask)
        r.step(a); o.step(a)
    r.close()
    r, o = ob.RefBalance(N), ob.OracleBalance(N)
    r.read()
    for t in range(60):
        for k in ("obs", "rewards", "done"):
            assert np.array_equal(getattr(r, k), getattr(o, k)), (k, t)
        a = rng.integers(0, 4, (2, N))
        r.step(a); o.step(a)
    r.close()
    r, o = ob.RefCartpole(N), ob.OracleCartpole(N)
    r.read()
    for t in range(150):
        for k in ("state", "rewards", "done"):
            assert np.array_equal(getattr(r, k), getattr(o, k)), (k, t)
        a = rng.integers(0, 2, (N,))
        r.step(a); o.step(a)
    r.close()
    for n_worlds in (7, 700):   # 700 worlds: the shared episode-length counter passes 500 inside every step
        r, o = ob.RefAcrobat(n_worlds), ob.OracleAcrobat(n_worlds)
        r.read()
        for t in range(400):
            for k in ("state", "rewards", "done"):
                assert np.array_equal(getattr(r, k), getattr(o, k)), (k, t)
            if t % 50 == 49:   # a few worlds near the upright position, so that some terminate
                s = o.state.copy()
                s[:: 5] = np.float32([2.9, 0.1, 0.5, -0.5])
                r.write_state(s); o.write_state(s)
            a = rng.integers(0, 3, (n_worlds,))
            r.step(a); o.step(a)
        r.close()


DUMMY_MDP_GOLDENS = ["cramped_room", "coordination_ring", "multiplayer_schelling", "asymmetric_advantages_tomato"]


@pytest.mark.parametrize("name", DUMMY_MDP_GOLDENS)
def test_overcooked_oracle_vs_dummy_mdp_golden(name):
    """Second, independent pin: the reference's numpy DummyMDP (envs/overcooked_reimplement.py:119-377), the
    oracle its own validate_step uses (envs/overcooked_env.py:478-527): exact obs / reward / done, resets included.
    Vectors from tests/golden/generate_dummy_mdp.py."""
    g = load_golden("dummy_mdp_" + name)
    params = overcooked_params_from_golden(g)
    N = g["actions"].shape[2]
    o = ob.OracleOvercooked(params, N)
    assert np.array_equal(o.obs, g["obs0"])
    for t in range(g["actions"].shape[0]):
        obs, rew, done = o.step(g["actions"][t])
        assert np.array_equal(obs, g["obs"][t]), (name, t)
        assert np.array_equal(rew, g["rewards"][t]), (name, t)
        assert np.array_equal(done, g["done"][t]), (name, t)
