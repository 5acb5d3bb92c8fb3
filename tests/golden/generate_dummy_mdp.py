"""Generate tests/golden/dummy_mdp_*.npz from the reference's SECOND, independent Overcooked implementation: the numpy
`DummyMDP` (/root/reference/envs/overcooked_reimplement.py:119-377), the oracle the reference's own validation uses
(`validate_step`, envs/overcooked_env.py:478-527, through `SimplifiedOvercooked`, :372-425).  Run in the build container
only (`python tests/golden/generate_dummy_mdp.py`; needs /root/reference; numpy only).

The driver below restates SimplifiedOvercooked.n_step / n_reset / get_obs and validate_step's reset-on-done (those
classes import gym): obs = lossless_state_encoding -> reshape (H, W, C) -> transpose (W, H, C); team reward = sum of the
per-player rewards; done = timestep >= horizon, after which the env is reset and the RESET observation is what the
step returns (envs/overcooked_env.py:512-514)."""
import importlib.util
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params   # noqa: E402  (pinned vs the reference's parser)

spec = importlib.util.spec_from_file_location("overcooked_reimplement", "/root/reference/envs/overcooked_reimplement.py")
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)


def obs_of(mdp, state):
    per_player = mdp.lossless_state_encoding(state)
    return np.stack([o[:mdp.size].reshape(mdp.height, mdp.width, -1).transpose(1, 0, 2) for o in per_player]).astype(np.uint8)


def generate(layout, npl, N, horizon, T, seed):
    params = get_base_layout_params(layout, horizon, npl)
    P = params["num_players"]
    mdps = [ref.DummyMDP(**params) for _ in range(N)]
    states = [m.get_standard_start_state() for m in mdps]
    rng = np.random.default_rng(seed)
    actions = rng.choice(6, size=(T, P, N), p=[0.14, 0.14, 0.14, 0.14, 0.04, 0.40]).astype(np.int8)
    obs0 = np.stack([obs_of(m, s) for m, s in zip(mdps, states)], axis=1)   # [P, N, W, H, C]
    obs, rew, done = [], [], []
    for t in range(T):
        o_t, r_t, d_t = [], [], []
        for n, m in enumerate(mdps):
            states[n], rewards = m.get_state_transition(states[n], [int(a) for a in actions[t, :, n]])
            d = states[n].timestep >= m.horizon
            if d:
                states[n] = m.get_standard_start_state()
            o_t.append(obs_of(m, states[n])); r_t.append(sum(rewards)); d_t.append(int(d))
        obs.append(np.stack(o_t, axis=1)); rew.append(np.tile(np.array(r_t, np.int32), (P, 1))); done.append(np.array(d_t, np.int32))
    name = f"dummy_mdp_{layout}.npz"
    np.savez_compressed(os.path.join(HERE, name), layout=layout, num_players=-1 if npl is None else npl, horizon=horizon,
                        actions=actions, obs0=obs0, obs=np.stack(obs), rewards=np.stack(rew), done=np.stack(done))
    print(name, "P", P, "reward steps", int((np.stack(rew)[:, 0] != 0).sum()), "resets", int(np.stack(done).sum()))


if __name__ == "__main__":
    # long enough for soups to be filled, cooked, plated and served by random play (rewards of every kind)
    generate("cramped_room", None, 8, 150, 330, 11)
    generate("coordination_ring", None, 6, 150, 320, 12)
    generate("multiplayer_schelling", None, 3, 90, 200, 13)
    generate("asymmetric_advantages_tomato", None, 4, 150, 320, 14)
