#!/usr/bin/env python
"""The reference's example scripts (scripts/{overcooked,hanabi,balance,cartpole}_example.py) on the B200 build: same
protocol — 5 warm-up steps, then --num-steps steps with fresh random actions each step through the env class's
n_step()/step(), wall clock around that call only, metric ``step * worlds / sec`` (scripts/overcooked_example.py:87-116)
— and the same pass criterion with --validation: ``Error rate: 0.0``.

    python tests/env_example.py --env overcooked --layout cramped_room --num-envs 1000 --num-steps 1000 [--validation]
    python tests/env_example.py --env hanabi --num-envs 1000 --validation

--validation steps the CPU oracle (oracle/, test infrastructure: the plain-C restatement pinned to the reference's
Madrona CPU exec mode) with the same actions and compares every observation / reward / done exactly (Cartpole: 1e-5
relative from the previous state, as envs/cartpole_env.py:246-288 does).  Hanabi picks uniformly random LEGAL moves by
masked argmax of random logits like scripts/hanabi_example.py:64-67.
"""
import argparse
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))   # lives under tests/: --validation loads the oracle
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--env", default="overcooked", choices=["overcooked", "hanabi", "balance", "cartpole"])
    ap.add_argument("--layout", default="cramped_room")
    ap.add_argument("--num-players", type=int, default=None)
    ap.add_argument("--num-envs", type=int, default=32)
    ap.add_argument("--num-steps", type=int, default=1000)
    ap.add_argument("--horizon", type=int, default=400)
    ap.add_argument("--validation", action="store_true")
    ap.add_argument("--asserts", action="store_true")
    args = ap.parse_args()
    from madrona_rl_envs_playground_b200.envs import (BalanceMadronaTorch, CartpoleMadronaTorch, HanabiMadrona,
                                                      OvercookedMadrona)
    from madrona_rl_envs_playground_b200.envs.hanabi_env import FULL_CONFIG
    N = args.num_envs
    orc = None
    if args.env == "overcooked":
        env = OvercookedMadrona(args.layout, N, 0, horizon=args.horizon, num_players=args.num_players)
        n_act, P = env.action_space.n, env.n_players
    elif args.env == "hanabi":
        env = HanabiMadrona(N, 0, config=FULL_CONFIG)
        n_act, P = env.action_space.n, 2
    elif args.env == "balance":
        env = BalanceMadronaTorch(N, 0)
        n_act, P = 4, 2
    else:
        env = CartpoleMadronaTorch(N, 0)
        n_act, P = 2, 1
    if args.validation:
        from oracle import bindings as ob
        orc = {"overcooked": lambda: ob.OracleOvercooked(env.base_layout_params, N),
               "hanabi": lambda: ob.OracleHanabi(N), "balance": lambda: ob.OracleBalance(N),
               "cartpole": lambda: ob.OracleCartpole(N)}[args.env]()
    dev = torch.device("cuda", 0)
    gen = torch.Generator(device=dev).manual_seed(0)

    def sample():
        if args.env == "cartpole":
            return torch.randint(0, 2, (N,), generator=gen, device=dev)
        if args.env == "hanabi":
            mask = env.sim.action_mask_tensor().to_torch()                      # [2, N, 20]
            logits = torch.rand(mask.shape, generator=gen, device=dev)
            logits[mask == 0] = -float("inf")
            return logits.argmax(dim=-1, keepdim=True)
        return torch.randint(0, n_act, (P, N, 1), generator=gen, device=dev)

    def validate(actions, out):
        a = actions.cpu().numpy().reshape(P, N) if args.env != "cartpole" else actions.cpu().numpy()
        if args.env == "overcooked":
            obs, rew, done = orc.step(a)
            got = np.stack([o.obs.cpu().numpy() for o in out[0]]).view(np.uint8)
            return (np.array_equal(got, obs) and np.array_equal(out[1].cpu().numpy(), rew)
                    and np.array_equal(out[2].cpu().numpy(), done))
        if args.env == "hanabi":
            orc.step(a)
            s = env.sim
            return (np.array_equal(s.observation_tensor().to_torch().cpu().numpy().view(np.uint8), orc.obs)
                    and np.array_equal(s.agent_state_tensor().to_torch().cpu().numpy().view(np.uint8), orc.state)
                    and np.array_equal(s.action_mask_tensor().to_torch().cpu().numpy(), orc.mask)
                    and np.array_equal(s.reward_tensor().to_torch().cpu().numpy(), orc.rewards)
                    and np.array_equal(s.done_tensor().to_torch().cpu().numpy(), orc.done))
        if args.env == "balance":
            orc.step(a)
            s = env.sim
            return (np.array_equal(s.observation_tensor().to_torch().cpu().numpy(), orc.obs)
                    and np.array_equal(s.reward_tensor().to_torch().cpu().numpy(), orc.rewards)
                    and np.array_equal(s.done_tensor().to_torch().cpu().numpy(), orc.done))
        orc.write_state(validate.prev)                 # cartpole: one step from the previous CUDA state
        orc.step(a)
        state = env.sim.observation_tensor().to_torch().cpu().numpy()
        alive = orc.done == 0
        ok = np.allclose(state[alive], orc.state[alive], rtol=1e-5, atol=1e-7)
        validate.prev = state.copy()
        return ok

    if args.env == "cartpole" and args.validation:
        validate.prev = env.sim.observation_tensor().to_torch().cpu().numpy().copy()
    step = env.step if args.env == "cartpole" else env.n_step
    errors = 0
    stamps = []
    for it in range(5 + args.num_steps):
        actions = sample()
        torch.cuda.synchronize()
        t0 = time.time()
        out = step(actions)
        torch.cuda.synchronize()      # the packaged env classes enqueue the step on the current stream: time it all
        t1 = time.time()
        if it >= 5:
            stamps.append(t1 - t0)
        if args.validation and not validate(actions, out):
            errors += 1
            print("MISMATCH at step", it)
            assert not args.asserts
    print("step * worlds / sec:", N / (sum(stamps) / len(stamps)))
    if args.validation:
        print("Error rate:", errors / (5 + args.num_steps))
    return errors


if __name__ == "__main__":
    sys.exit(1 if main() else 0)
