#!/usr/bin/env python
"""Time Overcooked rollouts on one GPU: fused T-step kernel vs one launch per step, for each lanes-per-world value.

    python tools/rollout_sweep.py [--worlds 32768] [--steps 2000] [--layout cramped_room]

CUDA events on the launch stream, obs written to a ring of 8 buffers (> L2).  Tuning aid; bench.py is the contract."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from madrona_rl_envs_playground_b200 import ExecMode, OvercookedSimulator, _lib  # noqa: E402
from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--worlds", type=int, default=32768)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--layout", default="cramped_room")
    ap.add_argument("--players", type=int, default=None)
    ap.add_argument("--lanes", type=int, nargs="*", default=[2, 4, 8])
    ap.add_argument("--ring", type=int, default=8)
    ap.add_argument("--launches", type=int, default=1, help="back-to-back rollout calls of --steps steps per timed region")
    ap.add_argument("--obs-dtype", default="int8", choices=["int8", "bfloat16", "float32"],
                    help="observation element type (bfloat16 / float32 = policy-ready observations)")
    args = ap.parse_args()
    dtype = getattr(torch, args.obs_dtype)
    es = {"int8": 1, "bfloat16": 2, "float32": 4}[args.obs_dtype]
    params = get_base_layout_params(args.layout, 400, args.players)
    P, W, H = params["num_players"], params["width"], params["height"]
    C, N = 5 * P + 16, args.worlds
    per_step = N * (P * W * H * C * es + 4 * P + 4 * P + 4)
    act = torch.randint(0, 6, (64, P, N), device="cuda", dtype=torch.int32)
    obs = torch.empty((args.ring, P, N, W, H, C), dtype=dtype, device="cuda")
    stream = torch.cuda.Stream()
    for lanes in args.lanes:
        sim = OvercookedSimulator(ExecMode.CUDA, 0, N, obs_dtype=dtype, **params)
        try:
            sim.set_lanes(lanes)
        except _lib.B200EnvError as e:
            print(f"lanes {lanes}: {e}")
            continue
        for fused in (False, True):
            sim.set_rollout_fused(fused)
            sim.rollout(stream, 50, act, obs)
            torch.cuda.synchronize()
            best = None
            for rep in range(3):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                for _ in range(args.launches):
                    sim.rollout(stream, args.steps, act, obs)
                e1.record(stream)
                torch.cuda.synchronize()
                us = 1e3 * e0.elapsed_time(e1) / (args.steps * args.launches)
                best = us if best is None else min(best, us)
            print(f"lanes {lanes} {'fused   ' if fused else 'per-step'}: {best:7.2f} us/step  "
                  f"{N / best / 1e3:7.3f} G env-steps/s  {per_step / best / 1e3:7.1f} GB/s (obs+act+rew+done)", flush=True)
        sim.close()


if __name__ == "__main__":
    main()
