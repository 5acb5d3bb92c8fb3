#!/usr/bin/env python
"""Device-only random-action rollouts of every BASELINE config on one GPU (b200env_random_rollout:
{sample legal actions -> step} x T, no host round trip), CUDA events on the launch stream.

    python tools/env_bench.py [--configs cartpole balance hanabi overcooked many_player] [--steps 2000]

Prints env-steps/s, us/step and algorithmic GB/s (bytes per env-step from SURVEY.md 8(d)) against the measured HBM
peak.  The sampler kernel is inside the timed loop (it is part of a random rollout); Overcooked's headline numbers
come from bench.py, which pre-generates actions.  Tuning / reporting aid for DESIGN.md section 3."""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from madrona_rl_envs_playground_b200 import (AcrobatSimulator, BalanceBeamSimulator, CartpoleSimulator, ExecMode, HanabiSimulator,  # noqa: E402
                                             OvercookedSimulator)
from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params  # noqa: E402


def make(name, worlds):
    if name in ("cartpole", "cartpole_per_step"):   # per_step: one launch per step instead of the cooperative rollout
        sim = CartpoleSimulator(ExecMode.CUDA, 0, worlds or 1024)
        sim.set_rollout_fused(name == "cartpole")
        return sim, worlds or 1024, 48
    if name == "acrobat":   # src/acrobat_env: state 16 R + 16 W, action 4 R, reward 4 W, done 4 W
        return AcrobatSimulator(ExecMode.CUDA, 0, worlds or 1024), worlds or 1024, 44
    if name in ("balance", "balance_per_step"):
        sim = BalanceBeamSimulator(ExecMode.CUDA, 0, worlds or 16384)
        sim.set_rollout_fused(name == "balance")
        return sim, worlds or 16384, 164
    if name in ("hanabi", "hanabi_per_step"):   # per_step: one launch per step instead of the fused cooperative rollout
        sim = HanabiSimulator(ExecMode.CUDA, 0, worlds or 65536, 5, 5, 2, 8, 3)
        sim.set_rollout_fused(name == "hanabi")
        return sim, worlds or 65536, 2065
    if name in ("overcooked", "overcooked_per_step"):
        # fused (default): the T-step kernel draws the moves itself — obs 1040 + rewards 8 + done 4 B per env-step, and
        # the observation tensor (34 MB) is overwritten in place every step, i.e. it stays in the 126 MB L2: the GB/s
        # printed is L2 write traffic, NOT an HBM figure (bench.py writes a ring larger than L2 for the HBM roofline).
        # per_step: sampler + step launch per step, state read and written every step (1 260 B)
        p = get_base_layout_params("cramped_room", 400)
        sim = OvercookedSimulator(ExecMode.CUDA, 0, worlds or 32768, **p)
        sim.set_rollout_fused(name == "overcooked")
        return sim, worlds or 32768, 1052 if name == "overcooked" else 1260
    if name == "simplecooked":   # src/overcooked2_env rules, layout `simple`: obs 2*20*20 = 800 B + 20 + state 100 R + W
        from madrona_rl_envs_playground_b200 import SimplecookedSimulator
        from madrona_rl_envs_playground_b200.overcooked2_layouts import get_base_layout_params as simple_params
        p = simple_params("simple", 200)   # fused like overcooked: obs 800 + rewards 8 + done 4, in place (L2-resident)
        return SimplecookedSimulator(ExecMode.CUDA, 0, worlds or 32768, **p), worlds or 32768, 812
    if name == "many_player":
        p = get_base_layout_params("many_player_layout", 400)
        return OvercookedSimulator(ExecMode.CUDA, 0, worlds or 8192, **p), worlds or 8192, 2206212
    raise SystemExit(f"unknown config {name}")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", nargs="*", default=["cartpole", "balance", "hanabi", "overcooked", "many_player"])
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--worlds", type=int, default=0)
    args = ap.parse_args()
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        peak = 6650.0
    stream = torch.cuda.Stream()
    for name in args.configs:
        sim, N, bytes_per = make(name, args.worlds)
        steps = args.steps if name != "many_player" else max(20, args.steps // 50)
        sim.random_rollout(max(5, steps // 20), 1, 0, stream)
        torch.cuda.synchronize()
        best = None
        for rep in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            sim.random_rollout(steps, 1, 1000 * (rep + 1), stream)
            e1.record(stream)
            torch.cuda.synchronize()
            us = 1e3 * e0.elapsed_time(e1) / steps
            best = us if best is None else min(best, us)
        gbs = bytes_per * N / best / 1e3
        where = (" [outputs overwritten in place: L2-resident, not HBM traffic]"
                 if name in ("overcooked", "simplecooked") and gbs > peak else "")
        print(f"{name:12s} N={N:6d}  {best:9.2f} us/step  {N / best:10.2f} M env-steps/s  {gbs:8.1f} GB/s algorithmic "
              f"({100 * gbs / peak:5.1f}% of {peak:.0f} measured){where}", flush=True)
        sim.close()


if __name__ == "__main__":
    main()
