#!/usr/bin/env python
"""Compact device-side timings of the step kernels for A/B runs (same gated, repeated event windows as bench.py):

    [B200ENVS_LIB=tmp_variants/lib_x.so] python tools/step_bench.py [overcooked] [hanabi] [many_player] [cartpole] [balance]

overcooked: fused rollout and one launch per step into output rings larger than L2 (bench.py's headline workload);
the others: bench.py's extra_configs legs.  Prints one line per measurement; not a bench line."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench   # noqa: E402


def overcooked(timer, stream, peak, K=int(os.environ.get("STEP_BENCH_K", 200)), ring=int(os.environ.get("STEP_BENCH_ACTION_RING", bench.ACTION_RING))):
    from madrona_rl_envs_playground_b200 import ExecMode, OvercookedSimulator
    from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
    N = bench.WORLDS_PER_GPU
    params = get_base_layout_params(bench.LAYOUT, bench.HORIZON)
    sim = OvercookedSimulator(ExecMode.CUDA, 0, N, **params)
    if os.environ.get("STEP_BENCH_LANES"):
        sim.set_lanes(int(os.environ["STEP_BENCH_LANES"]))
    gen = torch.Generator(device="cuda").manual_seed(1)
    act = torch.randint(0, 6, (ring, 2, N), generator=gen, device="cuda", dtype=torch.int32)
    obs = torch.empty((bench.OBS_RING, 2, N, 5, 4, 26), dtype=torch.int8, device="cuda")
    if stream is None:
        stream = torch.cuda.Stream()
        timer = bench.GatedTimer(0, stream)
    for fused, per in ((True, bench.ALGO_BYTES_PER_ENV_STEP_FUSED), (False, bench.ALGO_BYTES_PER_ENV_STEP)):
        sim.set_rollout_fused(fused)
        for _ in range(3):
            sim.rollout(stream, K, act, obs)
        stream.synchronize()
        L = int(os.environ.get("STEP_BENCH_LAUNCHES", 1))   # fused: back-to-back launches of K steps per timed region

        def region():
            for _ in range(L if fused else 1):
                sim.rollout(stream, K, act, obs)
        Kr = K * (L if fused else 1)
        ms = timer.run(region, max(6, 2400 // Kr), L if fused else K)
        t = bench.summarize(ms, Kr)
        us = t["us_per_step_median"]
        print(f"overcooked K={K} x{L if fused else 1} action ring {ring} {'fused' if fused else 'per-step'}: {us:.3f} us/step [{t['us_per_step_min']:.3f} .. "
              f"{t['us_per_step_max']:.3f}]  {per * N / us / 1e3 / peak:.3f} of peak", flush=True)
    sim.close()


def main():
    which = sys.argv[1:] or ["overcooked", "hanabi", "many_player"]
    peak, _ = bench.measured_peak_gbs()
    order = os.environ.get("STEP_BENCH_ORDER", "bench")   # "bench": torch owns the device first, stream made after the simulator
    setdev = os.environ.get("STEP_BENCH_SETDEV", "1" if order == "bench" else "0") == "1"
    late = os.environ.get("STEP_BENCH_STREAM_LATE", "1" if order == "bench" else "0") == "1"
    order = f"setdev={int(setdev)} stream_late={int(late)}"
    if setdev:
        torch.cuda.set_device(0)
    stream = timer = None
    if not late:
        stream = torch.cuda.Stream()
        timer = bench.GatedTimer(0, stream)
    print("library:", os.environ.get("B200ENVS_LIB", "in-tree"), "order:", order, flush=True)
    for name in which:
        if name == "overcooked":
            overcooked(timer, stream, peak)
            continue
        if stream is None:
            stream = torch.cuda.Stream()
            timer = bench.GatedTimer(0, stream)
        r = bench.run_extra_config(name, 0, timer, stream, 20 if name == "many_player" else 200, peak, cpu=False)
        for mode in ("fused_rollout", "per_step_launch"):
            if mode in r:
                m = r[mode]
                print(f"{name} {mode}: {m['us_per_step']:.3f} us/step [{m['us_per_step_min']:.3f} .. {m['us_per_step_max']:.3f}]  "
                      f"{m['frac']:.3f} of peak ({m['algorithmic_bytes_per_env_step']} B)", flush=True)


if __name__ == "__main__":
    main()
