"""us/step of the Hanabi kernels with int8 / bfloat16 / float32 rows (65 536 worlds, device-only random rollouts)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import bench
from madrona_rl_envs_playground_b200 import ExecMode, HanabiSimulator

N = 65536
stream = torch.cuda.Stream()
timer = bench.GatedTimer(0, stream)
peak, _ = bench.measured_peak_gbs()
for name, dt, es in (("int8", None, 1), ("bfloat16", torch.bfloat16, 2), ("float32", torch.float32, 4)):
    sim = HanabiSimulator(ExecMode.CUDA, 0, N, 5, 5, 2, 8, 3, obs_dtype=dt)
    for fused in (True, False):
        t = bench.time_random_rollout(sim, timer, stream, 100, fused, 1)
        byts = (658 + 783) * es + 104 + (0 if fused else 520)
        us = t["us_per_step_median"]
        print(f"hanabi {name:8s} {'fused' if fused else 'per-step'}: {us:7.2f} us/step  {byts} B/env-step  "
              f"{byts * N / us / 1e3:7.0f} GB/s = {byts * N / us / 1e3 / peak:.3f} of peak", flush=True)
    sim.close()
