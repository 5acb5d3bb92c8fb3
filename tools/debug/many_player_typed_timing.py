#!/usr/bin/env python
"""us per step of the CTA-per-world Overcooked kernel on many_player_layout (40 players) with int8 / bfloat16 / float32
observations, one launch per step with device-drawn random actions (argv[1] = worlds, default 2048)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from madrona_rl_envs_playground_b200 import ExecMode, OvercookedSimulator   # noqa: E402
from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params   # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
params = get_base_layout_params("many_player_layout", 400)
for dtype, es in ((torch.int8, 1), (torch.bfloat16, 2), (torch.float32, 4)):
    sim = OvercookedSimulator(ExecMode.CUDA, 0, N, obs_dtype=dtype, **params)
    sim.random_rollout(5, 1, first_step=0)
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    sim.random_rollout(20, 1, first_step=5)
    t1.record()
    torch.cuda.synchronize()
    us = t0.elapsed_time(t1) * 1e3 / 20
    gb = N * 40 * 255 * 216 * es / 1e9
    print(f"many_player {N} worlds {str(dtype):16s} {us:9.1f} us/step  {gb / us * 1e6 / 1e3:6.2f} TB/s of observation stores", flush=True)
    sim.close()
    del sim
    torch.cuda.empty_cache()
