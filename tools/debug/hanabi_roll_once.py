#!/usr/bin/env python
"""Three fused Hanabi rollouts of STEPS steps (argv[1], default 200) at 65 536 worlds, for `ncu -k regex:rollout_kernel -s 2 -c 1`;
argv[2] = "per_step" runs one launch per step instead (`-k regex:step_kernel -s 300 -c 1`)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from madrona_rl_envs_playground_b200 import ExecMode, HanabiSimulator   # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
sim = HanabiSimulator(ExecMode.CUDA, 0, 65536, 5, 5, 2, 8, 3)
if len(sys.argv) > 2 and sys.argv[2] == "per_step":
    sim.set_rollout_fused(False)
first = 0
for _ in range(3):
    sim.random_rollout(steps, 1, first_step=first)
    first += steps
torch.cuda.synchronize()
print("ok")
