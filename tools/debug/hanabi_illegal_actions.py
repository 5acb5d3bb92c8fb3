#!/usr/bin/env python
"""Hanabi per-step kernel fed arbitrary action values (-8 .. 39, mostly illegal).  The outputs are not compared with
anything (the reference's behaviour is undefined there); the run must end without a fault and with 0 / 1 rows.  (The same
logic runs under AddressSanitizer + UBSan on the CPU: tests/test_hanabi_game_host.py.)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from madrona_rl_envs_playground_b200 import ExecMode, HanabiSimulator   # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
for cfg in ((5, 5, 2, 8, 3), (3, 3, 2, 4, 2)):
    sim = HanabiSimulator(ExecMode.CUDA, 0, N, *cfg)
    act = sim.action_tensor().to_torch()
    gen = torch.Generator(device="cuda").manual_seed(1)
    for t in range(150):
        act.copy_(torch.randint(-8, 40, act.shape, generator=gen, device="cuda", dtype=torch.int32))
        sim.step()
    torch.cuda.synchronize()
    obs = sim.observation_tensor().to_torch()
    assert int(obs.max()) <= 1 and int(obs.min()) >= 0
    sim.close()
print("ok")
