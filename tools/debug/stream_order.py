"""Which stream runs the Overcooked per-step launch chain (programmatic dependent launches, 20-step regions) fastest?
torch streams made before / after the simulator, torch's default stream, a raw cudaStreamNonBlocking stream made by
cudart before / after.  Prints us/step (median of gated 20-step regions)."""
import ctypes
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench
from madrona_rl_envs_playground_b200 import ExecMode, OvercookedSimulator
from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params

torch.cuda.set_device(0)
early = torch.cuda.Stream()
early_hi = torch.cuda.Stream(priority=-1)
N, K = bench.WORLDS_PER_GPU, 20
sim = OvercookedSimulator(ExecMode.CUDA, 0, N, **get_base_layout_params(bench.LAYOUT, bench.HORIZON))
act = torch.randint(0, 6, (64, 2, N), device="cuda", dtype=torch.int32)
obs = torch.empty((8, 2, N, 5, 4, 26), dtype=torch.int8, device="cuda")
sim.set_rollout_fused(False)
late = torch.cuda.Stream()
late_hi = torch.cuda.Stream(priority=-1)
streams = {"torch early": early, "torch early high-priority": early_hi, "torch late": late, "torch late high-priority": late_hi,
           "torch default": torch.cuda.default_stream()}
for name, st in streams.items():
    timer = bench.GatedTimer(0, st)
    for _ in range(3):
        sim.rollout(st, K, act, obs)
    st.synchronize()
    t = bench.summarize(timer.run(lambda: sim.rollout(st, K, act, obs), 120, K), K)
    print(f"{name:28s} handle {st.cuda_stream:#x}: {t['us_per_step_median']:.3f} us/step [{t['us_per_step_min']:.3f} .. {t['us_per_step_max']:.3f}]", flush=True)
