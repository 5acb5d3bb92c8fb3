#!/usr/bin/env python
"""Benchmark of the env step hot path (BASELINE.json metric: Overcooked env-steps/sec @ 32K worlds per GPU).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

One "step" = one pass of the hot path (world update + observation featurisation) over this rank's 32 768 Overcooked
`cramped_room` worlds (2 agents, horizon 400, lossless obs) with fresh random actions.  Worlds are sharded over ranks
with no communication on the step path (weak scaling).  Prints ONE JSON line on rank 0.

  value        env-steps/s, whole job, inputs resident in HBM: the random-action rollout the metric names.  Actions of
               all steps are pre-generated on the device (ring of 64 step slices); the library runs 100 steps per
               launch with the world state kept on chip (fused rollout kernel), writing every step's observations to
               a ring of obs buffers larger than L2 so that every step streams to HBM (a rollout collector's
               [T, ...] buffer behaves the same way).  CUDA events on the launch stream.
  per_step_launch  the same steps as one launch of the step kernel per step (what env.step() does when a policy picks
               each action), for comparison.
  e2e          same metric through the public host-buffer call (C ABI b200env_step_host): per step, actions H2D from
               pinned memory, step, obs + rewards + done D2H into pinned memory.  Wall clock around synchronous calls.
  roofline     algorithmic bytes / kernel time against the measured HBM copy bandwidth in MEASURED_PEAKS.json.
               Fused rollout: 1 060 B per env-step (obs 1040 W + actions 8 R + rewards 8 W + done 4 W) plus 200 B of
               state per world once per launch; per-step launches: the 1 260 B of SURVEY.md 8(d).
  cpu_baseline the reference's own Madrona CPU exec mode (oracle/_ref) on all host cores, bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORLDS_PER_GPU = 32768
LAYOUT, HORIZON = "cramped_room", 400
ALGO_BYTES_PER_ENV_STEP = 1260      # obs 1040 W + actions 8 R + rewards 8 W + done 4 W + state 100 R + 100 W  
ALGO_BYTES_PER_ENV_STEP_FUSED = 1060   # fused rollout: obs + actions + rewards + done; state stays on chip ...
STATE_BYTES_PER_WORLD = 200            # ... and is read + written once per launch
# dram__bytes_read.sum + dram__bytes_write.sum of ONE 400-step rollout_tile<2> launch at 32 768 worlds, from the
# `ncu --set full` capture summarised in profiles/r01w_overcooked_rollout_tile_full.txt (19.3 MB read + 13 175.0 MB
# written; before the L2 eviction policies kept the action ring in L2, profiles/r01u_*: 123.4 + 13 162.8)
NCU_DRAM_BYTES_PER_400_STEP_LAUNCH = 13_194_311_000
METRIC = "overcooked_env_steps_per_sec"
UNIT = "env-steps/s"
ACTION_RING, OBS_RING = 64, 8       # 8 x 34 MB of obs > 2 x 126 MB L2


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region: NVML every 2 ms (nvidia-smi as a fallback)."""
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NVML_REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, gpu_index):
        self.gpu, self._stop, self._thread = gpu_index, threading.Event(), None
        self.sm, self.sm_max, self.reasons, self.source = [], [], set(), "nvml"
        try:
            import pynvml
            pynvml.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(visible.split(",")[gpu_index]) if visible and visible.split(",")[gpu_index].isdigit() else gpu_index
            self._nvml, self._h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(phys)
        except Exception:
            self._nvml, self.source = None, "nvidia-smi"

    def _sample_nvml(self):
        n = self._nvml
        self.sm.append(float(n.nvmlDeviceGetClockInfo(self._h, n.NVML_CLOCK_SM)))
        self.sm_max.append(float(n.nvmlDeviceGetMaxClockInfo(self._h, n.NVML_CLOCK_SM)))
        mask = n.nvmlDeviceGetCurrentClocksEventReasons(self._h) if hasattr(n, "nvmlDeviceGetCurrentClocksEventReasons") \
            else n.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
        for bit, name in self.NVML_REASONS.items():
            if mask & bit:
                self.reasons.add(name)

    def _sample_smi(self):
        out = subprocess.run(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                              "-i", str(self.gpu)], capture_output=True, text=True, timeout=5).stdout.strip()
        c = [x.strip() for x in out.split(",")]
        self.sm.append(float(c[1])); self.sm_max.append(float(c[2]))
        for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], c[5:9]):
            if v.lower().startswith("active"):
                self.reasons.add(name)

    def _run(self):
        while True:
            try:
                self._sample_nvml() if self._nvml else self._sample_smi()
            except Exception:
                pass
            if self._stop.wait(float(os.environ.get("BENCH_CLOCK_PERIOD", 0.002)) if self._nvml else 0.1):
                break

    def __enter__(self):
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._thread.join(timeout=6)

    def summary(self):
        sm = sorted(self.sm)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(self.sm_max) if self.sm_max else None,
                "reasons": sorted(self.reasons), "samples": len(sm), "source": self.source}


def dist_setup(n_gpus):
    import torch
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, world, local


def barrier(world):
    import torch
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(x, world):
    from madrona_rl_envs_playground_b200.sharding import max_over_ranks as _max
    return _max(x, device="cuda") if world > 1 else x


def cpu_reference_run(steps, warmup, worlds, seed=0):
    """Time the reference's Madrona CPU exec mode (oracle/_ref; the plain-C port if that build is absent)."""
    import numpy as np
    from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
    from oracle import bindings as ob
    params = get_base_layout_params(LAYOUT, HORIZON)
    rng = np.random.default_rng(seed)
    acts = rng.integers(0, 6, (16, 2, worlds)).astype(np.int32)
    if ob.ref_available("overcooked"):
        kind, cores = "reference", len(os.sched_getaffinity(0))
        sim = ob.RefOvercooked(params, worlds, num_workers=0)    # 0 = every core in the affinity mask
        acts_wm = np.ascontiguousarray(acts.transpose(0, 2, 1))  # the reference's buffers are world-major

        def step(i):
            sim.x_action[...] = acts_wm[i % 16]
            sim.raw_step()
    else:
        kind, cores = "port", 1
        ob.build(ref=False)
        sim = ob.OracleOvercooked(params, worlds)

        def step(i):
            sim.step(acts[i % 16], want_obs=True)
    for i in range(warmup):
        step(i)
    t0 = time.perf_counter()
    for i in range(steps):
        step(i)
    dt = time.perf_counter() - t0
    sim.close()
    return {"value": worlds * steps / dt, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": f"{steps} steps x {worlds} worlds of {LAYOUT} (horizon {HORIZON}), random actions, "
                      f"timed around exec.run() incl. the reference's export copy-in/out"}, dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # bounded sample: keep the whole run within ~2-3 minutes at ~1.3e5 env-steps/s
    worlds = WORLDS_PER_GPU
    while worlds > 256 and worlds * (args.steps + args.warmup) > 1.6e7:
        worlds //= 2
    base, dt = cpu_reference_run(args.steps, args.warmup, worlds)
    line = {"impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": f"overcooked {LAYOUT} 2-agent, horizon {HORIZON}, lossless obs; "
                                   f"{worlds} worlds per step on the host CPU", "worlds_per_step": worlds},
            "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_ours(args):
    import torch
    from madrona_rl_envs_playground_b200 import ExecMode, OvercookedSimulator, _lib
    from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params

    rank, world, local = dist_setup(args.gpus)
    N = args.worlds
    params = get_base_layout_params(LAYOUT, HORIZON)
    P, W, H = params["num_players"], params["width"], params["height"]
    C = 5 * P + 16
    sim = OvercookedSimulator(ExecMode.CUDA, local, N, **params)   # raises if libb200envs.so is missing
    if args.lanes:
        sim.set_lanes(args.lanes)

    gen = torch.Generator(device="cuda").manual_seed(1234 + rank)
    act_ring = torch.randint(0, 6, (ACTION_RING, P, N), generator=gen, device="cuda", dtype=torch.int32)
    obs_ring = torch.empty((OBS_RING, P, N, W, H, C), dtype=torch.int8, device="cuda")
    stream = torch.cuda.Stream()

    ROLLOUT_T = 400   # steps per fused launch: one full episode (horizon 400), the rollout length PPO collects here

    def run_steps(k, fused):
        # the rollout collector's loop, enqueued from C (b200env_overcooked_rollout).  fused: ONE launch per
        # ROLLOUT_T steps with the world state kept on chip; else one launch of the step kernel per step.
        sim.set_rollout_fused(fused)
        launches = 0
        while k > 0:
            t = min(k, ROLLOUT_T) if fused else k
            sim.rollout(stream, t, act_ring, obs_ring)
            launches += 1 if fused else t
            k -= t
        return launches

    def timed(fused):
        run_steps(args.warmup, fused)
        barrier(world)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with ClockSampler(local) as clk:
            barrier(world)
            e0.record(stream)
            launches = run_steps(args.steps, fused)
            e1.record(stream)
            barrier(world)
        return e0.elapsed_time(e1), launches, clk

    # ---- device-resident throughput ("value"): fused rollout; and the same steps as one launch per step ----
    local_ms, launches, clocks = timed(True)
    ms = max_over_ranks(local_ms, world)
    value = world * N * args.steps / (ms * 1e-3)
    peak, peak_src = measured_peak_gbs()
    algo_bytes = ALGO_BYTES_PER_ENV_STEP_FUSED * N * args.steps + STATE_BYTES_PER_WORLD * N * launches
    achieved = algo_bytes / (local_ms * 1e-3) / 1e9
    step_local_ms, step_launches, _ = timed(False)
    step_ms = max_over_ranks(step_local_ms, world)

    # fused rollout with the observation buffer fixed (L2-resident working set) for comparison
    sim.set_rollout_fused(True)
    ev2, ev3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev2.record(stream)
    sim.rollout(stream, args.steps, act_ring)
    ev3.record(stream)
    torch.cuda.synchronize()
    l2_ms = ev2.elapsed_time(ev3)

    # ---- end to end through the host-buffer C-ABI call ----
    e2e_steps = max(10, min(args.steps, 200))
    h_act = torch.randint(0, 6, (16, P, N), dtype=torch.int32).pin_memory()
    h_obs = torch.empty((P, N, W, H, C), dtype=torch.int8).pin_memory()
    h_rew = torch.empty((P, N), dtype=torch.int32).pin_memory()
    h_done = torch.empty((N,), dtype=torch.int32).pin_memory()
    outs = {_lib.T_OBSERVATION: h_obs, _lib.T_REWARD: h_rew, _lib.T_DONE: h_done}
    for i in range(3):
        sim.step_host(h_act[i % 16], outs)
    barrier(world)
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        sim.step_host(h_act[i % 16], outs)
    torch.cuda.synchronize()
    e2e_dt = max_over_ranks(time.perf_counter() - t0, world)
    e2e_value = world * N * e2e_steps / e2e_dt
    # ... and with only rewards + done read back (a GPU-resident policy consumes the obs tensor in place)
    outs_small = {_lib.T_REWARD: h_rew, _lib.T_DONE: h_done}
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        sim.step_host(h_act[i % 16], outs_small)
    torch.cuda.synchronize()
    e2e_small_dt = max_over_ranks(time.perf_counter() - t0, world)

    # optional global episode statistics: the only collective, off the step path
    from madrona_rl_envs_playground_b200.sharding import allreduce_episode_stats
    ep = allreduce_episode_stats(*sim.episode_stats(), device="cuda")

    cpu_base = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu_base, _ = cpu_reference_run(steps=50, warmup=3, worlds=WORLDS_PER_GPU)

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": f"overcooked {LAYOUT} 2-agent, {N} worlds per GPU, horizon {HORIZON}, lossless obs, "
                                   f"uniform random-action rollout: actions of all steps resident in HBM, fused "
                                   f"{ROLLOUT_T}-step launches (b200env_overcooked_rollout)",
                       "worlds_per_gpu": N, "steps_per_launch": ROLLOUT_T,
                       "parallelism": f"world-sharded x{world}, no step-path communication",
                       "l2": f"obs written to a ring of {OBS_RING} buffers ({OBS_RING * P * N * W * H * C / 1e6:.0f} MB) "
                             "> L2; actions from a 64-slice ring"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": NCU_DRAM_BYTES_PER_400_STEP_LAUNCH if (N == WORLDS_PER_GPU and ROLLOUT_T == 400) else None,
                         "traffic_source": "profiles/r01w_overcooked_rollout_tile_full.txt (ncu --set full, one 400-step launch)",
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": algo_bytes / launches,
                         "algorithmic_bytes_per_env_step": ALGO_BYTES_PER_ENV_STEP_FUSED,
                         "kernel": "b200::oc::rollout_tile<2>", "us_per_launch": 1e3 * local_ms / launches,
                         "us_per_step": 1e3 * local_ms / args.steps,
                         "l2_resident_us_per_step": 1e3 * l2_ms / args.steps},
            "per_step_launch": {"value": world * N * args.steps / (step_ms * 1e-3), "unit": UNIT,
                                "us_per_step": 1e3 * step_local_ms / args.steps, "launches": step_launches,
                                "kernel": "b200::oc::step_tile<2>",
                                "algorithmic_bytes_per_env_step": ALGO_BYTES_PER_ENV_STEP,
                                "roofline_frac": ALGO_BYTES_PER_ENV_STEP * N * args.steps / (step_local_ms * 1e-3) / 1e9 / peak,
                                "note": "same steps through one launch of the step kernel per step (what env.step() "
                                        "uses when a policy chooses each action); programmatic dependent launch"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": P * N * 4,
                    "d2h_bytes_per_step": P * N * W * H * C + P * N * 4 + N * 4, "steps": e2e_steps,
                    "api": "b200env_step_host (obs + rewards + done copied to pinned host memory)",
                    "rewards_done_only_value": world * N * e2e_steps / e2e_small_dt},
            "gpu_launches": launches,
            "clocks": clocks.summary(),
            "episodes_finished_all_ranks": ep[0],
        }
        if cpu_base is not None:
            line["cpu_baseline"] = cpu_base
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20000)
    ap.add_argument("--warmup", type=int, default=400)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--worlds", type=int, default=WORLDS_PER_GPU, help="worlds per GPU (default: the BASELINE config)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--lanes", type=int, default=0, help="lanes per world of the step kernel (0 = auto; tuning)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
