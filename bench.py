#!/usr/bin/env python
"""Benchmark of the env step hot path (BASELINE.json metric: Overcooked env-steps/sec @ 32K worlds per GPU).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload overcooked|many_player]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

One "step" = one pass of the hot path (world update + observation featurisation) over this rank's 32 768 Overcooked
`cramped_room` worlds (2 agents, horizon 400, lossless obs) with fresh random actions.  Worlds are sharded over ranks
with no communication on the step path (weak scaling).  Prints ONE JSON line on rank 0.

How the timed region is measured (all device-side, CUDA events on the launch stream):
  * the K-step region is enqueued `reps` times back to back BEHIND A GATE — a one-thread kernel that holds the stream
    until the host has finished enqueueing (b200env_debug_stream_gate) — with an event between consecutive regions,
    so no host launch latency, Python time or inter-process CPU contention is inside any event window;
  * `ms_per_step` = median over the reps of (event window / K), then the MAX over ranks; min / max over reps and the
    per-rank medians are kept in the line (`timing`); reps is chosen so that the timed windows add up to >= ~60 ms per
    rank (>= 25 reps), and is stated.

  value        env-steps/s, whole job, inputs resident in HBM: the random-action rollout the metric names.  Actions of
               all steps are pre-generated on the device (ring of 64 step slices); the library runs min(K, 400) steps
               per launch with the world state kept on chip (fused rollout kernel, b200env_overcooked_rollout), writing
               every step's observations to a ring of obs buffers larger than L2 so that every step streams to HBM (a
               rollout collector's [T, ...] buffer behaves the same way).
  per_step_launch  the same steps as one launch of the step kernel per step (what env.step() does when a policy picks
               each action), for comparison.
  e2e          same metric through the public host-buffer calls of the C ABI with pinned HOST buffers: per step,
               actions H2D, step, obs + rewards + done D2H.  `value` is the pipelined pair b200env_step_host_async /
               _wait (step t+1 runs while step t's outputs cross PCIe), `synchronous_value` is b200env_step_host.
               Wall clock around the calls.
  roofline     algorithmic bytes / kernel time against the measured HBM copy bandwidth in MEASURED_PEAKS.json.
               Fused rollout: 1 060 B per env-step (obs 1040 W + actions 8 R + rewards 8 W + done 4 W) plus 200 B of
               state per world once per launch; per-step launches: the 1 260 B of SURVEY.md 8(d).
  cpu_baseline the reference's own Madrona CPU exec mode (oracle/_ref) on all host cores, bounded sample.
  extra_configs (N = 1 only) the other BASELINE.json configs — Cartpole 1 024, Balance Beam 16 384, Hanabi 65 536,
               Overcooked many_player_layout 8 192 worlds — each as device-only random rollouts (fused and one launch
               per step), with the bytes per env-step that kernel actually moves, its fraction of the measured HBM
               peak, and the reference's CPU rate (all cores and one worker) on a bounded sample.
"""
import argparse
import json
import math
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
# dram__bytes_read.sum + dram__bytes_write.sum of ONE rollout_tile<2> launch at 32 768 worlds from `ncu --set full`
# captures, keyed by the steps of that launch: {steps: (bytes, summary under profiles/)}.  Reported only when the run
# times launches of exactly that shape.
NCU_DRAM_BYTES_PER_LAUNCH = {
    400: (13_178_776_000 + 19_688_448, "profiles/r02e_overcooked_rollout_tile_full.txt"),   # written + read
}
METRIC = "overcooked_env_steps_per_sec"
UNIT = "env-steps/s"
ACTION_RING, OBS_RING = 64, 8       # 8 x 34 MB of obs > 2 x 126 MB L2
ROLLOUT_T = 400                     # most steps per fused launch: one full episode (horizon 400)
MIN_TIMED_MS, MIN_REPS, MAX_REPS = 60.0, 25, 500
MAX_LAUNCHES_PER_GATE = 240         # launches + events queued behind one closed gate (well inside the launch queue)


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


def gather_ranks(x, world):
    """the per-rank values of a scalar, in rank order (on every rank)"""
    if world == 1:
        return [float(x)]
    import torch
    import torch.distributed as dist
    mine = torch.tensor([float(x)], dtype=torch.float64, device="cuda")
    out = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(out, mine)
    return [float(t.item()) for t in out]


class GatedTimer:
    """Times `reps` repetitions of a region enqueued on `stream`, each between two CUDA events, with the stream held by
    a gate kernel while the host enqueues (so that the event windows hold device time only)."""

    def __init__(self, local, stream):
        import torch
        from madrona_rl_envs_playground_b200 import stream_gate
        self.torch, self.gate, self.local, self.stream = torch, stream_gate, local, stream
        self.flag = torch.zeros(1, dtype=torch.int32).pin_memory()

    def run(self, enqueue, reps, launches_per_rep):
        """enqueue(): puts one region on the stream.  Returns the list of per-rep milliseconds."""
        torch = self.torch
        per_gate = max(1, MAX_LAUNCHES_PER_GATE // (launches_per_rep + 1))
        # No gate when one region alone holds more launches than may wait behind a closed gate (the host would block on a
        # full launch queue until the gate times out; such a region is long enough to amortise the host's head start), nor
        # under a profiler that runs kernels one at a time (BENCH_NO_GATE=1: the host could never open the gate).
        gated = launches_per_rep + 1 <= MAX_LAUNCHES_PER_GATE and not os.environ.get("BENCH_NO_GATE")
        out = []
        done = 0
        while done < reps:
            n = min(per_gate, reps - done)
            self.flag.zero_()
            if gated:
                self.gate(self.local, self.stream, self.flag, 4000)
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
            ev[0].record(self.stream)
            for i in range(n):
                enqueue()
                ev[i + 1].record(self.stream)
            self.flag.fill_(1)          # open the gate: everything above runs back to back from here
            self.stream.synchronize()
            out += [ev[i].elapsed_time(ev[i + 1]) for i in range(n)]
            done += n
        return out


def pick_reps(est_ms_per_rep):
    return int(min(MAX_REPS, max(MIN_REPS, math.ceil(MIN_TIMED_MS / max(est_ms_per_rep, 1e-3)))))


def summarize(ms_list, steps):
    s = sorted(ms_list)
    return {"median_ms": s[len(s) // 2], "min_ms": s[0], "max_ms": s[-1], "reps": len(s),
            "us_per_step_median": 1e3 * s[len(s) // 2] / steps, "us_per_step_min": 1e3 * s[0] / steps,
            "us_per_step_max": 1e3 * s[-1] / steps}


# --------------------------------------------------------------------------------------------------------------
# the reference's CPU path (oracle/_ref = its own sim.cpp + Madrona CPU runtime; the plain-C port if that is absent)
# --------------------------------------------------------------------------------------------------------------
def cpu_reference_run(steps, warmup, worlds, seed=0, layout=LAYOUT, num_workers=0):
    """Time the reference's Madrona CPU exec mode (oracle/_ref; the plain-C port if that build is absent)."""
    import numpy as np
    from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
    from oracle import bindings as ob
    params = get_base_layout_params(layout, HORIZON)
    P = params["num_players"]
    rng = np.random.default_rng(seed)
    acts = rng.integers(0, 6, (16, P, worlds)).astype(np.int32)
    if ob.ref_available("overcooked"):
        kind = "reference"
        cores = len(os.sched_getaffinity(0)) if num_workers == 0 else num_workers
        sim = ob.RefOvercooked(params, worlds, num_workers=num_workers)    # 0 = every core in the affinity mask
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
            "sample": f"{steps} steps x {worlds} worlds of {layout} (horizon {HORIZON}), random actions, "
                      f"timed around exec.run() incl. the reference's export copy-in/out"}, dt


def cpu_reference_simple(env, worlds, steps, num_workers):
    """The reference's CPU exec mode for Cartpole / Balance Beam / Hanabi (oracle/_ref), random (Hanabi: legal) actions.
    Only exec.run() is timed (as scripts/*_example.py time sim.step()); the action draw is outside."""
    import numpy as np
    from oracle import bindings as ob
    if not ob.ref_available(env):
        return None
    rng = np.random.default_rng(0)
    cores = len(os.sched_getaffinity(0)) if num_workers == 0 else num_workers
    if env == "cartpole":
        sim = ob.RefCartpole(worlds, num_workers=num_workers)
        draw = lambda: sim.x_action.__setitem__(Ellipsis, rng.integers(0, 2, worlds, dtype=np.int32))
    elif env == "balance":
        sim = ob.RefBalance(worlds, num_workers=num_workers)
        draw = lambda: sim.x_action.__setitem__(Ellipsis, rng.integers(0, 4, (worlds, 2), dtype=np.int32))
    else:
        sim = ob.RefHanabi(worlds, num_workers=num_workers)

        def draw():   # uniform over the legal moves: masked argmax of random logits (scripts/hanabi_example.py:64-67)
            logits = rng.random(sim.x_mask.shape, dtype=np.float32)
            logits[sim.x_mask == 0] = -1
            sim.x_action[...] = logits.argmax(-1)
    dt = 0.0
    for i in range(steps + 2):
        draw()
        t0 = time.perf_counter()
        sim.raw_step()
        if i >= 2:
            dt += time.perf_counter() - t0
    sim.close()
    return {"value": worlds * steps / dt, "unit": UNIT, "cores": cores, "kind": "reference",
            "sample": f"{steps} steps x {worlds} worlds, timed around exec.run()"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    layout = "many_player_layout" if args.workload == "many_player" else LAYOUT
    # bounded sample: keep the whole run within ~2-3 minutes at ~1.3e5 env-steps/s (many_player: ~500 env-steps/s,
    # and 3.4 MB of padded export per world)
    worlds = WORLDS_PER_GPU if layout == LAYOUT else 64
    budget = 1.6e7 if layout == LAYOUT else 3.0e4
    while worlds > 8 and worlds * (args.steps + args.warmup) > budget:
        worlds //= 2
    base, dt = cpu_reference_run(args.steps, args.warmup, worlds, layout=layout)
    line = {"impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": f"overcooked {layout}, horizon {HORIZON}, lossless obs; "
                                   f"{worlds} worlds per step on the host CPU", "worlds_per_step": worlds},
            "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------------------
# the other BASELINE.json configs (rows d1, d2, d4, d5 of SURVEY 8(d)): device-only random rollouts
# --------------------------------------------------------------------------------------------------------------
EXTRA = {
    # name: (worlds, bytes per env-step the FUSED kernel moves, bytes per env-step of one launch per step, kernels)
    # Cartpole: state 16 R + 16 W, action 4, reward 4 W, done 4 W, rng 4  = 48; the fused rollout
    # keeps the state in registers but still rewrites every exported tensor: 16 + 4 + 4 + 4 = 28
    "cartpole": (1024, 28, 48, ("b200::cartpole::rollout_kernel", "b200::cartpole::step_kernel")),
    # Balance Beam (config 2): obs 2x28 R + 2x28 W, actions 8, rewards 8 W, done 4 W, loc / time / rng 16 R + 16 W = 164;
    # fused: obs 56 W + actions 8 W + rewards 8 W + done 4 W = 76
    "balance": (16384, 76, 164, ("b200::balance::rollout_kernel", "b200::balance::step_kernel")),
    # Hanabi (config 4): obs 658 W + state 783 W + mask 80 W + active 8 W + reward 8 W + done 4 W + action 4 = 1 545 moved
    # by the fused rollout (game records stay on chip) — minus the rows it does not write: a game that ends is dealt
    # afresh a step later and the fresh game's first row (1 521 B) is only written in the last two steps of a launch
    # (HANABI_FRESH_ROW_BYTES x the measured reset rate is subtracted below); one launch per step also reads and writes
    # the record: SURVEY 8d's 2 065
    "hanabi": (65536, 1545, 2065, ("b200::hanabi::rollout_kernel", "b200::hanabi::step_kernel")),
    # Overcooked many_player_layout (config 5; 40 players, 15x17): obs 2 203 200 W + actions 160 + rewards 160 W + done 4
    # + state 1 344 R + W = 2 206 212; CTA-per-world kernel, one launch per step (+ the sampler launch)
    "many_player": (8192, None, 2206212, (None, "b200::oc::step_world_cta")),
}


def make_extra(name, local, worlds):
    from madrona_rl_envs_playground_b200 import (BalanceBeamSimulator, CartpoleSimulator, ExecMode, HanabiSimulator,
                                                 OvercookedSimulator)
    from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
    if name == "cartpole":
        return CartpoleSimulator(ExecMode.CUDA, local, worlds)
    if name == "balance":
        return BalanceBeamSimulator(ExecMode.CUDA, local, worlds)
    if name == "hanabi":
        return HanabiSimulator(ExecMode.CUDA, local, worlds, 5, 5, 2, 8, 3)
    return OvercookedSimulator(ExecMode.CUDA, local, worlds, **get_base_layout_params("many_player_layout", HORIZON))


HANABI_FRESH_ROW_BYTES = 658 + 783 + 80   # obs + state + mask row of a freshly dealt game


def time_random_rollout(sim, timer, stream, steps, fused, launches_per_step, target_ms=30.0):
    """median ms of `steps` x {random legal action -> step} (b200env_random_rollout), device-only; the summary also
    carries `resets_per_env_step` (episodes started per world and step over all the steps run here)"""
    import torch
    sim.set_rollout_fused(fused)
    counter = [0]
    episodes0 = sim.episode_stats()[0]

    def enqueue():
        sim.random_rollout(steps, 7, counter[0], stream)
        counter[0] += steps

    for _ in range(3):
        enqueue()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record(stream); enqueue(); t1.record(stream)
    stream.synchronize()
    est = t0.elapsed_time(t1)
    reps = int(min(100, max(5, math.ceil(target_ms / max(est, 1e-3)))))
    launches = 1 if fused else steps * launches_per_step
    out = summarize(timer.run(enqueue, reps, launches), steps)
    stream.synchronize()
    out["resets_per_env_step"] = (sim.episode_stats()[0] - episodes0) / max(1, counter[0] * sim.num_worlds)
    return out


def run_extra_config(name, local, timer, stream, steps, peak, worlds=None, cpu=True):
    worlds_default, fused_bytes, step_bytes, kernels = EXTRA[name]
    N = worlds or worlds_default
    sim = make_extra(name, local, N)
    out = {"worlds": N, "steps_per_region": steps}
    if fused_bytes is not None:
        t = time_random_rollout(sim, timer, stream, steps, True, 1)
        if name == "hanabi":   # rows the kernel really writes: all but the fresh games' first rows outside a launch's last two steps
            skipped = t["resets_per_env_step"] * max(0.0, 1.0 - 2.0 / steps)
            fused_bytes = round(fused_bytes - HANABI_FRESH_ROW_BYTES * skipped, 1)
            out["resets_per_env_step"] = t["resets_per_env_step"]
        gbs = fused_bytes * N / t["us_per_step_median"] / 1e3
        out["fused_rollout"] = {"us_per_step": t["us_per_step_median"], "env_steps_per_s": N / t["us_per_step_median"] * 1e6,
                                "algorithmic_bytes_per_env_step": fused_bytes, "achieved_gbs": gbs, "frac": gbs / peak,
                                "kernel": kernels[0], "reps": t["reps"], "us_per_step_min": t["us_per_step_min"],
                                "us_per_step_max": t["us_per_step_max"]}
    # many_player: sampler launch + step launch per step; the simple envs draw inside the step kernel; Hanabi: one sampler
    # launch for the first step of the region, then the step kernel draws the next move
    t = time_random_rollout(sim, timer, stream, steps, False, 2 if name == "many_player" else 1,
                            target_ms=100.0 if name == "many_player" else 30.0)
    gbs = step_bytes * N / t["us_per_step_median"] / 1e3
    if name == "many_player":
        out["note"] = ("a store-only stream (18 GB of observations per step, ~1 KB of reads per world) is not bounded by the "
                       "measured COPY bandwidth that `frac` is quoted against: frac > 1 is possible (HBM3e nominal 8 TB/s)")
    out["per_step_launch"] = {"us_per_step": t["us_per_step_median"], "env_steps_per_s": N / t["us_per_step_median"] * 1e6,
                              "algorithmic_bytes_per_env_step": step_bytes, "achieved_gbs": gbs, "frac": gbs / peak,
                              "kernel": kernels[1], "reps": t["reps"], "us_per_step_min": t["us_per_step_min"],
                              "us_per_step_max": t["us_per_step_max"]}
    sim.close()
    if cpu:
        try:
            if name == "many_player":
                allc, _ = cpu_reference_run(steps=8, warmup=2, worlds=64, layout="many_player_layout")
                one, _ = cpu_reference_run(steps=4, warmup=1, worlds=16, layout="many_player_layout", num_workers=1)
            else:
                sample = {"cartpole": (1024, 1000), "balance": (16384, 40), "hanabi": (65536, 6)}[name]
                allc = cpu_reference_simple(name, sample[0], sample[1], 0)
                one = cpu_reference_simple(name, sample[0] if name != "hanabi" else 16384, max(4, sample[1] // 2), 1)
            out["cpu_baseline"] = allc
            out["cpu_baseline_one_worker"] = one
        except Exception as e:   # the checker build is absent: say so, do not fail the GPU numbers
            out["cpu_baseline"] = {"unavailable": repr(e)}
    return out


def run_many_player(args):
    """BASELINE config 5 as the main workload: 8 192 many_player_layout worlds per GPU, sharded over the ranks."""
    import torch
    rank, world, local = dist_setup(args.gpus)
    peak, peak_src = measured_peak_gbs()
    stream = torch.cuda.Stream()
    timer = GatedTimer(local, stream)
    N = args.worlds or EXTRA["many_player"][0]
    barrier(world)
    with ClockSampler(local) as clk:
        res = run_extra_config("many_player", local, timer, stream, args.steps, peak, worlds=N,
                               cpu=(rank == 0 and not args.no_cpu_baseline))
    barrier(world)
    us_local = res["per_step_launch"]["us_per_step"]
    per_rank = gather_ranks(us_local, world)
    us = max(per_rank)
    if rank == 0:
        line = {"metric": "overcooked_many_player_env_steps_per_sec", "value": world * N / us * 1e6, "unit": UNIT,
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": us / 1e3,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                "config": {"workload": f"overcooked many_player_layout (40 agents, 15x17), {N} worlds per GPU, horizon "
                                       f"{HORIZON}, lossless obs, device-only random-action rollout, one step_world_cta "
                                       f"launch per step (BASELINE.json config 5)",
                           "worlds_per_gpu": N, "parallelism": f"world-sharded x{world}, no step-path communication",
                           "l2": "18 GB of observations per step per GPU, far larger than L2"},
                "roofline": {"bound": "hbm", "achieved": res["per_step_launch"]["achieved_gbs"], "peak": peak,
                             "unit": "GB/s", "frac": res["per_step_launch"]["frac"], "traffic": None,
                             "note": res.get("note"),
                             "peak_source": peak_src, "kernel": "b200::oc::step_world_cta",
                             "algorithmic_bytes_per_env_step": EXTRA["many_player"][2]},
                "timing": {"per_rank_us_per_step": per_rank, **res["per_step_launch"]},
                "gpu_launches": 2 * args.steps * res["per_step_launch"]["reps"],
                "clocks": clk.summary()}
        if "cpu_baseline" in res:
            line["cpu_baseline"] = res["cpu_baseline"]
            line["cpu_baseline_one_worker"] = res.get("cpu_baseline_one_worker")
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


def run_ours(args):
    import torch
    from madrona_rl_envs_playground_b200 import ExecMode, OvercookedSimulator, _lib
    from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params

    rank, world, local = dist_setup(args.gpus)
    N = args.worlds or WORLDS_PER_GPU
    K = args.steps
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
    timer = GatedTimer(local, stream)
    steps_per_launch = min(K, ROLLOUT_T)
    fused_launches = -(-K // ROLLOUT_T)

    def enqueue_steps(k, fused):
        # the rollout collector's loop, enqueued from C (b200env_overcooked_rollout).  fused: ONE launch per
        # ROLLOUT_T steps with the world state kept on chip; else one launch of the step kernel per step.
        while k > 0:
            t = min(k, ROLLOUT_T) if fused else k
            sim.rollout(stream, t, act_ring, obs_ring)
            k -= t

    def timed(fused):
        sim.set_rollout_fused(fused)
        enqueue_steps(args.warmup, fused)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); enqueue_steps(K, fused); e1.record(stream)     # sizes `reps`; also warm-up
        stream.synchronize()
        reps = pick_reps(e0.elapsed_time(e1))
        barrier(world)
        with ClockSampler(local) as clk:
            ms = timer.run(lambda: enqueue_steps(K, fused), reps, fused_launches if fused else K)
        barrier(world)
        return summarize(ms, K), clk

    # ---- device-resident throughput ("value"): fused rollout; and the same steps as one launch per step ----
    fused_t, clocks = timed(True)
    fused_ranks = gather_ranks(fused_t["median_ms"], world)
    ms = max(fused_ranks)
    value = world * N * K / (ms * 1e-3)
    peak, peak_src = measured_peak_gbs()
    algo_bytes = ALGO_BYTES_PER_ENV_STEP_FUSED * N * K + STATE_BYTES_PER_WORLD * N * fused_launches
    achieved = algo_bytes / (fused_t["median_ms"] * 1e-3) / 1e9
    step_t, _ = timed(False)
    step_ranks = gather_ranks(step_t["median_ms"], world)
    step_ms = max(step_ranks)

    # fused rollout with the observation buffer fixed (L2-resident working set) for comparison
    sim.set_rollout_fused(True)
    l2 = summarize(timer.run(lambda: sim.rollout(stream, steps_per_launch, act_ring), 10, 1), steps_per_launch)

    # ---- end to end through the host-buffer C-ABI calls ----
    e2e_steps = max(50, min(K, 200))
    h_act = torch.randint(0, 6, (16, P, N), dtype=torch.int32).pin_memory()
    h_obs = [torch.empty((P, N, W, H, C), dtype=torch.int8).pin_memory() for _ in range(2)]
    h_rew = [torch.empty((P, N), dtype=torch.int32).pin_memory() for _ in range(2)]
    h_done = [torch.empty((N,), dtype=torch.int32).pin_memory() for _ in range(2)]
    outs = [{_lib.T_OBSERVATION: h_obs[b], _lib.T_REWARD: h_rew[b], _lib.T_DONE: h_done[b]} for b in range(2)]
    for i in range(3):
        sim.step_host(h_act[i % 16], outs[0])
    barrier(world)
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        sim.step_host(h_act[i % 16], outs[0])
    torch.cuda.synchronize()
    e2e_sync_dt = max_over_ranks(time.perf_counter() - t0, world)
    # pipelined: step t+1 is enqueued before step t's outputs are waited for (two host buffer sets)
    def pipelined(n, outputs):
        sim.step_host_async(h_act[0], outputs[0])
        for i in range(1, n):
            sim.step_host_async(h_act[i % 16], outputs[i & 1])
            sim.step_host_wait()
        sim.step_host_wait()
    pipelined(4, outs)
    barrier(world)
    t0 = time.perf_counter()
    pipelined(e2e_steps, outs)
    torch.cuda.synchronize()
    e2e_dt = max_over_ranks(time.perf_counter() - t0, world)
    e2e_value = world * N * e2e_steps / e2e_dt
    # ... and with only rewards + done read back (a GPU-resident policy consumes the obs tensor in place)
    outs_small = [{_lib.T_REWARD: h_rew[b], _lib.T_DONE: h_done[b]} for b in range(2)]
    pipelined(4, outs_small)
    barrier(world)
    t0 = time.perf_counter()
    pipelined(e2e_steps, outs_small)
    torch.cuda.synchronize()
    e2e_small_dt = max_over_ranks(time.perf_counter() - t0, world)

    # optional global episode statistics: the only collective, off the step path
    from madrona_rl_envs_playground_b200.sharding import allreduce_episode_stats
    ep = allreduce_episode_stats(*sim.episode_stats(), device="cuda")
    sim.close()
    del obs_ring, act_ring

    cpu_base, extra = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu_base, _ = cpu_reference_run(steps=50, warmup=3, worlds=WORLDS_PER_GPU)
    if rank == 0 and world == 1 and not args.no_extra:
        extra = {}
        for name in ("cartpole", "balance", "hanabi", "many_player"):
            try:
                extra[name] = run_extra_config(name, local, timer, stream, max(K, 100) if name != "many_player" else 20,
                                               peak, cpu=not args.no_cpu_baseline)
            except Exception as e:
                extra[name] = {"error": repr(e)}

    if rank == 0:
        traffic = NCU_DRAM_BYTES_PER_LAUNCH.get(steps_per_launch) if (N == WORLDS_PER_GPU and K % steps_per_launch == 0) else None
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": args.warmup, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": f"overcooked {LAYOUT} 2-agent, {N} worlds per GPU, horizon {HORIZON}, lossless obs, "
                                   f"uniform random-action rollout: actions of all steps resident in HBM, fused "
                                   f"launches of {steps_per_launch} steps (b200env_overcooked_rollout)",
                       "worlds_per_gpu": N, "steps_per_launch": steps_per_launch,
                       "parallelism": f"world-sharded x{world}, no step-path communication",
                       "l2": f"obs written to a ring of {OBS_RING} buffers ({OBS_RING * P * N * W * H * C / 1e6:.0f} MB) "
                             "> L2; actions from a 64-slice ring",
                       "timing": f"the {K}-step region repeated {fused_t['reps']} times back to back behind a stream "
                                 "gate, one CUDA event window per repetition; ms_per_step = median window / steps, "
                                 "max over ranks"},
            "timing": {"reps": fused_t["reps"], "timed_total_ms": fused_t["median_ms"] * fused_t["reps"],
                       "us_per_step_median": fused_t["us_per_step_median"], "us_per_step_min": fused_t["us_per_step_min"],
                       "us_per_step_max": fused_t["us_per_step_max"],
                       "per_rank_us_per_step": [1e3 * m / K for m in fused_ranks]},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic[0] if traffic else None,
                         "traffic_source": (traffic[1] + f" (ncu --set full, one {steps_per_launch}-step launch)") if traffic else None,
                         "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": algo_bytes / fused_launches,
                         "algorithmic_bytes_per_env_step": ALGO_BYTES_PER_ENV_STEP_FUSED,
                         "kernel": "b200::oc::rollout_tile<2>", "us_per_launch": 1e3 * fused_t["median_ms"] / fused_launches,
                         "us_per_step": fused_t["us_per_step_median"],
                         "l2_resident_us_per_step": l2["us_per_step_median"]},
            "per_step_launch": {"value": world * N * K / (step_ms * 1e-3), "unit": UNIT,
                                "us_per_step": step_t["us_per_step_median"], "us_per_step_min": step_t["us_per_step_min"],
                                "us_per_step_max": step_t["us_per_step_max"], "reps": step_t["reps"],
                                "launches_per_region": K, "kernel": "b200::oc::step_tile<2>",
                                "algorithmic_bytes_per_env_step": ALGO_BYTES_PER_ENV_STEP,
                                "roofline_frac": ALGO_BYTES_PER_ENV_STEP * N / (step_t["us_per_step_median"] * 1e-6) / 1e9 / peak,
                                "per_rank_us_per_step": [1e3 * m / K for m in step_ranks],
                                "note": "same steps through one launch of the step kernel per step (what env.step() "
                                        "uses when a policy chooses each action); programmatic dependent launch"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": P * N * 4,
                    "d2h_bytes_per_step": P * N * W * H * C + P * N * 4 + N * 4, "steps": e2e_steps,
                    "api": "b200env_step_host_async + b200env_step_host_wait (pinned host buffers: actions in; obs + "
                           "rewards + done out, every step; step t+1 runs while step t's outputs cross PCIe)",
                    "synchronous_value": world * N * e2e_steps / e2e_sync_dt,
                    "synchronous_api": "b200env_step_host",
                    "d2h_gbs_per_gpu": (P * N * W * H * C + P * N * 4 + N * 4) * e2e_steps / e2e_dt / 1e9,
                    "rewards_done_only_value": world * N * e2e_steps / e2e_small_dt},
            "gpu_launches": fused_launches * fused_t["reps"],
            "clocks": clocks.summary(),
            "episodes_finished_all_ranks": ep[0],
        }
        if cpu_base is not None:
            line["cpu_baseline"] = cpu_base
        if extra is not None:
            line["extra_configs"] = extra
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=400)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="overcooked", choices=["overcooked", "many_player"],
                    help="overcooked = BASELINE config 3 (the metric); many_player = config 5 as the main workload")
    ap.add_argument("--worlds", type=int, default=0, help="worlds per GPU (default: the BASELINE config)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the extra_configs legs (the other BASELINE configs)")
    ap.add_argument("--lanes", type=int, default=0, help="lanes per world of the step kernel (0 = auto; tuning)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "many_player":
        run_many_player(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
