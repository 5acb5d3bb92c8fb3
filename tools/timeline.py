#!/usr/bin/env python
"""Per-CTA phase timeline of the Overcooked warp-tile step kernel under back-to-back (PDL) launches, or of the
Hanabi step kernel inside a device-only random rollout.

    python tools/timeline.py [--lanes G] [--steps 12] [--worlds 32768]
    python tools/timeline.py --env hanabi [--steps 12] [--worlds 65536]

Uses the library's debug trace (b200env_debug_trace): 8 %globaltimer stamps per CTA and step.  Prints, for a few
steady-state steps, when (relative to the first CTA entry of that step) each phase was reached — median / p5 / p95
over CTAs — plus the step-to-step period.  Diagnostic only; numbers taken with tracing on are not bench values."""
import argparse
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from madrona_rl_envs_playground_b200 import ExecMode, HanabiSimulator, OvercookedSimulator, _lib  # noqa: E402
from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params  # noqa: E402

NAMES = ["entry", "dep_resolved", "inputs_landed", "state_written", "obs0_queued", "obs0_read", "exit"]
# per-step Hanabi kernel stamps: game warp 0 ticket, 1 decks + records loaded, 2 played + count handed over, 3 play rows
# composed, 5 episode numbers known, 6 tail written (resets dealt, flags + records + reset rows written by the game warp);
# writing warp 4 play rows written, 7 exit
HANABI_NAMES = ["ticket", "loaded", "published", "composed", "rows_written", "numbered", "tail_written", "exit"]


def hanabi(args):
    N = args.worlds or 65536
    sim = HanabiSimulator(ExecMode.CUDA, 0, N, 5, 5, 2, 8, 3)
    tiles = (N + 31) // 32
    stream = torch.cuda.Stream()
    trace = torch.zeros((args.steps, tiles, 8), dtype=torch.int64, device="cuda")
    sim.random_rollout(200, seed=1, first_step=0, stream=stream)   # desynchronise the games
    torch.cuda.synchronize()
    _lib.check(sim._lib.b200env_debug_trace(sim._h, C.c_void_p(trace.data_ptr()), tiles * 8, args.steps))
    sim.random_rollout(args.steps, seed=1, first_step=200, stream=stream)
    torch.cuda.synchronize()
    _lib.check(sim._lib.b200env_debug_trace(sim._h, None, 0, 0))
    t = trace.cpu().numpy().astype(np.float64)
    start = t[:, :, 0].min(axis=1)
    print(f"hanabi, {N} worlds, {tiles} tiles; us relative to the first ticket of the step: median [p5 .. p95] over tiles")
    for s in range(2, args.steps):
        rel = (t[s] - start[s][None, None]) / 1e3
        line = f"step {s:2d} | " + "  ".join(
            f"{n} {np.median(rel[:, i]):5.2f} [{np.percentile(rel[:, i], 5):5.2f}..{np.percentile(rel[:, i], 95):5.2f}]"
            for i, n in enumerate(HANABI_NAMES))
        print(line)
    s = args.steps - 1
    rel = (t[s] - start[s]) / 1e3
    waited = rel[:, 5] - rel[:, 4]
    print(f"last step: tiles that waited > 0.5 us for episode numbers: {(waited > 0.5).sum()} of {tiles}; "
          f"their median wait {np.median(waited[waited > 0.5]) if (waited > 0.5).any() else 0:.2f} us")
    late = np.argsort(rel[:, 2])[-8:]
    print("the 8 tiles that published last: index, [ticket loaded published composed]")
    for i in late:
        print(f"  tile {i:5d}", np.round(rel[i, :4], 2))
    print("published, by tile index (us), every 128th tile:", np.round(rel[::128, 2], 2))
    print("numbered, by tile index (us), every 128th tile:", np.round(rel[::128, 5], 2))


ROLL_NAMES = {0: "begin", 1: "played", 2: "buf_free", 3: "composed", 4: "prev_known", 5: "lower_known", 6: "tail_composed",
              8: "w_rows", 9: "w_written", 10: "w_tail", 11: "w_done"}


def hanabi_rollout(args):
    """Fused rollout: per (step, tile) phase stamps; prints phase durations (median / p95 over tiles) for steady-state
    steps and the step-to-step period of each tile."""
    N = args.worlds or 65536
    sim = HanabiSimulator(ExecMode.CUDA, 0, N, 5, 5, 2, 8, 3)
    tiles = (N + 31) // 32
    stream = torch.cuda.Stream()
    S = args.steps
    trace = torch.zeros((S, tiles, 16), dtype=torch.int64, device="cuda")
    sim.random_rollout(200, seed=1, first_step=0, stream=stream)
    torch.cuda.synchronize()
    _lib.check(sim._lib.b200env_debug_trace(sim._h, C.c_void_p(trace.data_ptr()), tiles * 16, S))
    sim.random_rollout(S + 50, seed=1, first_step=200, stream=stream)
    torch.cuda.synchronize()
    _lib.check(sim._lib.b200env_debug_trace(sim._h, None, 0, 0))
    if args.save:
        np.save(args.save, trace.cpu().numpy())
    t = trace.cpu().numpy().astype(np.float64) / 1e3
    lo, hi = S // 2, S - 2
    period = (t[hi, :, 0] - t[lo, :, 0]) / (hi - lo)
    print(f"hanabi fused rollout, {N} worlds, {tiles} tiles; steps {lo}..{hi}: period per tile median {np.median(period):.2f} us "
          f"[p5 {np.percentile(period, 5):.2f} .. p95 {np.percentile(period, 95):.2f}]")
    # game warp stamps in program order: 0 step begins, 4 earlier steps' resets counted, 5 last step's ended games dealt
    # (lower tiles' counts known), 1 played + count published, 2 row buffer free, 3 rows composed
    seg = [("count earlier steps' resets", 0, 4), ("deal last step's ended games", 4, 5), ("play + publish", 5, 1),
           ("wait row buffer free", 1, 2), ("compose rows", 2, 3), ("writer: wait rows", 11, 8), ("writer: write rows", 8, 11)]
    for name, a, b in seg:
        if a == 11:   # from the previous pass's end to this pass's hand-over
            d = t[lo + 1:hi, :, b] - t[lo:hi - 1, :, a]
        else:
            d = t[lo:hi, :, b] - t[lo:hi, :, a]
        print(f"  {name:30s} mean {d.mean():6.2f}  median {np.median(d):6.2f}  p95 {np.percentile(d, 95):6.2f} us")
    skew = t[lo:hi, :, 0] - t[lo:hi, :, 0].min(axis=1, keepdims=True)
    print(f"  step-begin skew across tiles: median {np.median(skew):.2f}  p95 {np.percentile(skew, 95):.2f}  max {skew.max():.2f} us")
    wr = t[lo:hi, :, 11] - t[lo:hi, :, 8]
    print("  writer write-rows duration by tile index (every 256th tile, mean over steps):", np.round(wr.mean(axis=0)[::256], 1))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--lanes", type=int, default=0)
    ap.add_argument("--steps", type=int, default=12)
    ap.add_argument("--worlds", type=int, default=0)
    ap.add_argument("--layout", default="cramped_room")
    ap.add_argument("--save", default=None, help="hanabi_rollout: also save the raw stamps (.npy)")
    ap.add_argument("--env", default="overcooked", choices=["overcooked", "hanabi", "hanabi_rollout"])
    args = ap.parse_args()
    if args.env == "hanabi":
        return hanabi(args)
    if args.env == "hanabi_rollout":
        return hanabi_rollout(args)
    args.worlds = args.worlds or 32768
    params = get_base_layout_params(args.layout, 400)
    P, W, H = params["num_players"], params["width"], params["height"]
    Cc = 5 * P + 16
    N = args.worlds
    sim = OvercookedSimulator(ExecMode.CUDA, 0, N, **params)
    if args.lanes:
        sim.set_lanes(args.lanes)
    act = torch.randint(0, 6, (16, P, N), device="cuda", dtype=torch.int32)
    obs = torch.empty((8, P, N, W, H, Cc), dtype=torch.int8, device="cuda")
    stream = torch.cuda.Stream()
    max_ctas = N            # >= grid for any lanes value <= 4 ... (N*4/32)
    trace = torch.zeros((args.steps, max_ctas, 8), dtype=torch.int64, device="cuda")
    sim.rollout(stream, 50, act, obs)
    torch.cuda.synchronize()
    _lib.check(sim._lib.b200env_debug_trace(sim._h, C.c_void_p(trace.data_ptr()), max_ctas * 8, args.steps))
    sim.rollout(stream, args.steps, act, obs)
    torch.cuda.synchronize()
    _lib.check(sim._lib.b200env_debug_trace(sim._h, None, 0, 0))
    t = trace.cpu().numpy()
    used = t[:, :, 0] != 0
    grid = int(used[0].sum())
    t = t[:, :grid, :7].astype(np.float64)
    start = t[:, :, 0].min(axis=1)
    print(f"layout {args.layout}, {N} worlds, lanes {args.lanes or 'auto'}, grid {grid} CTAs; times in us relative to the "
          f"first CTA entry of the step: median [p5 .. p95] over CTAs")
    for s in range(2, args.steps):
        rel = (t[s] - start[s]) / 1e3
        period = (start[s] - start[s - 1]) / 1e3
        line = f"step {s:2d} period {period:6.2f} | "
        line += "  ".join(f"{n} {np.median(rel[:, i]):5.2f} [{np.percentile(rel[:, i], 5):5.2f}..{np.percentile(rel[:, i], 95):5.2f}]"
                          for i, n in enumerate(NAMES))
        print(line)
    last = (t[:, :, 6].max(axis=1) - start) / 1e3
    print("span first entry -> last exit per step (us):", np.round(last[2:], 2))
    prev_end = t[:-1, :, 6].max(axis=1)
    print("next step's first entry minus this step's last exit (us; negative = overlap):",
          np.round((start[1:] - prev_end)[2:] / 1e3, 2))


if __name__ == "__main__":
    main()
