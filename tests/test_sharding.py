"""Multi-process host logic on CPU (gloo, world_size 2): world slices, action sharding, the episode-statistics
reduction (the only collective of the package) and max-over-ranks timing.  Also checks, with the CPU oracle standing in
for one simulator per rank, that two shards stepped independently equal one job-wide run for Overcooked — the property
that lets worlds shard with no step-path communication."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from madrona_rl_envs_playground_b200.sharding import (allreduce_episode_stats, max_over_ranks, shard_actions,  # noqa: E402
                                                      world_slice)


def test_world_slice_partitions_every_world_once():
    for total in (1, 7, 32768, 65537):
        for ws in (1, 2, 3, 8):
            owned = np.zeros(total, dtype=np.int32)
            sizes = []
            for r in range(ws):
                a, b = world_slice(total, r, ws)
                owned[a:b] += 1
                sizes.append(b - a)
            assert (owned == 1).all()
            assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
    with pytest.raises(ValueError):
        world_slice(10, 2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world_size, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world_size))
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        from madrona_rl_envs_playground_b200.overcooked_layouts import get_base_layout_params
        from oracle.bindings import OracleOvercooked
        total, steps, horizon = 37, 60, 25
        params = get_base_layout_params("cramped_room", horizon)
        acts = np.random.default_rng(5).integers(0, 6, (steps, 2, total)).astype(np.int32)   # same stream on all ranks
        a, b = world_slice(total, rank, world_size)
        mine = shard_actions(torch.as_tensor(acts), rank, world_size)
        assert tuple(mine.shape) == (steps, 2, b - a)
        shard = OracleOvercooked(params, b - a)
        episodes = ret = 0
        running = np.zeros(b - a, dtype=np.int64)
        last = None
        for t in range(steps):
            obs, rew, done = shard.step(mine[t].numpy())
            running += rew[0]
            episodes += int(done.sum())
            ret += int(running[done != 0].sum())
            running[done != 0] = 0
            last = (obs, rew, done)
        job_eps, job_ret = allreduce_episode_stats(episodes, ret)
        slowest = max_over_ranks(1.0 + rank)
        if rank == 0:
            full = OracleOvercooked(params, total)
            f_eps = f_ret = 0
            f_run = np.zeros(total, dtype=np.int64)
            for t in range(steps):
                fo, fr, fd = full.step(acts[t])
                f_run += fr[0]
                f_eps += int(fd.sum())
                f_ret += int(f_run[fd != 0].sum())
                f_run[fd != 0] = 0
            ok = (np.array_equal(fo[:, a:b], last[0]) and np.array_equal(fr[:, a:b], last[1])
                  and np.array_equal(fd[a:b], last[2]) and (job_eps, job_ret) == (f_eps, f_ret)
                  and slowest == float(world_size))
            out.put(("ok" if ok else f"mismatch eps {job_eps}/{f_eps} ret {job_ret}/{f_ret} slowest {slowest}"))
    except Exception as e:   # pragma: no cover - reported to the parent
        if rank == 0:
            out.put(f"error: {e!r}")
        raise
    finally:
        dist.destroy_process_group()


def test_two_rank_sharding_gloo():
    from oracle import bindings
    bindings.build(ref=False)
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=120)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    assert out.get(timeout=5) == "ok"
