#!/usr/bin/env python
"""What is this box's device->host ceiling?  ("what's weak" 4: end-to-end Overcooked at 8 GPUs reached
96 GB/s aggregate; the wall or an accident?)

Every rank copies the per-step output of the end-to-end bench (34.5 MB: one observation tensor + rewards + done of
32 768 cramped_room worlds) from its GPU into pinned host memory with plain cudaMemcpyAsync, `iters` times back to
back, all ranks at once, and reports GB/s per rank and in aggregate.  No simulator involved.  Variants:

  affinity=inherit   whatever CPU affinity the launcher gave the rank (what bench.py runs under)
  affinity=spread    rank r pinned to its own slice of the visible cores before it allocates and touches the buffer
  chunks=1 | 3       one 34.5 MB copy, or three copies of the bench's tensor sizes (obs, rewards, done)
  h2d                the opposite direction, for comparison

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29512 \
        tools/probes/d2h_probe.py [--iters 200]

Prints one JSON line per variant on rank 0."""
import argparse
import json
import os
import time

import torch
import torch.distributed as dist

OBS, REW, DONE = 2 * 32768 * 5 * 4 * 26, 2 * 32768 * 4, 32768 * 4
VARIANTS = (("d2h", 1, "inherit"), ("d2h", 3, "inherit"), ("d2h", 1, "spread"), ("h2d", 1, "inherit"))


def run(args, rank, world, local, cores0, direction, chunks, affinity):
    if affinity == "spread":
        per = max(1, len(cores0) // world)
        os.sched_setaffinity(0, set(cores0[local * per:(local + 1) * per]) or set(cores0))
    else:
        os.sched_setaffinity(0, set(cores0))
    sizes = [OBS + REW + DONE] if chunks == 1 else [OBS, REW, DONE]
    dev = [torch.empty(n, dtype=torch.uint8, device="cuda") for n in sizes]
    host = [torch.empty(n, dtype=torch.uint8).pin_memory() for n in sizes]
    for h in host:
        h.zero_()          # first touch under the affinity chosen above
    stream = torch.cuda.Stream()

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def copy():
        with torch.cuda.stream(stream):
            for d, h in zip(dev, host):
                if direction == "d2h":
                    h.copy_(d, non_blocking=True)
                else:
                    d.copy_(h, non_blocking=True)

    for _ in range(5):
        copy()
    sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record(stream)
    for _ in range(args.iters):
        copy()
    e1.record(stream)
    stream.synchronize()
    wall = time.perf_counter() - t0
    gbs = sum(sizes) * args.iters / (e0.elapsed_time(e1) * 1e-3) / 1e9
    t = torch.tensor([gbs, wall], dtype=torch.float64, device="cuda")
    all_t = [torch.zeros_like(t) for _ in range(world)] if world > 1 else [t]
    if world > 1:
        dist.all_gather(all_t, t)
    sync_all()
    if rank == 0:
        per_rank = [float(x[0]) for x in all_t]
        wall_max = max(float(x[1]) for x in all_t)
        print(json.dumps({"probe": direction, "ranks": world, "chunks": chunks, "affinity": affinity,
                          "bytes_per_iter": sum(sizes), "iters": args.iters,
                          "gbs_per_rank": [round(g, 2) for g in per_rank],
                          "gbs_aggregate_sum_of_ranks": round(sum(per_rank), 2),
                          "gbs_aggregate_wall": round(world * sum(sizes) * args.iters / wall_max / 1e9, 2),
                          "visible_cores": len(cores0)}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=200)
    args = ap.parse_args()
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cores0 = sorted(os.sched_getaffinity(0))
    for direction, chunks, affinity in VARIANTS:
        run(args, rank, world, local, cores0, direction, chunks, affinity)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
