"""World sharding across the GPUs of one box (one process per GPU, launched by torchrun).

Worlds are independent units: rank g owns the contiguous slice ``[start, stop)`` of the job's worlds
and drives its own simulator instance; the step path has NO collective.  The reference has no multi-GPU mode — N GPUs
there are N independent Managers (`src/overcooked_env/mgr.cpp:97`, every script passes gpu_id 0) — so a shard here is
exactly one reference Manager of ``stop - start`` worlds, including its own episode counter for the RNG environments.

The only exchange is the optional reduction of episode statistics (a few integers), off the step path; it goes
through ``torch.distributed`` (NCCL over NVLink on GPUs, gloo in the CPU tests).
"""
import os


def rank_info():
    """(rank, world_size, local_rank) from the torchrun environment (1 process = 1 GPU)."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def world_slice(total_worlds, rank, world_size):
    """Contiguous slice of ``total_worlds`` owned by ``rank``; sizes differ by at most one (earlier ranks get the
    remainder), every world is owned exactly once."""
    if not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} outside world size {world_size}")
    base, extra = divmod(int(total_worlds), int(world_size))
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def shard_actions(actions, rank, world_size, world_dim=-1):
    """This rank's columns of a job-wide action tensor (worlds along ``world_dim``)."""
    start, stop = world_slice(actions.shape[world_dim], rank, world_size)
    return actions.narrow(world_dim % actions.dim(), start, stop - start)


def allreduce_episode_stats(episodes, return_sum, device=None, group=None):
    """Job-wide (finished episodes, sum of their returns) from the per-rank values of ``sim.episode_stats()``."""
    import torch
    import torch.distributed as dist
    t = torch.tensor([int(episodes), int(return_sum)], dtype=torch.int64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return int(t[0].item()), int(t[1].item())


def max_over_ranks(value, device=None, group=None):
    """Max of a per-rank scalar (device-side timing of a multi-GPU run is the max over ranks)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())
