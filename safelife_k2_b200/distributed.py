"""
Multi-GPU plumbing: environments shard by index across ranks (one process per GPU), with no
data-path collective — an env never reads another env's state and its spawn stream is keyed by its
GLOBAL env index, so results do not depend on placement.  The only exchange is an
all-reduce (NCCL over NVLink on GPUs, gloo in the CPU tests) of the small episode/reward statistics
vector that slb_accumulate_stats maintains.
"""
import os

import torch
import torch.distributed as dist

STAT_NAMES = ("sum_reward", "episodes_done", "sum_episode_length", "sum_episode_reward",
              "sum_completed", "sum_possible", "env_steps", "n_game_over")


def shard_range(n_total, rank, world_size):
    """Contiguous block of env indices for `rank`: env e -> rank floor(e * G / N) (inverse map)."""
    lo = (n_total * rank + world_size - 1) // world_size
    hi = (n_total * (rank + 1) + world_size - 1) // world_size
    return lo, hi


def init_from_env(backend=None):
    """Initialise torch.distributed from RANK/WORLD_SIZE/MASTER_* (torchrun).  Returns
    (rank, world_size, local_rank); world_size 1 needs no process group."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group(backend=backend, rank=rank, world_size=world)
    return rank, world, local


def all_reduce_stats(stats):
    """Sum the int64 statistics vector over all ranks (in place); returns it."""
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)
    return stats


def max_over_ranks(value, device=None):
    """Max of a python float over ranks (timing is reported as the slowest rank)."""
    if not (dist.is_initialized() and dist.get_world_size() > 1):
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def stats_dict(stats):
    v = stats.detach().cpu().tolist()
    return dict(zip(STAT_NAMES, v))
