"""Sharding of independent envs over the GPUs of one box and the one collective of the path.

The reference has no distributed code: its drivers fan
independent envs out over ``multiprocessing.Process`` (tests/test_rmsa_threads_us.py:130-256)
and join.  Here the envs of one job are a contiguous range of *global env ids*, split
into contiguous per-rank shards; a rank's engine is created with
``global_env_offset = shard start`` so that Philox request streams and replay-trace rows
are keyed by the global id and every env's trajectory is independent of the GPU count.
No collective is on the step path; the only exchange is a sum/max all-reduce of a small
episode-statistics vector (NCCL on GPUs, gloo in the CPU tests).
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

STAT_NAMES = (
    "services_processed", "services_accepted", "bit_rate_requested", "bit_rate_provisioned",
    "episodes", "reward_sum", "envs",
)


def shard_bounds(num_envs_global: int, world_size: int, rank: int) -> Tuple[int, int]:
    """[start, stop) of the global env ids owned by ``rank``; sizes differ by at most one
    and lower ranks take the remainder."""
    if world_size < 1 or not (0 <= rank < world_size) or num_envs_global < 0:
        raise ValueError("bad shard request")
    base, rem = divmod(int(num_envs_global), int(world_size))
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


def local_stats(counters, reward_sum=None):
    """counters: int64 tensor [n_env, 12] (orlg_get_counters layout, include/orlg.h).
    Returns (sums int64[7], fault int64[1]) on the counters' device."""
    import torch

    c = counters
    rs = torch.zeros((), dtype=torch.int64, device=c.device) if reward_sum is None else reward_sum.sum().to(torch.int64)
    sums = torch.stack([c[:, 0].sum(), c[:, 1].sum(), c[:, 4].sum(), c[:, 5].sum(), c[:, 11].sum(), rs,
                        torch.tensor(c.shape[0], dtype=torch.int64, device=c.device)]).to(torch.int64)
    fault = c[:, 10].max().reshape(1).to(torch.int64) if c.shape[0] else torch.zeros(1, dtype=torch.int64, device=c.device)
    return sums, fault


def reduce_episode_stats(counters, reward_sum=None, group=None) -> Dict[str, float]:
    """All-reduce the episode statistics of every rank's shard (sum; fault bits: max).

    Works with or without an initialised ``torch.distributed`` process group (without one
    it just summarises the local shard).  Returns plain Python numbers, identical on
    every rank, including the blocking rates the reference reports in its info dict
    (rmsa_env.py:293-307)."""
    import torch
    import torch.distributed as dist

    sums, fault = local_stats(counters, reward_sum)
    if dist.is_available() and dist.is_initialized():
        dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(fault, op=dist.ReduceOp.MAX, group=group)
    vals = [int(x) for x in sums.cpu()]
    out: Dict[str, float] = dict(zip(STAT_NAMES, vals))
    out["fault"] = int(fault.cpu()[0])
    proc, acc, brq, bpr = vals[0], vals[1], vals[2], vals[3]
    out["service_blocking_rate"] = (proc - acc) / proc if proc else 0.0
    out["bit_rate_blocking_rate"] = (brq - bpr) / brq if brq else 0.0
    return out


def init_from_env(backend: Optional[str] = None):
    """Join the torchrun-style rendezvous described by RANK/WORLD_SIZE/MASTER_* if any.
    Returns (rank, world_size, local_rank)."""
    import os

    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kw = {}
        if backend == "nccl":
            torch.cuda.set_device(local)
            kw["device_id"] = torch.device("cuda", local)
        dist.init_process_group(backend, **kw)
    return rank, world, local
