"""Episode-sharded data parallelism: one process per GPU, every rank runs the
same pipeline on its own shard of the global episode batch, the flat gradient buffer is summed with ONE
all-reduce (NCCL over NVLink/NVSwitch on GPUs, gloo in the CPU tests) and Adam applies it scaled by
1/world.  The reference has no multi-device code; the only shared state is the weights
(model/graph.py:57-85 touches one episode at a time), so no other collective exists on the path.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def env_world():
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


def init_from_env(backend=None):
    """Initialises torch.distributed from RANK/LOCAL_RANK/WORLD_SIZE/MASTER_* (torchrun).  Returns
    (rank, local_rank, world).  No-op for world == 1."""
    rank, local, world = env_world()
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kw = {}
        if backend == "nccl":
            torch.cuda.set_device(local)
            kw["device_id"] = torch.device("cuda", local)
        dist.init_process_group(backend, **kw)
    return rank, local, world


def episode_shard(global_batch: int, rank: int, world: int):
    """[lo, hi) of the global episode batch owned by ``rank``: contiguous, sizes differ by at most one."""
    if not 0 <= rank < world:
        raise ValueError("rank %d outside world %d" % (rank, world))
    base, extra = divmod(global_batch, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def allreduce_sum_(flat: torch.Tensor):
    """In-place sum of the flat gradient buffer over all ranks (enqueued on the current stream)."""
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)
    return flat


def grad_scale(shard_episodes: int, global_batch: int) -> float:
    """Each rank's loss is a mean over ITS episodes; the global loss is the mean over all of them, so
    a rank's gradient enters the sum with weight shard/global (= 1/world for equal shards)."""
    return float(shard_episodes) / float(global_batch)


def broadcast_(flat: torch.Tensor, src: int = 0):
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.broadcast(flat, src)
    return flat
