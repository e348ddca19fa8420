"""Multi-GPU plumbing for the hot path: one process per GPU, environments sharded by env index,
NO collective in the step itself; the only exchange is a sum all-reduce of the <= 16-float
episode-statistics vector once per horizon.

Reference behaviour kept: per-rank device from LOCAL_RANK (utils/rlgames_utils.py:89-107), per-rank
seed = seed + rank (utils/utils.py:87-94); every rank owns an independent env shard.
"""
import os

import torch

from . import _abi

STAT_NAMES = ["steps", "reward_sum", "done", "timeout", "done_length_sum"] + \
             [f"task{i}" for i in range(_abi.IGE_NUM_STATS - 5)]


def env_info():
    """(rank, local_rank, world_size) from the torchrun environment (defaults: single process)."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")),
            int(os.environ.get("WORLD_SIZE", "1")))


def init_process_group(device=None):
    """NCCL over NVLink for CUDA ranks, gloo for CPU ranks (tests).  No-op for world_size 1."""
    import torch.distributed as dist
    rank, local_rank, world = env_info()
    if world == 1 or dist.is_initialized():
        return rank, local_rank, world
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    use_cuda = device is not None and torch.device(device).type == "cuda"
    if use_cuda:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device(device))
    else:
        dist.init_process_group("gloo", rank=rank, world_size=world)
    return rank, local_rank, world


def env_shard(global_num_envs, rank, world):
    """Contiguous env-index range [start, start + count) owned by `rank` (agent rows env*A+a stay together)."""
    base, rem = divmod(int(global_num_envs), int(world))
    count = base + (1 if rank < rem else 0)
    start = rank * base + min(rank, rem)
    return start, count


def rank_seed(seed, rank):
    return int(seed) + int(rank)          # utils/utils.py:87-94


class StatsReducer:
    """Sum all-reduce of the episode-statistics vector, off the step's critical path: the step
    kernels keep adding into the env's accumulator while a snapshot is being reduced.

    On CUDA the snapshot and the NCCL all-reduce are issued on a SIDE stream that only waits for the step
    stream's work up to now; the step stream never waits for the collective (a `work.wait()` on the step
    stream put the ~15 us all-reduce latency into every horizon: 89 % weak-scaling efficiency at 2 GPUs).
    ``wait()`` joins the side stream when the reduced numbers are actually needed."""

    def __init__(self, device):
        self.device = torch.device(device)
        self.snap = torch.zeros(_abi.IGE_NUM_STATS, dtype=torch.float32, device=device)
        self.side = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None
        self._work = None

    def start(self, stats_vector):
        """`stats_vector`: the [IGE_NUM_STATS] tensor, or a callable returning it (evaluated on the side stream)."""
        import torch.distributed as dist
        multi = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
        if self.side is None:
            self.snap.copy_(stats_vector() if callable(stats_vector) else stats_vector)
            if multi:
                self._work = dist.all_reduce(self.snap, op=dist.ReduceOp.SUM, async_op=True)
            return
        self.side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(self.side):
            self.snap.copy_(stats_vector() if callable(stats_vector) else stats_vector)
            if multi:
                dist.all_reduce(self.snap, op=dist.ReduceOp.SUM, async_op=True).wait()   # side stream waits, not the step's

    def wait(self):
        if self._work is not None:
            self._work.wait()
            self._work = None
        if self.side is not None:
            torch.cuda.current_stream(self.device).wait_stream(self.side)
        return self.snap

    def result(self):
        v = self.wait().detach().cpu().tolist()
        out = dict(zip(STAT_NAMES, v))
        steps, done = max(out["steps"], 1.0), max(out["done"], 1.0)
        out["mean_reward"] = out["reward_sum"] / steps
        out["mean_episode_length"] = out["done_length_sum"] / done
        return out
