"""Multi-GPU plumbing: one process per GPU, environments sharded by contiguous blocks, and the one
collective of this path -- a sum all-reduce of the small fp64 statistics vector
[count, sum r, sum r^2, n_done, sum ep_return, sum ep_len, sum obs[D], sum obs^2[D]].
The reference has no collectives at all (SubprocVecEnv pipes only, src/sb3_ppo.py:275-278); the
normaliser statistics are new functionality ("parity unpinned"), checked against float64 numpy.
"""
from __future__ import annotations

import os

import numpy as np


def env_info():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def init(backend=None):
    """Initialise torch.distributed from the torchrun environment (no-op for world size 1)."""
    import torch
    import torch.distributed as dist
    rank, local_rank, world = env_info()
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local_rank)
        dist.init_process_group(backend=backend, rank=rank, world_size=world)
    return rank, local_rank, world


def shard(n_total, rank, world):
    """Contiguous block [lo, hi) of the global environment ids owned by `rank` (remainder to the low ranks)."""
    base, rem = divmod(int(n_total), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class _DevArray:
    """__cuda_array_interface__ view of a raw device pointer so torch can alias the library's accumulators."""

    def __init__(self, ptr, n, typestr="<f8"):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (int(ptr), False), "version": 2}


def stats_tensor(hotpath):
    """torch float64 tensor ALIASING the handle's device accumulators (no copy)."""
    import torch
    ptr, n = hotpath.stats_device_ptr()
    return torch.as_tensor(_DevArray(ptr, n), device=torch.device("cuda", hotpath.device))


def allreduce_(t):
    """In-place sum all-reduce (NCCL over NVLink on GPUs, gloo on CPU tensors); returns t."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t


def moments(stats, obs_dim):
    """Sums -> dict(count, reward_mean, reward_var, done_rate, ep_return_mean, ep_len_mean, obs_mean, obs_var)."""
    s = np.asarray(stats, dtype=np.float64)
    c = max(s[0], 1.0)
    nd = max(s[3], 1.0)
    om = s[6:6 + obs_dim] / c
    return dict(count=s[0], reward_mean=s[1] / c, reward_var=max(s[2] / c - (s[1] / c) ** 2, 0.0), done_rate=s[3] / c,
                ep_return_mean=s[4] / nd, ep_len_mean=s[5] / nd, obs_mean=om,
                obs_var=np.maximum(s[6 + obs_dim:6 + 2 * obs_dim] / c - om ** 2, 0.0))


class RunningNorm:
    """Running mean / variance merged with the parallel (Chan et al.) update; identical on every rank because the
    batch sums are all-reduced first."""

    def __init__(self, dim):
        self.n = 0.0
        self.mean = np.zeros(dim)
        self.m2 = np.zeros(dim)

    def update_from_sums(self, count, s, ss):
        if count <= 0:
            return
        bm = s / count
        bm2 = np.maximum(ss - count * bm * bm, 0.0)
        d = bm - self.mean
        tot = self.n + count
        self.mean = self.mean + d * (count / tot)
        self.m2 = self.m2 + bm2 + d * d * (self.n * count / tot)
        self.n = tot

    @property
    def var(self):
        return self.m2 / max(self.n, 1.0)
