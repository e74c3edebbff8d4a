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
    """torch float64 tensor ALIASING the handle's RUNNING device accumulators (no copy).  Never all-reduce this tensor in
    place: the step kernels keep adding into it, so every further collective would multiply what was already summed
    (round 1's bug: count 1.95e97 after 200 steps).  Use StatsCollective."""
    import torch
    ptr, n = hotpath.stats_device_ptr()
    return torch.as_tensor(_DevArray(ptr, n), device=torch.device("cuda", hotpath.device))


def allreduce_(t):
    """In-place sum all-reduce (NCCL over NVLink on GPUs, gloo on CPU tensors); returns t."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t


def snapshot_allreduce_(acc, total):
    """The protocol of the collective on plain torch tensors (any backend; the CPU / gloo tests run this):
    snapshot-and-zero the running sums, all-reduce the SNAPSHOT, add it to the global totals."""
    snap = acc.clone()
    acc.zero_()
    allreduce_(snap)
    total += snap
    return total


class StatsCollective:
    """The one collective of the path for a HotPath handle per rank.

    mode "peer"  : dm_stats_allreduce -- one single-CTA kernel per call that exchanges the snapshot through peer-mapped
                   memory over NVLink (csrc/dm_peer.cuh: tagged 16-byte lines, no flags, no fences); no host-enqueued
                   collective in the step loop.  auto(every) lets dm_step issue a lagged exchange itself.
    mode "nccl"  : checked fallback when the cudaIpc mapping is not available: dm_stats_snapshot, ncclAllReduce of the
                   snapshot through torch.distributed, dm_stats_commit.
    mode "local" : world size 1.
    Every rank must call reduce() the same number of times.  totals() = global running sums since the last stats_zero()."""

    def __init__(self, hotpath, force_nccl=False):
        import torch
        import torch.distributed as dist
        from . import _native as N
        self.hp = hotpath
        self.world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        self.rank = dist.get_rank() if self.world > 1 else 0
        self.mode = "local"
        self.why_not_peer = None
        if self.world > 1:
            ok = 0 if force_nccl else 1
            handle = b"\0" * N.IPC_HANDLE_BYTES
            if ok:
                try:
                    handle = hotpath.stats_peer_export(self.rank, self.world)
                except N.DmError as e:
                    ok, self.why_not_peer = 0, str(e)
            gathered = [None] * self.world
            dist.all_gather_object(gathered, (ok, handle))
            if all(g[0] for g in gathered):
                try:
                    hotpath.stats_peer_import([g[1] for g in gathered])
                except N.DmError as e:
                    ok, self.why_not_peer = 0, str(e)
            else:
                ok = 0
            flag = torch.tensor([ok], device=torch.device("cuda", hotpath.device), dtype=torch.int32)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)  # every rank takes the same route
            self.mode = "peer" if int(flag.item()) == 1 else "nccl"
            if self.mode == "nccl":
                ptr, n = hotpath.stats_snapshot()
                self._snap = torch.as_tensor(_DevArray(ptr, n), device=torch.device("cuda", hotpath.device))

    def reduce(self):
        if self.mode == "nccl":
            self.hp.stats_snapshot()
            allreduce_(self._snap)
            self.hp.stats_commit()
        else:
            self.hp.stats_allreduce()

    def auto(self, every):
        """Let dm_step issue a LAGGED exchange itself with every `every`-th step (small steps: a one-CTA kernel on a
        dependency-free side stream; steps that fill the GPU: one extra CTA of the step kernel).  The totals trail by one
        exchange until the next reduce().  Only the device-side route can do that; returns False in "nccl" mode, where the
        caller keeps calling reduce()."""
        if self.mode == "nccl":
            return False
        self.hp.stats_set_auto(every)
        return True

    def join(self):
        self.hp.stats_join()

    def totals(self):
        return self.hp.stats_read_total()


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
