"""Multi-GPU host logic on CPU: two `gloo` ranks shard the environments, reduce the fp64 statistics vector
with the same all-reduce the NCCL path uses, and merge it into running moments -- identically on every rank
and equal to float64 numpy over the concatenated batch."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from deepmimic_mujoco_b200 import dist as dmdist

D = 67


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _batch(step):
    rng = np.random.default_rng(100 + step)
    n = 1000 + 37 * step  # not divisible by the world size on purpose
    obs = rng.normal(0.3, 2.0, (n, D))
    rew = rng.uniform(0, 1, n)
    done = rng.random(n) < 0.05
    ep_ret = rng.uniform(0, 300, n)
    ep_len = rng.integers(1, 1000, n)
    return obs, rew, done, ep_ret, ep_len


def _sums(obs, rew, done, ep_ret, ep_len):
    return np.concatenate([[len(rew), rew.sum(), (rew ** 2).sum(), done.sum(), ep_ret[done].sum(), ep_len[done].sum()],
                           obs.sum(0), (obs ** 2).sum(0)])


def _worker(rank, world, port, q):
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    r, lr, w = dmdist.init("gloo")
    assert (r, w) == (rank, world)
    norm = dmdist.RunningNorm(D)
    last = None
    for step in range(3):
        full = _batch(step)
        lo, hi = dmdist.shard(len(full[1]), rank, world)
        t = torch.from_numpy(_sums(*[a[lo:hi] for a in full]))
        dmdist.allreduce_(t)
        last = t.numpy().copy()
        norm.update_from_sums(last[0], last[6:6 + D], last[6 + D:])
    # the collective's protocol on a RUNNING accumulator (reducing the running sums in place multiplied
    # them): four step -> reduce cycles, two of them with two steps between the reductions
    acc, total = torch.zeros(6 + 2 * D, dtype=torch.float64), torch.zeros(6 + 2 * D, dtype=torch.float64)
    step = 0
    for steps_between in (1, 2, 1, 2):
        for _ in range(steps_between):
            full = _batch(step)
            lo, hi = dmdist.shard(len(full[1]), rank, world)
            acc += torch.from_numpy(_sums(*[a[lo:hi] for a in full]))
            step += 1
        dmdist.snapshot_allreduce_(acc, total)
        assert not acc.any()
    q.put((rank, last, norm.mean, norm.var, norm.n, total.numpy().copy(), step))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_covers_every_env_once():
    for n, w in [(4096, 8), (1037, 2), (5, 8), (0, 4)]:
        spans = [dmdist.shard(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        assert max(h - l for l, h in spans) - min(h - l for l, h in spans) <= 1


@pytest.mark.timeout(120)
def test_two_rank_statistics_allreduce():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted([q.get(timeout=100) for _ in ps], key=lambda x: x[0])
    for p in ps:
        p.join(timeout=30)
        assert p.exitcode == 0
    full = _batch(2)
    np.testing.assert_allclose(res[0][1], _sums(*full), rtol=1e-12)
    np.testing.assert_array_equal(res[0][1], res[1][1])          # every rank holds the same reduced vector
    allobs = np.concatenate([_batch(s)[0] for s in range(3)])
    for r in res:
        np.testing.assert_allclose(r[2], allobs.mean(0), rtol=1e-10, atol=1e-12)
        np.testing.assert_allclose(r[3], allobs.var(0), rtol=1e-9)
        assert r[4] == len(allobs)
    np.testing.assert_array_equal(res[0][2], res[1][2])
    # running totals after four reductions == float64 numpy over every batch of every rank; the count is exact
    want = sum(_sums(*_batch(s)) for s in range(res[0][6]))
    for r in res:
        assert r[5][0] == want[0] == sum(len(_batch(s)[1]) for s in range(res[0][6]))
        np.testing.assert_allclose(r[5], want, rtol=1e-12)
    np.testing.assert_array_equal(res[0][5], res[1][5])
    m = dmdist.moments(res[0][1], D)
    np.testing.assert_allclose(m["reward_mean"], full[1].mean(), rtol=1e-12)
    np.testing.assert_allclose(m["done_rate"], full[2].mean(), rtol=1e-12)
    np.testing.assert_allclose(m["obs_var"], full[0].var(0), rtol=1e-9)
