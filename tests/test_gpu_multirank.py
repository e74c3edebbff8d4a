"""Two NCCL ranks on two GPUs (skipped on a one-GPU box; run with `gpurun --gpus 2`): BASELINE config 3's shape --
humanoid3d `spinkick` + `dance_b` by environment id, environments sharded by contiguous blocks, the statistics
collective after the steps -- against ONE GPU stepping the whole batch:
  * sharded observations / rewards / flags / counters == single-GPU outputs bit for bit;
  * the collective's global totals == float64 numpy over the concatenated batches (count exact), the same bits on both
    ranks, over several step -> reduce cycles on a running accumulator (one or two steps between reductions);
  * five routes: the device-side peer exchange launched between the steps (dm_stats_allreduce); the auto mode
    (dm_stats_set_auto, one explicit exchange at the end) with the lagged exchange as its own kernel on the side stream
    ("auto": what 768 environments per rank get), as the extra CTA of the step kernel ("fused"), and switching between the
    two from step to step ("mixed": the exchanges share one sequence number across two streams); the NCCL fallback on the
    snapshot."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

E = 128 * 6          # environments per rank: whole tiles, so every row comes from the tile kernel on both sides
CYCLES = (1, 2, 1)   # steps between reductions


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _states(hp, n, batches):
    """Seeded on the CPU so that every process holds the same numbers."""
    import torch
    import bench
    return bench.synth_states(hp, n, batches, 31, torch.device("cpu"))


def _step_all(hp, s, dev, lo, hi, batches, reduce=None, before_step=None):
    import torch
    d = hp.dims
    n = hi - lo
    res = []
    b = 0
    for k in CYCLES:
        for _ in range(k):
            if before_step is not None:
                before_step(b)
            t = {key: v[b, lo:hi].contiguous().to(dev) for key, v in s.items()}
            out = dict(obs=torch.zeros(n, d.obs_dim, device=dev), reward=torch.zeros(n, device=dev),
                       done=torch.zeros(n, dtype=torch.uint8, device=dev), ep_return=torch.zeros(n, device=dev))
            hp.step(n, t["qpos"], t["qvel"], t["action"], t["contact"], t["idx"], t["ep_len"], t["clip"], out["obs"], out["reward"],
                    out["done"], None, None, None, out["ep_return"], stats=True)
            out.update(idx=t["idx"], ep_len=t["ep_len"])
            res.append({key: v.cpu().numpy() for key, v in out.items()})
            b += 1
        if reduce is not None:
            reduce()
    assert b == batches
    return res


def _worker(rank, world, port, q, route):
    force_nccl = route == "nccl"
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch
    import torch.distributed as dist
    from deepmimic_mujoco_b200 import dist as dmdist
    from deepmimic_mujoco_b200.hotpath import HotPath
    dmdist.init("nccl")
    dev = torch.device("cuda", rank)
    hp = HotPath("humanoid3d", device=rank)
    hp.load_motions(["spinkick", "dance_b"])
    batches = sum(CYCLES)
    s = _states(hp, world * E, batches)
    lo, hi = dmdist.shard(world * E, rank, world)
    with torch.cuda.device(dev):
        coll = dmdist.StatsCollective(hp, force_nccl=force_nccl)
        hp.stats_zero()
        if route in ("auto", "fused", "mixed"):  # every step comes with an exchange; one explicit exchange afterwards makes the totals exact
            assert coll.auto(1)
            pick = {"auto": lambda b: None, "fused": lambda b: os.environ.update(DM_PEER_SIDE="0"),
                    "mixed": lambda b: os.environ.update(DM_PEER_SIDE=str(b % 2))}[route]
            res = _step_all(hp, s, dev, lo, hi, batches, lambda: None, pick)
            coll.reduce()
            coll.auto(0)
            os.environ.pop("DM_PEER_SIDE", None)
        else:
            res = _step_all(hp, s, dev, lo, hi, batches, coll.reduce)
        torch.cuda.synchronize()
        totals = coll.totals()
        local_left = hp.stats_read()  # the running accumulators after the last reduction: zero
        full = None
        if rank == 0:  # the same work on one GPU, one handle
            hp1 = HotPath("humanoid3d", device=0)
            hp1.load_motions(["spinkick", "dance_b"])
            full = _step_all(hp1, s, dev, 0, world * E, batches)
            hp1.close()
    q.put((rank, coll.mode, coll.why_not_peer, res, totals, local_left, full))
    dist.barrier()
    dist.destroy_process_group()
    hp.close()


@pytest.mark.timeout(300)
@pytest.mark.parametrize("route", ["peer", "auto", "fused", "mixed", "nccl"])
def test_two_ranks_equal_one_gpu(route):
    force_nccl = route == "nccl"
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, world, port, q, route)) for r in range(world)]
    for p in ps:
        p.start()
    got = sorted([q.get(timeout=240) for _ in ps], key=lambda x: x[0])
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    modes = {g[1] for g in got}
    assert modes == ({"nccl"} if force_nccl else {"peer"}), (modes, got[0][2], got[1][2])
    full = got[0][6]
    for b in range(sum(CYCLES)):
        for key in full[b]:  # concatenated shards == the one-GPU batch, bit for bit
            np.testing.assert_array_equal(np.concatenate([g[3][b][key] for g in got]), full[b][key], err_msg="%s batch %d" % (key, b))
    obs = np.concatenate([f["obs"] for f in full]).astype(np.float64)
    rew = np.concatenate([f["reward"] for f in full]).astype(np.float64)
    done = np.concatenate([f["done"] for f in full]).astype(bool)
    ep_ret = np.concatenate([f["ep_return"] for f in full]).astype(np.float64)
    ep_len = np.concatenate([f["ep_len"] for f in full]).astype(np.float64)
    want = np.concatenate([[len(rew), rew.sum(), (rew * rew).sum(), done.sum(), ep_ret[done].sum(), ep_len[done].sum()],
                           obs.sum(0), (obs * obs).sum(0)])
    for g in got:
        assert g[4][0] == world * E * sum(CYCLES)  # count exact after three reductions of a running accumulator
        assert g[4][3] == done.sum()
        np.testing.assert_allclose(g[4], want, rtol=2e-6, atol=1e-6)
        assert not g[5].any()
    np.testing.assert_array_equal(got[0][4], got[1][4])  # the same bits on every rank
