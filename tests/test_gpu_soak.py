"""Many consecutive steps from one seeded state at a size where every SM carries dozens of tiles (1 M environments, two
resident batches stepped in turn, statistics on).  A race between the groups of a CTA over the tenants of a ring slot
(mbarrier parity taken for the previous phase) produced an `unspecified launch failure` only after thousands of tiles -- ~20
steps at 1-3 M environments -- which no test of a few steps saw.  Checked: every launch completes, the statistics count is exact, the episode
counters advanced by exactly the number of steps, nothing non-finite appears."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("robot,combined,soa,n,steps", [("humanoid3d", False, False, 1 << 20, 60), ("humanoid3d", False, True, 1 << 20, 60),
                                                         ("unitree_g1", True, False, 1 << 20, 60),
                                                         # 3 M envs in the middle launch shape (3 groups, 4 ring slots): failed within
                                                         # 20 steps before the slot-tenant fix (dm_tile_kernel.cuh, s_issued)
                                                         ("humanoid3d", False, False, 3 << 20, 30)])
def test_consecutive_steps_at_scale(robot, combined, soa, n, steps):
    import torch
    import bench
    from deepmimic_mujoco_b200.hotpath import HotPath
    B = 2
    hp = HotPath(robot, combined=combined)
    hp.load_combined() if combined else hp.load_motions(["walk"])
    dev = torch.device("cuda", 0)
    s = bench.synth_states(hp, n, B, 99, dev)
    d = hp.dims
    if soa:
        for k in ("qpos", "qvel", "action"):
            s[k] = s[k].transpose(1, 2).contiguous()
    shape = (lambda dim: (B, dim, n)) if soa else (lambda dim: (B, n, dim))
    out = dict(obs=torch.zeros(shape(d.obs_dim), device=dev), torque=torch.zeros(shape(d.nu), device=dev),
               reward=torch.zeros(B, n, device=dev), done=torch.zeros(B, n, dtype=torch.uint8, device=dev),
               reason=torch.zeros(B, n, dtype=torch.uint8, device=dev), ep_return=torch.zeros(B, n, device=dev))
    ep0 = s["ep_len"].clone()
    hp.stats_zero()
    for i in range(steps):
        b = i % B
        hp.step(n, s["qpos"][b], s["qvel"][b], s["action"][b], s["contact"][b], s["idx"][b], s["ep_len"][b], s["clip"][b],
                out["obs"][b], out["reward"][b], out["done"][b], out["reason"][b], out["torque"][b], None, out["ep_return"][b],
                soa_ld=(n if soa else 0), stats=True)
    torch.cuda.synchronize()
    st = hp.stats_read()
    assert st[0] == n * steps
    assert torch.equal(s["ep_len"], ep0 + steps // B)
    assert torch.isfinite(out["reward"]).all() and torch.isfinite(out["obs"]).all()
    assert np.isfinite(st).all()
    hp.close()
