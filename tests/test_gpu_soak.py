"""Many consecutive steps from one seeded state at a size where every SM carries dozens of tiles (1 M environments, two
resident batches stepped in turn, statistics on): a variant of the step arithmetic once produced a launch failure after ~20
such steps that no test of a few steps saw.  Checked: every launch completes, the statistics count is exact, the episode
counters advanced by exactly the number of steps, nothing non-finite appears."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("robot,combined,soa", [("humanoid3d", False, False), ("humanoid3d", False, True), ("unitree_g1", True, False)])
def test_sixty_steps_at_one_million_envs(robot, combined, soa):
    import torch
    import bench
    from deepmimic_mujoco_b200.hotpath import HotPath
    n, B, steps = 1 << 20, 2, 60
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
