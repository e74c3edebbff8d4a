"""Batched GPU checks at sizes the golden files do not reach: tile kernel (K1b) == strided kernel (K1a)
(integers and flags bit for bit, floats to 2 ulp), both == the float64 oracle on a seeded random batch, SoA == AoS, fused statistics == numpy,
and size-independent properties at a BASELINE-sized batch."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CASES = [("humanoid3d", "walk", False), ("unitree_g1", "run", False), ("unitree_g1", None, True)]


def _make(robot, motion, combined, n, seed=3):
    import torch
    import bench
    from deepmimic_mujoco_b200.hotpath import HotPath
    hp = HotPath(robot, combined=combined)
    hp.load_combined() if combined else hp.load_motions([motion])
    s = bench.synth_states(hp, n, 1, seed, torch.device("cuda", 0))
    s = {k: v[0].contiguous() for k, v in s.items()}
    return hp, s


def _run(hp, s, variant, soa=False, stats=False):
    import torch
    n = s["qpos"].shape[0]
    d = hp.dims
    hp.set_kernel_variant(variant)
    dev = s["qpos"].device
    out = dict(obs=torch.zeros(n, d.obs_dim, device=dev), reward=torch.zeros(n, device=dev),
               done=torch.zeros(n, dtype=torch.uint8, device=dev), reason=torch.zeros(n, dtype=torch.uint8, device=dev),
               torque=torch.zeros(n, d.nu, device=dev), terms=torch.zeros(n, 8, device=dev), ep_return=torch.zeros(n, device=dev),
               idx=s["idx"].clone(), ep_len=s["ep_len"].clone(), clip=s["clip"].clone())
    q, v, a = s["qpos"], s["qvel"], s["action"]
    if soa:
        q, v, a = q.t().contiguous(), v.t().contiguous(), a.t().contiguous()
        for k in ("obs", "torque", "terms"):
            out[k] = out[k].t().contiguous()
    if stats:
        hp.stats_zero()
    hp.step(n, q, v, a, s["contact"], out["idx"], out["ep_len"], out["clip"], out["obs"], out["reward"], out["done"],
            out["reason"], out["torque"], out["terms"], out["ep_return"], soa_ld=(n if soa else 0), stats=stats)
    torch.cuda.synchronize()
    if soa:
        for k in ("obs", "torque", "terms"):
            out[k] = out[k].t().contiguous()
    res = {k: t.cpu().numpy() for k, t in out.items()}
    if stats:
        res["stats"] = hp.stats_read()
    return res


@pytest.mark.parametrize("robot,motion,combined", CASES)
def test_tile_kernel_equals_strided_kernel(robot, motion, combined):
    hp, s = _make(robot, motion, combined, 128 * 9 + 37)
    a, b = _run(hp, s, 0), _run(hp, s, 1)
    for k in a:  # two instantiations of the same arithmetic: nvcc may contract a*b+c differently, so floats agree to an
        if a[k].dtype.kind == "f":  # ulp or two, integers and flags exactly
            np.testing.assert_allclose(a[k], b[k], rtol=2e-6, atol=2e-7, err_msg=k)
        else:
            np.testing.assert_array_equal(a[k], b[k], err_msg=k)
    c = _run(hp, s, 0, soa=True)
    for k in a:
        np.testing.assert_array_equal(a[k], c[k], err_msg="soa " + k)


@pytest.mark.parametrize("robot,motion,combined", CASES)
def test_launch_shapes_agree(robot, motion, combined, monkeypatch):
    """The same batch through every launch shape of the tile kernel: the two-group instantiation small batches get (no
    register cap), the capped one large batches get (DM_WIDE=0), without programmatic dependent launch (DM_PDL=0), and without
    the per-environment row stage that tables too big for shared memory get (DM_ROW_STAGE=0; G1 combined);
    statistics included.  Same arithmetic, so integers and flags agree exactly and floats to an ulp or two."""
    n = 128 * 40
    res = []
    for env in ({}, {"DM_WIDE": "0"}, {"DM_PDL": "0"}, {"DM_ROW_STAGE": "0"}):
        for k in ("DM_WIDE", "DM_PDL", "DM_ROW_STAGE"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)  # read when the handle is created
        hp, s = _make(robot, motion, combined, n)
        res.append(_run(hp, s, 1, stats=True))
        hp.close()
    for other in res[1:]:
        for k in res[0]:
            if res[0][k].dtype.kind == "f":
                np.testing.assert_allclose(res[0][k], other[k], rtol=2e-6, atol=2e-7, err_msg=k)
            else:
                np.testing.assert_array_equal(res[0][k], other[k], err_msg=k)


@pytest.mark.parametrize("robot,motion,combined", CASES)
def test_random_batch_matches_oracle(robot, motion, combined):
    from oracle.envs import MultiClipEnv, SingleClipEnv
    n = 256
    hp, s = _make(robot, motion, combined, n, seed=11)
    got = _run(hp, s, 1)
    h = {k: v.cpu().numpy() for k, v in s.items()}
    env = MultiClipEnv() if combined else SingleClipEnv(motion, robot)
    if combined:
        env.reset(coin=0, frame=0)
    for i in range(n):
        if combined:
            env.current_motion_mocap = env.clips[int(h["clip"][i])]
            env.current_motion_n_steps = int(h["idx"][i])
        else:
            env.idx_curr = int(h["idx"][i])
        env.episode_length = int(h["ep_len"][i])
        obs, rew, done, info = env.step(h["action"][i].astype(np.float64), (h["qpos"][i].astype(np.float64), h["qvel"][i].astype(np.float64)),
                                        int(h["contact"][i]))
        tol = lambda b: 1e-5 * np.maximum(np.abs(b), 0.05)
        assert np.all(np.abs(got["obs"][i] - obs) <= tol(obs)), (i, np.abs(got["obs"][i] - obs).max())
        assert abs(got["reward"][i] - rew) <= tol(rew), (i, got["reward"][i], rew)
        assert bool(got["done"][i]) == bool(done), i
        assert np.all(np.abs(got["torque"][i] - env.torque) <= 1e-5 * np.maximum(np.abs(env.torque), 0.05))
        assert got["ep_len"][i] == env.episode_length
        if combined:
            assert got["idx"][i] == env.current_motion_n_steps and got["clip"][i] == env.clip_id
        else:
            assert got["idx"][i] == env.idx_curr


@pytest.mark.parametrize("variant", [0, 1])
def test_statistics_match_numpy(variant):
    hp, s = _make("humanoid3d", "walk", False, 128 * 20 + 11, seed=5)
    r = _run(hp, s, variant, stats=True)
    obs, rew, done = r["obs"].astype(np.float64), r["reward"].astype(np.float64), r["done"].astype(bool)
    want = np.concatenate([[len(rew), rew.sum(), (rew * rew).sum(), done.sum(), r["ep_return"][done].astype(np.float64).sum(),
                            r["ep_len"][done].astype(np.float64).sum()], obs.sum(0), (obs * obs).sum(0)])
    np.testing.assert_allclose(r["stats"], want, rtol=2e-6, atol=1e-6)


def test_full_size_properties():
    """BASELINE config sizes (65 536 envs, spinkick + dance_b by env id): structural properties."""
    import torch
    import bench
    from deepmimic_mujoco_b200.hotpath import HotPath
    hp = HotPath("humanoid3d")
    hp.load_motions(["spinkick", "dance_b"])
    n = 65536
    s = bench.synth_states(hp, n, 1, 21, torch.device("cuda", 0))
    s = {k: v[0].contiguous() for k, v in s.items()}
    r = _run(hp, s, 1)
    h = {k: v.cpu().numpy() for k, v in s.items()}
    lens = np.array(hp.clip_lens)[h["clip"]]
    np.testing.assert_array_equal(r["idx"], (h["idx"] + 1) % lens)              # frame index arithmetic
    np.testing.assert_array_equal(r["ep_len"], h["ep_len"] + 1)
    np.testing.assert_array_equal(r["obs"][:, :28], h["qpos"][:, 7:])            # joint block is a copy
    ph = np.clip(h["idx"] / lens, 0, 1)
    np.testing.assert_allclose(r["obs"][:, 66], ph, rtol=1e-6)
    guard = r["reason"] == 8
    assert np.all(r["reward"][~guard] <= 1.0 + 1e-6) and np.all(r["reward"][~guard] >= -0.1 - 1e-6)
    assert np.all(r["obs"][guard] == 0)
    # permutation equivariance: environments are independent
    perm = np.random.default_rng(0).permutation(n)
    s2 = {k: v[torch.from_numpy(perm).cuda()].contiguous() for k, v in s.items()}
    r2 = _run(hp, s2, 1)
    for k in ("obs", "reward", "done", "idx"):
        np.testing.assert_array_equal(r2[k], r[k][perm], err_msg=k)
