"""ENV_CFG.ADD_ABSPOS_OBS (deepmimic_env.py:133-137) on the GPU: the dm_obs_abspos post-pass behind the step / reset kernels
and its wiring into DPVecEnv / DPEnv / DPCombinedEnv, against tests/golden/abspos_steps.npz (generated from the UNMODIFIED
reference with the flag set, tools/gen_golden.py) and the float64 oracle.  Tolerance as everywhere: 1e-5 relative
(|a - b| <= 1e-5 * max(|b|, 0.05)); done flags, reasons and counters bit-exact."""
import numpy as np
import pytest

import replay

pytestmark = pytest.mark.gpu

AB = replay.load("abspos_steps.npz")
NAMES = ["h_walk_abspos", "g1_walk_abspos", "c_abspos"]


def _close(a, b, what, rtol=1e-5, floor=0.05):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    bad = np.abs(a - b) > rtol * np.maximum(np.abs(b), floor)
    bad &= ~(np.isnan(a) & np.isnan(b))
    assert not bad.any(), "%s: %d mismatches, worst |d| = %g at %s" % (what, bad.sum(), np.abs(a - b)[bad].max(), np.argwhere(bad)[0])


@pytest.mark.parametrize("name", NAMES)
def test_post_pass_rows_match_golden(name):
    """Every recorded step as one row of a batch: step kernel, then the post-pass in step mode."""
    import torch
    z = AB
    robot, kind, motions = replay.scenario_motions(z, name)
    be = replay.GpuBackend(robot, kind, motions)
    got = replay.replay_batched(z, name, be)
    n = got["obs"].shape[0]
    dev = {k: torch.from_numpy(np.ascontiguousarray(got[k])).cuda() for k in ("obs", "reward", "done", "reason")}
    qpos = torch.from_numpy(np.ascontiguousarray(z[name + "/qpos"], dtype=np.float32)).cuda()
    ext = torch.full((n, be.hp.obs_dim + be.hp.abspos_dim), 7.0, device="cuda")
    be.hp.obs_abspos(n, qpos, dev["obs"], ext, dev["reward"], dev["done"], dev["reason"])
    torch.cuda.synchronize()
    assert ext.shape[1] == z[name + "/obs"].shape[1]
    _close(ext.cpu().numpy(), z[name + "/obs"], name + " obs")
    _close(dev["reward"].cpu().numpy(), z[name + "/reward"], name + " reward")
    np.testing.assert_array_equal(dev["done"].cpu().numpy(), z[name + "/done"])
    np.testing.assert_array_equal(dev["reason"].cpu().numpy(), z[name + "/reason"])
    # the guard fired through the new block on a row the step kernel had let pass
    far = np.nonzero((z[name + "/reason"] == 8) & (got["reason"] != 8))[0]
    assert far.size >= 1 and not ext.cpu().numpy()[far].any()
    be.hp.close()


def test_mask_and_reset_mode():
    """Reset mode (no reward): rows are only widened; a mask leaves the other rows alone."""
    import torch
    from deepmimic_mujoco_b200.hotpath import HotPath
    from oracle import mjkin
    from oracle.envs import load_model
    hp = HotPath("unitree_g1", combined=True)
    hp.load_combined()
    rng = np.random.default_rng(5)
    n, d = 300, hp.dims
    q = rng.normal(0, 0.3, (n, d.nq)).astype(np.float32)
    q[:, 3:7] = rng.normal(0, 1, (n, 4))
    q[:, 2] += 0.8
    obs = rng.normal(0, 1, (n, d.obs_dim)).astype(np.float32)
    mask = (rng.random(n) < 0.5).astype(np.uint8)
    ext = torch.full((n, d.obs_dim + 3 * d.ngeom), -3.0, device="cuda")
    hp.obs_abspos(n, torch.from_numpy(q).cuda(), torch.from_numpy(obs).cuda(), ext, mask=torch.from_numpy(mask).cuda())
    torch.cuda.synchronize()
    e = ext.cpu().numpy()
    assert (e[mask == 0] == -3.0).all()
    model = load_model("unitree_g1")
    split = d.obs_dim - 8  # phase (1) + player action (7) follow the new block
    for i in np.nonzero(mask)[0][:40]:
        want = np.concatenate([obs[i, :split], np.asarray(mjkin.forward(model, q[i].astype(np.float64), np.zeros(d.nv)).geom_xpos).ravel(), obs[i, split:]])
        _close(e[i], want, "row %d" % i)
    with pytest.raises(Exception):
        hp.obs_abspos(n, torch.from_numpy(q).cuda(), ext, ext)  # not in place
    hp.close()


def _ref_steps(name):
    z = AB
    for t in range(z[name + "/qpos"].shape[0]):
        yield t, {k: z[name + "/" + k][t] for k in ("qpos", "qvel", "action", "contact_bits", "reset_before", "reset_obs", "eplen_before",
                                                     "obs", "reward", "done", "reason")}


@pytest.mark.parametrize("copy_outputs", [True, False])
def test_vecenv_sequence_matches_golden(copy_outputs):
    """DPVecEnv(1, add_abspos_obs=True) driven step by step like the reference run that made the golden: reset observations,
    step outputs, the guard through the new block, terminal_observation of the extended width."""
    from deepmimic_mujoco_b200.vec_env import DPVecEnv
    name = "h_walk_abspos"
    env = DPVecEnv(1, motion="walk", robot="humanoid3d", add_abspos_obs=True, copy_outputs=copy_outputs)
    width = AB[name + "/obs"].shape[1]
    assert env.observation_space.shape == (width,)
    held = []
    for t, r in _ref_steps(name):
        if int(r["reset_before"]) >= 0:
            ro = env.reset(frame=np.asarray([int(r["reset_before"])], np.int32))
            _close(ro[0], r["reset_obs"], "reset obs %d" % t)
        env.ep_len[0] = int(r["eplen_before"])
        obs, rew, done, infos = env.step(r["action"][None].astype(np.float32),
                                         force_state=(r["qpos"][None].astype(np.float32), r["qvel"][None].astype(np.float32),
                                                      np.asarray([int(r["contact_bits"])], np.uint32)))
        assert obs.shape == (1, width)
        assert bool(done[0]) == bool(r["done"]), t
        _close(rew[0], r["reward"], "reward %d" % t)
        if r["done"]:  # auto-reset: the step's own observation is the terminal one
            term = infos[0]["terminal_observation"]
            assert term.shape == (width,)
            _close(term, r["obs"], "terminal obs %d" % t)
            assert obs.shape == (1, width) and np.isfinite(obs).all()
        else:
            _close(obs[0], r["obs"], "obs %d" % t)
        if copy_outputs:
            held.append((obs, obs.copy()))
    for a, b in held:  # arrays handed out earlier are never rewritten
        np.testing.assert_array_equal(a, b)
    env.close()


def test_env_classes_take_the_flag():
    """DPEnv / DPCombinedEnv read ENV_CFG.ADD_ABSPOS_OBS like the reference (a class attribute set before construction);
    ADD_JOINT_FORCE_OBS is refused with the reason."""
    from deepmimic_mujoco_b200.combined_env import DPCombinedEnv
    from deepmimic_mujoco_b200.deepmimic_env import DPEnv
    z, name = AB, "g1_walk_abspos"
    DPEnv.ENV_CFG.ADD_ABSPOS_OBS = True
    DPCombinedEnv.ENV_CFG.ADD_ABSPOS_OBS = True
    try:
        env = DPEnv(motion="walk", robot="unitree_g1")
        assert env.observation_space.shape == (z[name + "/obs"].shape[1],)
        for t, r in _ref_steps(name):
            if int(r["reset_before"]) >= 0:
                _close(env.reset_model(idx_init=int(r["reset_before"])), r["reset_obs"], "DPEnv reset obs %d" % t)
            env._vec.ep_len[0] = int(r["eplen_before"])
            obs, rew, done, info = env.step(r["action"], force_state=(r["qpos"], r["qvel"], int(r["contact_bits"])))
            _close(obs, r["obs"], "DPEnv obs %d" % t)
            _close(rew, r["reward"], "DPEnv reward %d" % t)
            assert bool(done) == bool(r["done"]) and replay.reason_code(info) == int(r["reason"]), t
        cenv = DPCombinedEnv()
        assert cenv.observation_space.shape == (z["c_abspos/obs"].shape[1],)
        name = "c_abspos"
        for t, r in _ref_steps(name):
            code = int(r["reset_before"])
            if code >= 0:
                ro = cenv.reset(rsi=bool(code // 100000), coin=(code // 10000) % 10, frame=code % 10000)
                _close(ro, r["reset_obs"], "DPCombinedEnv reset obs %d" % t)
            cenv._vec.ep_len[0] = int(r["eplen_before"])
            obs, rew, done, info = cenv.step(r["action"], force_state=(r["qpos"], r["qvel"], int(r["contact_bits"])))
            _close(obs, r["obs"], "DPCombinedEnv obs %d" % t)
            _close(rew, r["reward"], "DPCombinedEnv reward %d" % t)
            assert bool(done) == bool(r["done"]), t
    finally:
        DPEnv.ENV_CFG.ADD_ABSPOS_OBS = False
        DPCombinedEnv.ENV_CFG.ADD_ABSPOS_OBS = False
    DPEnv.ENV_CFG.ADD_JOINT_FORCE_OBS = True
    try:
        with pytest.raises(NotImplementedError):
            DPEnv(motion="walk")
    finally:
        DPEnv.ENV_CFG.ADD_JOINT_FORCE_OBS = False


def test_tensor_api():
    """step_tensors / reset_tensors with the flag: extended rows on the device, guard and auto-reset included."""
    import torch
    from deepmimic_mujoco_b200.vec_env import DPVecEnv
    from oracle import mjkin
    from oracle.envs import load_model
    n = 256
    env = DPVecEnv(n, motion="walk", robot="humanoid3d", add_abspos_obs=True, seed=3)
    d = env.hp.dims
    obs = env.reset_tensors()
    assert obs.shape == (n, d.obs_dim + 3 * d.ngeom)
    b = env.device_buffers()
    model = load_model("humanoid3d")
    q = b["qpos"].cpu().numpy()
    o = obs.cpu().numpy()
    for i in range(0, n, 37):
        _close(o[i, d.obs_dim - 1:d.obs_dim - 1 + 3 * d.ngeom], np.asarray(mjkin.forward(model, q[i].astype(np.float64), np.zeros(d.nv)).geom_xpos).ravel(), "reset row %d" % i)
    qpos, qvel = b["qpos"].clone(), b["qvel"].clone()
    qpos[5, 0] = 500.0  # far away: the guard fires through the new block
    act = torch.zeros(n, d.na, device="cuda")
    obs, rew, done = env.step_tensors(act, qpos, qvel, contact=torch.zeros(n, dtype=torch.int32, device="cuda"))
    torch.cuda.synchronize()
    assert bool(done[5]) and float(rew[5]) == 0.0 and int(b["reason"][5]) == 8
    assert not b["terminal_obs_ext"][5].any()  # (obs * 0) was the terminal observation
    assert torch.isfinite(obs).all() and obs.shape == (n, d.obs_dim + 3 * d.ngeom)
    # env 5 was auto-reset: its row shows the reset state's geoms, not the far-away ones
    q5 = b["qpos"][5].cpu().numpy().astype(np.float64)
    _close(obs[5, d.obs_dim - 1:d.obs_dim - 1 + 3 * d.ngeom].cpu().numpy(), np.asarray(mjkin.forward(model, q5, np.zeros(d.nv)).geom_xpos).ravel(), "auto-reset row")
    env.close()
