"""The reference-facing Python surface on the GPU: DPVecEnv (SB3 VecEnv convention: auto-reset,
terminal_observation, infos), its pinned zero-copy inputs, and the single-env DPEnv / DPCombinedEnv mirrors,
all against the float64 oracle on the same seeded inputs.  Tolerance: 1e-5 relative (north_star), written as
|a - b| <= 1e-5 * max(|b|, 0.05); done flags, indices and counters bit-exact."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-5, 5e-7


def _states(env_or, rng, n, sig=1.0):
    """clip frame + Gaussian perturbation; a few envs are pushed below low_z so that they terminate."""
    m = env_or.mocap
    nq, nv = env_or.model.nq, env_or.model.nv
    f = rng.integers(0, m.get_length(), n)
    q = np.asarray(m.data_config)[f] + sig * np.concatenate([rng.normal(0, 0.02, (n, 7)), rng.normal(0, 0.05, (n, nq - 7))], 1)
    v = np.asarray(m.data_vel)[f] + sig * rng.normal(0, 0.5, (n, nv))
    q[::7, 2] -= 0.6  # fallen
    return q.astype(np.float32), v.astype(np.float32)


def test_vecenv_matches_oracle_with_autoreset():
    from deepmimic_mujoco_b200.vec_env import DPVecEnv, LazyInfos
    from oracle.envs import SingleClipEnv
    n, steps = 96, 4
    rng = np.random.default_rng(5)
    env = DPVecEnv(n, motion="walk", robot="humanoid3d", seed=11)
    frames = rng.integers(0, 76, n).astype(np.int32)
    obs = env.reset(frame=frames)
    ors = [SingleClipEnv("walk", "humanoid3d") for _ in range(n)]
    want = np.stack([o.reset(int(f)) for o, f in zip(ors, frames)])
    np.testing.assert_allclose(obs, want, rtol=RTOL, atol=ATOL)
    n_done = 0
    for t in range(steps):
        q, v = _states(ors[0], rng, n)
        contact = rng.integers(0, 4, n).astype(np.uint32)
        act = rng.uniform(-2, 2, (n, 28)).astype(np.float32)
        obs, rew, done, infos = env.step(act, force_state=(q, v, contact))
        assert isinstance(infos, LazyInfos) and len(infos) == n
        assert obs.dtype == np.float32 and done.dtype == np.bool_
        for i, o in enumerate(ors):
            wo, wr, wd, wi = o.step(act[i].astype(np.float64), (q[i].astype(np.float64), v[i].astype(np.float64)), int(contact[i]))
            assert bool(done[i]) == bool(wd), (t, i)
            np.testing.assert_allclose(rew[i], wr, rtol=RTOL, atol=ATOL)
            if wd:  # SB3: the returned row is the reset observation, the last one travels in infos
                n_done += 1
                np.testing.assert_allclose(infos[i]["terminal_observation"], wo, rtol=RTOL, atol=ATOL)
                assert infos[i]["done_reason"] == wi["done_reason"]
                assert infos[i]["TimeLimit.truncated"] is False
                assert infos[i]["episode"]["l"] == o.episode_length
                wo = o.reset(int(env.idx[i]))  # device RNG drew the frame: replay it in the oracle
                assert env.ep_len[i] == 0
            else:
                assert "terminal_observation" not in infos[i]
            assert env.idx[i] == o.idx_curr and env.ep_len[i] == o.episode_length
            np.testing.assert_allclose(obs[i], wo, rtol=RTOL, atol=ATOL)
    assert n_done >= steps * (n // 7)
    env.close()


def test_vecenv_pinned_inputs_are_zero_copy_and_equal():
    from deepmimic_mujoco_b200.vec_env import DPVecEnv
    from oracle.envs import SingleClipEnv
    n = 256
    rng = np.random.default_rng(9)
    o = SingleClipEnv("walk", "humanoid3d")
    q, v = _states(o, rng, n)
    q[::7, 2] += 0.6
    act = rng.uniform(-2, 2, (n, 28)).astype(np.float32)
    frames = rng.integers(0, 76, n).astype(np.int32)
    a = DPVecEnv(n, seed=1)
    b = DPVecEnv(n, seed=1, copy_outputs=False)
    a.reset(frame=frames)
    b.reset(frame=frames)
    ra = a.step(act, force_state=(q, v))
    qb, vb, cb, ab = b.input_buffers()
    qb[...] = q
    vb[...] = v
    cb[...] = 0
    ab[...] = act
    rb = b.step(ab, force_state=(qb, vb, cb))
    assert rb[0] is b.obs  # the pinned output array itself
    for x, y in zip(ra[:3], rb[:3]):
        np.testing.assert_array_equal(x, y)
    assert a.h2d_bytes == b.h2d_bytes and a.h2d_bytes >= n * 4 * (35 + 34 + 28)
    a.close()
    b.close()


@pytest.mark.parametrize("n,combined,chunk", [(128, False, 0), (4096 + 37, False, 0), (2048 + 5, False, 256), (1024 + 3, True, 512),
                                              (262144, True, 0)])  # BASELINE configs[3] at full size through the host API
def test_host_step_mapped_equals_staged(n, combined, chunk, monkeypatch):
    """dm_step_host on the pinned block directly (DM_HOST_MAPPED: the kernel's TMA copies cross PCIe, caller arrays staged
    in chunks of whole tiles over two streams) against the copy-engine form (DM_HOST_STAGED): the same kernel on other
    addresses, so every output is bit-identical -- pageable caller arrays and in-place pinned arrays alike, ragged tail
    (K1a on mapped memory) included."""
    from deepmimic_mujoco_b200.vec_env import DPVecEnv
    if chunk:
        monkeypatch.setenv("DM_HOST_CHUNK", str(chunk))
    rng = np.random.default_rng(n)
    import torch
    envs = [DPVecEnv(n, combined=combined, seed=3, copy_outputs=co) for co in (True, True, False, True)]
    hp0 = envs[0].hp
    m = hp0.mocaps[0]
    d = hp0.dims
    f = rng.integers(0, m.get_length(), n)
    q = (np.asarray(m.data_config)[f] + rng.normal(0, 0.03, (n, d.nq))).astype(np.float32)
    v = (np.asarray(m.data_vel)[f] + rng.normal(0, 0.4, (n, d.nv))).astype(np.float32)
    q[::11, 2] -= 0.6
    c = rng.integers(0, 256, n).astype(np.uint32)
    lo, hi = envs[0].action_space.low, envs[0].action_space.high
    act = rng.uniform(lo, hi, (n, d.na)).astype(np.float32)
    frames = rng.integers(0, 40, n).astype(np.int32)
    envs[0].hp.set_host_mode(False)
    envs[1].hp.set_host_mode(True)
    envs[2].hp.set_host_mode(True)
    envs[3].hp.set_host_mode(True)
    pin = lambda x: torch.from_numpy(x).pin_memory().numpy()  # caller arrays that are themselves pinned: used where they lie
    res = []
    for k, e in enumerate(envs):
        e.reset(frame=frames, coin=np.zeros(n, np.int32)) if combined else e.reset(frame=frames)
        if k == 2:  # everything in place: one launch on the pinned block
            qb, vb, cb, ab = e.input_buffers()
            qb[...] = q; vb[...] = v; cb[...] = c; ab[...] = act
            o, r, dn, _ = e.step(ab, force_state=(qb, vb, cb))
        elif k == 3:
            o, r, dn, _ = e.step(pin(act), force_state=(pin(q), pin(v), c))
        else:
            o, r, dn, _ = e.step(act, force_state=(q, v, c))
        res.append((np.array(o), np.array(r), np.array(dn), np.array(e.torque), np.array(e.idx), np.array(e.ep_len),
                    np.array(e.clip), np.array(e.reason)))
    assert combined or (res[0][2].any() and not res[0][2].all())
    for other in res[1:]:
        for x, y in zip(res[0], other):
            np.testing.assert_array_equal(x, y)
    assert envs[0].h2d_bytes == envs[1].h2d_bytes and envs[0].d2h_bytes == envs[1].d2h_bytes
    for e in envs:
        e.close()


def test_lazy_infos_behaves_like_a_list_of_dicts():
    from deepmimic_mujoco_b200.vec_env import LazyInfos
    x = LazyInfos(5)
    assert len(x) == 5 and x[3] == {} and x[-1] == {}
    x[2]["k"] = 1
    assert [d for d in x] == [{}, {}, {"k": 1}, {}, {}]
    assert x[1:3] == [{}, {"k": 1}]
    assert x == [{}, {}, {"k": 1}, {}, {}]
    for i, d in enumerate(x):
        d["i"] = i
    assert x[4]["i"] == 4


def test_dpenv_single_env_mirror():
    from deepmimic_mujoco_b200 import DPEnv
    from oracle.envs import SingleClipEnv
    rng = np.random.default_rng(2)
    env = DPEnv(motion="spinkick", robot="humanoid3d")
    o = SingleClipEnv("spinkick", "humanoid3d")
    assert env.mocap_data_len == o.mocap_data_len == 78
    np.testing.assert_allclose(env.reset_model(idx_init=5), o.reset(5), rtol=RTOL, atol=ATOL)
    for t in range(6):
        q, v = _states(o, rng, 1, sig=0.3)
        q[0, 2] += 0.6
        act = rng.uniform(-2, 2, 28)
        got = env.step(act, force_state=(q[0], v[0]))
        want = o.step(act.astype(np.float32).astype(np.float64), (q[0].astype(np.float64), v[0].astype(np.float64)))
        np.testing.assert_allclose(got[0], want[0], rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(got[1], want[1], rtol=RTOL, atol=ATOL)
        assert got[2] == want[2] and env.idx_curr == o.idx_curr and env.episode_length == o.episode_length
        for k in ("reward_config", "reward_qvel", "reward_end_eff", "reward_joint_limit"):
            if k in want[3]:
                np.testing.assert_allclose(got[3][k], want[3][k], rtol=RTOL, atol=ATOL)
    with pytest.raises(AssertionError):
        env.step(np.zeros(5), force_state=(q[0], v[0]))
    env.close()


def test_combined_env_single_env_mirror():
    from deepmimic_mujoco_b200 import DPCombinedEnv
    from oracle.envs import MultiClipEnv
    rng = np.random.default_rng(4)
    env, o = DPCombinedEnv(), MultiClipEnv()
    np.testing.assert_allclose(env.reset(coin=0, frame=7), o.reset(coin=0, frame=7), rtol=RTOL, atol=ATOL)
    assert env.current_motion == "walk" and env.current_motion_n_steps == o.current_motion_n_steps
    for t in range(5):
        q0, v0 = o.get_current_motion_state()
        q = (q0 + rng.normal(0, 0.01, q0.shape)).astype(np.float32)
        v = (v0 + rng.normal(0, 0.1, v0.shape)).astype(np.float32)
        act = rng.uniform(-0.5, 0.5, 23)
        contact = int(rng.integers(0, 1024)) << 2
        got = env.step(act, force_state=(q, v, contact))
        want = o.step(act.astype(np.float32).astype(np.float64), (q.astype(np.float64), v.astype(np.float64)), contact)
        np.testing.assert_allclose(got[0], want[0], rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(got[1], want[1], rtol=RTOL, atol=ATOL)
        assert got[2] == want[2] and env.current_motion_n_steps == o.current_motion_n_steps
    env.close()


def test_outputs_held_by_the_caller_never_change():
    """SAC keeps _last_obs, rollout buffers keep views; arrays handed out by step() must stay the caller's."""
    from deepmimic_mujoco_b200.vec_env import DPVecEnv
    from oracle.envs import SingleClipEnv
    n = 256
    rng = np.random.default_rng(3)
    env = DPVecEnv(n, motion="walk", robot="humanoid3d", seed=1)
    env.reset()
    o = SingleClipEnv("walk", "humanoid3d")
    held = []
    for t in range(9):
        q, v = _states(o, rng, n)
        out = env.step(rng.uniform(-1, 1, (n, 28)).astype(np.float32), force_state=(q, v))
        keep = (out[0][::2], out[1].reshape(-1, 1), out[2])  # views, not the arrays themselves
        held.append((keep, [x.copy() for x in keep]))
        del out
        if t % 3 == 2:
            env.reset(mask=np.arange(n) % 2 == 0)  # a masked reset must not touch arrays already handed out
        for views, copies in held:
            for a, b in zip(views, copies):
                np.testing.assert_array_equal(a, b)
    env.close()


def test_time_limit_is_a_plain_done_like_the_reference():
    """deepmimic_env.py:435-438: MAX_EP_LENGTH ends the episode inside step(); no TimeLimit wrapper, so SB3 must not see
    TimeLimit.truncated = True (it would bootstrap the terminal value)."""
    from deepmimic_mujoco_b200.vec_env import DPVecEnv
    from oracle.envs import SingleClipEnv
    o = SingleClipEnv("walk", "humanoid3d")
    for flag in (False, True):
        env = DPVecEnv(4, motion="walk", robot="humanoid3d", seed=1, timelimit_truncated=flag)
        env.reset(frame=np.zeros(4, np.int32))
        env.ep_len[:] = [1000, 5, 1000, 999]
        q = np.repeat(np.asarray(o.mocap.data_config[0], np.float32)[None], 4, 0)
        v = np.repeat(np.asarray(o.mocap.data_vel[0], np.float32)[None], 4, 0)
        obs, rew, done, infos = env.step(np.zeros((4, 28), np.float32), force_state=(q, v))
        assert done.tolist() == [True, False, True, False]
        for i in (0, 2):
            assert infos[i]["done_reason"] == "max_ep_len" and infos[i]["TimeLimit.truncated"] is flag
        env.close()


def test_failing_physics_ends_the_episode_like_a_diverged_simulation():
    """deepmimic_env.py:359-378: an exception out of do_simulation returns (obs * 0, 0, True, {})."""
    from deepmimic_mujoco_b200.vec_env import DPVecEnv
    n = 8
    state = {"mode": "ok"}

    def physics(qpos, qvel, action):
        if state["mode"] == "raise":
            raise FloatingPointError("mjWARN_BADQACC")
        bad = np.zeros(n, bool)
        q, v = qpos.copy(), qvel.copy()
        if state["mode"] == "mask":
            bad[[1, 6]] = True
            q[bad] = np.nan
        return q, v, None, bad

    env = DPVecEnv(n, motion="walk", robot="humanoid3d", seed=2, physics=physics)
    env.reset()
    obs, rew, done, infos = env.step(np.zeros((n, 28), np.float32))
    assert not done.any()
    state["mode"] = "mask"
    obs, rew, done, infos = env.step(np.zeros((n, 28), np.float32))
    assert done.tolist() == [i in (1, 6) for i in range(n)]
    for i in (1, 6):
        assert rew[i] == 0 and not infos[i]["terminal_observation"].any() and "done_reason" not in infos[i]
        assert env.ep_len[i] == 0  # auto-reset like any finished episode
    assert np.isfinite(obs).all()
    state["mode"] = "raise"
    obs, rew, done, infos = env.step(np.zeros((n, 28), np.float32))
    assert done.all() and not rew.any()
    env.close()


def test_clip_ids_and_reset_frames_are_validated_on_the_host():
    """an id / frame outside the table was a silent out-of-bounds read; the reference raises IndexError."""
    from deepmimic_mujoco_b200.vec_env import DPVecEnv
    env = DPVecEnv(4, motion=["spinkick", "dance_b"], robot="humanoid3d")
    env.set_clips([0, 1, 1, 0])
    with pytest.raises(IndexError):
        env.set_clips([0, 2, 0, 0])
    with pytest.raises(IndexError):
        env.set_clips([0, -1, 0, 0])
    env.reset(frame=[77, 152, 0, -1])  # last frames of both clips; -1 is Python's last element
    assert env.idx.tolist() == [77, 152, 0, 77]
    with pytest.raises(IndexError):
        env.reset(frame=[78, 0, 0, 0])
    env.close()


def test_dpenv_without_mocap_returns_the_observation_only():
    """deepmimic_env.py:276-291,394-395: DPEnv(load_mocap=False).step -> (obs, 0, False, {}) and no counter moves."""
    from deepmimic_mujoco_b200 import DPEnv
    from oracle.envs import SingleClipEnv
    o = SingleClipEnv("walk", "humanoid3d")
    env = DPEnv(motion="walk", load_mocap=False)
    assert env.mocap.data_config is None and env.mocap_data_len == 1
    q, v = np.asarray(o.mocap.data_config[9], np.float32), np.asarray(o.mocap.data_vel[9], np.float32)
    o.reset(0)
    want = o.step(np.zeros(28), (q.astype(np.float64), v.astype(np.float64)))[0]
    got = env.step(np.zeros(28), force_state=(q, v))
    assert got[1:] == (0, False, {})
    np.testing.assert_allclose(got[0][:-1], want[:-1], rtol=RTOL, atol=ATOL)
    assert got[0][-1] == 0.0 and env.episode_length == 1 and env.idx_curr == 0
    env.close()
