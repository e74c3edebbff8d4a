"""ENV_CFG.ADD_ABSPOS_OBS (deepmimic_env.py:133-137): the oracle's restatement against tests/golden/abspos_steps.npz, which
tools/gen_golden.py generated from the UNMODIFIED reference with the flag set (both env classes; a reset, a state whose
geoms leave the +-100 m box -- the bounds guard fires on the new block only -- and a state that trips it through a standard
entry).  CPU only."""
import numpy as np
import pytest

import replay

AB = replay.load("abspos_steps.npz")


def _close(a, b, what, rtol=1e-9, atol=1e-12):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    bad = np.abs(a - b) > atol + rtol * np.abs(b)
    bad &= ~(np.isnan(a) & np.isnan(b))
    assert not bad.any(), (what, int(bad.sum()), float(np.abs(a - b)[bad].max()))


@pytest.mark.parametrize("name", ["h_walk_abspos", "g1_walk_abspos"])
def test_single_clip_env(name):
    from oracle.envs import SingleClipEnv
    z = AB
    robot, motion, _ = [str(x) for x in z[name + "/meta"]]
    env = SingleClipEnv(motion, robot, add_abspos_obs=True)
    assert z[name + "/obs"].shape[1] == env.model.nq - 7 + env.model.nv - 6 + 8 + 2 + 3 * env.model.ngeom + 1
    for t in range(z[name + "/qpos"].shape[0]):
        r = int(z[name + "/reset_before"][t])
        if r >= 0:
            _close(env.reset(idx_init=r), z[name + "/reset_obs"][t], "reset obs %d" % t)
        env.episode_length = int(z[name + "/eplen_before"][t])
        obs, rew, done, info = env.step(z[name + "/action"][t].astype(np.float64),
                                        (z[name + "/qpos"][t].astype(np.float64), z[name + "/qvel"][t].astype(np.float64)),
                                        int(z[name + "/contact_bits"][t]))
        _close(obs, z[name + "/obs"][t], "obs %d" % t)
        _close(rew, z[name + "/reward"][t], "reward %d" % t)
        assert int(done) == int(z[name + "/done"][t]) and replay.reason_code(info) == int(z[name + "/reason"][t]), t
    assert 8 in z[name + "/reason"]  # the guard fired (once through the new block, once through a standard entry)


def test_multi_clip_env():
    from oracle.envs import MultiClipEnv
    z, name = AB, "c_abspos"
    env = MultiClipEnv(add_abspos_obs=True)
    assert z[name + "/obs"].shape[1] == 98 + 3 * env.model.ngeom
    for t in range(z[name + "/qpos"].shape[0]):
        code = int(z[name + "/reset_before"][t])
        if code >= 0:
            ro = env.reset(rsi=bool(code // 100000), coin=(code // 10000) % 10, frame=code % 10000)
            _close(ro, z[name + "/reset_obs"][t], "reset obs %d" % t)
        env.episode_length = int(z[name + "/eplen_before"][t])
        obs, rew, done, info = env.step(z[name + "/action"][t].astype(np.float64),
                                        (z[name + "/qpos"][t].astype(np.float64), z[name + "/qvel"][t].astype(np.float64)),
                                        int(z[name + "/contact_bits"][t]))
        _close(obs, z[name + "/obs"][t], "obs %d" % t)
        _close(rew, z[name + "/reward"][t], "reward %d" % t)
        assert int(done) == int(z[name + "/done"][t]) and replay.reason_code(info) == int(z[name + "/reason"][t]), t
