"""Recorded states in the reference's episode_debug_log JSON (src/deepmimic_env.py:458-476): reader / writer round
trip, the oracle reproduces the recorded rewards, and (GPU) the batched path replays a whole recording in ONE
launch -- env t = step t of the log -- within the 1e-5 tolerance; body_xpos / body_xvelp of the recorder's
device FK match the recorded ones."""
import json
import os

import numpy as np
import pytest

from deepmimic_mujoco_b200.debug_log import DebugLog, integrate_qpos, read_debug_log, write_debug_log

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
LOGS = [("humanoid3d", "walk"), ("unitree_g1", "run")]


def _path(robot, motion):
    return os.path.join(GOLD, "recorded_%s_%s.json" % (robot, motion))


@pytest.mark.parametrize("robot,motion", LOGS)
def test_round_trip_and_schema(robot, motion, tmp_path):
    log = read_debug_log(_path(robot, motion))
    assert (log.robot, log.motion) == (robot, motion) and len(log) > 40
    with open(_path(robot, motion)) as f:
        raw = json.load(f)
    for k in ("action", "body_xpos", "body_xvelp", "qpos", "qvel", "reward", "motion", "robot"):  # check_debug_log.py:12-28
        assert k in raw
    a = log.arrays()
    T = len(log)
    assert a["qpos"].shape[0] == a["qvel"].shape[0] == a["action"].shape[0] == a["reward"].shape[0] == T
    assert a["body_xpos"].shape == a["body_xvelp"].shape and a["body_xpos"].shape[2] == 3
    p = tmp_path / "x.json"
    write_debug_log(str(p), log, indent=4)  # the reference writes indent=4
    b = read_debug_log(str(p)).arrays()
    for k in a:
        np.testing.assert_array_equal(a[k], b[k])


def test_reader_rejects_other_json(tmp_path):
    p = tmp_path / "bad.json"
    p.write_text(json.dumps({"Frames": []}))
    with pytest.raises(ValueError):
        read_debug_log(str(p))
    p.write_text(json.dumps({"qpos": [[0.0]], "qvel": [[0.0]], "reward": [], "action": [[1], [2]], "motion": "walk", "robot": "humanoid3d"}))
    with pytest.raises(ValueError):
        read_debug_log(str(p))


def test_integrate_qpos_matches_first_order():
    rng = np.random.default_rng(0)
    q = rng.normal(0, 1, (5, 35))
    q[:, 3:7] /= np.linalg.norm(q[:, 3:7], axis=1, keepdims=True)
    v = rng.normal(0, 1, (5, 34))
    q1 = integrate_qpos(q, v, 1e-7)
    np.testing.assert_allclose(np.linalg.norm(q1[:, 3:7], axis=1), 1.0, atol=1e-12)
    np.testing.assert_allclose((q1[:, :3] - q[:, :3]) / 1e-7, v[:, :3], rtol=1e-6, atol=1e-8)
    np.testing.assert_allclose((q1[:, 7:] - q[:, 7:]) / 1e-7, v[:, 6:], rtol=1e-6, atol=1e-8)
    np.testing.assert_array_equal(integrate_qpos(q, v, 0.0), q)


@pytest.mark.parametrize("robot,motion", LOGS)
def test_oracle_reproduces_recorded_rewards(robot, motion):
    from oracle.envs import SingleClipEnv
    log = read_debug_log(_path(robot, motion))
    a = log.arrays()
    env = SingleClipEnv(motion, robot)
    env.reset(log.idx_init)
    for t in range(len(log)):
        _, r, _, _ = env.step(a["action"][t], (a["qpos"][t], a["qvel"][t]), 0)
        assert abs(r - a["reward"][t]) <= 1e-12
        np.testing.assert_allclose(env.data.xpos, a["body_xpos"][t], atol=1e-12)


@pytest.mark.gpu
@pytest.mark.parametrize("robot,motion", LOGS)
def test_gpu_replays_recording_in_one_launch(robot, motion):
    from deepmimic_mujoco_b200.debug_log import body_kinematics
    from deepmimic_mujoco_b200.vec_env import DPVecEnv
    log = read_debug_log(_path(robot, motion))
    a = log.arrays()
    T = len(log)
    env = DPVecEnv(T, motion=motion, robot=robot)
    L = env.hp.clip_lens[0]
    env.reset(frame=log.frame_indices(L).astype(np.int32))
    env.ep_len[:] = np.arange(T)
    obs, rew, done, infos = env.step(a["action"].astype(np.float32), force_state=(a["qpos"].astype(np.float32), a["qvel"].astype(np.float32)))
    np.testing.assert_allclose(rew, a["reward"], rtol=1e-5, atol=5e-7)
    assert not done.any()
    bx, bv = body_kinematics(env.hp, a["qpos"], a["qvel"])
    np.testing.assert_allclose(bx, a["body_xpos"], atol=1e-12)
    np.testing.assert_allclose(bv, a["body_xvelp"], atol=1e-8)
    env.close()
