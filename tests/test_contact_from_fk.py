"""floor-contact observation flags derived from forward kinematics (DM_CONTACT_FROM_FK) instead of
MuJoCo's contact list (get_foot_contact_obs / get_extra_contact_obs, src/deepmimic_env.py:78-121).  Oracle:
oracle/mjkin.floor_contact_bits, a float64 restatement of MuJoCo 2.1's plane-box / plane-sphere tests (PARITY
UNPINNED -- MuJoCo is not installable here); known answers from the MJCF geometry; kernel arithmetic on the CPU
harness and (gpu) through the product path.  Flags are exact away from the threshold: states whose signed distance is
within 1e-5 of the margin are skipped (fp32 FK cannot decide them)."""
import numpy as np
import pytest

import replay
from oracle import mjkin
from oracle.assets import load_model
from oracle.envs import ROBOTS, MultiClipEnv, SingleClipEnv


def _states(robot, motion, n, seed):
    rng = np.random.default_rng(seed)
    env = SingleClipEnv(motion, robot)
    m = env.mocap
    f = rng.integers(0, m.get_length(), n)
    q = np.asarray(m.data_config)[f] + np.concatenate([rng.normal(0, 0.02, (n, 7)), rng.normal(0, 0.05, (n, env.model.nq - 7))], 1)
    q[:, 2] += rng.normal(0, 0.03, n)  # feet in and out of the floor
    v = np.asarray(m.data_vel)[f] + rng.normal(0, 0.3, (n, env.model.nv))
    return f.astype(np.int32), q.astype(np.float32), v.astype(np.float32)


def _want_bits(robot, q, v):
    model, rc = load_model(robot), ROBOTS[robot]
    bits, ok = [], []
    for i in range(len(q)):
        d = mjkin.forward(model, q[i].astype(np.float64), v[i].astype(np.float64))
        bits.append(mjkin.floor_contact_bits(model, d, rc))
        dist = mjkin.floor_contact_dists(model, d, rc)
        marg = 0.001 if robot == "humanoid3d" else 0.0
        ok.append(np.all(np.abs(dist - marg) > 1e-5))
    return np.asarray(bits, dtype=np.uint32), np.asarray(ok)


def test_known_answers_standing_humanoid():
    """Upright default pose: the ankle box bottom sits 0.881416 below the root (MJCF body offsets 0.421546 + 0.40987,
    geom centre -0.0225, half height 0.0275); margin 0.001."""
    model, rc = load_model("humanoid3d"), ROBOTS["humanoid3d"]
    q = np.zeros(35)
    q[3] = 1.0
    for z, want in [(0.8824, 3), (0.8825, 0), (0.5, 3), (2.0, 0)]:
        q[2] = z
        assert mjkin.floor_contact_bits(model, mjkin.forward(model, q, np.zeros(34)), rc) == want, z
    q[2], q[28] = 0.8824, 0.5  # left hip rotated about x: the left foot lifts sideways, the right one stays down
    assert mjkin.floor_contact_bits(model, mjkin.forward(model, q, np.zeros(34)), rc) == 1


def test_known_answers_g1_visual_feet_never_touch():
    """right_foot / left_foot are class="visual" (contype = conaffinity = 0): bits 0 and 1 stay 0; the eight r = 0.001
    heel / toe spheres carry the contacts (margin 0)."""
    model, rc = load_model("unitree_g1"), ROBOTS["unitree_g1"]
    env = MultiClipEnv()
    q, v = np.asarray(env.walk_mocap.data_config[0], dtype=np.float64).copy(), np.zeros(43)
    q[2] -= 1.0
    w = mjkin.floor_contact_bits(model, mjkin.forward(model, q, v), rc)
    assert w == 0b1111111100
    q[2] += 3.0
    assert mjkin.floor_contact_bits(model, mjkin.forward(model, q, v), rc) == 0


@pytest.mark.parametrize("robot,motion,kind", [("humanoid3d", "walk", 0), ("unitree_g1", "walk", 0), ("unitree_g1", None, 1)])
def test_kernel_arithmetic_on_host_harness(robot, motion, kind):
    n = 200
    idx, q, v = _states(robot, motion or "walk", n, 4)
    be = replay.HarnessBackend(robot, kind, ["walk", "run", "getup_facedown_towalk"] if kind else [motion])
    be.L.hh_set_contact_mode(1)
    try:
        na = 28 if robot == "humanoid3d" else 23
        out = be.step(q, v, np.zeros((n, na), np.float32), np.full(n, 0xFFFFFFFF, np.uint32), idx, np.zeros(n, np.int32), np.zeros(n, np.int32))
    finally:
        be.L.hh_set_contact_mode(0)
    _check(robot, kind, q, v, out["obs"])


def _check(robot, kind, q, v, obs):
    want, ok = _want_bits(robot, q, v)
    nj = 28 if robot == "humanoid3d" else 37
    c0 = 2 * nj + 8
    guard = ~obs.any(axis=1)
    assert ok.sum() > 0.9 * len(ok)
    if kind == 0:
        got = obs[:, c0].astype(np.uint32) | (obs[:, c0 + 1].astype(np.uint32) << 1)
        want = want & 3
    else:
        got = sum(obs[:, c0 + k].astype(np.uint32) << (2 + k) for k in range(8))
        want = want & 0x3FC
    sel = ok & ~guard
    np.testing.assert_array_equal(got[sel], want[sel])
    assert 0.02 < (got[sel] != 0).mean() < 0.98 or robot == "unitree_g1" and kind == 0  # both outcomes are exercised


@pytest.mark.gpu
@pytest.mark.parametrize("robot,motion,combined", [("humanoid3d", "walk", False), ("unitree_g1", "walk", False), ("unitree_g1", None, True)])
def test_gpu_contact_from_fk(robot, motion, combined):
    from deepmimic_mujoco_b200.vec_env import DPVecEnv
    n = 384 + 17  # tile kernel + strided tail
    idx, q, v = _states(robot, motion or "walk", n, 6)
    env = DPVecEnv(n, motion=motion or "walk", robot=robot, combined=combined, contact_from_fk=True)
    env.reset(frame=idx, coin=np.zeros(n, np.int32) if combined else None)
    na = env.hp.na
    obs, rew, done, infos = env.step(np.zeros((n, na), np.float32), force_state=(q, v, np.full(n, 0xFFFFFFFF, np.uint32)))
    for i in np.nonzero(done)[0]:
        obs[i] = infos[i]["terminal_observation"]
    _check(robot, 1 if combined else 0, q, v, obs)
    # caller-supplied words are honoured again once the mode is switched back
    env.hp.set_contact_mode(False)
    env.reset(frame=idx, coin=np.zeros(n, np.int32) if combined else None)
    obs2, _, done2, infos2 = env.step(np.zeros((n, na), np.float32), force_state=(q, v, np.zeros(n, np.uint32)))
    for i in np.nonzero(done2)[0]:
        obs2[i] = infos2[i]["terminal_observation"]
    nj = 28 if robot == "humanoid3d" else 37
    assert not obs2[:, 2 * nj + 8:2 * nj + 8 + (8 if combined else 2)].any()
    env.close()
