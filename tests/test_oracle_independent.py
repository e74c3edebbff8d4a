"""Independent pins of the MuJoCo-side arithmetic the oracle restates (oracle/mjkin.py): the values the reference reads
from `sim.data` (src/deepmimic_env.py:29,51-54,232,236,421) and that no golden trajectory can pin because the reference's
Python was driven with the oracle's own kinematics (tools/ref_shims/gym/envs/mujoco/mujoco_env.py).  Three directions that
share no code with mjkin:

  * tests/golden/fk_independent.npz (tools/gen_fk_independent.py): FK recomputed from the reference's MJCF files with
    xml.etree and homogeneous matrices of the reference-held src/transformations.py;
  * central finite differences of FK for `cvel` (mj_comVel): angular part from R(t+h) R(t-h)^T, linear part = velocity
    of the body-fixed point that coincides with the robot COM;
  * the reference's own invariant `test_walk_hand_xpos_mocap` (src/deepmimic_env.py:540-554) on every loadable clip and
    every end-effector geom, against the independent FK.
CPU only."""
import math

import numpy as np
import pytest

import replay
from oracle import mjkin
from oracle import rotations as R
from oracle.assets import load_model
from oracle.envs import ROBOTS, tables_for

FK = replay.load("fk_independent.npz")
CLIPS = sorted({k.split("/")[1] for k in FK.files if k.startswith("clip/")})


def _robot_of(clip):
    return "unitree_g1" if clip.startswith("unitree_g1") else "humanoid3d"


@pytest.mark.parametrize("robot", ["humanoid3d", "unitree_g1"])
def test_fk_matches_homogeneous_matrix_chain(robot):
    m = load_model(robot)
    q = FK["rand/%s/qpos" % robot]
    for k in range(len(q)):
        d = mjkin.forward(m, q[k], np.zeros(m.nv))
        np.testing.assert_allclose(d.xpos, FK["rand/%s/xpos" % robot][k], rtol=0, atol=2e-13)
        np.testing.assert_allclose(d.xmat, FK["rand/%s/xmat" % robot][k], rtol=0, atol=2e-13)
        np.testing.assert_allclose(d.geom_xpos, FK["rand/%s/geom_xpos" % robot][k], rtol=0, atol=2e-13)


@pytest.mark.parametrize("clip", CLIPS)
def test_mocap_fk_invariant_every_clip_every_end_effector(clip):
    """FK of every table row == the independent FK; on key frames the STORED table positions agree as well (rows between
    key frames are lerp-of-FK, mocap_v2.py:309-336, so only key frames satisfy the reference's invariant)."""
    robot = _robot_of(clip)
    m = load_model(robot)
    t = tables_for(robot, clip[len(robot) + 1:])
    cfg = np.asarray(t.data_config)
    assert float(cfg.sum()) == float(FK["clip/%s/qpos_sum" % clip])
    stride = int(FK["clip/%s/stride" % clip])
    ee = [m.geom_name2id(g) for g in ROBOTS[robot]["ee"]]
    want_ee, want_x3 = FK["clip/%s/ee_xpos" % clip], FK["clip/%s/xpos3" % clip]
    assert len(want_ee) == len(cfg)
    for i in range(len(cfg)):
        d = mjkin.forward(m, cfg[i], t.data_vel[i])
        np.testing.assert_allclose(d.geom_xpos[ee], want_ee[i], rtol=0, atol=2e-13)
        if i % 3 == 0:
            np.testing.assert_allclose(d.xpos, want_x3[i // 3], rtol=0, atol=2e-13)
        if i % stride == 0:
            np.testing.assert_allclose(np.asarray(t.data_geom_xpos[i])[ee], want_ee[i], rtol=0, atol=2e-13)
            if i % 3 == 0:
                np.testing.assert_allclose(np.asarray(t.data_body_xpos[i]), want_x3[i // 3], rtol=0, atol=2e-13)


def _integrate(m, q, v, h):
    """mj_integratePos: free joint position += h v, quaternion <- q * exp(h w / 2) (w in the BODY frame); hinges += h v."""
    q = np.array(q, dtype=np.float64)
    q[:3] += h * v[:3]
    w = np.asarray(v[3:6]) * h
    ang = float(np.linalg.norm(w))
    dq = np.array([1.0, 0, 0, 0]) if ang == 0 else np.concatenate([[math.cos(ang / 2)], math.sin(ang / 2) * w / ang])
    q[3:7] = R.quat_mul(R.quat_normalize(q[3:7]), dq)
    q[7:] += h * np.asarray(v[6:])
    return q


@pytest.mark.parametrize("robot", ["humanoid3d", "unitree_g1"])
def test_cvel_matches_finite_differences_of_fk(robot):
    """cvel[b] = [world angular velocity; linear velocity of the point of body b that sits at the robot COM]
    (mj_comVel): check every body, the torso among them, against central differences of FK to 1e-6."""
    m = load_model(robot)
    rng = np.random.default_rng(7)
    h = 1e-6
    for q in FK["rand/%s/qpos" % robot][:12]:
        v = rng.normal(0, 1.5, m.nv)
        d0 = mjkin.forward(m, q, v)
        dp = mjkin.forward(m, _integrate(m, q, v, +h), v)
        dm = mjkin.forward(m, _integrate(m, q, v, -h), v)
        for b in range(1, m.nbody):
            r_body = d0.xmat[b].T @ (d0.subtree_com - d0.xpos[b])  # the COM as a body-fixed point
            pp = dp.xpos[b] + dp.xmat[b] @ r_body
            pm = dm.xpos[b] + dm.xmat[b] @ r_body
            lin = (pp - pm) / (2 * h)
            W = (dp.xmat[b] @ dm.xmat[b].T - dm.xmat[b] @ dp.xmat[b].T) / (4 * h)  # skew(w), first order in h
            ang = np.array([W[2, 1], W[0, 2], W[1, 0]])
            scale = max(1.0, float(np.abs(d0.cvel[b]).max()))
            np.testing.assert_allclose(d0.cvel[b][:3], ang, rtol=0, atol=1e-6 * scale)
            np.testing.assert_allclose(d0.cvel[b][3:], lin, rtol=0, atol=1e-6 * scale)


@pytest.mark.parametrize("robot", ["humanoid3d", "unitree_g1"])
def test_com_is_mass_weighted_inertial_frames(robot):
    """subtree_com of the root from the independent body frames: sum m (xpos + xmat ipos) / sum m (mj_comPos), the
    quantity behind z_com (deepmimic_env.py:421-422)."""
    m = load_model(robot)
    q = FK["rand/%s/qpos" % robot]
    for k in range(len(q)):
        xipos = FK["rand/%s/xpos" % robot][k] + np.einsum("bij,bj->bi", FK["rand/%s/xmat" % robot][k], m.body_ipos)
        com = (m.body_mass[:, None] * xipos).sum(0) / m.body_mass.sum()
        d = mjkin.forward(m, q[k], np.zeros(m.nv))
        np.testing.assert_allclose(d.subtree_com, com, rtol=0, atol=2e-13)


def test_to_rpy_matches_rotation_matrix_decomposition():
    """py3dtf to_rpy == the fixed-axis x-y-z angles of the same rotation read off its MATRIX (a second route: atan2 of
    matrix entries of the independent frames instead of the quaternion formula)."""
    for robot in ("humanoid3d", "unitree_g1"):
        m = load_model(robot)
        q = FK["rand/%s/qpos" % robot]
        for k in range(len(q)):
            d = mjkin.forward(m, q[k], np.zeros(m.nv))
            for b in (1, 2):
                M = FK["rand/%s/xmat" % robot][k][b]
                w, x, y, z = d.xquat[b]
                got = R.to_rpy(x, y, z, w)
                want = (math.atan2(M[2, 1], M[2, 2]), math.asin(max(-1.0, min(1.0, -M[2, 0]))), math.atan2(M[1, 0], M[0, 0]))
                np.testing.assert_allclose(got, want, rtol=0, atol=1e-9)
