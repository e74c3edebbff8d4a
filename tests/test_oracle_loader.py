"""Oracle clip loader and rotation helpers against fixtures recorded from the reference's
own mocap_v2.py / transformations.py (tools/gen_golden.py).  CPU only."""
import numpy as np
import pytest

import replay
from oracle import rotations as R
from oracle.assets import load_clip, load_model
from oracle.envs import tables_for
from oracle.mocap_tables import MocapTables

LD = replay.load("loader_tables.npz")
MG = replay.load("math_gohlke.npz")
CLIPS = sorted({k.split("/")[1] for k in LD.files if k.endswith("/data_config")})


@pytest.mark.parametrize("clip", CLIPS)
def test_tables_match_reference_loader(clip):
    robot = "unitree_g1" if clip.startswith("unitree_g1") else "humanoid3d"
    motion = clip[len(robot) + 1:]
    t = tables_for(robot, motion)
    n, dt, stride = LD["loader/%s/len_dt_stride" % clip]
    assert t.get_length() == int(n)
    assert t.dt == dt
    cfg = np.asarray(t.data_config)[::int(stride)]
    vel = np.asarray(t.data_vel)[::int(stride)]
    np.testing.assert_allclose(cfg, LD["loader/%s/data_config" % clip], rtol=1e-12, atol=1e-13)
    np.testing.assert_allclose(vel, LD["loader/%s/data_vel" % clip], rtol=1e-10, atol=1e-11)


@pytest.mark.parametrize("motion", ["backflip", "spin"])
def test_unloadable_clips_raise(motion):
    # reference mocap_v2.py:315-316: dt ratio 3.75 / 2.50 is not within 10 % of an integer
    assert int(LD["loader/humanoid3d_%s/raises" % motion]) == 1
    frames, fmt, name, loop = load_clip("humanoid3d", motion)
    with pytest.raises(Exception, match="Invalid dt ratio"):
        MocapTables(load_model("humanoid3d"), "humanoid3d").load(frames, fmt, name, loop)


def test_clip_lengths_match_survey():
    want = {("humanoid3d", "walk"): 76, ("humanoid3d", "run"): 48, ("humanoid3d", "spinkick"): 78,
            ("humanoid3d", "dance_b"): 153, ("humanoid3d", "kick"): 92, ("unitree_g1", "walk"): 76,
            ("unitree_g1", "run"): 48, ("unitree_g1", "getup_facedown_towalk"): 353}
    for (robot, motion), n in want.items():
        assert tables_for(robot, motion).get_length() == n


def test_gohlke_restatements_bit_exact():
    q = MG["math/quat_xyzw"]
    np.testing.assert_array_equal(np.asarray([R.gohlke_euler_rxyz(x) for x in q]), MG["math/euler_rxyz"])
    np.testing.assert_array_equal(np.asarray([R.gohlke_euler_sxyz(x) for x in q]), MG["math/euler_sxyz"])
    e = MG["math/euler_in"]
    np.testing.assert_array_equal(np.asarray([R.gohlke_quat_from_euler_rxyz(*x) for x in e]),
                                  MG["math/quat_from_euler_rxyz"])


def test_to_rpy_equals_static_xyz_of_vendored_transformations():
    q = MG["math/quat_xyzw"]
    got = np.asarray([R.to_rpy(*x) for x in q])
    np.testing.assert_allclose(got, MG["math/euler_sxyz"], rtol=0, atol=1e-12)


def test_fk_self_consistency_invariant():
    """reference deepmimic_env.py:540-554 (test_walk_hand_xpos_mocap): FK of table qpos reproduces
    the stored left-wrist height on every key frame (even rows of the x2-interpolated walk)."""
    from oracle import mjkin
    m = load_model("humanoid3d")
    t = tables_for("humanoid3d", "walk")
    g = m.geom_name2id("left_wrist")
    for i in range(0, t.get_length(), 2):
        d = mjkin.forward(m, t.data_config[i], t.data_vel[i])
        assert np.allclose(d.geom_xpos[g][2], t.data_geom_xpos[i][g][2])


def test_known_answer_mocap_frame_as_state():
    """reference deepmimic_env.py:570-653 (check_rewards_and_joint_limits): state := clip frame =>
    reward_config = reward_qvel = 1, end-effector / COM terms 1 to round-off."""
    from oracle.envs import SingleClipEnv
    env = SingleClipEnv("spinkick", "humanoid3d")
    for i in (0, 17, 77):
        env.reset(idx_init=i)
        _, r, _, info = env.step(np.zeros(28), (env.mocap.data_config[i], env.mocap.data_vel[i]))
        assert info["reward_config"] == 1.0 and info["reward_qvel"] == 1.0
        assert abs(info["reward_end_eff"] - 1) < 1e-12 and abs(info["reward_com"] - 1) < 1e-12
        assert abs(r - (1.0 - 0.1 * info["reward_joint_limit"])) < 1e-12
