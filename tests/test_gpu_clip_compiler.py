"""The device clip compiler (dm_compile_clips): raw "Frames" rows -> MocapDM tables on the GPU in float64,
against the oracle's restatement of the reference loader (oracle/mocap_tables.py, pinned to the reference's own
MocapDM by tools/gen_golden.py) for EVERY clip of the asset pack.  Tolerances: table lengths, dt and the discrete
outcome of the continuity fix exact; float64 values 1e-12 absolute (positions, angles) and 1e-9 (velocities, which
divide by dt) -- the only difference left is CUDA's libm against glibc's."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

UNLOADABLE = {("humanoid3d", "backflip"), ("humanoid3d", "spin")}  # dt / 0.01666 is 3.75 / 2.50 (mocap_v2.py:315-316)


def _motions(robot):
    from deepmimic_mujoco_b200 import assets
    return [m for m in assets.available_motions(robot) if (robot, m) not in UNLOADABLE]


def _cases():
    return [(r, m) for r in ("humanoid3d", "unitree_g1") for m in _motions(r)]


@pytest.mark.parametrize("robot,motion", _cases())
def test_compiled_tables_match_oracle(robot, motion):
    from deepmimic_mujoco_b200.mocap import MocapDM
    from oracle.envs import tables_for
    m = MocapDM(robot).load_motion(motion)
    t = tables_for(robot, motion)
    assert m.get_length() == t.get_length() and m.dt == t.dt
    got, want = np.asarray(m.data_config), np.asarray(t.data_config)
    np.testing.assert_allclose(got, want, rtol=0, atol=1e-12)
    np.testing.assert_allclose(np.asarray(m.data_vel), np.asarray(t.data_vel), rtol=0, atol=1e-9)
    np.testing.assert_allclose(np.asarray(m.data_body_xpos), np.asarray(t.data_body_xpos), rtol=0, atol=1e-12)
    np.testing.assert_allclose(np.asarray(m.data_geom_xpos), np.asarray(t.data_geom_xpos), rtol=0, atol=1e-12)


def test_continuity_fix_is_exercised_and_discrete_choice_matches():
    """spinkick has 97 (joint, frame) pairs the reference repairs; a wrong arg-min would land on another of the
    512 grid candidates, i.e. off by a whole grid step, not by round-off."""
    from deepmimic_mujoco_b200 import assets
    from deepmimic_mujoco_b200.mocap import MocapDM
    from oracle import mocap_tables as MT
    from oracle.envs import tables_for
    n_fixed = [0]
    orig = MT._continuity_fix

    def counting(joint, euler, prev, q, w):
        r = orig(joint, euler, prev, q, w)
        n_fixed[0] += tuple(r) != tuple(euler)
        return r

    MT._continuity_fix = counting
    try:
        t = MT.MocapTables(tables_for("humanoid3d", "walk").model, "humanoid3d")
        c = assets.load_clip("humanoid3d", "spinkick")
        t.load(c["frames"], c["format"], c["name"], c["loop"])
    finally:
        MT._continuity_fix = orig
    assert n_fixed[0] >= 50
    m = MocapDM("humanoid3d").load_motion("spinkick")
    assert np.abs(np.asarray(m.data_config) - np.asarray(t.data_config)).max() < 1e-12


@pytest.mark.parametrize("robot,motion", sorted(UNLOADABLE))
def test_bad_frame_duration_raises_like_the_reference(robot, motion):
    from deepmimic_mujoco_b200.mocap import MocapDM
    with pytest.raises(Exception, match="Invalid dt ratio"):
        MocapDM(robot).load_motion(motion)


@pytest.mark.parametrize("robot,combined", [("humanoid3d", False), ("unitree_g1", True)])
def test_device_packed_rows_equal_host_packed_rows(robot, combined):
    """The fp32 rows packed on the device (pack_rows_kernel) against dm_upload_clips' host packing of the same float64
    tables: identical up to the last bit of atan2 / asin (roll, pitch columns)."""
    from deepmimic_mujoco_b200.hotpath import HotPath
    a = HotPath(robot, combined=combined)
    a.load_combined() if combined else a.load_motions(["walk", "spinkick"])
    b = HotPath(robot, combined=combined)
    if combined:
        b.upload_mocaps(a.mocaps, [0, 0, 0], a.clip_names, transition=(2, 180, 1))
    else:
        b.upload_mocaps(a.mocaps, [0, 0], a.clip_names)
    ra, rb = a.packed_rows(), b.packed_rows()
    assert a.clip_lens == b.clip_lens
    if combined:  # the compiler stores the MotionTransition's pose source once; dm_upload_clips is handed (and keeps) a copy
        assert len(ra) == 76 + 48 + 353 and len(rb) == len(ra) + 353
        np.testing.assert_array_equal(rb[len(ra):], rb[76 + 48:76 + 48 + 353])
        rb = rb[:len(ra)]
    assert ra.shape == rb.shape
    np.testing.assert_allclose(ra, rb, rtol=2e-7, atol=1e-9)
    assert (ra != rb).mean() < 1e-3
    a.close()
    b.close()


def test_compile_clips_argument_errors():
    from deepmimic_mujoco_b200 import _native as N
    from deepmimic_mujoco_b200 import assets
    from deepmimic_mujoco_b200.hotpath import HotPath
    hp = HotPath("unitree_g1")
    h_walk = assets.load_clip("humanoid3d", "walk")  # 44-column DeepMimic rows: not defined for the G1
    with pytest.raises(N.DmError, match="humanoid3d only"):
        hp.compile_clips([h_walk], [0])
    g = assets.load_clip("unitree_g1", "walk")
    bad = dict(g, frames=g["frames"][:, :-1])
    with pytest.raises(N.DmError, match="nq \\+ 1 columns"):
        hp.compile_clips([bad], [0])
    with pytest.raises(N.DmError, match="same_as"):
        hp.compile_clips([], [], transition=(0, 10, 1))
    hp.close()
