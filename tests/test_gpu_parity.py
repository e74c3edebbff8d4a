"""GPU parity tests proper: the CUDA path, called through the C ABI (HotPath -> libdmhot.so), against
(1) the golden vectors recorded from the reference's own Python and (2) the float64 oracle on seeded
random batches.  Tolerance (BASELINE.json north_star): frame indices, counters, done flags and reasons
bit-exact; rewards, observations, torques within 1e-5 relative in fp32 (absolute floor 5e-7 = 1e-5 x 0.05
for values that are differences of O(1) quantities)."""
import numpy as np
import pytest

import replay

pytestmark = pytest.mark.gpu

DP = replay.load("dpenv_steps.npz")
CB = replay.load("combined_steps.npz")
_backends = {}


def backend(z, name, **kw):
    robot, kind, motions = replay.scenario_motions(z, name)
    key = (robot, kind, tuple(motions), tuple(sorted(kw.items())))
    if key not in _backends:
        _backends[key] = replay.GpuBackend(robot, kind, motions, **kw)
    return _backends[key]


@pytest.mark.parametrize("mode", ["aos", "soa", "host"])
@pytest.mark.parametrize("name", replay.scenarios(DP))
def test_dpenv_golden(name, mode):
    be = backend(DP, name, host=(mode == "host"), soa=(mode == "soa"))
    replay.assert_matches_golden(DP, name, replay.replay_batched(DP, name, be))


@pytest.mark.parametrize("mode", ["aos", "soa", "host"])
@pytest.mark.parametrize("name", replay.scenarios(CB))
def test_combined_golden(name, mode):
    be = backend(CB, name, host=(mode == "host"), soa=(mode == "soa"))
    replay.assert_matches_golden(CB, name, replay.replay_batched(CB, name, be))


def test_torque_matches_oracle():
    for name in ("g_walk", "h_walk_cycle"):
        got = replay.replay_batched(DP, name, backend(DP, name))
        want = replay.replay_dpenv_oracle(DP, name)["torque"]
        np.testing.assert_allclose(got["torque"], want, rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("robot,motion", [("humanoid3d", "walk"), ("humanoid3d", "getup_facedown"), ("unitree_g1", "run")])
def test_product_loader_tables_match_oracle(robot, motion):
    """MocapDM (host numpy + float64 FK kernel) == oracle restatement of the reference loader."""
    from deepmimic_mujoco_b200.mocap import MocapDM
    from oracle.envs import tables_for
    m = MocapDM(robot).load_motion(motion)
    t = tables_for(robot, motion)
    assert m.get_length() == t.get_length() and m.dt == t.dt
    np.testing.assert_allclose(np.asarray(m.data_config), np.asarray(t.data_config), rtol=0, atol=1e-12)
    np.testing.assert_allclose(np.asarray(m.data_vel), np.asarray(t.data_vel), rtol=0, atol=1e-9)
    np.testing.assert_allclose(np.asarray(m.data_body_xpos), np.asarray(t.data_body_xpos), rtol=0, atol=1e-12)
    np.testing.assert_allclose(np.asarray(m.data_geom_xpos), np.asarray(t.data_geom_xpos), rtol=0, atol=1e-12)


class _Tiled:
    """A golden file with every scenario's rows repeated up to a whole number of 32-environment tiles, so that the team
    kernel (variant 2, tiles of 32) produces every row and none falls to the strided tail."""

    def __init__(self, z, name, reps):
        self.z, self.name, self.reps = z, name, reps

    def __getitem__(self, key):
        a = self.z[key]
        if key.endswith("/meta"):
            return a
        return np.concatenate([a] * self.reps, 0)[:self.rows()]

    def rows(self):
        T = self.z[self.name + "/qpos"].shape[0]
        return T * self.reps // 32 * 32


@pytest.mark.parametrize("name", replay.scenarios(DP) + replay.scenarios(CB))
def test_team_kernel_golden(name):
    """K1c (dm_set_kernel_variant 2: one environment over four warps, rotation-matrix FK) against the same goldens."""
    z = DP if name in replay.scenarios(DP) else CB
    T = z[name + "/qpos"].shape[0]
    tz = _Tiled(z, name, reps=max(2, -(-64 // T)))
    assert tz.rows() >= 32 and tz.rows() % 32 == 0
    be = backend(z, name, variant=2)
    k0 = be.hp.launch_count()
    got = replay.replay_batched(tz, name, be)
    assert be.hp.launch_count() == k0 + 1  # one launch: the team kernel took every row
    replay.assert_matches_golden(tz, name, got)
