"""The kernel's per-environment arithmetic (csrc/dm_step.cuh + generated FK), compiled for the CPU by
tests/host_harness, against the golden vectors recorded from the reference.  This is a CPU-side unit
test of the SAME source the CUDA kernels compile -- the GPU parity tests proper are test_gpu_*.py."""
import os

import numpy as np
import pytest

import replay

DP = replay.load("dpenv_steps.npz")
CB = replay.load("combined_steps.npz")
_backends = {}


def backend(z, name, team=False):
    robot, kind, motions = replay.scenario_motions(z, name)
    key = (robot, kind, tuple(motions), team)
    if key not in _backends:
        _backends[key] = replay.HarnessBackend(robot, kind, motions, team=team)
    return _backends[key]


# team = False: the thread-per-environment step (dm_step.cuh + generated/robot_*.cuh: K1a, reset);
# team = True: the team kernel's arithmetic (dm_team_core.cuh + generated/parts_*.cuh: K1c), its four parts run in turn
@pytest.mark.parametrize("team", [False, True])
@pytest.mark.parametrize("name", replay.scenarios(DP))
def test_dpenv_kernel_math(name, team):
    replay.assert_matches_golden(DP, name, replay.replay_batched(DP, name, backend(DP, name, team)))


@pytest.mark.parametrize("team", [False, True])
@pytest.mark.parametrize("name", replay.scenarios(CB))
def test_combined_kernel_math(name, team):
    replay.assert_matches_golden(CB, name, replay.replay_batched(CB, name, backend(CB, name, team)))


def test_torque_matches_oracle():
    name = "g_walk"
    got = replay.replay_batched(DP, name, backend(DP, name))
    want = replay.replay_dpenv_oracle(DP, name)["torque"]
    np.testing.assert_allclose(got["torque"], want, rtol=1e-6, atol=1e-6)


def test_fk_tables_double_matches_oracle():
    import ctypes as C
    from deepmimic_mujoco_b200._native import np_ptr
    from oracle.envs import tables_for
    be = backend(CB, replay.scenarios(CB)[0])
    t = tables_for("unitree_g1", "run")
    q = np.ascontiguousarray(t.data_config, dtype=np.float64)
    v = np.ascontiguousarray(t.data_vel, dtype=np.float64)
    bx = np.zeros((len(q), 39, 3))
    gx = np.zeros((len(q), 94, 3))
    be.L.hh_fk_tables(1, len(q), np_ptr(q), np_ptr(v), np_ptr(bx), np_ptr(gx))
    np.testing.assert_allclose(bx, np.asarray(t.data_body_xpos), rtol=0, atol=1e-13)
    np.testing.assert_allclose(gx, np.asarray(t.data_geom_xpos), rtol=0, atol=1e-13)


@pytest.mark.parametrize("helpers", [0, 1, 3])
def test_copy_pool_matches_memcpy(helpers):
    """dm_copy_pool.h (the helper threads behind dm_step_host's staging copies): 400 copies of random size and alignment,
    each compared with memcpy including the bytes behind the payload."""
    import ctypes as C
    L = C.CDLL(os.path.join(os.path.dirname(os.path.abspath(__file__)), "host_harness", "libhostharness.so"))
    assert L.hh_copy_pool_selftest(helpers, 400, 7) == 0
