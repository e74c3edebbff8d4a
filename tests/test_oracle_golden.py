"""The oracle against the golden vectors recorded from the reference's own Python
(tools/gen_golden.py).  CPU only."""
import numpy as np
import pytest

import replay

DP = replay.load("dpenv_steps.npz")
CB = replay.load("combined_steps.npz")
TOL = dict(rtol=1e-11, atol=1e-12)


def _cmp(name, z, got, int_keys):
    np.testing.assert_allclose(got["obs"], z[name + "/obs"], err_msg=name + " obs", **TOL)
    np.testing.assert_allclose(got["reward"], z[name + "/reward"], err_msg=name + " reward", **TOL)
    np.testing.assert_allclose(got["terms"], z[name + "/terms"], err_msg=name + " terms", equal_nan=True, **TOL)
    np.testing.assert_allclose(got["reset_obs"], z[name + "/reset_obs"], err_msg=name + " reset obs", equal_nan=True, **TOL)
    np.testing.assert_array_equal(got["done"], z[name + "/done"], err_msg=name + " done")
    np.testing.assert_array_equal(got["reason"], z[name + "/reason"], err_msg=name + " reason")
    for k in int_keys:
        np.testing.assert_array_equal(got[k], z[name + "/" + k], err_msg=name + " " + k)


@pytest.mark.parametrize("name", replay.scenarios(DP))
def test_dpenv_oracle_matches_reference(name):
    _cmp(name, DP, replay.replay_dpenv_oracle(DP, name), ("idx_after", "eplen_after"))


@pytest.mark.parametrize("name", replay.scenarios(CB))
def test_combined_oracle_matches_reference(name):
    _cmp(name, CB, replay.replay_combined_oracle(CB, name), ("clip_after", "nsteps_after", "eplen_after"))


def test_scenarios_cover_every_branch():
    reasons = set()
    for n in replay.scenarios(DP):
        reasons |= set(DP[n + "/reason"][DP[n + "/done"] == 1].tolist())
    assert {1, 2, 3, 4, 5, 6, 8} <= reasons  # low_z high_z roll pitch max_ep acyclic guard
    clips, creasons = set(), set()
    for n in replay.scenarios(CB):
        clips |= set(CB[n + "/clip_after"].tolist())
        creasons |= set(CB[n + "/reason"][CB[n + "/done"] == 1].tolist())
    assert clips == {0, 1, 2, 3}
    assert {5, 7, 8} <= creasons
