"""CPU oracle for the DeepMimic_mujoco per-step environment hot path.

TEST INFRASTRUCTURE -- NOT PRODUCT CODE.  Only `tests/`, `__graft_entry__.smoke()`
and the `cpu_baseline` / `--impl reference` legs of `bench.py` may import this
package, and only as the checker or the timed CPU baseline.  The product
(`deepmimic_mujoco_b200`) never imports it and has no CPU fallback.

What it is: a float64 numpy restatement, one environment at a time, of what the
reference computes around MuJoCo on every control step
(reference src/deepmimic_env.py:33-256,335-510, src/combined_env.py:21-99,202-505,
src/config.py, src/mujoco/mocap_v2.py:33-362, src/mujoco/mocap_util.py:31-48),
plus restatements of the third-party arithmetic those files call and that is not
vendored in /root/reference:

  * MuJoCo 2.1 (mujoco-py 2.1.2.14): mj_kinematics / mj_comPos / mj_comVel,
    motor force = gear * clip(ctrl, ctrlrange)            -> oracle/mjkin.py
  * pyquaternion 0.9.9: Quaternion product / conjugate / axis / angle / from-matrix
                                                          -> oracle/rotations.py
  * py3dtf 0.2: Quaternion(x,y,z,w).to_rpy()              -> oracle/rotations.py
  * Gohlke transformations.py (vendored in the reference, src/transformations.py):
    euler_from_quaternion / quaternion_from_euler, axes 'rxyz' and 'sxyz'
                                                          -> oracle/rotations.py

Parity status.  The reference ships no golden vectors for this path (SURVEY.md
section 8c).  The oracle is pinned two ways: (1) against fixtures in
tests/golden/ produced by running the reference's OWN Python (deepmimic_env.py,
combined_env.py, mocap_v2.py, transformations.py) unmodified in the build
container with only the absent third-party modules shimmed (script:
tools/gen_golden.py); (2) against the reference's implied known answers
(mocap frame as state => reward_config = reward_qvel = 1).  At the third-party
edges -- MuJoCo FK/COM/cvel values, py3dtf.to_rpy, pyquaternion -- there is no
binary to compare with, so those restatements are "PARITY UNPINNED" at the value
level (they follow the published algorithms; see each module's docstring).
"""
