#!/usr/bin/env python3
"""Generate a SYNTHETIC-RECORDED episode in the reference's episode_debug_log JSON format
(src/deepmimic_env.py:458-476) by rolling the float64 oracle: no recordings ship with the reference and MuJoCo
is not installable here, so the states are a smooth, seeded perturbation of the clip ("a policy tracking the
reference imperfectly"), the velocities are the clip's plus smooth noise, and body_xpos / body_xvelp come from
the oracle's FK.  Output: tests/golden/recorded_<robot>_<motion>.json.

    python tools/make_recording.py [robot] [motion] [steps]
"""
import os
import sys

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)

from deepmimic_mujoco_b200.debug_log import DebugLog, integrate_qpos, write_debug_log  # noqa: E402
from oracle import mjkin  # noqa: E402
from oracle.envs import SingleClipEnv  # noqa: E402


def smooth_noise(rng, T, dim, scale, taps=9):
    x = rng.normal(0, 1, (T + taps, dim))
    k = np.hanning(taps + 2)[1:-1]
    k /= np.sqrt((k ** 2).sum())
    return scale * np.stack([np.convolve(x[:, j], k, mode="valid")[:T] for j in range(dim)], 1)


def main(robot="humanoid3d", motion="walk", steps=76, seed=7):
    rng = np.random.default_rng(seed)
    env = SingleClipEnv(motion, robot)
    env.reset(0)
    L = env.mocap_data_len
    nq, nv, na = env.model.nq, env.model.nv, env.action_dim
    nq_noise = np.concatenate([smooth_noise(rng, steps, 3, 0.01), smooth_noise(rng, steps, 4, 0.01), smooth_noise(rng, steps, nq - 7, 0.04)], 1)
    nv_noise = smooth_noise(rng, steps, nv, 0.3)
    act = np.clip(smooth_noise(rng, steps, na, 0.6), -2, 2) if robot == "humanoid3d" else smooth_noise(rng, steps, na, 0.3)
    log = DebugLog(motion, robot, 0)
    for t in range(steps):
        q = np.asarray(env.mocap.data_config[t % L]) + nq_noise[t]
        v = np.asarray(env.mocap.data_vel[t % L]) + nv_noise[t]
        obs, reward, done, info = env.step(act[t], (q, v), 0)
        eps = 1e-6
        xp = mjkin.forward(env.model, integrate_qpos(q, v, eps), v).xpos
        xm = mjkin.forward(env.model, integrate_qpos(q, v, -eps), v).xpos
        log.append(act[t], env.data.xpos, (xp - xm) / (2 * eps), q, v, reward)
        if done:
            break
    out = os.path.join(REPO, "tests", "golden", "recorded_%s_%s.json" % (robot, motion))
    log.extra["note"] = "synthetic-recorded by tools/make_recording.py (oracle roll-out, seed %d); not a MuJoCo episode" % seed
    write_debug_log(out, log)
    print(out, len(log), "steps", os.path.getsize(out), "bytes; mean reward %.4f" % np.mean(log.steps["reward"]))


if __name__ == "__main__":
    a = sys.argv[1:]
    main(*(a[:2]), *([int(a[2])] if len(a) > 2 else []))
