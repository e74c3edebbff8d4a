#!/usr/bin/env python3
"""DPVecEnv.step when a few environments finish every step (the normal case in training): the auto-reset goes through
dm_reset_host.  Staged = every array of the batch to the device and back; mapped = the reset kernel on pinned memory."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from deepmimic_mujoco_b200.vec_env import DPVecEnv

for n in [int(x) for x in (sys.argv[1:] or ["4096", "65536"])]:
    for mode in (0, 1):
        env = DPVecEnv(n, seed=0)
        env.hp.set_host_mode(mode)
        env.reset()
        rng = np.random.default_rng(0)
        m = env.hp.mocaps[0]
        f = rng.integers(0, 76, n)
        q = (np.asarray(m.data_config)[f] + rng.normal(0, 0.02, (n, 35))).astype(np.float32)
        v = (np.asarray(m.data_vel)[f] + rng.normal(0, 0.3, (n, 34))).astype(np.float32)
        a = rng.uniform(-1, 1, (n, 28)).astype(np.float32)
        c = np.zeros(n, np.uint32)
        q[:: max(n // 16, 1), 2] -= 0.6  # ~16 environments fall every step
        nd = 0
        for _ in range(20):
            env.step(a, force_state=(q, v, c))
        k = 200 if n <= 16384 else 40
        t = time.perf_counter()
        for _ in range(k):
            nd += int(env.step(a, force_state=(q, v, c))[2].sum())
        dt = (time.perf_counter() - t) / k * 1e6
        print("n=%d  %s host mode: %.1f us per step with %.1f auto-resets per step" % (n, "mapped" if mode else "staged", dt, nd / k), flush=True)
        env.close()
