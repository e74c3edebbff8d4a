#!/usr/bin/env python3
"""Where the host step goes at scale: DPVecEnv.step at a large batch (default G1 combined, 262144 envs; BASELINE
configs[3]) in mapped and staged mode, pageable and in-place inputs, with the library's own phase trace.
Usage: DM_HOST_TRACE=1 python tools/e2e_scale_probe.py [envs] [combined(1/0)]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from deepmimic_mujoco_b200.vec_env import DPVecEnv

n = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
combined = (sys.argv[2] if len(sys.argv) > 2 else "1") == "1"
rng = np.random.default_rng(0)


def T(fn, k=8):
    for _ in range(3):
        fn()
    t = time.perf_counter()
    for _ in range(k):
        fn()
    return (time.perf_counter() - t) / k * 1e3


for mapped in (1, 0):
    env = DPVecEnv(n, combined=True, seed=0) if combined else DPVecEnv(n, seed=0)
    env.hp.set_host_mode(bool(mapped))
    env.reset()
    d = env.hp.dims
    q = env.qpos.copy() + rng.normal(0, 0.01, env.qpos.shape).astype(np.float32)
    v = env.qvel.copy()
    a = rng.uniform(-0.3, 0.3, (n, d.na)).astype(np.float32)
    c = np.zeros(n, np.uint32)
    qb, vb, cb, ab = env.input_buffers()
    qb[:], vb[:], ab[:] = q, v, a
    mb = (q.nbytes + v.nbytes + a.nbytes + 4 * n + env.obs.nbytes + 4 * n * d.nu) / 1e6
    t1 = T(lambda: env.step(a, force_state=(q, v, c)))
    env.copy_outputs = False
    t2 = T(lambda: env.step(a, force_state=(q, v, c)))
    t3 = T(lambda: env.step(ab, force_state=(qb, vb, cb)))
    print("mode=%s n=%d  %.0f MB/step | pageable in, owned out %.2f ms | pageable in, views out %.2f ms | in place %.2f ms"
          % ("mapped" if mapped else "staged", n, mb, t1, t2, t3), flush=True)
    env.close()
    del env
