#!/usr/bin/env python3
"""Host-buffer step: copy engines (DM_HOST_STAGED: H2D, kernel, D2H) against the kernel working straight on the pinned
host block (DM_HOST_MAPPED: TMA bulk loads / stores and plain stores cross PCIe themselves, both directions at once),
for arrays already in the pinned block and for ordinary (pageable) caller arrays.
usage: zerocopy_probe.py [n ...]      (DM_HOST_CHUNK=<envs> overrides the chunk size of the mapped form)"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from deepmimic_mujoco_b200.vec_env import DPVecEnv


def T(fn, k=300):
    for _ in range(20):
        fn()
    t = time.perf_counter()
    for _ in range(k):
        fn()
    return (time.perf_counter() - t) / k * 1e6


for n in [int(x) for x in (sys.argv[1:] or ["4096", "16384", "65536"])]:
    env = DPVecEnv(n, seed=0, copy_outputs=False)
    env.reset()
    hp = env.hp
    rng = np.random.default_rng(0)
    m = hp.mocaps[0]
    f = rng.integers(0, 76, n)
    q = (np.asarray(m.data_config)[f] + rng.normal(0, 0.02, (n, 35))).astype(np.float32)
    v = (np.asarray(m.data_vel)[f] + rng.normal(0, 0.3, (n, 34))).astype(np.float32)
    a = rng.uniform(-1, 1, (n, 28)).astype(np.float32)
    c = np.zeros(n, np.uint32)
    qb, vb, cb, ab = env.input_buffers()
    qb[:] = q; vb[:] = v; ab[:] = a; cb[:] = c
    k = 300 if n <= 65536 else 50
    out = {}
    for mode in (0, 1):
        hp.set_host_mode(mode)
        env.copy_outputs = False
        out["pinned", mode] = T(lambda: env.step(ab, force_state=(qb, vb, cb)), k)
        env.copy_outputs = True
        out["pageable", mode] = T(lambda: env.step(a, force_state=(q, v, c)), k)
    obs_out = np.empty((n, 67), np.float32)
    out["raw"] = T(lambda: hp.step_host(n, q, v, a, c, env.idx, env.ep_len, env.clip, obs_out, env.reward, env.done, env.reason,
                                        env.torque, None, env.ep_return), k)
    print("n=%d  step_host alone, pageable in, obs copied out (mapped): %.1f us" % (n, out["raw"]))
    print("n=%d  pinned in/out: staged %.1f us, mapped %.1f us   pageable in, fresh out: staged %.1f us, mapped %.1f us  (chunk %s)"
          % (n, out["pinned", 0], out["pinned", 1], out["pageable", 0], out["pageable", 1], os.environ.get("DM_HOST_CHUNK", "auto")),
          flush=True)
    env.close()
