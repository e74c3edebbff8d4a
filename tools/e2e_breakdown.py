#!/usr/bin/env python3
"""Where the end-to-end step time goes at 4096 environments (host side)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from deepmimic_mujoco_b200.vec_env import DPVecEnv, LazyInfos

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
env = DPVecEnv(n, seed=0)
env.reset()
rng = np.random.default_rng(0)
m = env.hp.mocaps[0]
f = rng.integers(0, 76, n)
q = (np.asarray(m.data_config)[f] + rng.normal(0, 0.02, (n, 35))).astype(np.float32)
v = (np.asarray(m.data_vel)[f] + rng.normal(0, 0.3, (n, 34))).astype(np.float32)
a = rng.uniform(-1, 1, (n, 28)).astype(np.float32)
c = np.zeros(n, np.uint32)

def T(fn, k=300):
    for _ in range(20): fn()
    t = time.perf_counter()
    for _ in range(k): fn()
    return (time.perf_counter() - t) / k * 1e6

qb, vb, cb, ab = env.input_buffers()
hp = env.hp
print("full step (pageable in, copies out)   %.1f us" % T(lambda: env.step(a, force_state=(q, v, c))))
env.copy_outputs = False
print("full step (pageable in, views out)    %.1f us" % T(lambda: env.step(a, force_state=(q, v, c))))
print("full step (pinned in, views out)      %.1f us" % T(lambda: env.step(ab, force_state=(qb, vb, cb))))
print("dm_step_host only (pinned)            %.1f us" % T(lambda: hp.step_host(n, qb, vb, ab, cb, env.idx, env.ep_len, env.clip, env.obs, env.reward, env.done, env.reason, env.torque, None, env.ep_return)))
print("input copies                          %.1f us" % T(lambda: (np.copyto(qb, q), np.copyto(vb, v), np.copyto(ab, a), np.copyto(cb, c))))
print("obs copy                              %.1f us" % T(lambda: env.obs.copy()))
print("LazyInfos                             %.1f us" % T(lambda: LazyInfos(n)))
import torch
x = torch.empty(1671168, dtype=torch.uint8).pin_memory(); y = torch.empty(1671168, dtype=torch.uint8, device="cuda")
def h2d():
    y.copy_(x, non_blocking=True); torch.cuda.synchronize()
def d2h():
    x.copy_(y, non_blocking=True); torch.cuda.synchronize()
print("torch H2D 1.67MB + sync               %.1f us" % T(h2d))
print("torch D2H 1.67MB + sync               %.1f us" % T(d2h))
