import sys, os, time
sys.path.insert(0, "/root/repo")
import numpy as np
from deepmimic_mujoco_b200.vec_env import DPVecEnv
n = int(sys.argv[1])
env = DPVecEnv(n, seed=0, copy_outputs=True)
env.reset()
hp = env.hp
rng = np.random.default_rng(0)
m = hp.mocaps[0]
f = rng.integers(0, 76, n)
q = (np.asarray(m.data_config)[f] + rng.normal(0, 0.02, (n, 35))).astype(np.float32)
v = (np.asarray(m.data_vel)[f] + rng.normal(0, 0.3, (n, 34))).astype(np.float32)
a = rng.uniform(-1, 1, (n, 28)).astype(np.float32)
c = np.zeros(n, np.uint32)
for _ in range(30):
    r = env.step(a, force_state=(q, v, c))
print("dones per step", int(r[2].sum()))
import cProfile, pstats
pr = cProfile.Profile()
pr.enable()
t = time.perf_counter()
for _ in range(50):
    r = env.step(a, force_state=(q, v, c))
dt = (time.perf_counter() - t) / 50 * 1e6
pr.disable()
print("env.step %.1f us" % dt)
pstats.Stats(pr).sort_stats("cumulative").print_stats(12)
