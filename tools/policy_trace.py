#!/usr/bin/env python3
"""Per-tile event timeline of the tensor-core policy kernel (CTA 0, epilogue thread 0): clock64 marks
0/1 layer-0 halves ready, 2 H1 handed over, 3 next X0 staged, 4 layer 1 ready, 5 H2 handed over, 6 head ready."""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deepmimic_mujoco_b200.hotpath import HotPath  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1048576
dims = (67, 256, 128, 28)
rng = np.random.default_rng(0)
W = [rng.normal(0, 0.1, sh).astype(np.float32) for sh in ((dims[0], dims[1]), (dims[1],), (dims[1], dims[2]), (dims[2],),
                                                          (dims[2], dims[3]), (dims[3],))]
hp = HotPath("humanoid3d")
hp.policy_upload(*W)
hp.set_policy_variant(1)
dev = torch.device("cuda", 0)
obs, act = torch.randn(n, dims[0], device=dev), torch.empty(n, dims[3], device=dev)
trace = torch.zeros(64 * 16, dtype=torch.int64, device=dev)
hp.policy_act(n, obs, act)
cta = int(sys.argv[2]) if len(sys.argv) > 2 else 0
hp.L.dm_policy_set_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
hp.L.dm_policy_set_trace(hp.h, C.c_void_p(trace.data_ptr()), cta)
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
hp.policy_act(n, obs, act)
b.record()
torch.cuda.synchronize()
print("CTA %d; kernel %.3f ms" % (cta, a.elapsed_time(b)))
t = trace.cpu().numpy().reshape(64, 16)
t0 = t[0, 7]
names = ["D0a", "D0b", "H1>", "X0'", "D1", "H2>", "D2"]
print("tile  " + "  ".join("%7s" % x for x in names) + "   (cycles since kernel start; then deltas)")
prev = t0
for i in range(12):
    row = t[i, :7] - t0
    print("%4d  " % i + "  ".join("%7d" % x for x in row))
print("issuer: tile  L0 start/issued   L1 start/issued   head start/issued  | L1 run = D1 - L1 start, head run = D2 - head start")
for i in range(2, 10):
    r = t[i] - t0
    print("        %4d  %7d %7d   %7d %7d   %7d %7d  | L0a run %5d  L1 run %5d  head run %5d" % (
        i, r[8], r[11], r[9], r[12], r[10], r[13], r[0] - r[8], r[4] - r[9], r[6] - r[10]))
print("stage_x0 (during tile i): H1> -> raw ready %d -> X0 built %d -> done %d" % (
    np.mean(t[2:20, 14] - t[2:20, 2]), np.mean(t[2:20, 15] - t[2:20, 14]), np.mean(t[2:20, 3] - t[2:20, 15])))
last = max(i for i in range(64) if t[i, 6] > 0)
print("tiles traced: %d; last head ready at %d cycles" % (last + 1, t[last, 6] - t0))
d = np.diff(t[:20, 0])
print("tile period (cycles):", d)
seg = t[2:20]
print("mean: D0a->D0b %d  D0b->H1 %d  H1->X0' %d  X0'->D1 %d  D1->H2 %d  H2->D2 %d  D2->next D0a %d" % tuple(
    [np.mean(seg[:, 1] - seg[:, 0]), np.mean(seg[:, 2] - seg[:, 1]), np.mean(seg[:, 3] - seg[:, 2]), np.mean(seg[:, 4] - seg[:, 3]),
     np.mean(seg[:, 5] - seg[:, 4]), np.mean(seg[:, 6] - seg[:, 5]), np.mean(seg[1:, 0] - seg[:-1, 6])]))
