#!/usr/bin/env python3
"""Device time of the fused normaliser + policy kernel (K6) with CUDA events; prints one JSON line per batch size.
fp32 FMA roofline: 148 SMs x 128 lanes x SM clock (read from nvidia-smi under load)."""
import json
import os
import subprocess
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deepmimic_mujoco_b200.hotpath import HotPath  # noqa: E402


def main():
    sizes = [int(x) for x in (sys.argv[1].split(",") if len(sys.argv) > 1 else ["4096", "65536", "1048576"])]
    dims = (67, 256, 128, 28)
    rng = np.random.default_rng(0)
    W = [rng.normal(0, 0.1, sh).astype(np.float32) for sh in ((dims[0], dims[1]), (dims[1],), (dims[1], dims[2]), (dims[2],),
                                                              (dims[2], dims[3]), (dims[3],))]
    hp = HotPath("humanoid3d")
    hp.policy_upload(*W)
    hp.policy_set_normalizer(np.zeros(67), np.ones(67))
    dev = torch.device("cuda", 0)
    fma_per_env = dims[0] * dims[1] + dims[1] * dims[2] + dims[2] * dims[3]
    for variant, n in [(v, n) for v in (0, 1) for n in sizes]:
        hp.set_policy_variant(variant)
        obs = torch.randn(n, dims[0], device=dev)
        act = torch.empty(n, dims[3], device=dev)
        for _ in range(5):
            hp.policy_act(n, obs, act, normalize=True)
        torch.cuda.synchronize()
        reps = 20
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            hp.policy_act(n, obs, act, normalize=True)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / reps
        clk = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm", "--format=csv,noheader,nounits", "-i", "0"],
                             capture_output=True, text=True).stdout.strip()
        mhz = float(clk or 1965)
        peak = 148 * 128 * mhz * 1e6
        fma = fma_per_env * n / (ms * 1e-3)
        print(json.dumps({"kernel": "policy_mlp_tc_kernel" if variant else "policy_mlp_kernel", "envs": n, "ms": ms, "env_per_s": n / (ms * 1e-3), "fma_per_env": fma_per_env,
                          "TFLOPs": 2 * fma / 1e12, "fp32_peak_TFLOPs": 2 * peak / 1e12, "frac": fma / peak, "sm_mhz": mhz}))


if __name__ == "__main__":
    main()
