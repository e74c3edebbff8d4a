#!/usr/bin/env python3
"""Kernel-time sweep over batch sizes / robots / options (GPU box).  Prints one JSON line per point."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", default="4096,16384,65536,262144,1048576")
    ap.add_argument("--cases", default="h,g,c")
    ap.add_argument("--stats", default="1,0")
    ap.add_argument("--layouts", default="aos")
    ap.add_argument("--variants", default="-1")
    ap.add_argument("--steps", type=int, default=30)
    a = ap.parse_args()
    peak, _ = bench.peaks()
    dev = torch.device("cuda", 0)
    for case in a.cases.split(","):
        for layout in a.layouts.split(","):
            for variant in a.variants.split(","):
                for st in a.stats.split(","):
                    for n in [int(x) for x in a.envs.split(",")]:
                        argv = ["--envs", str(n), "--layout", layout, "--variant", variant]
                        if case == "g":
                            argv += ["--robot", "unitree_g1", "--motion", "walk"]
                        if case == "c":
                            argv += ["--combined"]
                        if st == "0":
                            argv += ["--no-stats"]
                        sys.argv = ["bench.py"] + argv
                        args = bench.parse()
                        wl = bench.Workload(args, dev, n, 5, 0)
                        ms, _ = bench.kernel_time(wl, a.steps)
                        gb = wl.bpe * n / (ms * 1e-3) / 1e9
                        print(json.dumps({"case": case, "layout": layout, "variant": int(variant), "stats": st == "1", "envs": n,
                                          "ms": round(ms, 5), "env_steps_per_s": round(n / (ms * 1e-3)), "GBps": round(gb, 1),
                                          "frac": round(gb / peak, 4)}), flush=True)
                        del wl
                        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
