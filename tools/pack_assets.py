#!/usr/bin/env python3
"""Build the asset pack the package ships with (run in a container that has the reference).

Inputs (read-only, never copied verbatim):
  <ref>/src/mujoco/humanoid_deepmimic/envs/asset/deepmimic_{humanoid3d,unitree_g1}.xml
  <ref>/src/mujoco/motions/*.txt     (DeepMimic JSON clips, reference loader mocap_v2.py:40-51)
Outputs:
  deepmimic_mujoco_b200/assets/models.json   compiled kinematic models (mjcf.RobotModel)
  deepmimic_mujoco_b200/assets/motions.npz   per clip: raw "Frames" (float64), "Loop", "Format"

The GPU box has no /root/reference, so everything that needs the robot models or
the clips at run time reads this pack (or a reference checkout named by
$DEEPMIMIC_MUJOCO_SRC).
"""
import argparse
import glob
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from deepmimic_mujoco_b200.mjcf import compile_mjcf, dump_models  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default="/root/reference")
    ap.add_argument("--out", default=os.path.join(os.path.dirname(HERE), "deepmimic_mujoco_b200", "assets"))
    a = ap.parse_args()
    asset = os.path.join(a.ref, "src/mujoco/humanoid_deepmimic/envs/asset")
    models = {r: compile_mjcf(os.path.join(asset, "deepmimic_%s.xml" % r), r) for r in ("humanoid3d", "unitree_g1")}
    os.makedirs(a.out, exist_ok=True)
    dump_models(models, os.path.join(a.out, "models.json"))
    pack = {}
    for p in sorted(glob.glob(os.path.join(a.ref, "src/mujoco/motions/*.txt"))):
        name = os.path.splitext(os.path.basename(p))[0]
        with open(p) as f:
            d = json.load(f)
        pack[name + "/Frames"] = np.asarray(d["Frames"], dtype=np.float64)
        pack[name + "/Loop"] = np.asarray(d.get("Loop", ""))
        pack[name + "/HasLoop"] = np.asarray("Loop" in d)
        pack[name + "/Format"] = np.asarray(d.get("Format", ""))
    np.savez_compressed(os.path.join(a.out, "motions.npz"), **pack)
    print("models:", {k: (m.nq, m.nv, m.nu, m.nbody) for k, m in models.items()})
    print("clips:", len(pack) // 4, "->", os.path.getsize(os.path.join(a.out, "motions.npz")), "bytes")


if __name__ == "__main__":
    main()
