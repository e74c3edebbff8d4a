"""Access to the robot models and motion clips.

Default source is the pack built by tools/pack_assets.py (assets/models.json, assets/motions.npz).
If $DEEPMIMIC_MUJOCO_SRC names a checkout of the reference's `src/` directory, models are compiled
from its MJCF files and clips parsed from its motion files instead (reference paths:
src/mujoco/humanoid_deepmimic/envs/asset/*.xml, src/mujoco/motions/*.txt)."""
import json
import os

import numpy as np

from . import mjcf

ASSET_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "assets")
_models = {}
_motions = None


def _src():
    p = os.environ.get("DEEPMIMIC_MUJOCO_SRC")
    return p if p and os.path.isdir(p) else None


def load_model(robot) -> mjcf.RobotModel:
    if robot not in _models:
        src = _src()
        if src:
            path = os.path.join(src, "mujoco/humanoid_deepmimic/envs/asset/deepmimic_%s.xml" % robot)
            _models[robot] = mjcf.compile_mjcf(path, robot)
        else:
            _models.update(mjcf.load_models(os.path.join(ASSET_DIR, "models.json")))
        if robot not in _models:
            raise Exception("Unknown robot: %s" % robot)
    return _models[robot]


def clip_name(robot, motion):
    return "%s_%s" % (robot, motion)


def load_clip(robot, motion):
    """-> dict(frames=float64 [n, cols], format=str or None, loop=str or None, name=str)"""
    global _motions
    name = clip_name(robot, motion)
    src = _src()
    if src:
        return parse_clip_file(os.path.join(src, "mujoco/motions", name + ".txt"))
    if _motions is None:
        _motions = np.load(os.path.join(ASSET_DIR, "motions.npz"))
    if name + "/Frames" not in _motions.files:
        raise FileNotFoundError("no motion clip %r in the asset pack" % name)
    return dict(frames=np.array(_motions[name + "/Frames"], dtype=np.float64),
                format=str(_motions[name + "/Format"]) or None,
                loop=str(_motions[name + "/Loop"]) if bool(_motions[name + "/HasLoop"]) else None, name=name)


def parse_clip_file(path):
    """DeepMimic JSON clip (mocap_v2.py:40-51): "Frames", optional "Loop", optional "Format"."""
    with open(path) as f:
        d = json.load(f)
    return dict(frames=np.array(d["Frames"], dtype=np.float64), format=d.get("Format") or None, loop=d.get("Loop"),
                name=os.path.splitext(os.path.basename(path))[0])


def available_motions(robot):
    global _motions
    if _motions is None:
        _motions = np.load(os.path.join(ASSET_DIR, "motions.npz"))
    pre = robot + "_"
    return sorted(k[len(pre):-len("/Frames")] for k in _motions.files if k.startswith(pre) and k.endswith("/Frames"))
