"""Reads the packed model constants and clips for the oracle (test infrastructure).

Data only: deepmimic_mujoco_b200/assets/{models.json,motions.npz} are produced by
tools/pack_assets.py from the reference's MJCF files and motion clips.
"""
import json
import os
from types import SimpleNamespace

import numpy as np

ASSET_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                         "deepmimic_mujoco_b200", "assets")
_INT = {"body_parentid", "body_jntadr", "body_jntnum", "jnt_type", "jnt_bodyid", "jnt_qposadr",
        "jnt_dofadr", "geom_bodyid", "geom_contype", "geom_conaffinity", "actuator_jntid", "actuator_dofadr"}
_BOOL = {"jnt_limited", "actuator_ctrllimited"}
_STR = {"body_names", "jnt_names", "geom_names", "actuator_names"}
_cache = {}


def load_model(robot):
    if ("m", robot) in _cache:
        return _cache[("m", robot)]
    with open(os.path.join(ASSET_DIR, "models.json")) as f:
        d = json.load(f)[robot]
    m = SimpleNamespace()
    for k, v in d.items():
        if isinstance(v, list) and k not in _STR:
            v = np.asarray(v, dtype=np.int64 if k in _INT else bool if k in _BOOL else np.float64)
        setattr(m, k, v)
    m.geom_pos = m.geom_pos.reshape(-1, 3)
    m.body_name2id = lambda name: m.body_names.index(name)
    m.geom_name2id = lambda name: m.geom_names.index(name)
    m.geom_id2name = lambda i: m.geom_names[i] or None
    m.joint_names = tuple(m.jnt_names)
    _cache[("m", robot)] = m
    return m


def load_clip(robot, motion):
    """-> (frames float64 [n, cols], format or None, clip name, loop or None)"""
    key = ("z",)
    if key not in _cache:
        _cache[key] = np.load(os.path.join(ASSET_DIR, "motions.npz"))
    z = _cache[key]
    name = "%s_%s" % (robot, motion)
    fmt = str(z[name + "/Format"]) or None
    loop = str(z[name + "/Loop"]) if bool(z[name + "/HasLoop"]) else None
    return z[name + "/Frames"], fmt, name, loop
