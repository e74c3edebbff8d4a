#!/usr/bin/env python3
"""Independent forward kinematics fixtures: tests/golden/fk_independent.npz.

A SECOND code path for the pose quantities the reference reads from MuJoCo
(`body_xpos`, `geom_xpos`, `xmat`; reference src/deepmimic_env.py:29,51-54,232,236,421 and
src/mujoco/mocap_v2.py:292-307), sharing nothing with oracle/mjkin.py or the package's MJCF
compiler:

  * the kinematic tree, joint axes, body / geom offsets are read here straight from the
    reference's MJCF files with xml.etree (not through deepmimic_mujoco_b200/mjcf.py or
    assets/models.json);
  * every body frame is a product of 4x4 homogeneous matrices built with the reference-held
    src/transformations.py (`translation_matrix` :195, `rotation_matrix` :275,
    `quaternion_matrix` :1174, `concatenate_matrices` :1649) -- rotation matrices all the way,
    where mjkin chains quaternions.

Runs only in the build container (needs /root/reference).  Recorded: random states of both
robots (inputs rounded to float32 like every other fixture) and EVERY key frame of every clip in
the pack (the reference's `test_walk_hand_xpos_mocap` invariant, src/deepmimic_env.py:540-554,
needs exactly those), each with body_xpos, geom_xpos and the body rotation matrices.

Usage: python tools/gen_fk_independent.py
"""
import os
import sys
import xml.etree.ElementTree as ET

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("DM_REFERENCE", "/root/reference")
sys.path.insert(0, os.path.join(REF, "src"))
sys.path.insert(0, REPO)
import transformations as T  # noqa: E402  (reference, vendored Gohlke; quaternions are x, y, z, w)

ASSET = os.path.join(REF, "src", "mujoco", "humanoid_deepmimic", "envs", "asset")
# RobotConfig.endeffector_geom_names (reference src/config.py:7-22)
EE = {"humanoid3d": ["left_ankle", "right_ankle", "left_wrist", "right_wrist"],
      "unitree_g1": ["left_foot", "right_foot", "left_hand", "right_hand"]}
XML = {"humanoid3d": "deepmimic_humanoid3d.xml", "unitree_g1": "deepmimic_unitree_g1.xml"}


def _floats(s, default):
    return np.asarray([float(x) for x in (s if s is not None else default).split()], dtype=np.float64)


def _wxyz_matrix(q):
    return T.quaternion_matrix([q[1], q[2], q[3], q[0]])


class Tree:
    """Bodies, joints and geoms in MJCF document order (depth first), which is MuJoCo's id order."""

    def __init__(self, robot):
        root = ET.parse(os.path.join(ASSET, XML[robot])).getroot()
        self.bodies = []  # (name, parent index, pos, quat wxyz, [joint ...], [geom ...])
        self.geom_names = []
        self.body_names = ["world"]
        self.nq = 0
        wb = root.find("worldbody")
        self.world_geoms = []
        for g in wb.findall("geom"):
            self.geom_names.append(g.get("name") or "")
            self.world_geoms.append(self._geom_pos(g))
        self._walk(wb, 0)

    @staticmethod
    def _geom_pos(g):
        if g.get("fromto") is not None:  # capsule given by its end points: the frame sits at the midpoint
            ft = _floats(g.get("fromto"), "")
            return 0.5 * (ft[:3] + ft[3:])
        return _floats(g.get("pos"), "0 0 0")

    def _walk(self, elem, parent):
        for b in elem.findall("body"):
            joints = []
            for j in b.findall("joint") + b.findall("freejoint"):
                typ = "free" if j.tag == "freejoint" else j.get("type", "hinge")
                adr = self.nq
                self.nq += 7 if typ == "free" else 1
                joints.append((typ, adr, _floats(j.get("axis"), "0 0 1"), _floats(j.get("pos"), "0 0 0")))
            geoms = []
            for g in b.findall("geom"):
                geoms.append((len(self.geom_names), self._geom_pos(g)))
                self.geom_names.append(g.get("name") or "")
            me = len(self.bodies) + 1
            self.body_names.append(b.get("name"))
            self.bodies.append((b.get("name"), parent, _floats(b.get("pos"), "0 0 0"), _floats(b.get("quat"), "1 0 0 0"),
                                joints, geoms))
            self._walk(b, me)

    def forward(self, qpos):
        nb = len(self.bodies) + 1
        frames = [np.identity(4)] * nb
        xpos = np.zeros((nb, 3))
        xmat = np.zeros((nb, 3, 3))
        xmat[0] = np.identity(3)
        gx = np.zeros((len(self.geom_names), 3))
        for k, p in enumerate(self.world_geoms):
            gx[k] = p
        for i, (_, parent, pos, quat, joints, geoms) in enumerate(self.bodies, start=1):
            M = T.concatenate_matrices(frames[parent], T.translation_matrix(pos), _wxyz_matrix(quat))
            for typ, adr, axis, jpos in joints:
                if typ == "free":  # the free joint replaces the body's frame by (qpos[0:3], qpos[3:7])
                    M = T.concatenate_matrices(T.translation_matrix(qpos[adr:adr + 3]), _wxyz_matrix(qpos[adr + 3:adr + 7]))
                else:  # rotation by the joint angle about the axis through the joint anchor
                    M = T.concatenate_matrices(M, T.rotation_matrix(qpos[adr], axis, point=jpos))
            frames[i] = M
            xpos[i] = M[:3, 3]
            xmat[i] = M[:3, :3]
            for gid, gpos in geoms:
                gx[gid] = (M @ np.append(gpos, 1.0))[:3]
        return xpos, xmat, gx


def random_states(robot, tree, n, rng):
    from oracle.assets import load_model
    m = load_model(robot)
    assert tree.nq == m.nq and tree.body_names == list(m.body_names) and tree.geom_names == list(m.geom_names)
    qs = np.zeros((n, m.nq))
    for k in range(n):
        q = np.zeros(m.nq)
        q[:3] = rng.normal(0, 0.5, 3) + [0, 0, 0.9]
        q[3:7] = rng.normal(0, 1, 4)
        lo, hi = m.jnt_range[1:, 0], m.jnt_range[1:, 1]
        q[7:] = rng.uniform(lo - 0.2, hi + 0.2)
        qs[k] = np.asarray(q, dtype=np.float32)
    return qs


def main():
    from oracle.assets import ASSET_DIR, load_clip
    from oracle.envs import tables_for
    rng = np.random.default_rng(20261105)
    out = {}
    z = np.load(os.path.join(ASSET_DIR, "motions.npz"))
    clips = sorted({k.split("/")[0] for k in z.files})
    for robot in ("humanoid3d", "unitree_g1"):
        tree = Tree(robot)
        qs = random_states(robot, tree, 48, rng)
        res = [tree.forward(q) for q in qs]
        out["rand/%s/qpos" % robot] = qs
        out["rand/%s/xpos" % robot] = np.asarray([r[0] for r in res])
        out["rand/%s/xmat" % robot] = np.asarray([r[1] for r in res])
        out["rand/%s/geom_xpos" % robot] = np.asarray([r[2] for r in res])
        for name in clips:
            if not name.startswith(robot + "_"):
                continue
            motion = name[len(robot) + 1:]
            try:
                t = tables_for(robot, motion)
            except Exception:  # backflip / spin: "Invalid dt ratio" (mocap_v2.py:315-316)
                continue
            cfg = np.asarray(t.data_config)
            n_key = len(load_clip(robot, motion)[0])
            stride = 1 if len(cfg) == n_key else len(cfg) // (n_key - 1)  # x k resampling (mocap_v2.py:309-336)
            out["clip/%s/stride" % name] = np.int64(stride)
            res = [tree.forward(q) for q in cfg]
            # inputs = the oracle loader's data_config rows (pinned to the reference's MocapDM by loader_tables.npz);
            # a checksum guards against the two drifting apart.  Geoms: the four end effectors only (size).
            ee = [tree.geom_names.index(g) for g in EE[robot]]
            out["clip/%s/qpos_sum" % name] = np.float64(cfg.sum())
            out["clip/%s/xpos3" % name] = np.asarray([r[0] for r in res[::3]])  # all bodies, every third row
            out["clip/%s/ee_xpos" % name] = np.asarray([r[2][ee] for r in res])
    path = os.path.join(REPO, "tests", "golden", "fk_independent.npz")
    np.savez_compressed(path, **out)
    print("wrote %s (%.1f kB, %d arrays)" % (path, os.path.getsize(path) / 1e3, len(out)))


if __name__ == "__main__":
    main()
