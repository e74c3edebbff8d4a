"""Clip-table builder restated from the reference's MocapDM (oracle; test infrastructure).

Follows reference src/mujoco/mocap_v2.py:33-362 (read_raw_data, calc_rot_vel, accessors)
and src/mujoco/mocap_util.py:5-48 (joint orders, align_rotation, align_position).
The FK stage (mocap_v2.py:301-307: a throw-away DPEnv + set_state per key-frame) uses
oracle.mjkin.forward.  Input is the parsed clip ("Frames" array, optional "Format"),
so the same function serves the JSON files and the packed assets.
"""
import json

import numpy as np

from . import mjkin
from . import rotations as R

# mocap_util.py:5-20
MJCF_JOINT_ORDER = ["chest", "neck", "right_shoulder", "right_elbow", "left_shoulder", "left_elbow",
                    "right_hip", "right_knee", "right_ankle", "left_hip", "left_knee", "left_ankle"]
FILE_JOINT_ORDER = ["chest", "neck", "right_hip", "right_knee", "right_ankle", "right_shoulder",
                    "right_elbow", "left_hip", "left_knee", "left_ankle", "left_shoulder", "left_elbow"]
JOINT_DOF = {"chest": 3, "neck": 3, "right_shoulder": 3, "right_elbow": 1, "left_shoulder": 3,
             "left_elbow": 1, "right_hip": 3, "right_knee": 1, "right_ankle": 3, "left_hip": 3,
             "left_knee": 1, "left_ankle": 3}

# mocap_v2.py:148-154 (hard windows of the "continuity" fix)
BALL_JOINTS = ("left_shoulder", "right_shoulder", "left_hip", "right_hip")
E_LIM = {  # joint: ((xmin, xmax), (ymin, ymax), (zmin, zmax))
    "left_shoulder": ((-0.50, 3.14), (-3.14, 0.70), (-1.50, 1.50)),
    "right_shoulder": ((-3.14, 0.50), (-3.14, 0.70), (-1.50, 1.50)),
    "left_hip": ((-1.2, 1.2), (-2.57, 1.57), (-1.0, 1.0)),
    "right_hip": ((-1.2, 1.2), (-2.57, 1.57), (-1.0, 1.0)),
}

TARGET_DT = 0.01666  # mocap_v2.py:311


def _align_quats():
    # mocap_util.py:33-38: Quaternion(matrix=...) of the two constant matrices
    right = R.pq_from_matrix(np.array([[1.0, 0.0, 0.0], [0.0, 0.0, 1.0], [0.0, -1.0, 0.0]]))
    left = R.pq_from_matrix(np.array([[1.0, 0.0, 0.0], [0.0, 0.0, -1.0], [0.0, 1.0, 0.0]]))
    return left, right


_QL, _QR = _align_quats()


def align_rotation(q_wxyz):
    """mocap_util.py:31-40: q_left * q * q_right (y-up -> z-up)."""
    return R.pq_mul(R.pq_mul(_QL, np.asarray(q_wxyz, dtype=np.float64)), _QR)


def align_position(p):
    """mocap_util.py:42-48."""
    return np.array([[1.0, 0.0, 0.0], [0.0, 0.0, -1.0], [0.0, 1.0, 0.0]]) @ np.asarray(p, dtype=np.float64)


def calc_rot_vel(q0, q1, dt):
    """mocap_v2.py:350-362: angle/dt * axis of conj(q0) * q1 (pyquaternion semantics)."""
    qd = R.pq_mul(R.pq_conj(np.asarray(q0, dtype=np.float64)), np.asarray(q1, dtype=np.float64))
    return (R.pq_angle(qd) / dt) * R.pq_axis(qd)


def _continuity_fix(joint, euler, prev, quat_xyzw, vmx_dt):
    """mocap_v2.py:196-222, FIX_SINGULARITY_MODE == 'continuity'."""
    lo, hi, tgt = [0.0] * 3, [0.0] * 3, [0.0] * 3
    for a in range(3):
        lo[a] = max(E_LIM[joint][a][0], prev[a] - vmx_dt)
        hi[a] = min(E_LIM[joint][a][1], prev[a] + vmx_dt)
        tgt[a] = np.clip(euler[a], lo[a], hi[a])
    if np.allclose(list(euler), tgt):
        return tuple(euler)
    best, best_err = None, np.inf
    cands = [[tgt[a], prev[a]] + list(np.linspace(lo[a], hi[a], 6)) for a in range(3)]
    for cx in cands[0]:
        for cy in cands[1]:
            for cz in cands[2]:
                qn = R.gohlke_quat_from_euler_rxyz(cx, cy, cz)
                err = min(np.linalg.norm(qn - quat_xyzw), np.linalg.norm(-qn - quat_xyzw)) ** 2
                if err < best_err:
                    best_err, best = err, (cx, cy, cz)
    return best


class MocapTables:
    """Restated MocapDM: data_config / data_vel / data_body_xpos / data_geom_xpos lists."""

    def __init__(self, model, robot):
        self.model = model
        self.robot = robot
        self.dt = None
        self.loop = None
        self.data_config = None
        self.data_vel = None
        self.data_body_xpos = None
        self.data_geom_xpos = None

    # accessors, mocap_v2.py:25-28,338-348
    def get_length(self):
        return 0 if self.data_config is None else len(self.data_config)

    def get_qpos(self, i):
        return self.data_config[i]

    def get_qvel(self, i):
        return self.data_vel[i]

    def get_geom_xpos(self, i):
        return self.data_geom_xpos[i]

    def get_body_xpos(self, i):
        return self.data_body_xpos[i]

    def load_file(self, path):
        with open(path) as f:
            d = json.load(f)
        import os
        name = os.path.splitext(os.path.basename(path))[0]
        return self.load(np.array(d["Frames"], dtype=np.float64), d.get("Format") or None, name, d.get("Loop"))

    def load(self, frames, fmt, clip_name, loop=None):
        frames = np.array(frames, dtype=np.float64)
        self.loop = loop
        self.dt = float(frames[0][0])  # mocap_v2.py:48
        self.motion_name = clip_name
        if not fmt:
            self.data_config = self._convert_deepmimic(frames, clip_name)
        else:  # "direct_qpos", mocap_v2.py:271-272
            self.data_config = [row[1:].copy() for row in frames]
        self._velocities()
        self._fk_tables()
        self._interpolate()
        return self

    def _convert_deepmimic(self, frames, clip_name):
        """mocap_v2.py:56-251: align, per-joint Euler 'rxyz', singularity fix, MJCF joint order."""
        vmx = 5.0 if "getup" in clip_name else 10.0  # mocap_v2.py:156-158
        states = []
        for row in frames:
            st = {"root_pos": align_position(row[1:4]), "root_rot": align_rotation(row[4:8])}
            c = 8
            for jn in FILE_JOINT_ORDER:
                if JOINT_DOF[jn] == 1:
                    st[jn] = row[c:c + 1].copy()
                    c += 1
                else:
                    st[jn] = align_rotation(row[c:c + 4])
                    c += 4
            states.append(st)
        fixed_prev = {}
        rows = []
        for k, st in enumerate(states):
            ang = list(st["root_pos"]) + list(st["root_rot"])
            for jn in MJCF_JOINT_ORDER:
                if JOINT_DOF[jn] == 1:
                    ang += list(st[jn])
                    continue
                w, x, y, z = st[jn]
                q_xyzw = np.array([x, y, z, w])
                e = R.gohlke_euler_rxyz(q_xyzw)
                if jn in BALL_JOINTS:
                    prev = e if k == 0 else fixed_prev[jn]  # mocap_v2.py:163-169
                    e = _continuity_fix(jn, e, prev, q_xyzw, vmx * self.dt)
                    fixed_prev[jn] = e
                ang += list(e)
            rows.append(np.array(ang, dtype=np.float64))
        return rows

    def _velocities(self):
        """mocap_v2.py:274-289 (frame 0 differences itself -> zeros)."""
        self.data_vel = []
        for k in range(len(self.data_config)):
            a, b = self.data_config[max(k - 1, 0)], self.data_config[k]
            self.data_vel.append(np.concatenate([
                (b[:3] - a[:3]) / self.dt, calc_rot_vel(a[3:7], b[3:7], self.dt), (b[7:] - a[7:]) / self.dt]))

    def _fk_tables(self):
        """mocap_v2.py:292-307."""
        self.data_body_xpos, self.data_geom_xpos = [], []
        for qpos, qvel in zip(self.data_config, self.data_vel):
            d = mjkin.forward(self.model, qpos, qvel)
            self.data_body_xpos.append(d.xpos * 1.0)
            self.data_geom_xpos.append(d.geom_xpos * 1.0)

    def _interpolate(self):
        """mocap_v2.py:309-336: x k linear blend to the 60 Hz simulator step."""
        ratio = self.dt / TARGET_DT
        k = int(ratio)
        if abs(ratio - k) > 0.1:
            raise Exception("Invalid dt ratio, cannot interpolate mocap frames: %f" % ratio)
        if k <= 1:
            return
        out = ([], [], [], [])
        src = (self.data_config, self.data_vel, self.data_body_xpos, self.data_geom_xpos)
        for ia in range(len(self.data_config) - 1):
            for j in range(k):
                B = j * 1.0 / k
                A = 1.0 - B
                for o, s in zip(out, src):
                    o.append(A * s[ia] + B * s[ia + 1])
        self.dt = TARGET_DT
        self.data_config, self.data_vel, self.data_body_xpos, self.data_geom_xpos = out
