"""MocapDM: the clip-table builder, same public surface as the reference's
src/mujoco/mocap_v2.py:12-362 (`load_mocap`, `get_length`, `get_qpos/get_qvel/get_body_xpos/
get_geom_xpos`, `data_config`, `data_vel`, `data_body_xpos`, `data_geom_xpos`, `dt`, `loop`).

What differs from the reference, by design:
  * the per-key-frame forward kinematics (mocap_v2.py:292-307: a throw-away MuJoCo env and one
    `set_state` per frame) is ONE launch of the float64 FK kernel (`dm_fk_tables`, K3);
  * the 8x8x8 candidate search of the shoulder / hip "continuity" fix (mocap_v2.py:196-222) is
    evaluated as one numpy batch per joint and frame instead of three nested Python loops;
  * the third-party helpers (pyquaternion, transformations.py) are replaced by the few closed-form
    expressions below.
Everything is float64 on the host, as in the reference.
"""
import math
import os

import numpy as np

from . import assets

BODY_JOINTS = ["chest", "neck", "right_shoulder", "right_elbow", "left_shoulder", "left_elbow",
               "right_hip", "right_knee", "right_ankle", "left_hip", "left_knee", "left_ankle"]  # MJCF order
BODY_JOINTS_IN_DP_ORDER = ["chest", "neck", "right_hip", "right_knee", "right_ankle", "right_shoulder",
                           "right_elbow", "left_hip", "left_knee", "left_ankle", "left_shoulder", "left_elbow"]
DOF_DEF = {"root": 3, "chest": 3, "neck": 3, "right_shoulder": 3, "right_elbow": 1, "right_wrist": 0,
           "left_shoulder": 3, "left_elbow": 1, "left_wrist": 0, "right_hip": 3, "right_knee": 1,
           "right_ankle": 3, "left_hip": 3, "left_knee": 1, "left_ankle": 3}
_BALL = {  # hard Euler windows of the continuity fix, mocap_v2.py:148-154: (x, y, z) x (min, max)
    "left_shoulder": np.array([[-0.50, 3.14], [-3.14, 0.70], [-1.50, 1.50]]),
    "right_shoulder": np.array([[-3.14, 0.50], [-3.14, 0.70], [-1.50, 1.50]]),
    "left_hip": np.array([[-1.2, 1.2], [-2.57, 1.57], [-1.0, 1.0]]),
    "right_hip": np.array([[-1.2, 1.2], [-2.57, 1.57], [-1.0, 1.0]]),
}
_SIM_DT = 0.01666
_H = math.sqrt(0.5)
_EPS = np.finfo(float).eps * 4.0


def _qmul(a, b):
    return np.array([a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3],
                     a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
                     a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1],
                     a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0]])


_Q_LEFT = np.array([2.0, 2.0, 0.0, 0.0]) * (0.5 / math.sqrt(2.0))    # Rx(+90 deg), as pyquaternion's trace method gives it
_Q_RIGHT = np.array([2.0, -2.0, 0.0, 0.0]) * (0.5 / math.sqrt(2.0))  # Rx(-90 deg)


def align_rotation(rot):
    """y-up -> z-up: q_left * q * q_right (mocap_util.py:31-40)."""
    return _qmul(_qmul(_Q_LEFT, np.asarray(rot, dtype=np.float64)), _Q_RIGHT)


def align_position(pos):
    """(x, y, z) -> (x, -z, y) (mocap_util.py:42-48)."""
    x, y, z = pos
    return np.array([1.0 * x + 0.0 * y + 0.0 * z, 0.0 * x + 0.0 * y + -1.0 * z, 0.0 * x + 1.0 * y + 0.0 * z])


def _euler_rxyz(q_wxyz):
    """Intrinsic x-y-z Euler angles of a quaternion = euler_from_quaternion(xyzw, 'rxyz') of the
    vendored transformations.py (mocap_v2.py:140-142); the matrix is self-normalising."""
    w, x, y, z = q_wxyz
    n = x * x + y * y + z * z + w * w
    if n < _EPS:
        m00 = m11 = m22 = 1.0
        m01 = m02 = m10 = m12 = 0.0
    else:
        s = math.sqrt(2.0 / n)
        x, y, z, w = x * s, y * s, z * s, w * s
        m00, m01, m02 = 1.0 - y * y - z * z, x * y - z * w, x * z + y * w
        m10, m11, m12 = x * y + z * w, 1.0 - x * x - z * z, y * z - x * w
        m22 = 1.0 - x * x - y * y
    cy = math.sqrt(m22 * m22 + m12 * m12)
    if cy > _EPS:
        return -math.atan2(m12, m22), -math.atan2(-m02, cy), -math.atan2(m01, m00)
    return -0.0, -math.atan2(-m02, cy), -math.atan2(-m10, m11)


def _quat_xyzw_from_euler_rxyz(ex, ey, ez):
    """Batched quaternion_from_euler(ex, ey, ez, 'rxyz') -> [..., 4] xyzw."""
    ai, aj, ak = ez / 2.0, -ey / 2.0, ex / 2.0
    ci, si, cj, sj, ck, sk = np.cos(ai), np.sin(ai), np.cos(aj), np.sin(aj), np.cos(ak), np.sin(ak)
    cc, cs, sc, ss = ci * ck, ci * sk, si * ck, si * sk
    return np.stack([cj * cs - sj * sc, -(cj * ss + sj * cc), cj * sc - sj * cs, cj * cc + sj * ss], axis=-1)


def _continuity_fix(joint, e, prev, q_xyzw, window):
    """Keep ball-joint Euler angles continuous and inside their limits (mocap_v2.py:196-222)."""
    lim = _BALL[joint]
    lo = np.maximum(lim[:, 0], np.asarray(prev) - window)
    hi = np.minimum(lim[:, 1], np.asarray(prev) + window)
    tgt = np.array([np.clip(e[a], lo[a], hi[a]) for a in range(3)])
    if np.allclose(list(e), tgt):
        return tuple(e)
    cand = [np.concatenate(([tgt[a], prev[a]], np.linspace(lo[a], hi[a], 6))) for a in range(3)]
    gx, gy, gz = np.meshgrid(cand[0], cand[1], cand[2], indexing="ij")
    qn = _quat_xyzw_from_euler_rxyz(gx, gy, gz)
    d_pos = np.sqrt(((qn - q_xyzw) ** 2).sum(-1))
    d_neg = np.sqrt(((-qn - q_xyzw) ** 2).sum(-1))
    err = np.minimum(d_pos, d_neg) ** 2
    i, j, k = np.unravel_index(int(np.argmin(err)), err.shape)  # first minimum = the loops' strict '<'
    return float(cand[0][i]), float(cand[1][j]), float(cand[2][k])


def calc_rot_vel(q0, q1, dt):
    """Body-frame angular velocity between two orientations (mocap_v2.py:350-362): axis * angle / dt of
    conj(q0) * q1, angle wrapped into (-pi, pi] (pyquaternion semantics)."""
    q0 = np.asarray(q0, dtype=np.float64)
    qd = _qmul(np.array([q0[0], -q0[1], -q0[2], -q0[3]]), np.asarray(q1, dtype=np.float64))
    ss = float(np.dot(qd, qd))
    if abs(1.0 - ss) >= 1e-14 and ss > 0:
        qd = qd / math.sqrt(ss)
    vn = float(np.linalg.norm(qd[1:]))
    axis = np.zeros(3) if vn < 1e-17 else qd[1:] / vn
    theta = 2.0 * math.atan2(vn, qd[0])
    ang = ((theta + math.pi) % (2 * math.pi)) - math.pi
    if ang == -math.pi:
        ang = math.pi
    return (ang / dt) * axis


class MocapDM(object):
    def __init__(self, robot, hotpath=None):
        self.robot = robot
        self.model = assets.load_model(robot)
        self._hotpath = hotpath  # provides fk_tables(); created lazily
        self.dt = None
        self.loop = None
        self.motion_name = None
        self.data_config = None
        self.data_vel = None
        self.data_body_xpos = None
        self.data_geom_xpos = None

    def get_length(self):
        return 0 if self.data_config is None else len(self.data_config)

    def get_qpos(self, idx):
        return self.data_config[idx]

    def get_qvel(self, idx):
        return self.data_vel[idx]

    def get_geom_xpos(self, idx):
        return self.data_geom_xpos[idx]

    def get_body_xpos(self, idx):
        return self.data_body_xpos[idx]

    # ------------------------------------------------------------------ loading
    def load_mocap(self, filepath):
        """`filepath` may be a DeepMimic JSON file or "<robot>_<motion>" / a MotionConfig.mocap_path
        whose basename names a clip of the asset pack."""
        if os.path.isfile(filepath):
            clip = assets.parse_clip_file(filepath)
        else:
            name = os.path.splitext(os.path.basename(filepath))[0]
            if not name.startswith(self.robot + "_"):
                raise FileNotFoundError(filepath)
            clip = assets.load_clip(self.robot, name[len(self.robot) + 1:])
        self.read_raw_data(clip)

    def load_motion(self, motion):
        self.read_raw_data(assets.load_clip(self.robot, motion))
        return self

    def read_raw_data(self, clip):
        frames = np.array(clip["frames"], dtype=np.float64)
        self.loop = clip.get("loop")
        if self.loop is None:
            print("Warning: No loop information found in mocap file")
        self.dt = float(frames[0][0])
        self.motion_name = clip["name"]
        if not clip.get("format"):
            cfg = self._convert_deepmimic(frames)
        else:  # "direct_qpos"
            cfg = frames[:, 1:].copy()
        vel = self._finite_difference(cfg)
        body_xpos, geom_xpos = self._fk(cfg, vel)
        cfg, vel, body_xpos, geom_xpos = self._resample(cfg, vel, body_xpos, geom_xpos)
        self.data_config = list(cfg)
        self.data_vel = list(vel)
        self.data_body_xpos = list(body_xpos)
        self.data_geom_xpos = list(geom_xpos)

    def _convert_deepmimic(self, frames):
        window = (5.0 if "getup" in self.motion_name else 10.0) * self.dt  # VMX * dt, mocap_v2.py:156-158,197
        cols, c = {}, 8
        for jn in BODY_JOINTS_IN_DP_ORDER:
            w = 1 if DOF_DEF[jn] == 1 else 4
            cols[jn] = (c, c + w)
            c += w
        out = np.zeros((len(frames), 7 + sum(DOF_DEF[j] for j in BODY_JOINTS)))
        prev_fixed = {}
        for k, row in enumerate(frames):
            out[k, 0:3] = align_position(row[1:4])
            out[k, 3:7] = align_rotation(row[4:8])
            c = 7
            for jn in BODY_JOINTS:
                a, b = cols[jn]
                if DOF_DEF[jn] == 1:
                    out[k, c] = row[a]
                    c += 1
                    continue
                q = align_rotation(row[a:b])
                e = _euler_rxyz(q)
                if jn in _BALL:
                    prev = e if k == 0 else prev_fixed[jn]
                    e = _continuity_fix(jn, e, prev, np.array([q[1], q[2], q[3], q[0]]), window)
                    prev_fixed[jn] = e
                out[k, c:c + 3] = e
                c += 3
        return out

    def _finite_difference(self, cfg):
        """mocap_v2.py:274-289; frame 0 differences itself."""
        vel = np.zeros((len(cfg), cfg.shape[1] - 1))
        for k in range(len(cfg)):
            a, b = cfg[max(k - 1, 0)], cfg[k]
            vel[k, 0:3] = (b[:3] - a[:3]) / self.dt
            vel[k, 3:6] = calc_rot_vel(a[3:7], b[3:7], self.dt)
            vel[k, 6:] = (b[7:] - a[7:]) / self.dt
        return vel

    def _fk(self, cfg, vel):
        if self._hotpath is None:
            from .hotpath import HotPath
            self._hotpath = HotPath.shared(self.robot)
        return self._hotpath.fk_tables(cfg, vel)

    def _resample(self, cfg, vel, body_xpos, geom_xpos):
        """x k linear blend down to the 60 Hz simulator step (mocap_v2.py:309-336); the last key frame
        is dropped and the root quaternion is blended component-wise without re-normalising."""
        ratio = self.dt / _SIM_DT
        k = int(ratio)
        if abs(ratio - k) > 0.1:
            raise Exception("Invalid dt ratio, cannot interpolate mocap frames: %f" % ratio)
        if k <= 1:
            return cfg, vel, body_xpos, geom_xpos
        B = (np.arange(k) * 1.0 / k)
        A = 1.0 - B

        def blend(x):
            a, b = x[:-1], x[1:]
            shape = (len(a), k) + (1,) * (x.ndim - 1)
            y = A.reshape((1, k) + (1,) * (x.ndim - 1)) * a[:, None] + B.reshape((1, k) + (1,) * (x.ndim - 1)) * b[:, None]
            return y.reshape((len(a) * k,) + x.shape[1:])

        self.dt = _SIM_DT
        return blend(cfg), blend(vel), blend(body_xpos), blend(geom_xpos)
