"""Rotation arithmetic the reference borrows from third-party packages (oracle; test infrastructure).

MuJoCo helpers use (w, x, y, z).  `pq_*` restate pyquaternion 0.9.9 (used by the
reference at src/mujoco/mocap_util.py:31-40 and src/mujoco/mocap_v2.py:350-362);
`to_rpy` restates py3dtf.Quaternion(x,y,z,w).to_rpy() (used at
src/deepmimic_env.py:56,163,217,219); `gohlke_*` restate the two functions of the
vendored src/transformations.py the loader calls (mocap_v2.py:142,165,216).
"""
import math

import numpy as np


# ---------------------------------------------------------------- MuJoCo (mju_*)
def quat_mul(a, b):
    """Hamilton product, wxyz (mju_mulQuat)."""
    aw, ax, ay, az = a
    bw, bx, by, bz = b
    return np.array([
        aw * bw - ax * bx - ay * by - az * bz,
        aw * bx + ax * bw + ay * bz - az * by,
        aw * by - ax * bz + ay * bw + az * bx,
        aw * bz + ax * by - ay * bx + az * bw,
    ])


def quat_normalize(q):
    """mju_normalize4: unit quaternion; a (near-)zero quaternion becomes identity."""
    q = np.asarray(q, dtype=np.float64)
    n = math.sqrt(float(np.dot(q, q)))
    if n < 1e-15:  # mjMINVAL
        return np.array([1.0, 0.0, 0.0, 0.0])
    return q / n


def quat_to_mat(q):
    """mju_quat2Mat: 3x3 rotation matrix of a unit wxyz quaternion."""
    w, x, y, z = q
    return np.array([
        [w * w + x * x - y * y - z * z, 2 * (x * y - w * z), 2 * (x * z + w * y)],
        [2 * (x * y + w * z), w * w - x * x + y * y - z * z, 2 * (y * z - w * x)],
        [2 * (x * z - w * y), 2 * (y * z + w * x), w * w - x * x - y * y + z * z],
    ])


def rot_vec_quat(v, q):
    """mju_rotVecQuat."""
    return quat_to_mat(q) @ np.asarray(v, dtype=np.float64)


def axis_angle_to_quat(axis, angle):
    """mju_axisAngle2Quat (axis is unit length)."""
    s = math.sin(0.5 * angle)
    return np.array([math.cos(0.5 * angle), axis[0] * s, axis[1] * s, axis[2] * s])


# ---------------------------------------------------------------- pyquaternion 0.9.9
def pq_from_matrix(m):
    """Quaternion(matrix=m): trace method on the transposed matrix; returns wxyz."""
    m = np.asarray(m, dtype=np.float64).T
    if m[2, 2] < 0:
        if m[0, 0] > m[1, 1]:
            t = 1 + m[0, 0] - m[1, 1] - m[2, 2]
            q = [m[1, 2] - m[2, 1], t, m[0, 1] + m[1, 0], m[2, 0] + m[0, 2]]
        else:
            t = 1 - m[0, 0] + m[1, 1] - m[2, 2]
            q = [m[2, 0] - m[0, 2], m[0, 1] + m[1, 0], t, m[1, 2] + m[2, 1]]
    else:
        if m[0, 0] < -m[1, 1]:
            t = 1 - m[0, 0] - m[1, 1] + m[2, 2]
            q = [m[0, 1] - m[1, 0], m[2, 0] + m[0, 2], m[1, 2] + m[2, 1], t]
        else:
            t = 1 + m[0, 0] + m[1, 1] + m[2, 2]
            q = [t, m[1, 2] - m[2, 1], m[2, 0] - m[0, 2], m[0, 1] - m[1, 0]]
    return np.array(q, dtype=np.float64) * (0.5 / math.sqrt(t))


def pq_mul(a, b):
    """Quaternion.__mul__ (Hamilton, wxyz, no normalisation)."""
    return quat_mul(a, b)


def pq_conj(q):
    return np.array([q[0], -q[1], -q[2], -q[3]])


def _pq_unit(q):
    # Quaternion._normalise(): rescale unless |1 - |q|^2| < 1e-14
    q = np.asarray(q, dtype=np.float64)
    ss = float(np.dot(q, q))
    if abs(1.0 - ss) < 1e-14:
        return q
    n = math.sqrt(ss)
    return q / n if n > 0 else q


def pq_axis(q):
    """Quaternion.axis: unit rotation axis, zeros when the vector part vanishes (< 1e-17)."""
    q = _pq_unit(q)
    n = float(np.linalg.norm(q[1:]))
    if n < 1e-17:
        return np.zeros(3)
    return q[1:] / n


def pq_angle(q):
    """Quaternion.angle: 2*atan2(|v|, w) wrapped into (-pi, pi]."""
    q = _pq_unit(q)
    theta = 2.0 * math.atan2(float(np.linalg.norm(q[1:])), q[0])
    r = ((theta + math.pi) % (2 * math.pi)) - math.pi
    return math.pi if r == -math.pi else r


# ---------------------------------------------------------------- py3dtf
def to_rpy(x, y, z, w):
    """py3dtf.Quaternion(x, y, z, w).to_rpy() -> (roll, pitch, yaw), static-xyz convention.

    PARITY UNPINNED: py3dtf is not installed and not vendored.  Restated as the
    standard roll-pitch-yaw of the NORMALISED quaternion; for unit quaternions this
    equals transformations.euler_from_quaternion([x,y,z,w], 'sxyz') away from gimbal
    lock (checked in tests against the vendored transformations.py fixtures).
    """
    n = math.sqrt(x * x + y * y + z * z + w * w)
    if n > 0:
        x, y, z, w = x / n, y / n, z / n, w / n
    roll = math.atan2(2 * (w * x + y * z), 1 - 2 * (x * x + y * y))
    sp = 2 * (w * y - z * x)
    pitch = math.asin(max(-1.0, min(1.0, sp)))
    yaw = math.atan2(2 * (w * z + x * y), 1 - 2 * (y * y + z * z))
    return roll, pitch, yaw


# ---------------------------------------------------------------- Gohlke transformations.py
_EPS = np.finfo(float).eps * 4.0


def _gohlke_matrix(q_xyzw):
    """quaternion_matrix (transformations.py:1174-1193): xyzw in, 3x3 out, self-normalising."""
    q = np.array(q_xyzw[:4], dtype=np.float64)
    nq = float(np.dot(q, q))
    if nq < _EPS:
        return np.identity(3)
    q = q * math.sqrt(2.0 / nq)
    o = np.outer(q, q)
    return np.array([
        [1.0 - o[1, 1] - o[2, 2], o[0, 1] - o[2, 3], o[0, 2] + o[1, 3]],
        [o[0, 1] + o[2, 3], 1.0 - o[0, 0] - o[2, 2], o[1, 2] - o[0, 3]],
        [o[0, 2] - o[1, 3], o[1, 2] + o[0, 3], 1.0 - o[0, 0] - o[1, 1]],
    ])


def gohlke_euler_rxyz(q_xyzw):
    """euler_from_quaternion(q, axes='rxyz') (transformations.py:1031-1097 with
    firstaxis=2, parity=1, repetition=0, frame=1): angles (ex, ey, ez) with R = Rx Ry Rz."""
    M = _gohlke_matrix(q_xyzw)
    cy = math.sqrt(M[2, 2] * M[2, 2] + M[1, 2] * M[1, 2])
    if cy > _EPS:
        ax = math.atan2(M[0, 1], M[0, 0])
        ay = math.atan2(-M[0, 2], cy)
        az = math.atan2(M[1, 2], M[2, 2])
    else:
        ax = math.atan2(-M[1, 0], M[1, 1])
        ay = math.atan2(-M[0, 2], cy)
        az = 0.0
    ax, ay, az = -ax, -ay, -az  # parity
    return az, ay, ax  # frame: swap first and last


def gohlke_euler_sxyz(q_xyzw):
    """euler_from_quaternion(q, axes='sxyz'): static roll, pitch, yaw."""
    M = _gohlke_matrix(q_xyzw)
    cy = math.sqrt(M[0, 0] * M[0, 0] + M[1, 0] * M[1, 0])
    if cy > _EPS:
        return math.atan2(M[2, 1], M[2, 2]), math.atan2(-M[2, 0], cy), math.atan2(M[1, 0], M[0, 0])
    return math.atan2(-M[1, 2], M[1, 1]), math.atan2(-M[2, 0], cy), 0.0


def gohlke_quat_from_euler_rxyz(ex, ey, ez):
    """quaternion_from_euler(ex, ey, ez, axes='rxyz') -> xyzw (transformations.py:1100-1154)."""
    # i=2, j=1, k=0; frame swaps (ai, ak); parity negates aj
    ai, aj, ak = ez / 2.0, -ey / 2.0, ex / 2.0
    ci, si = math.cos(ai), math.sin(ai)
    cj, sj = math.cos(aj), math.sin(aj)
    ck, sk = math.cos(ak), math.sin(ak)
    cc, cs, sc, ss = ci * ck, ci * sk, si * ck, si * sk
    q = np.empty(4)
    q[2] = cj * sc - sj * cs
    q[1] = -(cj * ss + sj * cc)
    q[0] = cj * cs - sj * sc
    q[3] = cj * cc + sj * ss
    return q
