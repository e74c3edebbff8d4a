"""Reader / writer of the reference's `episode_debug_log` JSON -- the only on-disk format for recorded
states the reference has (written by src/deepmimic_env.py:458-476 and :366-378, read by src/check_debug_log.py:9-28):

    {"action": [[..na..], ...], "body_xpos": [[[x,y,z] * nbody], ...], "body_xvelp": [[[..3..] * nbody], ...],
     "qpos": [[..nq..], ...], "qvel": [[..nv..], ...], "reward": [r, ...], "motion": str, "robot": str,
     "full_traceback": str (optional), "last_action": [...] (optional)}

one list entry per control step.  The log does not store the mocap index: check_debug_log.py replays from
`reset_model(idx_init=0)`, so step t of a log sits at frame `t % mocap_len`; `DebugLog.idx_init` carries a
different start for logs produced here.
"""
from __future__ import annotations

import json

import numpy as np

STEP_KEYS = ("action", "body_xpos", "body_xvelp", "qpos", "qvel", "reward")


class DebugLog:
    def __init__(self, motion, robot, idx_init=0):
        self.motion, self.robot, self.idx_init = motion, robot, int(idx_init)
        self.steps = {k: [] for k in STEP_KEYS}
        self.extra = {}

    def __len__(self):
        return len(self.steps["qpos"])

    def append(self, action, body_xpos, body_xvelp, qpos, qvel, reward):
        """Same six appends as deepmimic_env.py:458-463."""
        for k, v in zip(STEP_KEYS, (action, body_xpos, body_xvelp, qpos, qvel, reward)):
            self.steps[k].append(float(v) if k == "reward" else np.asarray(v, dtype=np.float64).tolist())

    def arrays(self):
        """float64 arrays: action [T,na], body_xpos [T,nbody,3], body_xvelp [T,nbody,3], qpos [T,nq], qvel [T,nv], reward [T]."""
        return {k: np.asarray(self.steps[k], dtype=np.float64) for k in STEP_KEYS}

    def frame_indices(self, mocap_len):
        return (self.idx_init + np.arange(len(self))) % int(mocap_len)

    def to_json(self, indent=None):
        d = dict(self.steps)
        d.update(self.extra)
        d["motion"], d["robot"] = self.motion, self.robot
        if self.idx_init:
            d["idx_init"] = self.idx_init
        return json.dumps(d, indent=indent)


def write_debug_log(path, log, indent=None):
    with open(path, "w") as f:
        f.write(log.to_json(indent))


def read_debug_log(path):
    with open(path) as f:
        d = json.load(f)
    for k in ("qpos", "qvel", "motion", "robot"):
        if k not in d:
            raise ValueError("%s: not an episode_debug_log (missing %r)" % (path, k))
    log = DebugLog(d["motion"], d["robot"], d.get("idx_init", 0))
    T = len(d["qpos"])
    for k in STEP_KEYS:
        v = d.get(k, [])
        if len(v) not in (0, T):  # the step that raised inside do_simulation has no entries of its own
            raise ValueError("%s: %r has %d entries, qpos has %d" % (path, k, len(v), T))
        log.steps[k] = v
    log.extra = {k: d[k] for k in d if k not in STEP_KEYS + ("motion", "robot", "idx_init")}
    return log


def integrate_qpos(qpos, qvel, eps):
    """qpos (+) eps * qvel for a free joint followed by hinges (MuJoCo mj_integratePos: the free-joint angular
    velocity is in the body frame).  float64, batched over the leading axis."""
    q = np.array(qpos, dtype=np.float64, copy=True)
    v = np.asarray(qvel, dtype=np.float64)
    q[..., :3] += eps * v[..., :3]
    w = v[..., 3:6] * eps
    ang = np.linalg.norm(w, axis=-1, keepdims=True)
    half = 0.5 * ang
    k = np.where(ang > 1e-12, np.sin(half) / np.maximum(ang, 1e-300), 0.5)
    dq = np.concatenate([np.cos(half), k * w], axis=-1)
    a, b = q[..., 3:7].copy(), dq
    q[..., 3] = a[..., 0] * b[..., 0] - a[..., 1] * b[..., 1] - a[..., 2] * b[..., 2] - a[..., 3] * b[..., 3]
    q[..., 4] = a[..., 0] * b[..., 1] + a[..., 1] * b[..., 0] + a[..., 2] * b[..., 3] - a[..., 3] * b[..., 2]
    q[..., 5] = a[..., 0] * b[..., 2] - a[..., 1] * b[..., 3] + a[..., 2] * b[..., 0] + a[..., 3] * b[..., 1]
    q[..., 6] = a[..., 0] * b[..., 3] + a[..., 1] * b[..., 2] - a[..., 2] * b[..., 1] + a[..., 3] * b[..., 0]
    q[..., 7:] += eps * v[..., 6:]
    return q


def body_kinematics(hotpath, qpos, qvel, eps=1e-6):
    """body_xpos [T,nbody,3] and body_xvelp [T,nbody,3] (MuJoCo data.body_xpos / data.body_xvelp) of recorded
    states: float64 FK on the device (dm_fk_tables), velocities as the central difference of the FK along qvel."""
    q = np.asarray(qpos, dtype=np.float64)
    v = np.asarray(qvel, dtype=np.float64)
    T = len(q)
    qs = np.concatenate([q, integrate_qpos(q, v, eps), integrate_qpos(q, v, -eps)])
    bx, _ = hotpath.fk_tables(qs, np.concatenate([v, v, v]))
    return bx[:T], (bx[T:2 * T] - bx[2 * T:]) / (2 * eps)


class EpisodeRecorder:
    """Collects the debug log of ONE environment of a DPVecEnv (what each reference env process keeps in
    self.episode_debug_log); body_xpos / body_xvelp are filled in by finish() in one batched FK launch."""

    def __init__(self, env, index=0):
        self.env, self.i = env, int(index)
        self.log = DebugLog(env.motions[int(env.clip[index])] if not env.combined else "combined", env.robot, int(env.idx[index]))
        self._rows = []

    def record(self, actions, rewards):
        e, i = self.env, self.i
        self._rows.append((np.array(actions[i], dtype=np.float64), e.qpos[i].astype(np.float64), e.qvel[i].astype(np.float64),
                           float(rewards[i])))

    def finish(self):
        if self._rows:
            q = np.stack([r[1] for r in self._rows])
            v = np.stack([r[2] for r in self._rows])
            bx, bv = body_kinematics(self.env.hp, q, v)
            for t, (a, qq, vv, r) in enumerate(self._rows):
                self.log.append(a, bx[t], bv[t], qq, vv, r)
            self._rows = []
        return self.log
