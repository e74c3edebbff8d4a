"""Per-environment step logic restated from the reference (oracle; test infrastructure).

One environment at a time, float64, plain Python/numpy -- the way the reference runs.
Restates reference src/deepmimic_env.py:33-173 (observation), :193-256 (imitation reward),
:258-270,335-510 (DPEnv step / reset), src/combined_env.py:21-99,202-244,246-505,529-533
(DPCombinedEnv, PlayerAction, MotionTransition) and src/config.py:3-37 (robot / motion
constants).  MuJoCo's part of a step is replaced by `force_state` semantics (reference
deepmimic_env.py:355-357): the pose is given and kinematics are recomputed from it
(oracle.mjkin.forward); mj_step itself is outside the hot path.

Contacts: the reference scans `mjdata.contact` and compares geom names
(deepmimic_env.py:88-101,113-120).  Here `contacts` is either such a list of
(geom1_id, geom2_id) pairs or an int bit mask -- bit 0 right foot & floor, bit 1 left
foot & floor, bits 2..9 the eight extra-contact geoms & floor -- which is the form the
CUDA path takes.
"""
import math
import random

import numpy as np

from . import mjkin
from .assets import load_clip, load_model
from .mocap_tables import MocapTables
from .rotations import to_rpy

# ------------------------------------------------------------------ constants (config.py)
ROBOTS = {
    "humanoid3d": dict(torso="chest", lfoot="left_ankle", rfoot="right_ankle", floor="floor", extra=None,
                       ee=["left_ankle", "right_ankle", "left_wrist", "right_wrist"], low_z=0.7),
    "unitree_g1": dict(torso="pelvis", lfoot="left_foot", rfoot="right_foot", floor="floor",
                       extra=["left_foot_lheel", "left_foot_rheel", "left_foot_ltoe", "left_foot_rtoe",
                              "right_foot_lheel", "right_foot_rheel", "right_foot_ltoe", "right_foot_rtoe"],
                       ee=["left_foot", "right_foot", "left_hand", "right_hand"], low_z=0.4),
}
ACYCLICAL_MOTIONS = FLOOR_MOTIONS = ("getup_faceup", "getup_facedown", "getup_facedown_slow",
                                     "getup_facedown_slow_FSI", "getup_facedown_towalk")
# deepmimic_env.py:206-207 (G1: root and hand joints excluded)
G1_QPOS_SEL = list(range(7, 25)) + list(range(32, 37))
G1_QVEL_SEL = list(range(6, 24)) + list(range(31, 36))

DONE_REASONS = ["", "low_z", "high_z", "run roll limit", "run pitch limit", "max_ep_len", "acyclical_end",
                "fallen without amnesty", "obs_bounds"]
R_NONE, R_LOW_Z, R_HIGH_Z, R_RUN_ROLL, R_RUN_PITCH, R_MAX_EP, R_ACYC_END, R_FALLEN, R_OBS_BOUNDS = range(9)


class ObsFlags:
    """DPEnvConfig (deepmimic_env.py:258-270) / DPCombinedEnvConfig (combined_env.py:21-35)."""

    def __init__(self, combined=False):
        self.MAX_EP_LENGTH = 2000 if combined else 1000
        self.VEL_OBS_SCALE = 0.1
        self.FRC_OBS_SCALE = 0.001
        self.ADD_FOOT_CONTACT_OBS = not combined
        self.ADD_EXTRA_CONTACT_OBS = combined
        self.ADD_TORSO_OBS = True
        self.ADD_JOINT_FORCE_OBS = False
        self.ADD_ABSPOS_OBS = False
        self.ADD_PHASE_OBS = True
        self.ADD_PLAYER_ACTION_OBS = combined
        self.MAX_PLAYER_ACTIONS = 3
        if combined:
            self.ACT_SCALE = 20.0
            self.AMNESTY_STEPS = 150


def contact_flags(model, rc, contacts):
    """(rfoot_floor, lfoot_floor, extra[8]) from a contact list or a bit mask."""
    n_extra = len(rc["extra"]) if rc["extra"] else 0
    if isinstance(contacts, (int, np.integer)):
        c = int(contacts)
        return float(c & 1), float((c >> 1) & 1), [float((c >> (2 + i)) & 1) for i in range(n_extra)]
    rf = lf = 0.0
    extra = [0.0] * n_extra
    for g1, g2 in contacts or []:
        n1, n2 = model.geom_id2name(g1), model.geom_id2name(g2)
        floor = n1 == rc["floor"] or n2 == rc["floor"]
        if floor and rc["rfoot"] in (n1, n2):
            rf = 1.0
        if floor and rc["lfoot"] in (n1, n2):
            lf = 1.0
        if floor:
            for i in range(n_extra):
                if rc["extra"][i] in (n1, n2):
                    extra[i] = 1.0
    return rf, lf, extra


# ------------------------------------------------------------------ observation
def observation(d, model, idx, motion_len, player_action, getup_state, cfg, rc):
    """get_obs (deepmimic_env.py:33-45); the joint-force block (:123-131, default off) needs MuJoCo's solver and is omitted."""
    S = cfg.VEL_OBS_SCALE
    parts = [d.qpos[7:].copy(), d.qvel[6:] * S]
    b = model.body_name2id(rc["torso"])
    if cfg.ADD_TORSO_OBS:  # :47-76
        w, x, y, z = d.xquat[b]
        roll, pitch, yaw = to_rpy(x, y, z, w)
        v = d.cvel[b][3:]
        c, s = math.cos(-yaw), math.sin(-yaw)
        vx = c * v[0] + (-s) * v[1] + 0 * v[2]
        vy = s * v[0] + c * v[1] + 0 * v[2]
        vz = 0 * v[0] + 0 * v[1] + 1 * v[2]
        parts.append(np.array([roll * S, pitch * S, vx * S, vy * S, vz * S,
                               d.cvel[b][0] * S, d.cvel[b][1] * S, d.cvel[b][2] * S]))
    rf, lf, extra = contact_flags(model, rc, d.contact)
    if cfg.ADD_FOOT_CONTACT_OBS:  # :78-105
        parts.append(np.array([rf, lf]))
    if cfg.ADD_EXTRA_CONTACT_OBS:  # :107-121
        parts.append(np.array(extra))
    if cfg.ADD_JOINT_FORCE_OBS:
        raise NotImplementedError("get_joint_force_obs (deepmimic_env.py:123-131): qfrc_unc / qfrc_constraint need MuJoCo's solver")
    if cfg.ADD_ABSPOS_OBS:  # :133-137
        parts.append(np.array(d.geom_xpos).flatten() * 1.0)
    if cfg.ADD_PHASE_OBS:  # :139-143
        parts.append(np.array([np.clip(1.0 * idx / motion_len, 0.0, 1.0)]))
    if cfg.ADD_PLAYER_ACTION_OBS:  # :145-173
        gs = np.zeros(2) if getup_state is None else np.asarray(getup_state, dtype=np.float64)
        if player_action is None:
            oh, head = np.zeros(cfg.MAX_PLAYER_ACTIONS), np.zeros(3)
        else:
            oh, head = player_action.onehot(cfg.MAX_PLAYER_ACTIONS), player_action.heading_in_world()
        w, x, y, z = d.xquat[b]
        yaw = to_rpy(x, y, z, w)[2]
        hx = head[0] * math.cos(-yaw) - head[1] * math.sin(-yaw)
        hy = head[0] * math.sin(-yaw) + head[1] * math.cos(-yaw)
        parts.append(np.concatenate(([hx, hy], oh, gs)))
    return np.concatenate(parts)


# ------------------------------------------------------------------ imitation reward
def imitation_reward(wp, wv, we, wc, wj, d, model, mocap, i, robot, rc, info):
    """calc_imitation_reward (deepmimic_env.py:193-256)."""
    if mocap is None:
        return 0.0, [0.0] * 6
    mass = model.body_mass[:, None]
    if robot == "unitree_g1":
        qs, vs = G1_QPOS_SEL, G1_QVEL_SEL
    else:
        qs, vs = list(range(7, model.nq)), list(range(6, model.nv))
    q_ref, v_ref = mocap.get_qpos(i), mocap.get_qvel(i)
    dq = np.abs(d.qpos[qs] - q_ref[qs])
    w, x, y, z = d.qpos[3:7]
    roll, pitch, _ = to_rpy(x, y, z, w)
    w, x, y, z = q_ref[3:7]
    roll_ref, pitch_ref, _ = to_rpy(x, y, z, w)
    r_pose = math.exp(-(np.sum(dq) + abs(pitch - pitch_ref)))
    r_vel = math.exp(-0.1 * np.sum(np.abs(v_ref[vs] - d.qvel[vs])))
    e_ee = 0.0
    g_ref = mocap.get_geom_xpos(i)
    for name in rc["ee"]:
        g = model.geom_name2id(name)
        e_ee += np.linalg.norm(d.geom_xpos[g] - g_ref[g]) ** 2
    r_ee = math.exp(-40 * e_ee)
    com_ref = np.sum(mocap.get_body_xpos(i) * mass, 0) / np.sum(mass)
    com = np.sum(d.xpos * mass, 0) / np.sum(mass)
    r_com = math.exp(-10 * np.linalg.norm(com_ref - com) ** 2)
    tol = model.jnt_range[1:] * 0.99
    jp = d.qpos[7:]
    if robot == "unitree_g1":
        sel = np.array(qs) - 7
        tol, jp = tol[sel], jp[sel]
    qlim = np.sum((jp <= tol[:, 0]) + (jp >= tol[:, 1])) / len(jp)
    reward = wp * r_pose + wv * r_vel + we * r_ee + wc * r_com + wj * qlim
    info["reward_config"] = r_pose
    info["reward_qvel"] = r_vel
    info["reward_end_eff"] = r_ee
    info["reward_com"] = r_com
    info["reward_joint_limit"] = qlim
    return reward, [mass, roll, roll_ref, pitch, pitch_ref, dq]


W_POSE, W_VEL, W_EE, W_COM, W_JLIM = 0.75, 0.1, 0.15, 0.0, -0.1  # deepmimic_env.py:400-404


def _com_height(d, mass):
    return (np.sum(mass * d.xipos, 0) / np.sum(mass))[2]


def _build_tables(robot, motion, model):
    frames, fmt, name, loop = load_clip(robot, motion)
    return MocapTables(model, robot).load(frames, fmt, name, loop)


_TABLE_CACHE = {}


def tables_for(robot, motion):
    key = (robot, motion)
    if key not in _TABLE_CACHE:
        _TABLE_CACHE[key] = _build_tables(robot, motion, load_model(robot))
    return _TABLE_CACHE[key]


# ------------------------------------------------------------------ DPEnv
class SingleClipEnv:
    """DPEnv (deepmimic_env.py:273-510) driven by force_state."""

    def __init__(self, motion=None, robot="humanoid3d", tables=None, add_abspos_obs=False):
        self.motion = "walk" if motion is None else motion
        self.robot = robot
        self.rc = ROBOTS[robot]
        self.model = load_model(robot)
        self.cfg = ObsFlags(False)
        self.cfg.ADD_ABSPOS_OBS = bool(add_abspos_obs)
        self.mocap = tables if tables is not None else tables_for(robot, self.motion)
        self.mocap_data_len = self.mocap.get_length()
        self.act_scale = 20.0 if robot == "unitree_g1" else 1.0  # :347-349
        self.n_hand = 14 if robot == "unitree_g1" else 0
        self.action_dim = self.model.nu - self.n_hand  # :304-307
        # state after gym's init-time step (Q4: idx_curr -1 -> 0, one episode step counted)
        self.idx_curr = 0
        self.episode_reward = 0
        self.episode_length = 1
        self.data = mjkin.forward(self.model, self._default_qpos(), np.zeros(self.model.nv))
        self.torque = np.zeros(self.model.nu)

    def _default_qpos(self):
        q = np.zeros(self.model.nq)
        q[:3] = self.model.body_pos[1]
        q[3] = 1.0
        return q

    def _obs(self):
        return observation(self.data, self.model, self.idx_curr, self.mocap_data_len, None, None, self.cfg, self.rc)

    def reset(self, idx_init=None):
        """reset + reset_model + reference_state_init (:312-316,496-510)."""
        self.episode_reward = 0
        self.episode_length = 0
        self.idx_init = random.randint(0, self.mocap_data_len - 1) if idx_init is None else idx_init
        self.idx_curr = self.idx_init
        self.data = mjkin.forward(self.model, self.mocap.data_config[self.idx_init], self.mocap.data_vel[self.idx_init])
        return self._obs()

    def step(self, action, force_state, contacts=0):
        info = {}
        ctrl = np.asarray(action, dtype=np.float64) * self.act_scale  # :347-352
        if len(ctrl) == self.model.nu - self.n_hand and self.n_hand:
            ctrl = np.concatenate((ctrl, np.zeros(self.n_hand)))
        assert len(ctrl) == self.model.nu
        self.torque = mjkin.motor_forces(self.model, ctrl)
        qpos, qvel = force_state  # :355-357
        self.data = mjkin.forward(self.model, qpos, qvel, contacts)
        obs = self._obs()  # :389 (pre-increment idx_curr)
        reward, inter = imitation_reward(W_POSE, W_VEL, W_EE, W_COM, W_JLIM, self.data, self.model, self.mocap,
                                         self.idx_curr, self.robot, self.rc, info)
        mass, roll, roll_ref, pitch, pitch_ref, _ = inter
        done = False  # :418-442
        if self.motion not in FLOOR_MOTIONS:
            z = _com_height(self.data, mass)
            done = bool(z < self.rc["low_z"] or z > 2.0)
            info["done_reason"] = "low_z" if z < self.rc["low_z"] else "high_z"
            info["z_com"] = z
        if self.motion == "run" and self.robot == "unitree_g1":
            lim = np.deg2rad(60.0)
            if abs(roll - roll_ref) > lim:
                done = True
                info["done_reason"] = "run roll limit"
            if abs(pitch - pitch_ref) > lim:
                done = True
                info["done_reason"] = "run pitch limit"
        if self.episode_length >= self.cfg.MAX_EP_LENGTH:
            done = True
            info["done_reason"] = "max_ep_len"
        if self.idx_curr + 1 == self.mocap_data_len and self.motion in ACYCLICAL_MOTIONS:
            done = True
            info["done_reason"] = "acyclical_end"
        self.idx_curr = (self.idx_curr + 1) % self.mocap_data_len  # :452-455
        self.episode_reward += reward
        self.episode_length += 1
        if np.max(obs) > 100.0 or np.min(obs) < -100.0:  # :465-476
            return self._obs() * 0.0, 0, True, {}
        return obs, reward, done, info


# ------------------------------------------------------------------ DPCombinedEnv
class PlayerAction:
    """combined_env.py:38-64."""
    IDXS = {"walk": 0, "run": 1, "action": 2}

    def __init__(self, name, vx, vy):
        self.name, self.vx, self.vy = name, vx, vy

    def onehot(self, n):
        v = np.zeros(n)
        v[PlayerAction.IDXS[self.name]] = 1.0
        return v

    def heading_in_world(self):
        n = np.linalg.norm([self.vx, self.vy])
        return np.array([self.vx, self.vy, 0]) / n if n != 0 else np.array([0, 0, 0])


def pa_walk():
    return PlayerAction("walk", 1.0, 0.0)


def pa_run():
    return PlayerAction("run", 3.0, 0.0)


class ConstantPoseClip:
    """MotionTransition / MTToGetup (combined_env.py:67-99): always frame 1 of the target clip."""

    def __init__(self, target, length, name):
        self.target, self.length, self.motion_name = target, length, name

    def get_qpos(self, i):
        return self.target.get_qpos(1)

    def get_qvel(self, i):
        return self.target.get_qvel(1)

    def get_body_xpos(self, i):
        return self.target.get_body_xpos(1)

    def get_geom_xpos(self, i):
        return self.target.get_geom_xpos(1)

    def get_length(self):
        return self.length


CLIP_WALK, CLIP_RUN, CLIP_GETUP, CLIP_TO_GETUP = range(4)


class MultiClipEnv:
    """DPCombinedEnv (combined_env.py:102-533) driven by force_state."""

    def __init__(self, add_abspos_obs=False):
        self.robot = "unitree_g1"
        self.rc = ROBOTS[self.robot]
        self.model = load_model(self.robot)
        self.cfg = ObsFlags(True)
        self.cfg.ADD_ABSPOS_OBS = bool(add_abspos_obs)
        self.walk_mocap = tables_for(self.robot, "walk")
        self.run_mocap = tables_for(self.robot, "run")
        self.getup_mocap = tables_for(self.robot, "getup_facedown_towalk")
        self.to_getup_mocap = ConstantPoseClip(self.getup_mocap, 180, "to_getup")
        self.clips = [self.walk_mocap, self.run_mocap, self.getup_mocap, self.to_getup_mocap]
        self.action_dim = self.model.nu - 14
        self.episode_reward = 0
        self.episode_length = 1  # init-time step of gym's MujocoEnv
        self.current_motion_n_steps = None
        self.current_motion_mocap = None
        self.current_player_action = None
        self.data = None
        self.torque = np.zeros(self.model.nu)

    @property
    def clip_id(self):
        return self.clips.index(self.current_motion_mocap)

    def get_current_motion_state(self):  # :202-206
        i = self.current_motion_n_steps % self.current_motion_mocap.get_length()
        return self.current_motion_mocap.get_qpos(i) * 1.0, self.current_motion_mocap.get_qvel(i) * 1.0

    def change_to_motion(self, clip):  # :529-533
        self.current_motion_mocap = clip
        self.current_motion_n_steps = 0

    def reset(self, rsi=True, coin=None, frame=None):
        """:208-244.  `coin` / `frame` inject the two random draws (parity tests)."""
        if rsi:
            c = random.randint(0, 1) if coin is None else coin
            if c == 0:
                self.current_motion_mocap = self.walk_mocap
                L = self.walk_mocap.get_length()
                r = random.randint(0, L - 1) if frame is None else frame
                self.current_motion_n_steps = self.cfg.AMNESTY_STEPS + 10 + r
            else:
                self.current_motion_mocap = self.getup_mocap
                L = self.getup_mocap.get_length()
                self.current_motion_n_steps = random.randint(0, L - 1) if frame is None else frame
        else:
            self.current_motion_mocap = self.getup_mocap
            self.current_motion_n_steps = 0
        self.current_player_action = pa_walk()
        self.episode_reward = 0
        self.episode_length = 0
        qpos, qvel = self.get_current_motion_state()
        self.data = mjkin.forward(self.model, qpos, qvel)
        return self._obs()

    def _obs(self):  # :495-505
        L = self.current_motion_mocap.get_length() if self.current_motion_mocap is not None else 1
        i = (self.current_motion_n_steps % L) if self.current_motion_n_steps is not None else 0
        if self.current_motion_mocap is self.getup_mocap:
            gs = [0, 1]
        elif self.current_motion_mocap is self.to_getup_mocap:
            gs = [1, 0]
        else:
            gs = [0, 0]
        return observation(self.data, self.model, i, L, self.current_player_action, gs, self.cfg, self.rc)

    def step(self, action, force_state, contacts=0):
        info = {}
        ctrl = np.asarray(action, dtype=np.float64) * self.cfg.ACT_SCALE  # :255-260
        if len(ctrl) == self.model.nu - 14:
            ctrl = np.concatenate((ctrl, np.zeros(14)))
        assert len(ctrl) == self.model.nu
        self.torque = mjkin.motor_forces(self.model, ctrl)
        qpos, qvel = force_state
        self.data = mjkin.forward(self.model, qpos, qvel, contacts)
        obs = self._obs() * 1.0  # :315
        cur = self.current_motion_mocap
        L = cur.get_length() if cur is not None else 1
        i = (self.current_motion_n_steps % L) if self.current_motion_n_steps is not None else 0
        imit, inter = imitation_reward(W_POSE, W_VEL, W_EE, W_COM, W_JLIM, self.data, self.model, cur, i,
                                       self.robot, self.rc, info)
        mass, roll, roll_ref, pitch, pitch_ref, dq = inter
        task = 0.0  # :339-352
        if cur is self.walk_mocap or cur is self.run_mocap:
            err = np.linalg.norm(cur.get_qvel(i)[:2] - self.data.qvel[:2])
            task = np.exp(-err * 10.0)
        if cur is self.to_getup_mocap:
            imit = 0.0
            cfg_err = np.sum(np.abs(dq)) + abs(pitch - pitch_ref) + abs(roll - roll_ref)
            task = np.exp(-cfg_err / 5.0) / 3.0
        reward = imit * 0.7 + task * 0.3
        info["imitation_reward"] = imit
        info["task_reward"] = task
        done = False
        if cur is not None:  # :398-445
            if self.current_motion_n_steps >= cur.get_length() - 1:
                if self.current_motion_mocap is self.getup_mocap:
                    # Q3: `current_player_action == PAWalk()` compares identities -> always run
                    self.change_to_motion(self.run_mocap)
                if self.current_motion_mocap is self.to_getup_mocap:
                    self.change_to_motion(self.getup_mocap)
            alim = np.deg2rad(15)
            ok = (abs(pitch - pitch_ref) < alim) and (abs(roll - roll_ref) < alim) and bool(np.all(np.abs(dq) < alim))
            if ok and self.current_motion_mocap is self.to_getup_mocap:
                self.change_to_motion(self.getup_mocap)
            if self.current_motion_mocap is self.walk_mocap or self.current_motion_mocap is self.run_mocap:
                z = _com_height(self.data, mass)
                info["z_com"] = z
                fallen = bool(z < self.rc["low_z"] or z > 2.0)
                lim = np.deg2rad(60.0)
                if abs(roll - roll_ref) > lim or abs(pitch - pitch_ref) > lim:
                    fallen = True
                if fallen:
                    if not self.current_motion_n_steps > self.cfg.AMNESTY_STEPS:
                        done = True
                        info["done_reason"] = "fallen without amnesty"
                    self.change_to_motion(self.to_getup_mocap)
            if self.episode_length >= self.cfg.MAX_EP_LENGTH:
                done = True
                info["done_reason"] = "max_ep_len"
        if self.current_motion_n_steps is not None:  # :456-462
            self.current_motion_n_steps += 1
        self.episode_reward += reward
        self.episode_length += 1
        if np.max(obs) > 100.0 or np.min(obs) < -100.0:  # :474-485
            return self._obs() * 0.0, 0, True, {}
        return obs, reward, done, info
