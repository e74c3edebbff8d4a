#!/usr/bin/env python3
"""Generate tests/golden/*.npz by running the REFERENCE'S OWN Python, unmodified.

Runs only in the build container (needs /root/reference).  The reference modules
  src/deepmimic_env.py  (DPEnv, get_obs, calc_imitation_reward)
  src/combined_env.py   (DPCombinedEnv and its motion state machine)
  src/mujoco/mocap_v2.py, src/mujoco/mocap_util.py (MocapDM clip loader)
  src/config.py, src/transformations.py
are imported from where they lie and driven through their public API
(`reset()`, `step(action, force_state=(qpos, qvel))`).  The third-party modules they
import and this image lacks (gym/mujoco_py, py3dtf, pyquaternion) are stood in for by
tools/ref_shims/ (kinematics = oracle.mjkin).  So the fixtures pin everything the
reference itself owns -- observation layout and scaling, reward terms and weights,
termination, counters, the obs-bounds guard, the combined env's state machine and
quirks, the loader's alignment / Euler / continuity fix / velocity / interpolation --
and leave MuJoCo, py3dtf and pyquaternion VALUES unpinned (see oracle/__init__.py).

All recorded inputs are rounded to float32 first, so the CUDA path (fp32) and the
float64 oracle see bit-identical inputs.

Usage: python tools/gen_golden.py            (writes tests/golden/*.npz)
"""
import builtins
import io
import os
import random
import sys

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("DM_REFERENCE", "/root/reference")
OUT = os.path.join(REPO, "tests", "golden")


def _setup():
    home = os.path.join(REPO, "build", "refhome")
    os.makedirs(os.path.join(home, "Code"), exist_ok=True)
    link = os.path.join(home, "Code", "DeepMimic_mujoco")
    if not os.path.islink(link):
        os.symlink(REF, link)
    os.environ["HOME"] = home  # config.py:26,38 hard-codes ~/Code/DeepMimic_mujoco/src
    sys.path.insert(0, REPO)
    sys.path.insert(0, os.path.join(REPO, "tools", "ref_shims"))
    sys.path.insert(0, os.path.join(REF, "src"))
    # the obs-bounds guard dumps a debug log to /tmp (deepmimic_env.py:468-473): swallow it
    real_open = builtins.open

    def guarded_open(path, *a, **k):
        if isinstance(path, str) and path.startswith("/tmp/deepmimic_episode_"):
            return io.StringIO()
        return real_open(path, *a, **k)

    builtins.open = guarded_open


_setup()
import transformations as T  # noqa: E402  (reference, vendored Gohlke)
from config import MotionConfig  # noqa: E402  (reference)
from mujoco.mocap_v2 import MocapDM  # noqa: E402  (reference)
import deepmimic_env as ref_dp  # noqa: E402  (reference)
import combined_env as ref_cb  # noqa: E402  (reference)

REASON_CODES = {"": 0, "low_z": 1, "high_z": 2, "run roll limit": 3, "run pitch limit": 4, "max_ep_len": 5,
                "acyclical_end": 6, "fallen without amnesty": 7}
TERM_KEYS = ["reward_config", "reward_qvel", "reward_end_eff", "reward_com", "reward_joint_limit",
             "imitation_reward", "task_reward"]


def f32(x):
    return np.asarray(x, dtype=np.float32).astype(np.float64)


def reason_code(info, empty_info_is_guard):
    if not info and empty_info_is_guard:
        return 8
    r = info.get("done_reason", "")
    for k, v in REASON_CODES.items():
        if k and r.startswith(k):
            return v
    return 0


def contact_pairs(model, robot, bits, rng):
    """bit mask -> a MuJoCo-like contact list (plus distractor pairs that must not set flags)."""
    pairs = []
    floor = model.geom_name2id("floor")
    if robot == "humanoid3d":
        r, l = model.geom_name2id("right_ankle"), model.geom_name2id("left_ankle")
        other = model.geom_name2id("right_knee")
    else:
        r, l = model.geom_name2id("right_foot"), model.geom_name2id("left_foot")
        other = model.geom_name2id("left_hand")
    if bits & 1:
        pairs.append((floor, r))
    if bits & 2:
        pairs.append((l, floor))
    if robot == "unitree_g1":
        names = ["left_foot_lheel", "left_foot_rheel", "left_foot_ltoe", "left_foot_rtoe",
                 "right_foot_lheel", "right_foot_rheel", "right_foot_ltoe", "right_foot_rtoe"]
        for i, n in enumerate(names):
            if bits & (4 << i):
                g = model.geom_name2id(n)
                pairs.append((floor, g) if i % 2 else (g, floor))
    pairs.append((r, other))  # foot touching something that is not the floor
    pairs.append((other, l))
    rng.shuffle(pairs)
    return [tuple(p) for p in pairs]


class Recorder:
    def __init__(self):
        self.rows = {}

    def add(self, **kw):
        for k, v in kw.items():
            self.rows.setdefault(k, []).append(v)

    def arrays(self, prefix):
        out = {}
        for k, v in self.rows.items():
            a = np.asarray(v)
            if k in ("qpos", "qvel", "action"):
                a = a.astype(np.float32)
            out[prefix + "/" + k] = a
        return out


def noisy_state(rng, mocap, idx, sq=0.05, sroot=0.02, sv=0.5):
    q = np.array(mocap.get_qpos(idx), dtype=np.float64)
    v = np.array(mocap.get_qvel(idx), dtype=np.float64)
    q[7:] += rng.normal(0, sq, q.size - 7)
    q[:3] += rng.normal(0, sroot, 3)
    q[3:7] += rng.normal(0, sroot, 4)  # left un-normalised on purpose (mj_kinematics normalises)
    v += rng.normal(0, sv, v.size)
    return f32(q), f32(v)


# ----------------------------------------------------------------------------- DPEnv
def run_dpenv(name, robot, motion, script, seed, act_hi):
    """script: list of dicts per step: optional reset=idx, optional edit(q, v) callable, optional set_eplen."""
    rng = np.random.default_rng(seed)
    env = ref_dp.DPEnv(motion=motion, robot=robot)
    rec = Recorder()
    na = env.action_space.shape[0]
    for st in script:
        reset_idx = st.get("reset", -1)
        reset_obs = np.full(env.observation_space.shape[0], np.nan)
        if reset_idx >= 0:
            real = random.randint
            random.randint = lambda a, b: reset_idx
            env.set_contacts([])  # sim.forward() in reset recomputes contacts; the reset pose has none here
            try:
                reset_obs = env.reset()
            finally:
                random.randint = real
        if "set_eplen" in st:
            env.episode_length = st["set_eplen"]
        idx_before, eplen_before = env.idx_curr, env.episode_length
        q, v = noisy_state(rng, env.mocap, env.idx_curr, *st.get("noise", (0.05, 0.02, 0.5)))
        if "edit" in st:
            st["edit"](q, v)
            q, v = f32(q), f32(v)
        a = f32(rng.uniform(-act_hi, act_hi, na))
        bits = int(rng.integers(0, 4))
        env.set_contacts(contact_pairs(env.model, robot, bits, rng))
        obs, rew, done, info = env.step(a, force_state=(q, v))
        rec.add(qpos=q, qvel=v, action=a, contact_bits=bits, reset_before=reset_idx, reset_obs=reset_obs,
                idx_before=idx_before, eplen_before=eplen_before,
                obs=np.asarray(obs, dtype=np.float64), reward=float(rew), done=int(bool(done)),
                terms=[float(info.get(k, np.nan)) for k in TERM_KEYS],
                reason=reason_code(info, True), idx_after=env.idx_curr, eplen_after=env.episode_length)
    out = rec.arrays(name)
    out[name + "/meta"] = np.asarray([robot, motion, "dpenv"])
    return out


def dpenv_scenarios():
    out = {}
    steps = [{"reset": 70}] + [{} for _ in range(199)]
    out.update(run_dpenv("h_walk_cycle", "humanoid3d", "walk", steps, 1, 2.0))

    def low(q, v):
        q[2] = 0.45

    def high(q, v):
        q[2] = 2.4

    def fast(q, v):
        v[10] = 1500.0

    def fast_neg(q, v):
        v[20] = -1200.0

    def at_limit(q, v):
        q[7:] = 3.2

    steps = [{"reset": 3}, {}, {"edit": low}, {"reset": 10}, {}, {"edit": high}, {"reset": 0}, {"edit": fast},
             {"reset": 75}, {}, {"edit": fast_neg}, {"reset": 40}, {"edit": at_limit}, {},
             {"set_eplen": 998}, {}, {}, {"reset": 5}, {}]
    out.update(run_dpenv("h_walk_events", "humanoid3d", "walk", steps, 2, 2.0))
    steps = [{"reset": 176}] + [{} for _ in range(8)] + [{"reset": 0}, {}, {}]
    out.update(run_dpenv("h_getup_acyclic", "humanoid3d", "getup_facedown", steps, 3, 2.0))
    for i, m in enumerate(["run", "spinkick", "dance_b", "kick"]):
        steps = [{"reset": 7 * (i + 1)}] + [{} for _ in range(29)]
        out.update(run_dpenv("h_" + m, "humanoid3d", m, steps, 10 + i, 2.0))

    def roll_over(q, v):
        q[3:7] = [0.7, 0.7, 0.0, 0.0]

    def pitch_over(q, v):
        q[3:7] = [0.75, 0.0, 0.66, 0.0]

    steps = [{"reset": 4}, {}, {"edit": roll_over}, {"reset": 20}, {"edit": pitch_over}, {"reset": 47}, {}, {}, {"edit": low}]
    out.update(run_dpenv("g_run_limits", "unitree_g1", "run", steps, 20, 4.0))
    steps = [{"reset": 60}] + [{} for _ in range(39)]
    out.update(run_dpenv("g_walk", "unitree_g1", "walk", steps, 21, 4.0))
    steps = [{"reset": 325}] + [{} for _ in range(9)]
    out.update(run_dpenv("g_getup_fsi", "unitree_g1", "getup_facedown_slow_FSI", steps, 22, 4.0))
    return out


# ----------------------------------------------------------------------------- DPCombinedEnv
def clip_code(env):
    m = env.current_motion_mocap
    return [env.walk_mocap, env.run_mocap, env.getup_mocap, env.to_getup_mocap].index(m)


def run_combined(name, script, seed):
    rng = np.random.default_rng(seed)
    env = ref_cb.DPCombinedEnv()
    rec = Recorder()
    na = env.action_space.shape[0]
    hi = env.action_space.high
    for st in script:
        reset = st.get("reset")  # (rsi, coin, frame)
        code = -1
        reset_obs = np.full(env.observation_space.shape[0], np.nan)
        if reset is not None:
            rsi, coin, frame = reset
            draws = [coin, frame]
            real = random.randint
            random.randint = lambda a, b: draws.pop(0)
            env.set_contacts([])
            try:
                reset_obs = env.reset(rsi=rsi)
            finally:
                random.randint = real
            code = (1 if rsi else 0) * 100000 + coin * 10000 + frame
        if "set_eplen" in st:
            env.episode_length = st["set_eplen"]
        clip_before, n_before, eplen_before = clip_code(env), env.current_motion_n_steps, env.episode_length
        mocap = env.current_motion_mocap
        i = env.current_motion_n_steps % mocap.get_length()
        q, v = noisy_state(rng, mocap, i, *st.get("noise", (0.03, 0.01, 0.3)))
        if "edit" in st:
            st["edit"](q, v)
            q, v = f32(q), f32(v)
        a = f32(rng.uniform(-hi, hi))
        bits = int(rng.integers(0, 256)) << 2
        env.set_contacts(contact_pairs(env.model, "unitree_g1", bits, rng))
        obs, rew, done, info = env.step(a, force_state=(q, v))
        rec.add(qpos=q, qvel=v, action=a, contact_bits=bits, reset_before=code, reset_obs=reset_obs,
                clip_before=clip_before, nsteps_before=n_before, eplen_before=eplen_before,
                obs=np.asarray(obs, dtype=np.float64), reward=float(rew), done=int(bool(done)),
                terms=[float(info.get(k, np.nan)) for k in TERM_KEYS],
                reason=reason_code(info, True), clip_after=clip_code(env), nsteps_after=env.current_motion_n_steps,
                eplen_after=env.episode_length)
    out = rec.arrays(name)
    out[name + "/meta"] = np.asarray(["unitree_g1", "combined", "combined"])
    return out


def combined_scenarios():
    out = {}

    def low(q, v):
        q[2] = 0.2

    def tilt(q, v):
        q[3:7] = [0.8, 0.0, 0.6, 0.0]

    def fast(q, v):
        v[8] = 1300.0

    far = (0.6, 0.05, 0.3)
    exact = (0.0, 0.0, 0.0)
    # getup clip runs out of time -> run (Q3), then a fall without amnesty ends the episode
    steps = [{"reset": (True, 1, 300)}] + [{} for _ in range(60)] + [{"edit": low}, {}, {}]
    out.update(run_combined("c_getup_to_run_fall", steps, 30))
    # walk with amnesty: fall -> to_getup without done; far poses; exact pose -> success -> getup
    steps = ([{"reset": (True, 0, 20)}] + [{} for _ in range(20)] + [{"edit": tilt}] + [{"noise": far} for _ in range(8)]
             + [{"noise": exact}] + [{} for _ in range(10)])
    out.update(run_combined("c_walk_amnesty_getup", steps, 31))
    # to_getup never succeeds: times out after 180 steps -> getup
    steps = [{"reset": (True, 0, 75)}, {}, {"edit": low}] + [{"noise": far} for _ in range(185)]
    out.update(run_combined("c_to_getup_timeout", steps, 32))
    # walk wraps its clip for a long time; max episode length; obs guard; rsi=False reset
    steps = ([{"reset": (True, 0, 0)}] + [{} for _ in range(90)] + [{"set_eplen": 1999}, {}, {"reset": (False, 0, 0)}, {}, {},
             {"edit": fast}, {"reset": (True, 1, 352)}, {}, {}])
    out.update(run_combined("c_misc", steps, 33))
    return out


# ----------------------------------------------------------------------------- loader and math
def abspos_goldens():
    """ENV_CFG.ADD_ABSPOS_OBS = True (deepmimic_env.py:133-137; off in every configuration the reference ships): the
    observation grows by geom_xpos.flatten() in front of the phase entry.  Short trajectories of both env classes with a
    reset, a state whose geoms leave the +-100 m box (the observation-bounds guard fires on the NEW block only) and a
    state that trips the guard through a standard entry."""
    out = {}

    def far(q, v):
        q[0] = 150.0   # root x: every geom is > 100 m from the origin, no standard entry is out of range

    def fast(q, v):
        v[8] = 1500.0  # a joint velocity: 0.1 * 1500 > 100 in the standard block

    ref_dp.DPEnv.ENV_CFG.ADD_ABSPOS_OBS = True
    ref_cb.DPCombinedEnv.ENV_CFG.ADD_ABSPOS_OBS = True
    try:
        steps = [{"reset": 12}] + [{} for _ in range(9)] + [{"edit": far}, {"reset": 3}, {}, {"edit": fast}, {"reset": 20}] + [{} for _ in range(5)]
        out.update(run_dpenv("h_walk_abspos", "humanoid3d", "walk", steps, 71, 1.0))
        out.update(run_dpenv("g1_walk_abspos", "unitree_g1", "walk", steps, 72, 0.5))
        csteps = [{"reset": (True, 0, 40)}] + [{} for _ in range(7)] + [{"edit": far}, {"reset": (True, 1, 10)}, {}, {}]
        out.update(run_combined("c_abspos", csteps, 73))
    finally:
        ref_dp.DPEnv.ENV_CFG.ADD_ABSPOS_OBS = False
        ref_cb.DPCombinedEnv.ENV_CFG.ADD_ABSPOS_OBS = False
    return out


def loader_goldens():
    out = {}
    clips = [("humanoid3d", "walk", 1), ("humanoid3d", "run", 1), ("humanoid3d", "spinkick", 1),
             ("humanoid3d", "dance_b", 3), ("humanoid3d", "getup_facedown", 1), ("humanoid3d", "kick", 2),
             ("humanoid3d", "cartwheel", 4), ("humanoid3d", "getup_faceup", 4), ("humanoid3d", "roll", 4),
             ("unitree_g1", "walk", 2), ("unitree_g1", "run", 2), ("unitree_g1", "getup_facedown_towalk", 8)]
    for robot, motion, stride in clips:
        m = MocapDM(robot=robot)
        m.load_mocap(MotionConfig(motion=motion, robot=robot).mocap_path)
        key = "loader/%s_%s" % (robot, motion)
        out[key + "/len_dt_stride"] = np.asarray([len(m.data_config), m.dt, stride], dtype=np.float64)
        out[key + "/data_config"] = np.asarray(m.data_config, dtype=np.float64)[::stride]
        out[key + "/data_vel"] = np.asarray(m.data_vel, dtype=np.float64)[::stride]
    for motion in ("backflip", "spin"):  # mocap_v2.py:315-316 must raise
        try:
            MocapDM(robot="humanoid3d").load_mocap(MotionConfig(motion=motion, robot="humanoid3d").mocap_path)
            out["loader/humanoid3d_%s/raises" % motion] = np.asarray(0)
        except Exception as e:  # noqa: BLE001
            out["loader/humanoid3d_%s/raises" % motion] = np.asarray(1)
            out["loader/humanoid3d_%s/message" % motion] = np.asarray(str(e))
    return out


def math_goldens():
    rng = np.random.default_rng(99)
    q = rng.normal(size=(96, 4))
    q[::2] /= np.linalg.norm(q[::2], axis=1, keepdims=True)
    e = rng.uniform(-3.1, 3.1, size=(96, 3))
    return {
        "math/quat_xyzw": q,
        "math/euler_rxyz": np.asarray([T.euler_from_quaternion(x, axes="rxyz") for x in q]),
        "math/euler_sxyz": np.asarray([T.euler_from_quaternion(x, axes="sxyz") for x in q]),
        "math/euler_in": e,
        "math/quat_from_euler_rxyz": np.asarray([T.quaternion_from_euler(a, b, c, axes="rxyz") for a, b, c in e]),
    }


def main():
    os.makedirs(OUT, exist_ok=True)
    import contextlib
    sink = io.StringIO()
    with contextlib.redirect_stdout(sink):  # the reference prints warnings / debug-log notices
        dp = dpenv_scenarios()
        cb = combined_scenarios()
        ld = loader_goldens()
        ab = abspos_goldens()
    mt = math_goldens()
    np.savez_compressed(os.path.join(OUT, "dpenv_steps.npz"), **dp)
    np.savez_compressed(os.path.join(OUT, "combined_steps.npz"), **cb)
    np.savez_compressed(os.path.join(OUT, "loader_tables.npz"), **ld)
    np.savez_compressed(os.path.join(OUT, "math_gohlke.npz"), **mt)
    np.savez_compressed(os.path.join(OUT, "abspos_steps.npz"), **ab)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))
    for k in sorted(dp):
        if k.endswith("/done"):
            print(k, int(dp[k].sum()), "dones of", dp[k].size, "reasons", sorted(set(dp[k[:-4] + "reason"].tolist())))
    for k in sorted(cb):
        if k.endswith("/clip_after"):
            print(k, "clips", cb[k].tolist()[:0], sorted(set(cb[k].tolist())), "dones", int(cb[k[:-10] + "done"].sum()))


if __name__ == "__main__":
    main()
