"""Shared helpers: replay a golden scenario (tests/golden/*.npz, made by tools/gen_golden.py
from the reference's own code) through the oracle."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TERM_KEYS = ["reward_config", "reward_qvel", "reward_end_eff", "reward_com", "reward_joint_limit",
             "imitation_reward", "task_reward"]
REASON_CODES = {"": 0, "low_z": 1, "high_z": 2, "run roll limit": 3, "run pitch limit": 4, "max_ep_len": 5,
                "acyclical_end": 6, "fallen without amnesty": 7}


def load(name):
    return np.load(os.path.join(GOLDEN, name))


def scenarios(z):
    return sorted({k.split("/")[0] for k in z.files if k.endswith("/meta")})


def reason_code(info):
    if not info:
        return 8
    return REASON_CODES[info.get("done_reason", "")]


def replay_dpenv_oracle(z, name):
    from oracle.envs import SingleClipEnv
    robot, motion, _ = [str(x) for x in z[name + "/meta"]]
    env = SingleClipEnv(motion, robot)
    T = z[name + "/qpos"].shape[0]
    out = {k: [] for k in ("obs", "reward", "done", "terms", "reason", "idx_after", "eplen_after", "reset_obs", "torque")}
    for t in range(T):
        r = int(z[name + "/reset_before"][t])
        ro = np.full(z[name + "/obs"].shape[1], np.nan)
        if r >= 0:
            ro = env.reset(idx_init=r)
        env.episode_length = int(z[name + "/eplen_before"][t])  # scenarios may poke the counter
        q = z[name + "/qpos"][t].astype(np.float64)
        v = z[name + "/qvel"][t].astype(np.float64)
        a = z[name + "/action"][t].astype(np.float64)
        obs, rew, done, info = env.step(a, (q, v), int(z[name + "/contact_bits"][t]))
        out["obs"].append(obs)
        out["reward"].append(float(rew))
        out["done"].append(int(done))
        out["terms"].append([float(info.get(k, np.nan)) for k in TERM_KEYS])
        out["reason"].append(reason_code(info))
        out["idx_after"].append(env.idx_curr)
        out["eplen_after"].append(env.episode_length)
        out["reset_obs"].append(ro)
        out["torque"].append(env.torque.copy())
    return {k: np.asarray(v) for k, v in out.items()}


def replay_combined_oracle(z, name):
    from oracle.envs import MultiClipEnv
    env = MultiClipEnv()
    T = z[name + "/qpos"].shape[0]
    out = {k: [] for k in ("obs", "reward", "done", "terms", "reason", "clip_after", "nsteps_after", "eplen_after",
                           "reset_obs", "torque")}
    for t in range(T):
        code = int(z[name + "/reset_before"][t])
        ro = np.full(98, np.nan)
        if code >= 0:
            ro = env.reset(rsi=bool(code // 100000), coin=(code // 10000) % 10, frame=code % 10000)
        env.episode_length = int(z[name + "/eplen_before"][t])
        q = z[name + "/qpos"][t].astype(np.float64)
        v = z[name + "/qvel"][t].astype(np.float64)
        a = z[name + "/action"][t].astype(np.float64)
        obs, rew, done, info = env.step(a, (q, v), int(z[name + "/contact_bits"][t]))
        out["obs"].append(obs)
        out["reward"].append(float(rew))
        out["done"].append(int(done))
        out["terms"].append([float(info.get(k, np.nan)) for k in TERM_KEYS])
        out["reason"].append(reason_code(info))
        out["clip_after"].append(env.clip_id)
        out["nsteps_after"].append(env.current_motion_n_steps)
        out["eplen_after"].append(env.episode_length)
        out["reset_obs"].append(ro)
        out["torque"].append(env.torque.copy())
    return {k: np.asarray(v) for k, v in out.items()}
