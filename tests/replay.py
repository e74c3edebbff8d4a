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


# --------------------------------------------------------------------------- fp32 back ends
def oracle_clip_structs(robot, motions, combined=False):
    """dm_clip_t array (+ keep-alive list) built from ORACLE tables -- for tests only."""
    import ctypes as C
    from deepmimic_mujoco_b200 import _native as N
    from oracle.envs import ACYCLICAL_MOTIONS, FLOOR_MOTIONS, tables_for
    keep, clips = [], (N.Clip * (len(motions) + (1 if combined else 0)))()
    for i, m in enumerate(motions):
        t = tables_for(robot, m)
        arrs = [np.ascontiguousarray(np.asarray(x, dtype=np.float64)) for x in
                (t.data_config, t.data_vel, t.data_body_xpos, t.data_geom_xpos)]
        keep.append(arrs)
        flags = 0
        if not combined:
            flags |= N.CLIP_FLOOR if m in FLOOR_MOTIONS else 0
            flags |= N.CLIP_ACYCLIC if m in ACYCLICAL_MOTIONS else 0
            flags |= N.CLIP_G1_RUN if (m == "run" and robot == "unitree_g1") else 0
        clips[i] = N.Clip(t.get_length(), t.get_length(), flags, -1, *[N.np_ptr(a) for a in arrs])
    if combined:  # MTToGetup: constant frame 1 of the getup clip, virtual length 180
        a = keep[2]
        clips[3] = N.Clip(180, len(a[0]), 0, 1, *[N.np_ptr(x) for x in a])
    return clips, keep


class HarnessBackend:
    """tests/host_harness: the kernel's per-env arithmetic compiled for the CPU (not the product)."""

    def __init__(self, robot, kind, motions, team=False):
        import ctypes as C
        import subprocess
        self.team = bool(team)
        d = os.path.join(os.path.dirname(os.path.abspath(__file__)), "host_harness")
        so = os.path.join(d, "libhostharness.so")
        srcs = [os.path.join(d, "harness.cpp")]
        csrc = os.path.join(os.path.dirname(d), "..", "deepmimic_mujoco_b200", "csrc")
        for root, _, files in os.walk(csrc):
            srcs += [os.path.join(root, f) for f in files if f.endswith((".cuh", ".h"))]
        if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
            subprocess.check_call(["g++", "-O1", "-std=c++17", "-fPIC", "-shared", "-w", "-o", so, srcs[0]])
        self.L = C.CDLL(so)
        self.robot_id = {"humanoid3d": 0, "unitree_g1": 1}[robot]
        self.kind = kind
        self.clips, self._keep = oracle_clip_structs(robot, motions, combined=(kind == 1))
        self.obs_dim = self.L.hh_obs_dim(self.robot_id, kind)
        self.nu = 28 if robot == "humanoid3d" else 37

    def step(self, qpos, qvel, action, contact, idx, ep_len, clip):
        from deepmimic_mujoco_b200._native import np_ptr
        n = qpos.shape[0]
        f = lambda a: np.ascontiguousarray(a, dtype=np.float32)
        qpos, qvel, action = f(qpos), f(qvel), f(action)
        contact = np.ascontiguousarray(contact, dtype=np.uint32)
        idx, ep_len, clip = (np.ascontiguousarray(x, dtype=np.int32).copy() for x in (idx, ep_len, clip))
        obs = np.zeros((n, self.obs_dim), np.float32)
        reward = np.zeros(n, np.float32)
        done = np.zeros(n, np.uint8)
        reason = np.zeros(n, np.uint8)
        torque = np.zeros((n, self.nu), np.float32)
        terms = np.zeros((n, 8), np.float32)
        self.L.hh_set_team(1 if self.team else 0)
        rc = self.L.hh_step(self.robot_id, self.kind, len(self.clips), self.clips, n, np_ptr(qpos), np_ptr(qvel),
                            np_ptr(action), np_ptr(contact), np_ptr(idx), np_ptr(ep_len), np_ptr(clip), np_ptr(obs),
                            np_ptr(reward), np_ptr(done), np_ptr(reason), np_ptr(torque), np_ptr(terms))
        assert rc == 0
        return dict(obs=obs, reward=reward, done=done, reason=reason, torque=torque, terms=terms, idx=idx,
                    ep_len=ep_len, clip=clip)


def scenario_motions(z, name):
    robot, motion, kind = [str(x) for x in z[name + "/meta"]]
    if kind == "combined":
        return robot, 1, ["walk", "run", "getup_facedown_towalk"]
    return robot, 0, [motion]


def replay_batched(z, name, backend):
    """All T recorded steps as ONE batch of T independent environments: the recorded pre-step
    counters are the per-env state, so a batched back end must reproduce every row."""
    robot, kind, _ = scenario_motions(z, name)
    T = z[name + "/qpos"].shape[0]
    if kind == 0:
        idx, clip = z[name + "/idx_before"], np.zeros(T, np.int32)
    else:
        idx, clip = z[name + "/nsteps_before"], z[name + "/clip_before"]
    return backend.step(z[name + "/qpos"], z[name + "/qvel"], z[name + "/action"], z[name + "/contact_bits"],
                        idx, z[name + "/eplen_before"], clip)


def assert_matches_golden(z, name, got, rtol=1e-5, floor=0.05):
    """fp32 path vs the float64 reference outputs: |a-b| <= rtol * max(|b|, floor) for floats
    (BASELINE.json: 1e-5 relative in fp32); integers and flags bit-exact."""
    robot, kind, _ = scenario_motions(z, name)

    def close(a, b, what, rt=rtol):
        a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
        bad = np.abs(a - b) > rt * np.maximum(np.abs(b), floor)
        bad &= ~(np.isnan(a) & np.isnan(b))
        assert not bad.any(), "%s %s: %d mismatches, worst |d|=%g at %s (got %r want %r)" % (
            name, what, bad.sum(), np.abs(a - b)[bad].max(), np.argwhere(bad)[0], a[bad][0], b[bad][0])

    close(got["obs"], z[name + "/obs"], "obs")
    close(got["reward"], z[name + "/reward"], "reward")
    np.testing.assert_array_equal(got["done"], z[name + "/done"], err_msg=name + " done")
    np.testing.assert_array_equal(got["reason"], z[name + "/reason"], err_msg=name + " reason")
    np.testing.assert_array_equal(got["ep_len"], z[name + "/eplen_after"], err_msg=name + " ep_len")
    if kind == 0:
        np.testing.assert_array_equal(got["idx"], z[name + "/idx_after"], err_msg=name + " idx")
    else:
        np.testing.assert_array_equal(got["idx"], z[name + "/nsteps_after"], err_msg=name + " n_steps")
        np.testing.assert_array_equal(got["clip"], z[name + "/clip_after"], err_msg=name + " clip")
    want_terms = z[name + "/terms"]
    guard = z[name + "/reason"] == 8
    nt = 5 if kind == 0 else 7
    close(got["terms"][~guard][:, :nt], want_terms[~guard][:, :nt], "terms")


class GpuBackend:
    """The product path: deepmimic_mujoco_b200.HotPath -> libdmhot.so -> CUDA kernels, with clip tables
    built by the PRODUCT loader (MocapDM + device FK), not by the oracle."""

    def __init__(self, robot, kind, motions, host=False, soa=False, variant=-1):
        import torch
        from deepmimic_mujoco_b200.hotpath import HotPath
        self.torch = torch
        self.hp = HotPath(robot, combined=(kind == 1))
        if kind == 1:
            self.hp.load_combined()
        else:
            self.hp.load_motions(motions)
        self.hp.set_kernel_variant(variant)
        self.kind, self.host, self.soa = kind, host, soa
        self.obs_dim, self.nu = self.hp.obs_dim, self.hp.nu

    def step(self, qpos, qvel, action, contact, idx, ep_len, clip):
        n = qpos.shape[0]
        f = lambda a: np.ascontiguousarray(a, dtype=np.float32)
        host = dict(qpos=f(qpos), qvel=f(qvel), action=f(action), contact=np.ascontiguousarray(contact, dtype=np.uint32),
                    idx=np.ascontiguousarray(idx, dtype=np.int32).copy(), ep_len=np.ascontiguousarray(ep_len, dtype=np.int32).copy(),
                    clip=np.ascontiguousarray(clip, dtype=np.int32).copy(),
                    obs=np.zeros((n, self.obs_dim), np.float32), reward=np.zeros(n, np.float32), done=np.zeros(n, np.uint8),
                    reason=np.zeros(n, np.uint8), torque=np.zeros((n, self.nu), np.float32), terms=np.zeros((n, 8), np.float32),
                    ep_return=np.zeros(n, np.float32))
        order = ("qpos", "qvel", "action", "contact", "idx", "ep_len", "clip", "obs", "reward", "done", "reason", "torque",
                 "terms", "ep_return")
        if self.host:
            self.hp.step_host(n, *[host[k] for k in order])
            return host
        torch = self.torch
        dev = {}
        for k in order:
            a = host[k]
            if a.dtype == np.uint32:
                t = torch.from_numpy(a.view(np.int32)).cuda()
            else:
                t = torch.from_numpy(a).cuda()
            if self.soa and a.ndim == 2:
                t = t.t().contiguous()  # [dim, n]
            dev[k] = t
        self.hp.step(n, *[dev[k] for k in order], soa_ld=(n if self.soa else 0))
        torch.cuda.synchronize()
        out = {}
        for k in order:
            t = dev[k]
            if self.soa and host[k].ndim == 2:
                t = t.t().contiguous()
            out[k] = t.cpu().numpy()
        return out
