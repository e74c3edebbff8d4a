"""DPCombinedEnv: single-environment view with the reference's signature
(src/combined_env.py:102-533), backed by the batched CUDA path with N = 1."""
from types import SimpleNamespace

import numpy as np

from . import _native as N
from .assets import load_model
from .config import DPCombinedEnvConfig, RobotConfig
from .vec_env import DPVecEnv

MOTION_NAMES = ["walk", "run", "getup", "to_getup"]


class DPCombinedEnv:
    version = "v0.2.up"
    ENV_CFG = DPCombinedEnvConfig()

    def __init__(self, verbose=0, _profile=False, device=0, physics=None):
        self.PROFILE = _profile
        self.verbose = verbose
        self.robot = "unitree_g1"
        self.robot_config = RobotConfig(robot=self.robot)
        self.model = load_model(self.robot)
        if self.ENV_CFG.ADD_JOINT_FORCE_OBS:
            raise NotImplementedError("ADD_JOINT_FORCE_OBS (deepmimic_env.py:123-131) needs qfrc_unc / qfrc_constraint of MuJoCo's solver, "
                                      "which is outside this package")
        self._vec = DPVecEnv(1, combined=True, device=device, physics=physics, info_terms=True,
                             add_abspos_obs=self.ENV_CFG.ADD_ABSPOS_OBS)
        self.walk_mocap, self.run_mocap, self.getup_mocap = self._vec.hp.mocaps
        self.action_space = self._vec.action_space
        self.observation_space = self._vec.observation_space
        self.sim = SimpleNamespace(data=SimpleNamespace(qpos=np.zeros(self.model.nq), qvel=np.zeros(self.model.nv), time=0.0))
        self.episode_reward = 0
        self.episode_length = 1
        self.current_motion_n_steps = None
        self.current_motion = None

    def _pull(self):
        v = self._vec
        self.current_motion_n_steps = int(v.idx[0])
        self.current_motion = MOTION_NAMES[int(v.clip[0])]
        self.episode_length = int(v.ep_len[0])
        self.episode_reward = float(v.ep_return[0])
        self.sim.data.qpos = v.qpos[0].astype(np.float64)
        self.sim.data.qvel = v.qvel[0].astype(np.float64)

    def get_current_motion_state(self):
        """combined_env.py:202-206"""
        c = int(self._vec.clip[0])
        m = [self.walk_mocap, self.run_mocap, self.getup_mocap, self.getup_mocap][c]
        i = 1 if c == 3 else self.current_motion_n_steps % m.get_length()
        return m.get_qpos(i) * 1.0, m.get_qvel(i) * 1.0

    def reset(self, rsi=True, coin=None, frame=None):
        """`coin` / `frame` inject the two random.randint draws of combined_env.py:223-231."""
        if not rsi:
            coin, frame = 2, 0
        obs = self._vec.reset(frame=None if frame is None else [frame], coin=None if coin is None else [coin])
        self._pull()
        return obs[0].astype(np.float64)

    def step(self, action, force_state=None):
        v = self._vec
        a = np.asarray(action, dtype=np.float32)
        assert a.shape == (v.hp.na,)
        if force_state is None:
            if v.physics is None:
                raise N.DmError("DPCombinedEnv.step needs force_state=(qpos, qvel) or a physics callback")
            q, qd, c = v.physics(v.qpos, v.qvel, a[None])
        else:
            q, qd = np.asarray(force_state[0])[None], np.asarray(force_state[1])[None]
            c = np.asarray([force_state[2]], dtype=np.uint32) if len(force_state) > 2 else None
        v.qpos = np.ascontiguousarray(q, dtype=np.float32).copy()
        v.qvel = np.ascontiguousarray(qd, dtype=np.float32).copy()
        c = None if c is None else np.ascontiguousarray(c, dtype=np.uint32)
        v.hp.step_host(1, v.qpos, v.qvel, np.ascontiguousarray(a[None]), c, v.idx, v.ep_len, v.clip, v._std_obs(), v.reward, v.done,
                       v.reason, v.torque, v.terms, v.ep_return)
        v._after_direct_step()
        self._pull()
        reason = int(v.reason[0])
        if reason == 8:
            return v.obs[0].astype(np.float64), 0, True, {}
        info = {k: float(v.terms[0, j]) for j, k in enumerate(N.TERM_NAMES[:7])}
        if reason:
            info["done_reason"] = N.DONE_REASONS[reason]
        return v.obs[0].astype(np.float64), float(v.reward[0]), bool(v.done[0]), info

    def render(self, mode=None):
        return None

    def close(self):
        self._vec.close()
