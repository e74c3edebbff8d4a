"""DPEnv: single-environment view with the reference's signature and attributes
(src/deepmimic_env.py:273-538), backed by the batched CUDA path with N = 1.

    env = DPEnv(motion="walk", robot="humanoid3d")
    obs = env.reset()
    obs, reward, done, info = env.step(action, force_state=(qpos, qvel))

`force_state` is the reference's own hook (deepmimic_env.py:355-357).  Without it the env needs a
`physics` callback (see DPVecEnv): MuJoCo's mj_step is not part of this package.
"""
from types import SimpleNamespace

import numpy as np

from . import _native as N
from .assets import load_model
from .config import DPEnvConfig, MotionConfig, RobotConfig
from .vec_env import DPVecEnv


class DPEnv:
    version = "v1.0"
    ENV_CFG = DPEnvConfig()

    def __init__(self, motion=None, load_mocap=True, robot="humanoid3d", _profile=False, device=0, physics=None):
        self.PROFILE = _profile
        self.motion_config = MotionConfig(motion=motion, robot=robot)
        self.robot_config = RobotConfig(robot=robot)
        self.model = load_model(robot)
        if self.ENV_CFG.ADD_JOINT_FORCE_OBS:
            raise NotImplementedError("ADD_JOINT_FORCE_OBS (deepmimic_env.py:123-131) needs qfrc_unc / qfrc_constraint of MuJoCo's solver, "
                                      "which is outside this package")
        self._vec = DPVecEnv(1, self.motion_config.motion, robot, device=device, physics=physics, info_terms=True,
                             add_abspos_obs=self.ENV_CFG.ADD_ABSPOS_OBS)
        self.mocap = self._vec.hp.mocaps[0]
        self.mocap_dt = self.mocap.dt
        self.mocap_data_len = self.mocap.get_length()
        # load_mocap=False (deepmimic_env.py:276-291, the loader's internal FK env): no tables, mocap_data_len = 1 and
        # step() returns (obs, 0, False, {}) before any reward / termination / counter logic (:394-395).  The kernel still
        # needs a table to index, so the motion's clip stays loaded underneath; only this view hides it.
        self._no_mocap = not load_mocap
        if self._no_mocap and self._vec.add_abspos_obs:
            raise NotImplementedError("load_mocap=False together with ADD_ABSPOS_OBS is not provided")
        if self._no_mocap:
            self.mocap = SimpleNamespace(data_config=None, data_vel=None, data_body_xpos=None, data_geom_xpos=None, dt=None)
            self.mocap_dt = None
            self.mocap_data_len = 1
        self.action_space = self._vec.action_space
        self.observation_space = self._vec.observation_space
        self.sim = SimpleNamespace(data=SimpleNamespace(qpos=np.zeros(self.model.nq), qvel=np.zeros(self.model.nv), time=0.0))
        self.episode_reward = 0
        self.episode_length = 1  # state after gym's init-time step (Q4), until reset()
        self.idx_curr = 0
        self.idx_init = 0

    def _pull(self):
        v = self._vec
        self.idx_curr = int(v.idx[0])
        self.episode_length = int(v.ep_len[0])
        self.episode_reward = float(v.ep_return[0])
        self.sim.data.qpos = v.qpos[0].astype(np.float64)
        self.sim.data.qvel = v.qvel[0].astype(np.float64)

    def reference_state_init(self, idx_init=None):
        self.reset_model(idx_init)

    def reset(self):
        self.episode_reward = 0  # deepmimic_env.py:496-500: only reset() zeroes the episode counters
        self.episode_length = 0
        self._vec.ep_len[0] = 0
        self._vec.ep_return[0] = 0.0
        return self.reset_model()

    def reset_model(self, idx_init=None):
        """reference_state_init + set_state + observation (deepmimic_env.py:502-510); the episode counters are left alone."""
        v = self._vec
        keep = (int(v.ep_len[0]), float(v.ep_return[0]))
        if self._no_mocap:  # :504-505: the observation of the current state, idx_curr = idx_init
            self.idx_init = self.idx_curr = 0 if idx_init is None else int(idx_init)
            return self._obs_only()
        obs = v.reset(frame=None if idx_init is None else [idx_init])
        v.ep_len[0], v.ep_return[0] = keep
        self._pull()
        self.idx_init = self.idx_curr
        return obs[0].astype(np.float64)

    def _obs_only(self, force_state=None):
        """get_obs of the current (or given) state without touching reward, termination or counters."""
        v = self._vec
        keep = (int(v.idx[0]), int(v.ep_len[0]), float(v.ep_return[0]))
        if force_state is not None:
            v.qpos[0], v.qvel[0] = force_state[0], force_state[1]
        v.idx[0] = 0  # phase = clip(idx_curr / 1, 0, 1) with idx_curr in {-1, 0} -> 0
        contact = np.asarray([force_state[2]], dtype=np.uint32) if force_state is not None and len(force_state) > 2 else None
        v.hp.step_host(1, v.qpos, v.qvel, np.zeros((1, v.hp.na), np.float32), contact, v.idx, v.ep_len, v.clip, v.obs,
                       v.reward, v.done, v.reason, v.torque, v.terms, v.ep_return)
        obs = v.obs[0].astype(np.float64)
        ph = self.ENV_CFG.ADD_PHASE_OBS and v.hp.obs_dim - 1
        if ph:
            obs[ph] = 0.0
        v.idx[0], v.ep_len[0], v.ep_return[0] = keep
        self.sim.data.qpos = v.qpos[0].astype(np.float64)
        self.sim.data.qvel = v.qvel[0].astype(np.float64)
        return obs

    def step(self, action, force_state=None):
        a = np.asarray(action, dtype=np.float32)
        assert a.shape == (self._vec.hp.na,), "action must have %d entries" % self._vec.hp.na
        v = self._vec
        if self._no_mocap:  # deepmimic_env.py:394-395
            if force_state is None:
                raise N.DmError("DPEnv.step needs force_state=(qpos, qvel): MuJoCo's mj_step is outside this package")
            return self._obs_only(force_state), 0, False, {}
        fs = None
        if force_state is not None:
            fs = (np.asarray(force_state[0], dtype=np.float32)[None], np.asarray(force_state[1], dtype=np.float32)[None]) + \
                 ((np.asarray([force_state[2]], dtype=np.uint32),) if len(force_state) > 2 else ())
        # single-env semantics: no auto-reset (gym env, not VecEnv) -- step the kernel directly
        v.step_async(a[None], fs)
        n = 1
        if fs is None and v.physics is None:
            raise N.DmError("DPEnv.step needs force_state=(qpos, qvel) or a physics callback")
        if fs is None:
            q, qd, c = v.physics(v.qpos, v.qvel, v._actions)
            fs = (np.ascontiguousarray(q, dtype=np.float32), np.ascontiguousarray(qd, dtype=np.float32),
                  None if c is None else np.ascontiguousarray(c, dtype=np.uint32))
        v.qpos, v.qvel = np.ascontiguousarray(fs[0]).copy(), np.ascontiguousarray(fs[1]).copy()
        contact = fs[2] if len(fs) > 2 else None
        v.hp.step_host(n, v.qpos, v.qvel, v._actions, contact, v.idx, v.ep_len, v.clip, v._std_obs(), v.reward, v.done, v.reason,
                       v.torque, v.terms, v.ep_return)
        v._after_direct_step()
        self._pull()
        self.sim.data.time += self.model.timestep
        reason = int(v.reason[0])
        if reason == 8:  # obs-bounds guard: (obs * 0, 0, True, {})
            return v.obs[0].astype(np.float64), 0, True, {}
        info = {k: float(v.terms[0, j]) for j, k in enumerate(N.TERM_NAMES[:5])}
        if reason:
            info["done_reason"] = N.DONE_REASONS[reason]
        return v.obs[0].astype(np.float64), float(v.reward[0]), bool(v.done[0]), info

    @property
    def torque(self):
        return self._vec.torque[0].copy()

    def get_time(self):
        return self.sim.data.time

    def render(self, mode=None):
        return None  # no graphics stack on this box

    def close(self):
        self._vec.close()
