"""gym 0.15.4 `MujocoEnv` stand-in: same constructor protocol (builds model + sim, takes one
step with a sampled action, asserts it is not done), `set_state` = state + forward,
`do_simulation` = ctrl + "step".  Kinematics come from oracle.mjkin; dynamics do not exist:
`sim.step()` leaves qpos/qvel unchanged and re-runs forward (golden runs always pass
force_state)."""
import numpy as np

from deepmimic_mujoco_b200.mjcf import compile_mjcf
from gym.spaces import Box
from oracle import mjkin


class _Contact(object):
    def __init__(self, g1, g2):
        self.geom1, self.geom2 = g1, g2


class _Data(object):
    pass


class _Sim(object):
    def __init__(self, model):
        self.model = model
        self.data = _Data()
        q = np.zeros(model.nq)
        q[:3] = model.body_pos[1]
        q[3] = 1.0
        self.data.qpos = q
        self.data.qvel = np.zeros(model.nv)
        self.data.ctrl = np.zeros(model.nu)
        self.data.contact = []
        self.data.time = 0.0
        self.forward()

    def forward(self):
        d = mjkin.forward(self.model, self.data.qpos, self.data.qvel)
        self.data.qpos[:] = d.qpos  # mj_kinematics normalises the free-joint quaternion in place
        for k in ("xpos", "xquat", "xmat", "xipos", "geom_xpos", "cvel", "subtree_com"):
            setattr(self.data, k, getattr(d, k))
        self.data.body_xpos = d.xpos
        self.data.body_xquat = d.xquat
        self.data.body_xvelp = np.zeros_like(d.xpos)

    def step(self):
        self.forward()
        self.data.time += self.model.timestep


class MujocoEnv(object):
    def __init__(self, model_path, frame_skip):
        self.frame_skip = frame_skip
        self.model = compile_mjcf(model_path)
        self.sim = _Sim(self.model)
        self.data = self.sim.data
        self.viewer = None
        self.init_qpos = self.sim.data.qpos.ravel().copy()
        self.init_qvel = self.sim.data.qvel.ravel().copy()
        b = self.model.actuator_ctrlrange.copy()
        self.action_space = Box(low=b[:, 0], high=b[:, 1], dtype=np.float32)
        action = self.action_space.sample()
        observation, _reward, done, _info = self.step(action)
        assert not done
        self.observation_space = Box(low=-np.inf * np.ones(observation.size), high=np.inf * np.ones(observation.size))

    def set_state(self, qpos, qvel):
        assert qpos.shape == (self.model.nq,) and qvel.shape == (self.model.nv,)
        self.sim.data.qpos = np.array(qpos, dtype=np.float64)
        self.sim.data.qvel = np.array(qvel, dtype=np.float64)
        self.sim.forward()

    def do_simulation(self, ctrl, n_frames):
        self.sim.data.ctrl[:] = ctrl
        for _ in range(n_frames):
            self.sim.step()

    def set_contacts(self, pairs):
        """test hook: the contact list MuJoCo's collision stage would have produced"""
        self.sim.data.contact = [_Contact(a, b) for a, b in pairs]
