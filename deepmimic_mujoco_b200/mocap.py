"""MocapDM: the clip-table builder, same public surface as the reference's
src/mujoco/mocap_v2.py:12-362 (`load_mocap`, `read_raw_data`, `get_length`, `get_qpos / get_qvel / get_body_xpos /
get_geom_xpos`, `data_config`, `data_vel`, `data_body_xpos`, `data_geom_xpos`, `dt`, `loop`, `calc_rot_vel`).

What differs from the reference, by design: the whole of read_raw_data -- y-up -> z-up alignment, quaternion ->
Euler conversion, the shoulder / hip "continuity" fix (three nested Python loops over 8 x 8 x 8 candidates per joint
and frame, mocap_v2.py:196-222), finite-difference velocities, one MuJoCo `set_state` per key frame for the FK tables
(:292-307) and the x k resampling (:309-336) -- runs on the GPU in float64 (dm_compile_clips, csrc/dm_clip_compiler.cu).
This class only parses the motion file and reads the finished tables back; there is no host implementation of the
conversion in this package (the float64 numpy restatement lives in oracle/mocap_tables.py, for the tests).
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


class MocapDM(object):
    def __init__(self, robot, hotpath=None):
        self.robot = robot
        self.model = assets.load_model(robot)
        self._hotpath = hotpath  # owner of the device tables; a shared load-time handle is created lazily
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
        """Compile one parsed clip on the device (on a handle kept for load-time work) and read its tables back."""
        from .hotpath import HotPath
        hp = HotPath.shared(self.robot)
        try:
            hp.compile_clips([clip], [0])
        except Exception as e:  # the reference raises a bare Exception with this text (mocap_v2.py:315-316)
            if "Invalid dt ratio" in str(e):
                raise Exception("Invalid dt ratio, cannot interpolate mocap frames: %f" % (float(clip["frames"][0][0]) / 0.01666))
            raise
        return self._adopt(clip, hp, 0)

    def _adopt(self, clip, hotpath, index):
        """Take over clip `index` of a handle that has already compiled `clip`."""
        self.loop = clip.get("loop")
        if self.loop is None:
            print("Warning: No loop information found in mocap file")
        self.motion_name = clip["name"]
        t = hotpath.compiled_clip(index)
        self.dt = t["dt"]
        self.data_config = list(t["qpos"])
        self.data_vel = list(t["qvel"])
        self.data_body_xpos = list(t["body_xpos"])
        self.data_geom_xpos = list(t["geom_xpos"])
        return self

    @staticmethod
    def calc_rot_vel(seg_0, seg_1, dura):
        """Body-frame angular velocity between two wxyz orientations (mocap_v2.py:350-362).  Kept for callers that
        use it as a utility; the tables' velocities come from the device compiler."""
        a, b = np.asarray(seg_0, dtype=np.float64), np.asarray(seg_1, dtype=np.float64)
        c = np.array([a[0], -a[1], -a[2], -a[3]])
        qd = np.array([c[0] * b[0] - c[1] * b[1] - c[2] * b[2] - c[3] * b[3],
                       c[0] * b[1] + c[1] * b[0] + c[2] * b[3] - c[3] * b[2],
                       c[0] * b[2] - c[1] * b[3] + c[2] * b[0] + c[3] * b[1],
                       c[0] * b[3] + c[1] * b[2] - c[2] * b[1] + c[3] * b[0]])
        ss = float(np.dot(qd, qd))
        if abs(1.0 - ss) >= 1e-14 and ss > 0:
            qd = qd / math.sqrt(ss)
        vn = float(np.linalg.norm(qd[1:]))
        axis = np.zeros(3) if vn < 1e-17 else qd[1:] / vn
        ang = ((2.0 * math.atan2(vn, qd[0]) + math.pi) % (2 * math.pi)) - math.pi
        return ((math.pi if ang == -math.pi else ang) / dura) * axis
