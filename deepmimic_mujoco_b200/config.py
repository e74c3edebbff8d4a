"""Robot / motion / environment constants, same names and values as the reference's
src/config.py:3-49 (RobotConfig, MotionConfig), src/deepmimic_env.py:258-270 (DPEnvConfig) and
src/combined_env.py:21-35 (DPCombinedEnvConfig).  The reference hard-codes
~/Code/DeepMimic_mujoco/src for its asset paths (config.py:26,38); here assets come from the
packed copy under deepmimic_mujoco_b200/assets or from $DEEPMIMIC_MUJOCO_SRC."""
import os

ROBOTS = ("humanoid3d", "unitree_g1")


class RobotConfig:
    def __init__(self, robot="humanoid3d"):
        self.robot = robot
        self.floor_geom_name = "floor"
        if robot == "humanoid3d":
            self.torso_body_name = "chest"
            self.lfoot_geom_name = "left_ankle"
            self.rfoot_geom_name = "right_ankle"
            self.extra_contact_geom_names = None
            self.endeffector_geom_names = ["left_ankle", "right_ankle", "left_wrist", "right_wrist"]
            self.low_z = 0.7
        elif robot == "unitree_g1":
            self.torso_body_name = "pelvis"
            self.lfoot_geom_name = "left_foot"
            self.rfoot_geom_name = "right_foot"
            self.extra_contact_geom_names = [side + part for side in ("left_foot_", "right_foot_")
                                             for part in ("lheel", "rheel", "ltoe", "rtoe")]
            self.endeffector_geom_names = ["left_foot", "right_foot", "left_hand", "right_hand"]
            self.low_z = 0.4
        else:
            raise Exception("Unknown robot: %s" % robot)
        self.env_name = "deepmimic_" + robot
        self.curr_path = source_root()
        self.xml_folder = "/mujoco/humanoid_deepmimic/envs/asset"
        self.xml_path = "%s%s/%s.xml" % (self.curr_path, self.xml_folder, self.env_name)


class MotionConfig(object):
    def __init__(self, motion=None, robot="humanoid3d"):
        getups = ["getup_faceup", "getup_facedown", "getup_facedown_slow", "getup_facedown_slow_FSI",
                  "getup_facedown_towalk"]
        self.acyclical_motions = list(getups)
        self.floor_motions = list(getups)
        self.curr_path = source_root()
        self.motion = "walk" if motion is None else motion
        self.robot = robot
        self.env_name = "deepmimic_" + robot
        self.motion_folder = "/mujoco/motions"
        self.xml_folder = "/mujoco/humanoid_deepmimic/envs/asset"
        self.mocap_path = "%s%s/%s_%s.txt" % (self.curr_path, self.motion_folder, robot, self.motion)
        self.xml_path = "%s%s/%s.xml" % (self.curr_path, self.xml_folder, self.env_name)


class DPEnvConfig:
    def __init__(self):
        self.MAX_EP_LENGTH = 1000
        self.VEL_OBS_SCALE = 0.1
        self.FRC_OBS_SCALE = 0.001
        self.ADD_FOOT_CONTACT_OBS = True
        self.ADD_EXTRA_CONTACT_OBS = False
        self.ADD_TORSO_OBS = True
        self.ADD_JOINT_FORCE_OBS = False
        self.ADD_ABSPOS_OBS = False
        self.ADD_PHASE_OBS = True
        self.ADD_PLAYER_ACTION_OBS = False
        self.MAX_PLAYER_ACTIONS = 3


class DPCombinedEnvConfig:
    def __init__(self):
        self.MAX_EP_LENGTH = 2000
        self.VEL_OBS_SCALE = 0.1
        self.FRC_OBS_SCALE = 0.001
        self.ADD_FOOT_CONTACT_OBS = False
        self.ADD_EXTRA_CONTACT_OBS = True
        self.ACT_SCALE = 20.0
        self.ADD_TORSO_OBS = True
        self.ADD_JOINT_FORCE_OBS = False
        self.ADD_ABSPOS_OBS = False
        self.ADD_PHASE_OBS = True
        self.ADD_PLAYER_ACTION_OBS = True
        self.MAX_PLAYER_ACTIONS = 3
        self.AMNESTY_STEPS = 150


def source_root():
    """Directory standing in for the reference's `src/` (only used to form display paths)."""
    return os.environ.get("DEEPMIMIC_MUJOCO_SRC", os.path.expanduser("~/Code/DeepMimic_mujoco/src"))


def clip_flags(robot, motion):
    """DM_CLIP_* bits of a (robot, motion) pair for DPEnv semantics (deepmimic_env.py:420,426,440)."""
    mc = MotionConfig(motion, robot)
    f = 0
    if mc.motion in mc.floor_motions:
        f |= 1
    if mc.motion in mc.acyclical_motions:
        f |= 2
    if mc.motion == "run" and robot == "unitree_g1":
        f |= 4
    return f
