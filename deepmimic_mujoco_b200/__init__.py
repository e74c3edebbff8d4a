"""deepmimic_mujoco_b200 -- B200-native batched DeepMimic_mujoco environment hot path.

Host side (Python, mirrors the reference's gym / VecEnv surface):
  DPEnv, DPCombinedEnv   single-environment shims   (reference src/deepmimic_env.py, src/combined_env.py)
  DPVecEnv               N environments per call     (SB3 VecEnv duck type)
  MocapDM                clip-table builder          (reference src/mujoco/mocap_v2.py)
  HotPath                owner of one libdmhot handle (include/dmhot.h)
Device side: deepmimic_mujoco_b200/csrc (CUDA, sm_100a) built into libdmhot.so.  No CPU fallback.
"""
from .config import DPCombinedEnvConfig, DPEnvConfig, MotionConfig, RobotConfig  # noqa: F401

__all__ = ["DPEnv", "DPCombinedEnv", "DPVecEnv", "MocapDM", "HotPath", "RobotConfig", "MotionConfig", "DPEnvConfig",
           "DPCombinedEnvConfig"]


def __getattr__(name):  # lazy: importing the package must not require the built library
    if name == "HotPath":
        from .hotpath import HotPath
        return HotPath
    if name == "MocapDM":
        from .mocap import MocapDM
        return MocapDM
    if name == "DPVecEnv":
        from .vec_env import DPVecEnv
        return DPVecEnv
    if name == "DPEnv":
        from .deepmimic_env import DPEnv
        return DPEnv
    if name == "DPCombinedEnv":
        from .combined_env import DPCombinedEnv
        return DPCombinedEnv
    raise AttributeError(name)
