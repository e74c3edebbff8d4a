"""ctypes binding of libdmhot.so (include/dmhot.h).  Fails loudly: there is no CPU fallback."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libdmhot.so")

ROBOT_IDS = {"humanoid3d": 0, "unitree_g1": 1}
ENV_SINGLE, ENV_COMBINED = 0, 1
LAYOUT_AOS, LAYOUT_SOA = 0, 1
CLIP_FLOOR, CLIP_ACYCLIC, CLIP_G1_RUN = 1, 2, 4
NTERMS = 8
DONE_REASONS = ["", "low_z", "high_z", "run roll limit", "run pitch limit", "max_ep_len", "acyclical_end",
                "fallen without amnesty", "obs_bounds"]
TERM_NAMES = ["reward_config", "reward_qvel", "reward_end_eff", "reward_com", "reward_joint_limit",
              "imitation_reward", "task_reward", "z_com"]


class DmError(RuntimeError):
    pass


class Dims(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("nq", "nv", "nu", "na", "obs_dim", "nbody", "ngeom", "nsel", "row_stride",
                                         "n_stats", "robot", "env_kind")]


class Clip(C.Structure):
    _fields_ = [("len", C.c_int32), ("n_rows", C.c_int32), ("flags", C.c_int32), ("fixed_row", C.c_int32),
                ("qpos", C.c_void_p), ("qvel", C.c_void_p), ("body_xpos", C.c_void_p), ("geom_xpos", C.c_void_p)]


class RawClip(C.Structure):
    _fields_ = [("n_frames", C.c_int32), ("n_cols", C.c_int32), ("direct_qpos", C.c_int32), ("getup", C.c_int32),
                ("flags", C.c_int32), ("fixed_row", C.c_int32), ("virtual_len", C.c_int32), ("same_as", C.c_int32),
                ("frames", C.c_void_p)]


class Policy(C.Structure):
    _fields_ = [("in_dim", C.c_int32), ("h0", C.c_int32), ("h1", C.c_int32), ("out_dim", C.c_int32)] + \
               [(k, C.c_void_p) for k in ("W0", "b0", "W1", "b1", "W2", "b2")]


class PolicyArgs(C.Structure):
    _fields_ = [("n", C.c_int32), ("normalize", C.c_int32), ("obs", C.c_void_p), ("action", C.c_void_p),
                ("obs_norm", C.c_void_p), ("reward", C.c_void_p), ("reward_norm", C.c_void_p),
                ("act_lo", C.c_float), ("act_hi", C.c_float), ("stream", C.c_void_p)]


class StepArgs(C.Structure):
    _fields_ = [("n", C.c_int32), ("layout", C.c_int32), ("ld", C.c_int64),
                ("qpos", C.c_void_p), ("qvel", C.c_void_p), ("action", C.c_void_p), ("contact", C.c_void_p),
                ("idx", C.c_void_p), ("ep_len", C.c_void_p), ("clip", C.c_void_p),
                ("obs", C.c_void_p), ("reward", C.c_void_p), ("done", C.c_void_p), ("reason", C.c_void_p),
                ("torque", C.c_void_p), ("terms", C.c_void_p), ("ep_return", C.c_void_p),
                ("accumulate_stats", C.c_int32), ("stream", C.c_void_p)]


class ResetArgs(C.Structure):
    _fields_ = [("n", C.c_int32), ("layout", C.c_int32), ("ld", C.c_int64),
                ("mask", C.c_void_p), ("frame", C.c_void_p), ("coin", C.c_void_p), ("contact", C.c_void_p),
                ("seed", C.c_uint64), ("counter", C.c_uint64),
                ("qpos", C.c_void_p), ("qvel", C.c_void_p), ("idx", C.c_void_p), ("ep_len", C.c_void_p),
                ("clip", C.c_void_p), ("obs", C.c_void_p), ("ep_return", C.c_void_p), ("terminal_obs", C.c_void_p),
                ("stream", C.c_void_p)]


# every symbol include/dmhot.h declares (tests check that the library exports each one)
SYMBOLS = ["dm_abi_version", "dm_create", "dm_destroy", "dm_last_error", "dm_get_dims", "dm_get_model", "dm_fk_tables",
           "dm_upload_clips", "dm_compile_clips", "dm_get_compiled_clip", "dm_get_packed_rows", "dm_step", "dm_reset", "dm_step_host", "dm_reset_host", "dm_host_buffers",
           "dm_stats_zero", "dm_stats_device_ptr", "dm_stats_read", "dm_launch_count", "dm_step_kernel_time",
           "dm_set_timing", "dm_set_kernel_variant", "dm_set_contact_mode", "dm_set_host_mode", "dm_policy_upload", "dm_policy_set_normalizer",
           "dm_policy_normalizer_from_stats", "dm_policy_act", "dm_set_policy_variant", "dm_batch_stats_noalloc",
           "dm_stats_peer_export", "dm_stats_peer_import", "dm_stats_allreduce", "dm_stats_snapshot", "dm_stats_commit", "dm_stats_read_total",
           "dm_stats_set_auto", "dm_stats_join", "dm_obs_abspos"]
IPC_HANDLE_BYTES = 64
NOALLOC_MAX_OBS_DIM, NOALLOC_MAX_ENVS = 125, 1 << 22

_lib = None


def lib():
    """Load libdmhot.so (built in-tree by __graft_entry__.build()); raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise DmError("libdmhot.so not found at %s -- build it with `python -c 'import __graft_entry__ as g; g.build()'`; "
                      "deepmimic_mujoco_b200 has no CPU fallback" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    L.dm_last_error.restype = C.c_char_p
    L.dm_last_error.argtypes = [C.c_void_p]
    L.dm_create.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
    L.dm_destroy.argtypes = [C.c_void_p]
    L.dm_get_dims.argtypes = [C.c_void_p, C.POINTER(Dims)]
    L.dm_get_model.argtypes = [C.c_void_p] + [C.c_void_p] * 7
    L.dm_fk_tables.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.dm_upload_clips.argtypes = [C.c_void_p, C.c_int32, C.POINTER(Clip)]
    L.dm_compile_clips.argtypes = [C.c_void_p, C.c_int32, C.POINTER(RawClip)]
    L.dm_get_compiled_clip.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_double)] + [C.c_void_p] * 4
    L.dm_get_packed_rows.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_int32)]
    L.dm_step.argtypes = [C.c_void_p, C.POINTER(StepArgs)]
    L.dm_reset.argtypes = [C.c_void_p, C.POINTER(ResetArgs)]
    L.dm_step_host.argtypes = [C.c_void_p, C.POINTER(StepArgs), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    L.dm_reset_host.argtypes = [C.c_void_p, C.POINTER(ResetArgs)]
    L.dm_host_buffers.argtypes = [C.c_void_p, C.c_int32, C.POINTER(StepArgs)]
    L.dm_stats_zero.argtypes = [C.c_void_p, C.c_void_p]
    L.dm_stats_device_ptr.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int32)]
    L.dm_stats_read.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.dm_launch_count.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]
    L.dm_step_kernel_time.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int64)]
    L.dm_set_timing.argtypes = [C.c_void_p, C.c_int]
    L.dm_set_kernel_variant.argtypes = [C.c_void_p, C.c_int]
    L.dm_set_contact_mode.argtypes = [C.c_void_p, C.c_int]
    L.dm_set_host_mode.argtypes = [C.c_void_p, C.c_int]
    L.dm_policy_upload.argtypes = [C.c_void_p, C.POINTER(Policy)]
    L.dm_policy_set_normalizer.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_float]
    L.dm_policy_normalizer_from_stats.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_void_p]
    L.dm_policy_act.argtypes = [C.c_void_p, C.POINTER(PolicyArgs)]
    L.dm_set_policy_variant.argtypes = [C.c_void_p, C.c_int]
    L.dm_stats_peer_export.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]
    L.dm_stats_peer_import.argtypes = [C.c_void_p, C.c_int32, C.c_void_p]
    L.dm_stats_allreduce.argtypes = [C.c_void_p, C.c_void_p]
    L.dm_stats_snapshot.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.c_void_p]
    L.dm_stats_commit.argtypes = [C.c_void_p, C.c_void_p]
    L.dm_stats_read_total.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.dm_stats_set_auto.argtypes = [C.c_void_p, C.c_int32]
    L.dm_stats_join.argtypes = [C.c_void_p, C.c_void_p]
    L.dm_batch_stats_noalloc.argtypes = [C.c_int, C.c_int32, C.c_int32] + [C.c_void_p] * 7
    L.dm_obs_abspos.argtypes = [C.c_void_p, C.c_int32] + [C.c_void_p] * 8
    _lib = L
    return L


def check(rc, handle=None, what=""):
    if rc != 0:
        msg = lib().dm_last_error(handle)
        raise DmError("%s failed (%d): %s" % (what or "dmhot call", rc, (msg or b"").decode()))


def np_ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)
