"""DPVecEnv: N DeepMimic environments per call behind the Stable-Baselines3 VecEnv surface.

Replaces `SubprocVecEnv([lambda: DPEnv(...)] * N)` / `SubprocVecEnv([DPCombinedEnv] * N)`
(reference src/sb3_ppo.py:275-278, src/sac_sb3.py:77, src/ppo.py:32, src/profile_subproc_dpenv.py:10):
`reset() -> obs[N, D]`, `step_async(actions)`, `step_wait() -> (obs, rews, dones, infos)`, with SB3's
auto-reset convention (a done env returns its reset observation and
`infos[i]["terminal_observation"]` holds the last one).

MuJoCo's mj_step is outside this package.  Each step therefore needs the post-physics state:
either `step(actions, force_state=(qpos[N,nq], qvel[N,nv]))` -- the reference's own test hook
(src/deepmimic_env.py:355-357) -- or a `physics` callback given at construction,
`physics(qpos, qvel, torque) -> (qpos, qvel, contact_bits)`, which is where a batched simulator plugs in.

Two data paths:
  host   numpy in / numpy out through `dm_step_host` (pinned staging, H2D + kernel + D2H per step);
  device `step_tensors(...)` on torch CUDA tensors, no host synchronisation (the resident path).
"""
from __future__ import annotations

import ctypes as C
import weakref
from collections.abc import Sequence

import numpy as np

from . import _native as N
from .hotpath import HotPath


def _vecenv_base():
    """stable_baselines3's VecEnv ABC when it is importable, else `object`.  SB3 wraps anything that is not an
    `isinstance(env, VecEnv)` in a DummyVecEnv (the reference hands a SubprocVecEnv to PPO, src/sb3_ppo.py:275-278,307),
    so the drop-in has to BE one."""
    try:
        from stable_baselines3.common.vec_env import VecEnv
        return VecEnv
    except Exception:
        return object


def _gym_box():
    for mod in ("gymnasium", "gym"):
        try:
            return __import__(mod + ".spaces", fromlist=["Box"]).Box
        except Exception:
            pass
    return None


class Box:
    """Minimal gym.spaces.Box stand-in (gym is not a dependency)."""

    def __init__(self, low, high, dtype=np.float32):
        self.low = np.asarray(low, dtype=dtype)
        self.high = np.asarray(high, dtype=dtype)
        self.shape = self.low.shape
        self.dtype = np.dtype(dtype)
        self._rng = np.random.default_rng(0)

    def seed(self, seed=None):
        self._rng = np.random.default_rng(seed)

    def sample(self):
        lo = np.where(np.isfinite(self.low), self.low, -1.0)
        hi = np.where(np.isfinite(self.high), self.high, 1.0)
        return self._rng.uniform(lo, hi).astype(self.dtype)

    def contains(self, x):
        x = np.asarray(x)
        return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))


class LazyInfos(Sequence):
    """The `infos` of a VecEnv step: a sequence of num_envs per-env dicts whose entries are created on first access.
    (SB3's SubprocVecEnv returns a tuple here; consumers index, iterate and take len().)  Construction is O(1): building
    num_envs empty dicts -- or even a list of num_envs None entries -- eagerly costs more than the whole GPU step at 4096
    environments (17 us; 1.1 ms at 262144)."""

    __slots__ = ("_n", "_d", "__weakref__")

    def __init__(self, n):
        self._n = int(n)
        self._d = {}

    def __len__(self):
        return self._n

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[j] for j in range(*i.indices(self._n))]
        i = int(i)
        if i < 0:
            i += self._n
        if not 0 <= i < self._n:
            raise IndexError("info index out of range")
        v = self._d.get(i)
        if v is None:
            v = self._d[i] = {}
        return v

    def __setitem__(self, i, v):
        i = int(i)
        if i < 0:
            i += self._n
        if not 0 <= i < self._n:
            raise IndexError("info index out of range")
        self._d[i] = v

    def __iter__(self):
        for i in range(self._n):
            yield self[i]

    def __eq__(self, other):
        try:
            return len(other) == self._n and all(a == b for a, b in zip(self, other))
        except TypeError:
            return NotImplemented

    __hash__ = None

    def touched(self):
        """Indices whose dict exists (every other entry is an empty dict)."""
        return sorted(self._d)


class _OutputRing:
    """Caller-owned output arrays without per-step allocation.  A ring slot is a block of (pinned) memory; every hand-out
    wraps it in a NEW 1-D ndarray `f` (whose base is the slot's buffer object, not an ndarray) and gives the caller
    `f.reshape(shape)`.  numpy collapses the base of every further view -- slices, reshapes, torch.from_numpy -- onto `f`,
    so `f` dies exactly when nothing of the caller's aliases the slot any more; the ring keeps only a weak reference to it
    and reuses a slot whose reference is dead.  When the caller holds every slot (SAC's replay buffer keeping observations,
    a rollout buffer of views) a plain fresh array is handed out instead: never an array somebody may still read."""

    def __init__(self, shape, dtype, slots, pinned):
        self.shape, self.dtype = tuple(shape), np.dtype(dtype)
        self._slots = []  # [buffer object, weakref to the live 1-D array or None]
        self._keep = []
        count = int(np.prod(self.shape))
        for _ in range(slots):
            if pinned:
                import torch
                t = torch.empty(max(count * self.dtype.itemsize, 1), dtype=torch.uint8, pin_memory=True)
                self._keep.append(t)
                buf = (C.c_uint8 * (count * self.dtype.itemsize)).from_address(t.data_ptr())
            else:
                buf = (C.c_uint8 * (count * self.dtype.itemsize))()
            self._slots.append([buf, None])

    def get(self):
        for s in self._slots:
            if s[1] is None or s[1]() is None:
                f = np.frombuffer(s[0], self.dtype)
                s[1] = weakref.ref(f)
                return f.reshape(self.shape)
        return np.empty(self.shape, self.dtype)


class DPVecEnv(_vecenv_base()):
    def __init__(self, num_envs, motion="walk", robot="humanoid3d", combined=False, device=0, seed=0, physics=None,
                 info_terms=False, copy_outputs=True, contact_from_fk=False, timelimit_truncated=False, pin_cpus=None,
                 add_abspos_obs=False):
        self.num_envs = int(num_envs)
        # ENV_CFG.ADD_ABSPOS_OBS (deepmimic_env.py:133-137; False in every configuration the reference ships): the
        # observation grows by geom_xpos.flatten() in front of the phase entry.  An opt-in post-pass (dm_obs_abspos) behind
        # the step / reset kernels; the default path is untouched.
        self.add_abspos_obs = bool(add_abspos_obs)
        # pin_cpus: the CPUs this process (and the copy helper threads the library starts later, which inherit the mask)
        # may run on.  "auto" under a one-process-per-GPU launcher (LOCAL_RANK / LOCAL_WORLD_SIZE) gives every rank its own
        # contiguous slice of the cores, so that N ranks staging their batches at the same time do not share cores
        # (8 ranks x 4 copy threads on one box: 0.47 e2e scaling efficiency in round 1); a list of CPU ids is taken as is.
        self.cpus = self._pin_cpus(pin_cpus)
        # The reference's envs end on MAX_EP_LENGTH inside step() as a plain `done` (deepmimic_env.py:435-438,
        # combined_env.py:442-445): no gym TimeLimit wrapper, so SubprocVecEnv reports TimeLimit.truncated = False and SB3
        # does not bootstrap the terminal value.  True is an opt-in deviation.
        self.timelimit_truncated = bool(timelimit_truncated)
        self.robot = "unitree_g1" if combined else robot
        self.combined = bool(combined)
        self.device = device
        self.physics = physics
        self.info_terms = info_terms
        self.hp = HotPath(self.robot, combined=self.combined, device=device)
        if self.combined:
            self.hp.load_combined()
            self.motions = list(self.hp.clip_names)
        else:
            self.motions = [motion] if isinstance(motion, str) else list(motion)
            self.hp.load_motions(self.motions)
        self.hp.set_contact_mode(contact_from_fk)
        d = self.hp.dims
        self.obs_dim = d.obs_dim + (3 * d.ngeom if self.add_abspos_obs else 0)
        mc = self.hp.model_constants()
        lo, hi = mc["ctrl_lo"][:d.na].copy(), mc["ctrl_hi"][:d.na].copy()
        if self.combined:  # combined_env.py:194-200: ctrlrange / ACT_SCALE
            lo, hi = lo / 20.0, hi / 20.0
        GymBox = _gym_box()
        if type(self).__mro__[1] is not object and GymBox is not None:  # a real SB3 VecEnv wants real gym spaces
            act_space = GymBox(low=lo, high=hi, dtype=np.float32)
            obs_space = GymBox(low=-np.inf, high=np.inf, shape=(self.obs_dim,), dtype=np.float32)
        else:
            act_space = Box(lo, hi)
            obs_space = Box(np.full(self.obs_dim, -np.inf), np.full(self.obs_dim, np.inf))
        if type(self).__mro__[1] is not object:
            super().__init__(self.num_envs, obs_space, act_space)
        self.action_space, self.observation_space = act_space, obs_space
        n = self.num_envs
        self._seed, self._counter = int(seed), 0
        self.copy_outputs = bool(copy_outputs)
        # Per-env state and the step's input / output arrays live in PINNED host memory owned by the handle, laid
        # out as dm_step_host stages them: a step is one H2D transfer, the kernel, one D2H transfer, no staging
        # copies.  (The reference keeps these as Python attributes of each env process.)
        pin = self.hp.host_buffers(n, terms=info_terms)
        self._pin = pin
        self.idx = pin["idx"]              # idx_curr / current_motion_n_steps
        self.ep_len = pin["ep_len"]        # episode_length
        self.clip = pin["clip"]            # clip id / motion state
        self.ep_return = pin["ep_return"]
        self.qpos, self.qvel = pin["qpos"], pin["qvel"]
        self.obs, self.reward, self.done, self.reason = pin["obs"], pin["reward"], pin["done"], pin["reason"]
        self.torque = pin["torque"]
        self.terms = pin.get("terms")
        self.terminal_obs = np.zeros((n, self.obs_dim), np.float32)
        self._actions = None
        self._force = None
        self.h2d_bytes = self.d2h_bytes = 0
        self._dev = None
        # the argument block of dm_step_host, built once: the pinned slots never move, only the caller's arrays do
        self._args = self.hp._step_args(n, pin["qpos"], pin["qvel"], pin["action"], pin["contact"], self.idx, self.ep_len, self.clip,
                                        pin["obs"], self.reward, self.done, self.reason, self.torque, self.terms, self.ep_return,
                                        0, False, None)
        self._slot_ptr = {k: self._args_ptr(pin[k]) for k in ("qpos", "qvel", "action", "contact", "obs")}
        # ... and of the auto-reset (dm_reset_host): only the mask, the observation array and the RNG counter change
        self._rargs = self.hp._reset_args(n, None, None, None, None, pin["qpos"], pin["qvel"], self.idx, self.ep_len, self.clip,
                                          pin["obs"], self.ep_return, 0, self._seed, 0, None)
        self._h2d, self._d2h = C.c_int64(), C.c_int64()
        self._call = (self.hp.L.dm_step_host, self.hp.h, C.byref(self._args), C.byref(self._h2d), C.byref(self._d2h))
        # caller-owned outputs (copy_outputs=True) come from small rings that are reused only once the caller has
        # dropped every view of a slot (see _OutputRing).  The observation slots are PINNED host memory: dm_step_host
        # then lets the kernel write the observations straight into the array the caller receives.
        if self.add_abspos_obs:  # the extended rows and the reset mask have to be device-accessible: pinned
            import torch
            self._ext_keep = [torch.empty(n * self.obs_dim, dtype=torch.float32, pin_memory=True), torch.zeros(max(n, 1), dtype=torch.uint8, pin_memory=True)]
            self._ext = self._ext_keep[0].numpy().reshape(n, self.obs_dim)
            self._ext_mask = self._ext_keep[1].numpy()[:n]
        self._rings = {"obs": _OutputRing((n, self.obs_dim), np.float32, 4, True),
                       "rew": _OutputRing((n,), np.float32, 4, False), "done": _OutputRing((n,), np.bool_, 4, False)}

    @staticmethod
    def _pin_cpus(spec):
        import os
        if spec is None or not hasattr(os, "sched_setaffinity"):
            return None
        if spec == "auto":
            lw = int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1")))
            lr = int(os.environ.get("LOCAL_RANK", "0"))
            cores = sorted(os.sched_getaffinity(0))
            if lw <= 1 or len(cores) < 2 * lw:
                return None
            per = len(cores) // lw
            spec = cores[lr * per:(lr + 1) * per]
        cpus = sorted(int(c) for c in spec)
        os.sched_setaffinity(0, cpus)
        return cpus

    @staticmethod
    def _args_ptr(x):
        return x.__array_interface__["data"][0]

    def _fresh(self, kind):
        return self._rings[kind].get()

    def input_buffers(self):
        """The pinned (qpos, qvel, contact, action) arrays of the next step.  A simulator that writes its
        post-physics state straight into them and then calls step(actions_view, force_state=(qpos, qvel, contact))
        with these very arrays skips every host-side copy."""
        p = self._pin
        return p["qpos"], p["qvel"], p["contact"], p["action"]

    # ------------------------------------------------------------------ VecEnv surface
    def set_clips(self, clip_ids):
        """DPEnv semantics with several motions: which clip each env imitates."""
        ids = np.asarray(clip_ids, dtype=np.int64)
        n_clips = len(self.hp.clip_lens)
        if ids.size and (ids.min() < 0 or ids.max() >= n_clips):  # the reference indexes a Python list: IndexError
            raise IndexError("clip id out of range: this env has %d clip(s) %s" % (n_clips, self.motions))
        self.clip[:] = ids.astype(np.int32)

    def seed(self, seed=None):
        self._seed = int(seed or 0)
        return [self._seed + i for i in range(self.num_envs)]

    def reset(self, frame=None, coin=None, mask=None):
        """All envs (or those in `mask`): reference-state initialisation + reset observation."""
        self._counter += 1
        if frame is not None:
            frame = np.ascontiguousarray(frame, dtype=np.int32)
            if not self.combined:  # data_config[idx_init] of the reference raises IndexError (deepmimic_env.py:506)
                lens = np.asarray(self.hp.clip_lens, dtype=np.int64)[self.clip]
                bad = (frame < -lens) | (frame >= lens)
                if mask is not None:
                    bad &= np.asarray(mask, dtype=bool)
                if bad.any():
                    raise IndexError("reset frame out of range for env %d (clip length %d)" % (int(np.nonzero(bad)[0][0]), int(lens[np.nonzero(bad)[0][0]])))
                frame = np.where(frame < 0, frame + lens, frame).astype(np.int32)  # Python's negative indexing
        # always into the handle's own pinned slot: an array handed to the caller by an earlier step() is never rewritten
        obs = self._pin["obs"]
        if self.add_abspos_obs:  # the pinned slot always holds the last standard rows; keep the untouched extended rows
            if mask is not None and self.obs is not self._ext and getattr(self.obs, "shape", None) == self._ext.shape:
                np.copyto(self._ext, self.obs)
        elif mask is not None and self.obs is not obs:
            np.copyto(obs, self.obs)
        self.obs = obs
        self.hp.reset_host(self.num_envs, self.idx, self.ep_len, self.clip, obs,
                           mask=None if mask is None else np.ascontiguousarray(mask, dtype=np.uint8),
                           frame=frame,
                           coin=None if coin is None else np.ascontiguousarray(coin, dtype=np.int32),
                           qpos=self.qpos, qvel=self.qvel, ep_return=self.ep_return, seed=self._seed, counter=self._counter)
        if self.add_abspos_obs:
            self._extend_obs(obs, self._ext, mask=mask)
            self.obs = self._ext
            return self._ext.copy()
        return obs.copy()

    def _std_obs(self):
        """Where a direct dm_step_host call (the N = 1 env classes) has to write the standard observation rows."""
        return self._pin["obs"] if self.add_abspos_obs else self.obs

    def _after_direct_step(self):
        """... and what follows it: with ADD_ABSPOS_OBS the rows are widened and the bounds guard re-run over the new block."""
        if self.add_abspos_obs:
            self._extend_obs(self._pin["obs"], self._ext, step=True)
            self.obs = self._ext

    def _extend_obs(self, std_obs, ext, mask=None, step=False):
        """dm_obs_abspos over the pinned block: ext rows = [std[:phase], geom_xpos(qpos), std[phase:]] (+ the bounds guard
        over the new block in step mode).  Synchronous: the rows are read on the host right after."""
        import torch
        if self.qpos is not self._pin["qpos"]:  # the N = 1 env classes keep their own copy of the state
            np.copyto(self._pin["qpos"], self.qpos)
        m = None
        if mask is not None:
            np.copyto(self._ext_mask, np.asarray(mask).astype(np.uint8, copy=False))
            m = self._ext_mask
        self.hp.obs_abspos(self.num_envs, self._pin["qpos"], std_obs, ext, self.reward if step else None,
                           self.done if step else None, self.reason if step else None, m)
        torch.cuda.synchronize(self.device)

    def step_async(self, actions, force_state=None):
        actions = np.asarray(actions)
        assert actions.shape == (self.num_envs, self.hp.na), "actions must be [num_envs, %d]" % self.hp.na
        self._actions = self._into(self._pin["action"], actions)
        self._force = force_state

    @staticmethod
    def _into(dst, src):
        """The array to hand to dm_step_host for the pinned slot dst: dst itself when the caller already wrote into it;
        the caller's array when the library can read it as it is (C-contiguous, right dtype and shape -- the copy into
        the pinned block then happens inside dm_step_host, pipelined with the H2D transfers); else a converted copy in dst."""
        if src is dst:
            return dst
        if isinstance(src, np.ndarray) and src.dtype == dst.dtype and src.shape == dst.shape and src.flags.c_contiguous:
            return src
        np.copyto(dst, src, casting="unsafe")
        return dst

    def step_wait(self):
        n = self.num_envs
        p = self._pin
        contact = None
        failed = None
        if self._force is not None:
            fs = self._force
        elif self.physics is not None:
            # the torque belongs to the kernel (gear * clip(scale * a, ctrlrange)), so physics gets the action.
            # A simulation that diverges (deepmimic_env.py:359-378: any exception out of do_simulation) ends the episode
            # with (obs * 0, 0, True, {}): the callback either raises -- every env of the batch failed -- or returns a
            # fourth element, a per-env boolean mask of the environments whose simulation failed.
            try:
                fs = self.physics(self.qpos, self.qvel, self._actions)
            except Exception as exc:  # noqa: BLE001  (the reference catches everything)
                print("Error in step: physics callback raised %r; all %d episodes end" % (exc, n))
                fs, failed = (self.qpos, self.qvel, None), np.ones(n, np.bool_)
            if len(fs) > 3 and fs[3] is not None and np.any(fs[3]):
                failed = np.asarray(fs[3], dtype=np.bool_)
                # failed rows keep their last valid state so the kernel sees finite numbers; their outputs are overwritten below
                q, v = np.array(fs[0], dtype=np.float32), np.array(fs[1], dtype=np.float32)
                q[failed], v[failed] = self.qpos[failed], self.qvel[failed]
                fs = (q, v, fs[2])
        else:
            raise N.DmError("DPVecEnv.step needs force_state=(qpos, qvel[, contact]) or a physics callback: MuJoCo's "
                            "mj_step is outside this package")
        assert np.shape(fs[0]) == p["qpos"].shape and np.shape(fs[1]) == p["qvel"].shape
        qpos_in, qvel_in = self._into(p["qpos"], fs[0]), self._into(p["qvel"], fs[1])
        if len(fs) > 2 and fs[2] is not None:
            contact = self._into(p["contact"], fs[2])
        # copy_outputs: the observation goes straight into a fresh caller-owned array (copied inside dm_step_host while
        # the remaining outputs are still in flight); self.obs then aliases it until the next step
        self.obs = None  # drop last step's array: if the caller did too, _fresh hands it out again
        ext = None
        if self.add_abspos_obs:  # the kernel writes the standard rows into the pinned slot; the post-pass below widens them
            ext = self._fresh("obs") if self.copy_outputs else self._ext
            out_obs = p["obs"]
        else:
            out_obs = self._fresh("obs") if self.copy_outputs else p["obs"]
        self.obs = out_obs
        a, sp, ptr = self._args, self._slot_ptr, self._args_ptr
        a.qpos = sp["qpos"] if qpos_in is p["qpos"] else ptr(qpos_in)
        a.qvel = sp["qvel"] if qvel_in is p["qvel"] else ptr(qvel_in)
        a.action = sp["action"] if self._actions is p["action"] else ptr(self._actions)
        a.contact = None if contact is None else sp["contact"] if contact is p["contact"] else ptr(contact)
        a.obs = sp["obs"] if out_obs is p["obs"] else ptr(out_obs)
        fn, h, pa, ph, pd = self._call
        rc = fn(h, pa, ph, pd)
        if rc:
            N.check(rc, h, "dm_step_host")
        self.qpos, self.qvel, self._actions = p["qpos"], p["qvel"], p["action"]  # the pinned copies, never the caller's arrays
        self.h2d_bytes, self.d2h_bytes = self._h2d.value, self._d2h.value
        if ext is not None:
            self._extend_obs(p["obs"], self._ext, step=True)  # always into the pinned rows: a ring that is fully held by the
            if ext is not self._ext:                          # caller hands out ordinary (pageable) arrays
                np.copyto(ext, self._ext)
            self.obs = out_obs = ext
        if failed is not None:  # (obs * 0, 0, True, {}) and no counter advanced past the failure (deepmimic_env.py:377-378)
            self.obs[failed] = 0.0
            self.reward[failed] = 0.0
            self.done[failed] = 1
            self.reason[failed] = 8
        dones = self.done.view(np.bool_)
        infos = LazyInfos(n)
        if self.info_terms:
            for i in range(n):
                if self.reason[i] != 8:
                    infos[i].update({k: float(self.terms[i, j]) for j, k in enumerate(N.TERM_NAMES[:5])})
        if self.copy_outputs:
            rews = self._fresh("rew")
            np.copyto(rews, self.reward)
        else:
            rews = self.reward
        obs_out = out_obs if self.copy_outputs else self.obs
        if dones.any():
            d2 = self._fresh("done")
            np.copyto(d2, dones)
            dones = d2
            ids = np.nonzero(dones)[0]
            ep_ret, ep_len, reasons = self.ep_return[ids].copy(), self.ep_len[ids].copy(), self.reason[ids].copy()
            self.terminal_obs[ids] = self.obs[ids]
            self._counter += 1
            ra = self._rargs
            ra.mask, ra.obs = self._args_ptr(dones), self._args_ptr(p["obs"] if ext is not None else self.obs)
            ra.seed, ra.counter = self._seed, self._counter
            rc = self.hp.L.dm_reset_host(self.hp.h, C.byref(ra))
            if rc:
                N.check(rc, self.hp.h, "dm_reset_host")
            if ext is not None:
                self._extend_obs(p["obs"], self._ext, mask=dones)
                if ext is not self._ext:
                    ext[ids] = self._ext[ids]
            for k, i in enumerate(ids):
                info = infos[i]
                info["terminal_observation"] = self.terminal_obs[i].copy()
                info["TimeLimit.truncated"] = bool(self.timelimit_truncated and reasons[k] == 5)
                if reasons[k] not in (0, 8):
                    info["done_reason"] = N.DONE_REASONS[reasons[k]]
                info["episode"] = {"r": float(ep_ret[k]), "l": int(ep_len[k])}
        elif self.copy_outputs:
            d2 = self._fresh("done")
            np.copyto(d2, dones)
            dones = d2
        return obs_out, rews, dones, infos

    def step(self, actions, force_state=None):
        self.step_async(actions, force_state)
        return self.step_wait()

    def close(self):
        self.hp.close()

    def get_attr(self, name, indices=None):
        idx = range(self.num_envs) if indices is None else ([indices] if isinstance(indices, int) else indices)
        return [getattr(self, name) for _ in idx]

    def set_attr(self, name, value, indices=None):
        setattr(self, name, value)

    def env_method(self, method_name, *args, indices=None, **kwargs):
        idx = range(self.num_envs) if indices is None else ([indices] if isinstance(indices, int) else indices)
        return [getattr(self, method_name)(*args, **kwargs) for _ in idx]

    def env_is_wrapped(self, wrapper_class, indices=None):
        idx = range(self.num_envs) if indices is None else ([indices] if isinstance(indices, int) else indices)
        return [False for _ in idx]

    def render(self, mode=None):
        return None  # graphics stack unavailable on this box; the reference renders through mujoco_py (deepmimic_env.py:527-538)

    # ------------------------------------------------------------------ device-resident path
    def device_buffers(self):
        """Allocate (once) torch CUDA tensors mirroring the per-env state for step_tensors()."""
        if self._dev is None:
            import torch
            dev = torch.device("cuda", self.device)
            d, n = self.hp.dims, self.num_envs
            t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
            self._dev = dict(idx=t(self.idx), ep_len=t(self.ep_len), clip=t(self.clip), ep_return=t(self.ep_return),
                             obs=torch.zeros(n, d.obs_dim, device=dev), reward=torch.zeros(n, device=dev),
                             done=torch.zeros(n, dtype=torch.uint8, device=dev),
                             reason=torch.zeros(n, dtype=torch.uint8, device=dev),
                             torque=torch.zeros(n, d.nu, device=dev), terminal_obs=torch.zeros(n, d.obs_dim, device=dev),
                             qpos=t(self.qpos), qvel=t(self.qvel))
            if self.add_abspos_obs:
                self._dev.update(obs_ext=torch.zeros(n, self.obs_dim, device=dev), terminal_obs_ext=torch.zeros(n, self.obs_dim, device=dev))
        return self._dev

    def reset_tensors(self):
        """Resident reset of every environment (device RNG reference-state init); returns the device observation
        tensor.  The reference states are in device_buffers()['qpos'] / ['qvel']."""
        b = self.device_buffers()
        self._counter += 1
        self.hp.reset(self.num_envs, b["idx"], b["ep_len"], b["clip"], b["obs"], mask=None, qpos=b["qpos"], qvel=b["qvel"],
                      ep_return=b["ep_return"], seed=self._seed, counter=self._counter)
        if self.add_abspos_obs:
            self.hp.obs_abspos(self.num_envs, b["qpos"], b["obs"], b["obs_ext"])
            return b["obs_ext"]
        return b["obs"]

    def step_tensors(self, actions, qpos, qvel, contact=None, auto_reset=True, stats=False):
        """Resident path: everything stays on the GPU; returns (obs, reward, done) tensors owned by the env.
        With auto_reset, done envs are re-initialised by a masked reset launch (their previous observation goes to
        device_buffers()['terminal_obs'], their reference state to ['qpos'] / ['qvel'])."""
        b = self.device_buffers()
        n = self.num_envs
        self.hp.step(n, qpos, qvel, actions, contact, b["idx"], b["ep_len"], b["clip"], b["obs"], b["reward"], b["done"],
                     b["reason"], b["torque"], None, b["ep_return"], stats=stats)
        if self.add_abspos_obs:  # widen the rows (and re-run the bounds guard over the new block) before anybody reads `done`
            self.hp.obs_abspos(n, qpos, b["obs"], b["obs_ext"], b["reward"], b["done"], b["reason"])
        if auto_reset:
            self._counter += 1
            if self.add_abspos_obs:
                import torch
                torch.where(b["done"].bool()[:, None], b["obs_ext"], b["terminal_obs_ext"], out=b["terminal_obs_ext"])
            self.hp.reset(n, b["idx"], b["ep_len"], b["clip"], b["obs"], mask=b["done"], qpos=b["qpos"], qvel=b["qvel"],
                          ep_return=b["ep_return"], seed=self._seed, counter=self._counter, terminal_obs=b["terminal_obs"])
            if self.add_abspos_obs:  # the reset rows: their reference state is in b['qpos']
                self.hp.obs_abspos(n, b["qpos"], b["obs"], b["obs_ext"], mask=b["done"])
        if self.add_abspos_obs:
            return b["obs_ext"], b["reward"], b["done"]
        return b["obs"], b["reward"], b["done"]
