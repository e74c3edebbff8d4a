"""HotPath: Python owner of one libdmhot handle (one robot x env kind on one GPU).

It uploads clip tables and launches the batched step / reset kernels on buffers the caller owns:
torch CUDA tensors (device pointers + the current torch stream; the resident path) or numpy arrays
(host pointers; the library stages them through pinned memory -- the end-to-end path).
There is no CPU implementation behind this class: without libdmhot.so and a CUDA device it raises.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _native as N
from . import assets
from .config import clip_flags

_shared = {}


class HotPath:
    def __init__(self, robot="humanoid3d", combined=False, device=0):
        self.L = N.lib()
        self.robot = robot
        self.combined = bool(combined)
        self.device = int(device)
        h = C.c_void_p()
        N.check(self.L.dm_create(self.device, N.ROBOT_IDS[robot], N.ENV_COMBINED if combined else N.ENV_SINGLE,
                                 C.byref(h)), None, "dm_create")
        self.h = h
        d = N.Dims()
        N.check(self.L.dm_get_dims(self.h, C.byref(d)), self.h, "dm_get_dims")
        self.dims = d
        self.nq, self.nv, self.nu, self.na, self.obs_dim = d.nq, d.nv, d.nu, d.na, d.obs_dim
        self.clip_names = []
        self.clip_lens = []
        self._keep = None

    @classmethod
    def shared(cls, robot, device=0):
        """A handle used only for load-time FK (MocapDM)."""
        key = (robot, device)
        if key not in _shared:
            _shared[key] = cls(robot, False, device)
        return _shared[key]

    def close(self):
        if getattr(self, "h", None):
            self.L.dm_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ model / tables
    def model_constants(self):
        d = self.dims
        out = dict(body_mass=np.zeros(d.nbody), lim_lo=np.zeros(d.nsel, np.float32), lim_hi=np.zeros(d.nsel, np.float32),
                   gear=np.zeros(d.nu, np.float32), ctrl_lo=np.zeros(d.nu, np.float32), ctrl_hi=np.zeros(d.nu, np.float32),
                   ee_geom=np.zeros(4, np.int32))
        N.check(self.L.dm_get_model(self.h, *[N.np_ptr(out[k]) for k in
                                              ("body_mass", "lim_lo", "lim_hi", "gear", "ctrl_lo", "ctrl_hi", "ee_geom")]),
                self.h, "dm_get_model")
        return out

    def fk_tables(self, qpos, qvel):
        """float64 FK of key frames on the device -> (body_xpos [n,nbody,3], geom_xpos [n,ngeom,3])."""
        q = np.ascontiguousarray(qpos, dtype=np.float64)
        v = np.ascontiguousarray(qvel, dtype=np.float64)
        n = q.shape[0]
        assert q.shape == (n, self.nq) and v.shape == (n, self.nv)
        bx = np.zeros((n, self.dims.nbody, 3))
        gx = np.zeros((n, self.dims.ngeom, 3))
        N.check(self.L.dm_fk_tables(self.h, n, N.np_ptr(q), N.np_ptr(v), N.np_ptr(bx), N.np_ptr(gx)), self.h, "dm_fk_tables")
        return bx, gx

    def upload_mocaps(self, mocaps, flags, names=None, transition=None):
        """mocaps: MocapDM-like objects (data_config / data_vel / data_body_xpos / data_geom_xpos).
        transition = (target_index, virtual_len, frame): appends a MotionTransition clip."""
        n = len(mocaps) + (1 if transition else 0)
        clips = (N.Clip * n)()
        keep = []
        for i, m in enumerate(mocaps):
            arrs = [np.ascontiguousarray(np.asarray(x, dtype=np.float64)) for x in
                    (m.data_config, m.data_vel, m.data_body_xpos, m.data_geom_xpos)]
            keep.append(arrs)
            L = len(arrs[0])
            clips[i] = N.Clip(L, L, int(flags[i]), -1, *[N.np_ptr(a) for a in arrs])
        if transition:
            t, vlen, frame = transition
            a = keep[t]
            clips[n - 1] = N.Clip(int(vlen), len(a[0]), 0, int(frame), *[N.np_ptr(x) for x in a])
        N.check(self.L.dm_upload_clips(self.h, n, clips), self.h, "dm_upload_clips")
        self.clip_lens = [int(clips[i].len) for i in range(n)]
        self.clip_names = list(names) if names else ["clip%d" % i for i in range(n)]
        self._keep = keep
        return self

    def compile_clips(self, clips, flags, names=None, transition=None):
        """The device clip compiler (dm_compile_clips): `clips` are parsed motion files as assets.load_clip returns
        them (dict(frames=float64 [n, cols], format, loop, name)); the tables are built and packed on the GPU.
        transition = (target_index, virtual_len, frame): appends a MotionTransition clip sharing the target's tables."""
        n = len(clips) + (1 if transition else 0)
        raw = (N.RawClip * n)()
        keep = []
        for i, c in enumerate(clips):
            fr = np.ascontiguousarray(c["frames"], dtype=np.float64)
            keep.append(fr)
            raw[i] = N.RawClip(fr.shape[0], fr.shape[1], 1 if c.get("format") else 0, 1 if "getup" in c["name"] else 0,
                               int(flags[i]), -1, 0, -1, N.np_ptr(fr))
        if transition:
            t, vlen, frame = transition
            raw[n - 1] = N.RawClip(0, 0, 0, 0, 0, int(frame), int(vlen), int(t), None)
        N.check(self.L.dm_compile_clips(self.h, n, raw), self.h, "dm_compile_clips")
        self.clip_names = list(names) if names else [c["name"] for c in clips] + (["transition"] if transition else [])
        self.clip_lens = []
        for i in range(n):
            ln = C.c_int32()
            N.check(self.L.dm_get_compiled_clip(self.h, i, C.byref(ln), None, None, None, None, None), self.h, "dm_get_compiled_clip")
            self.clip_lens.append(int(transition[1]) if transition and i == n - 1 else ln.value)
        self._keep = None
        return self

    def compiled_clip(self, i):
        """float64 tables of clip i as built by the device compiler -> dict(dt, qpos, qvel, body_xpos, geom_xpos)."""
        ln, dt = C.c_int32(), C.c_double()
        N.check(self.L.dm_get_compiled_clip(self.h, i, C.byref(ln), C.byref(dt), None, None, None, None), self.h, "dm_get_compiled_clip")
        d, L = self.dims, ln.value
        out = dict(qpos=np.zeros((L, d.nq)), qvel=np.zeros((L, d.nv)), body_xpos=np.zeros((L, d.nbody, 3)),
                   geom_xpos=np.zeros((L, d.ngeom, 3)))
        N.check(self.L.dm_get_compiled_clip(self.h, i, None, None, *[N.np_ptr(out[k]) for k in
                                                                       ("qpos", "qvel", "body_xpos", "geom_xpos")]),
                self.h, "dm_get_compiled_clip")
        out["dt"] = dt.value
        return out

    def _load(self, motions, flags, names, transition=None):
        from . import assets
        from .mocap import MocapDM
        clips = [assets.load_clip(self.robot, m) for m in motions]
        self.compile_clips(clips, flags, names, transition)
        self.mocaps = [MocapDM(self.robot, hotpath=self)._adopt(c, self, i) for i, c in enumerate(clips)]
        return self

    def load_motions(self, motions):
        """DPEnv semantics: one clip per motion name (per-env clip ids select among them)."""
        return self._load(motions, [clip_flags(self.robot, m) for m in motions], list(motions))

    def load_combined(self):
        """DPCombinedEnv: walk, run, getup_facedown_towalk + MTToGetup (combined_env.py:167-178)."""
        motions = ["walk", "run", "getup_facedown_towalk"]
        return self._load(motions, [0, 0, 0], motions + ["to_getup"], transition=(2, 180, 1))

    def packed_rows(self):
        n = C.c_int32()
        N.check(self.L.dm_get_packed_rows(self.h, None, C.byref(n)), self.h, "dm_get_packed_rows")
        rows = np.zeros((n.value, self.dims.row_stride), np.float32)
        N.check(self.L.dm_get_packed_rows(self.h, N.np_ptr(rows), C.byref(n)), self.h, "dm_get_packed_rows")
        return rows

    # ------------------------------------------------------------------ policy (K6)
    def policy_upload(self, W0, b0, W1, b1, W2, b2):
        """Two tanh layers + linear head, weights as np.matmul(x, W) + b uses them (extracted_policy.py:471-478)."""
        arrs = [np.ascontiguousarray(a, dtype=np.float32) for a in (W0, b0, W1, b1, W2, b2)]
        W0, b0, W1, b1, W2, b2 = arrs
        assert W0.ndim == 2 and W1.shape[0] == W0.shape[1] and W2.shape[0] == W1.shape[1]
        assert b0.shape == (W0.shape[1],) and b1.shape == (W1.shape[1],) and b2.shape == (W2.shape[1],)
        pol = N.Policy(W0.shape[0], W0.shape[1], W1.shape[1], W2.shape[1], *[N.np_ptr(a) for a in arrs])
        N.check(self.L.dm_policy_upload(self.h, C.byref(pol)), self.h, "dm_policy_upload")
        self.policy_dims = (W0.shape[0], W0.shape[1], W1.shape[1], W2.shape[1])

    def set_policy_variant(self, variant):
        """0 = fp32 CUDA-core kernel, 1 = tcgen05 tensor-core kernel (default)."""
        N.check(self.L.dm_set_policy_variant(self.h, int(variant)), self.h, "dm_set_policy_variant")

    def policy_set_normalizer(self, mean=None, var=None, eps=1e-8, clip_obs=10.0, ret_var=1.0, clip_reward=10.0):
        m = None if mean is None else np.ascontiguousarray(mean, dtype=np.float32)
        v = None if var is None else np.ascontiguousarray(var, dtype=np.float32)
        N.check(self.L.dm_policy_set_normalizer(self.h, 0 if m is None else len(m), N.np_ptr(m), N.np_ptr(v), eps, clip_obs,
                                                ret_var, clip_reward), self.h, "dm_policy_set_normalizer")

    def policy_normalizer_from_stats(self, eps=1e-8, clip_obs=10.0, clip_reward=10.0, stream=None):
        """mean / variance from the statistics the step kernels accumulated (all-reduce them first on several GPUs)."""
        if stream is None:
            import torch
            stream = torch.cuda.current_stream(self.device).cuda_stream
        N.check(self.L.dm_policy_normalizer_from_stats(self.h, eps, clip_obs, clip_reward, stream), self.h,
                "dm_policy_normalizer_from_stats")

    def policy_act(self, n, obs, action, normalize=False, obs_norm=None, reward=None, reward_norm=None,
                   act_lo=-np.inf, act_hi=np.inf, stream=None):
        """Device buffers (torch CUDA tensors); asynchronous on `stream` (default: torch's current stream)."""
        if stream is None:
            import torch
            stream = torch.cuda.current_stream(self.device).cuda_stream
        a = N.PolicyArgs(int(n), 1 if normalize else 0, self._ptr(obs), self._ptr(action), self._ptr(obs_norm),
                         self._ptr(reward), self._ptr(reward_norm), float(act_lo), float(act_hi), stream)
        N.check(self.L.dm_policy_act(self.h, C.byref(a)), self.h, "dm_policy_act")

    # ------------------------------------------------------------------ step / reset
    @staticmethod
    def _ptr(x):
        if x is None:
            return None
        if isinstance(x, np.ndarray):
            return x.ctypes.data
        return x.data_ptr()  # torch tensor

    def _step_args(self, n, qpos, qvel, action, contact, idx, ep_len, clip, obs, reward, done, reason, torque, terms,
                   ep_return, soa_ld, stats, stream):
        a = N.StepArgs()
        a.n = n
        a.layout = N.LAYOUT_SOA if soa_ld else N.LAYOUT_AOS
        a.ld = int(soa_ld or 0)
        for k, v in (("qpos", qpos), ("qvel", qvel), ("action", action), ("contact", contact), ("idx", idx),
                     ("ep_len", ep_len), ("clip", clip), ("obs", obs), ("reward", reward), ("done", done),
                     ("reason", reason), ("torque", torque), ("terms", terms), ("ep_return", ep_return)):
            setattr(a, k, self._ptr(v))
        a.accumulate_stats = 1 if stats else 0
        a.stream = stream
        return a

    def step(self, n, qpos, qvel, action, contact, idx, ep_len, clip, obs, reward, done, reason=None, torque=None,
             terms=None, ep_return=None, soa_ld=0, stats=False, stream=None):
        """Device buffers (torch CUDA tensors).  Asynchronous on `stream` (default: torch's current stream)."""
        if stream is None:
            import torch
            stream = torch.cuda.current_stream(self.device).cuda_stream
        a = self._step_args(n, qpos, qvel, action, contact, idx, ep_len, clip, obs, reward, done, reason, torque, terms,
                            ep_return, soa_ld, stats, stream)
        N.check(self.L.dm_step(self.h, C.byref(a)), self.h, "dm_step")

    def prepare_step(self, n, qpos, qvel, action, contact, idx, ep_len, clip, obs, reward, done, reason=None, torque=None,
                     terms=None, ep_return=None, soa_ld=0, stats=False, stream=None):
        """Build the argument block once for buffers that are reused every step (keeps the per-step host cost to a
        single foreign call; see step_prepared)."""
        if stream is None:
            import torch
            stream = torch.cuda.current_stream(self.device).cuda_stream
        return self._step_args(n, qpos, qvel, action, contact, idx, ep_len, clip, obs, reward, done, reason, torque, terms,
                               ep_return, soa_ld, stats, stream)

    def step_prepared(self, args):
        rc = self.L.dm_step(self.h, C.byref(args))
        if rc:
            N.check(rc, self.h, "dm_step")

    def step_host(self, n, qpos, qvel, action, contact, idx, ep_len, clip, obs, reward, done, reason=None, torque=None,
                  terms=None, ep_return=None, stats=False):
        """Host buffers (C-contiguous numpy).  Returns (h2d_bytes, d2h_bytes)."""
        a = self._step_args(n, qpos, qvel, action, contact, idx, ep_len, clip, obs, reward, done, reason, torque, terms,
                            ep_return, 0, stats, None)
        hb, db = C.c_int64(), C.c_int64()
        N.check(self.L.dm_step_host(self.h, C.byref(a), C.byref(hb), C.byref(db)), self.h, "dm_step_host")
        return hb.value, db.value

    def host_buffers(self, n, terms=False):
        """Pinned host arrays (numpy views of memory owned by the handle) laid out as dm_step_host stages them.
        Passing exactly these arrays to step_host() makes the step zero-copy on the host: one H2D transfer, the
        kernel, one D2H transfer.  Valid until host_buffers / step_host is called with a larger n, or close()."""
        a = N.StepArgs()
        N.check(self.L.dm_host_buffers(self.h, int(n), C.byref(a)), self.h, "dm_host_buffers")
        d = self.dims

        def view(ptr, shape, ctype, dtype):
            cnt = int(np.prod(shape))
            arr = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ctype)), shape=(cnt,))
            return arr.view(dtype).reshape(shape)

        f, i32, u32, u8 = (C.c_float, np.float32), (C.c_int32, np.int32), (C.c_uint32, np.uint32), (C.c_uint8, np.uint8)
        spec = dict(qpos=((n, d.nq), f), qvel=((n, d.nv), f), action=((n, d.na), f), contact=((n,), u32),
                    idx=((n,), i32), ep_len=((n,), i32), clip=((n,), i32), ep_return=((n,), f),
                    obs=((n, d.obs_dim), f), reward=((n,), f), torque=((n, d.nu), f), done=((n,), u8), reason=((n,), u8))
        if terms:
            spec["terms"] = ((n, N.NTERMS), f)
        out = {k: view(getattr(a, k), shape, ct, dt) for k, (shape, (ct, dt)) in spec.items()}
        for v in out.values():
            v[...] = 0
        return out

    def _reset_args(self, n, mask, frame, coin, contact, qpos, qvel, idx, ep_len, clip, obs, ep_return, soa_ld, seed,
                    counter, stream, terminal_obs=None):
        a = N.ResetArgs()
        a.n = n
        a.layout = N.LAYOUT_SOA if soa_ld else N.LAYOUT_AOS
        a.ld = int(soa_ld or 0)
        for k, v in (("mask", mask), ("frame", frame), ("coin", coin), ("contact", contact), ("qpos", qpos),
                     ("qvel", qvel), ("idx", idx), ("ep_len", ep_len), ("clip", clip), ("obs", obs),
                     ("ep_return", ep_return), ("terminal_obs", terminal_obs)):
            setattr(a, k, self._ptr(v))
        a.seed, a.counter, a.stream = int(seed), int(counter), stream
        return a

    def reset(self, n, idx, ep_len, clip, obs, mask=None, frame=None, coin=None, contact=None, qpos=None, qvel=None,
              ep_return=None, soa_ld=0, seed=0, counter=0, stream=None, terminal_obs=None):
        if stream is None:
            import torch
            stream = torch.cuda.current_stream(self.device).cuda_stream
        a = self._reset_args(n, mask, frame, coin, contact, qpos, qvel, idx, ep_len, clip, obs, ep_return, soa_ld, seed,
                             counter, stream, terminal_obs)
        N.check(self.L.dm_reset(self.h, C.byref(a)), self.h, "dm_reset")

    def reset_host(self, n, idx, ep_len, clip, obs, mask=None, frame=None, coin=None, contact=None, qpos=None, qvel=None,
                   ep_return=None, seed=0, counter=0, terminal_obs=None):
        a = self._reset_args(n, mask, frame, coin, contact, qpos, qvel, idx, ep_len, clip, obs, ep_return, 0, seed,
                             counter, None, terminal_obs)
        N.check(self.L.dm_reset_host(self.h, C.byref(a)), self.h, "dm_reset_host")

    # ------------------------------------------------------------------ statistics / bookkeeping
    def stats_zero(self, stream=None):
        N.check(self.L.dm_stats_zero(self.h, stream), self.h, "dm_stats_zero")

    def stats_read(self, stream=None):
        out = np.zeros(self.dims.n_stats)
        N.check(self.L.dm_stats_read(self.h, N.np_ptr(out), stream), self.h, "dm_stats_read")
        return out

    def stats_device_ptr(self):
        p, n = C.c_void_p(), C.c_int32()
        N.check(self.L.dm_stats_device_ptr(self.h, C.byref(p), C.byref(n)), self.h, "dm_stats_device_ptr")
        return p.value, n.value

    # the collective (include/dmhot.h: dm_stats_allreduce)
    def stats_peer_export(self, rank, world):
        buf = C.create_string_buffer(N.IPC_HANDLE_BYTES)
        N.check(self.L.dm_stats_peer_export(self.h, int(rank), int(world), buf), self.h, "dm_stats_peer_export")
        return buf.raw

    def stats_peer_import(self, handles):
        """handles: the `world` exported handles in rank order (bytes objects)."""
        blob = b"".join(handles)
        N.check(self.L.dm_stats_peer_import(self.h, len(handles), C.c_char_p(blob)), self.h, "dm_stats_peer_import")

    def stats_allreduce(self, stream=None):
        """Snapshot-and-zero this rank's sums, exchange with the other ranks on the device, add to the global totals."""
        if stream is None:
            import torch
            stream = torch.cuda.current_stream(self.device).cuda_stream
        rc = self.L.dm_stats_allreduce(self.h, stream)
        if rc:
            N.check(rc, self.h, "dm_stats_allreduce")

    def stats_snapshot(self, stream=None):
        """Fallback route, step 1: acc -> snapshot (zeroing acc); returns (device pointer, n) of the snapshot."""
        if stream is None:
            import torch
            stream = torch.cuda.current_stream(self.device).cuda_stream
        p = C.c_void_p()
        N.check(self.L.dm_stats_snapshot(self.h, C.byref(p), stream), self.h, "dm_stats_snapshot")
        return p.value, self.dims.n_stats

    def stats_commit(self, stream=None):
        """Fallback route, step 3 (after the all-reduce of the snapshot): totals += snapshot."""
        if stream is None:
            import torch
            stream = torch.cuda.current_stream(self.device).cuda_stream
        N.check(self.L.dm_stats_commit(self.h, stream), self.h, "dm_stats_commit")

    def stats_set_auto(self, every):
        """every > 0: dm_step issues a lagged exchange itself with every `every`-th step with statistics (dm_stats_set_auto)."""
        N.check(self.L.dm_stats_set_auto(self.h, int(every)), self.h, "dm_stats_set_auto")

    def stats_join(self, stream=None):
        """Make `stream` wait for every exchange issued so far."""
        if stream is None:
            import torch
            stream = torch.cuda.current_stream(self.device).cuda_stream
        N.check(self.L.dm_stats_join(self.h, stream), self.h, "dm_stats_join")

    def stats_read_total(self, stream=None):
        out = np.zeros(self.dims.n_stats)
        N.check(self.L.dm_stats_read_total(self.h, N.np_ptr(out), stream), self.h, "dm_stats_read_total")
        return out

    def obs_abspos(self, n, qpos, obs_in, obs_out, reward=None, done=None, reason=None, mask=None, stream=None):
        """get_abspos_obs (deepmimic_env.py:133-137) as a post-pass over DEVICE-accessible buffers (torch CUDA tensors, or
        the handle's pinned host block): obs_out[e] = [obs_in[e][:phase], geom_xpos(qpos[e]).flatten(), obs_in[e][phase:]].
        With reward / done (/ reason) the observation-bounds guard is applied to the new block (step mode)."""
        p = self._ptr
        N.check(self.L.dm_obs_abspos(self.h, int(n), p(qpos), p(obs_in), p(obs_out), p(reward), p(done), p(reason), p(mask), stream),
                self.h, "dm_obs_abspos")

    @property
    def abspos_dim(self):
        return 3 * self.dims.ngeom

    def launch_count(self):
        k = C.c_int64()
        N.check(self.L.dm_launch_count(self.h, C.byref(k)), self.h, "dm_launch_count")
        return k.value

    def set_timing(self, on):
        N.check(self.L.dm_set_timing(self.h, 1 if on else 0), self.h, "dm_set_timing")

    def kernel_time(self, reset=True):
        ms, n = C.c_double(), C.c_int64()
        N.check(self.L.dm_step_kernel_time(self.h, 1 if reset else 0, C.byref(ms), C.byref(n)), self.h, "dm_step_kernel_time")
        return ms.value, n.value

    def set_contact_mode(self, from_fk):
        """True: the floor-contact observation bits are derived from FK inside the kernels (include/dmhot.h)."""
        N.check(self.L.dm_set_contact_mode(self.h, 1 if from_fk else 0), self.h, "dm_set_contact_mode")

    def set_host_mode(self, mapped):
        """How step_host moves data: True = the kernel works on the pinned host block directly (default), False = copy
        engines (H2D, kernel, D2H).  Same results."""
        N.check(self.L.dm_set_host_mode(self.h, 1 if mapped else 0), self.h, "dm_set_host_mode")

    def set_kernel_variant(self, v):
        N.check(self.L.dm_set_kernel_variant(self.h, int(v)), self.h, "dm_set_kernel_variant")
