#!/usr/bin/env python3
"""bench.py -- env-steps/s of the DeepMimic_mujoco per-step hot path (lookup + FK + reward + termination +
obs + action->torque) on N B200s, with the HBM roofline of the step kernel and a CPU baseline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--envs E] [--impl reference]

Workload (BASELINE.json configs[1]): humanoid3d 'walk', E = 4096 environments per GPU, DPEnv semantics.
A "step" is one control step of all E environments = one launch of the step kernel (plus the statistics
launch).  Inputs are synthetic: reference-clip frame + Gaussian perturbation, resident in HBM
as B distinct state batches that are cycled through, B chosen so that the batches together exceed the
126 MB L2 (no L2 reuse between timed iterations).  N > 1: every rank owns its own E environments (weak
scaling); the fp64 statistics vector is all-reduced over NCCL every --allreduce-every steps.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

BYTES_PER_ENV_STEP = {("humanoid3d", 0): 793, ("unitree_g1", 0): 953, ("unitree_g1", 1): 1021}
L2_BYTES = 126e6
# dram__bytes_read.sum + dram__bytes_write.sum of ONE step-kernel launch from `ncu --set full` captures of this very
# command (profiles/r01_step_tile_s18_*_ncu_full_summary.txt), keyed by (robot, env kind, envs per GPU); at 4096 envs the
# 1.7 MB of outputs are still dirty in L2 when the kernel ends, so only the reads reach DRAM inside the launch
NCU_TRAFFIC = {("humanoid3d", 0, 4096): (1.764096e6, "profiles/r01_step_tile_s18_h4096_ncu_full_summary.txt"),
               ("humanoid3d", 0, 1048576): (427.913216e6 + 362.767616e6, "profiles/r01_step_tile_s18_h1M_ncu_full_summary.txt"),
               ("unitree_g1", 1, 1048576): (484.756480e6 + 535.972096e6, "profiles/r01_step_tile_s18_g1c_1M_ncu_full_summary.txt"),
               ("unitree_g1", 1, 262144): (120.824832e6 + 95.611392e6, "profiles/r01_step_tile_s18_g1c_262144_ncu_full_summary.txt")}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--envs", type=int, default=4096, help="environments per GPU")
    ap.add_argument("--robot", default="humanoid3d")
    ap.add_argument("--motion", default="walk")
    ap.add_argument("--combined", action="store_true")
    ap.add_argument("--layout", default="aos", choices=["aos", "soa"])
    ap.add_argument("--variant", type=int, default=-1)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--allreduce-every", type=int, default=64)
    ap.add_argument("--no-stats", action="store_true")
    ap.add_argument("--no-sweep", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--sweep-envs", default="65536,1048576")
    ap.add_argument("--states", default="synthetic", choices=["synthetic", "recorded"],
                    help="recorded: tile an episode_debug_log JSON (reference src/deepmimic_env.py:458-476) over the envs")
    ap.add_argument("--recording", default=None, help="path of the episode_debug_log (default: tests/golden/recorded_<robot>_<motion>.json)")
    return ap.parse_args()


def peaks():
    p = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        super().__init__(daemon=True)
        self.gpu, self.rows, self._halt = gpu, [], threading.Event()

    def run(self):
        while not self._halt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self._halt.wait(0.05)

    def stop(self):
        self._halt.set()
        self.join(timeout=6)
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.rows)}


def synth_states(hp, n_envs, n_batches, seed, device, clip_ids=None):
    """Reference-clip frame + Gaussian perturbation (sigma_q 0.05 rad, sigma_root 0.02, sigma_v 0.5), on the device."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    d = hp.dims
    tabs_q = [torch.tensor(np.asarray(m.data_config), dtype=torch.float32, device=device) for m in hp.mocaps]
    tabs_v = [torch.tensor(np.asarray(m.data_vel), dtype=torch.float32, device=device) for m in hp.mocaps]
    B, N = n_batches, n_envs
    if hp.combined:
        clip = torch.randint(0, 3, (B, N), generator=g, device=device, dtype=torch.int32)
        lens = torch.tensor(hp.clip_lens, device=device)
        frame = (torch.rand(B, N, generator=g, device=device) * lens[clip.long()]).long()
        nsteps = frame + torch.where(clip == 0, 160, 0)
        qpos = torch.empty(B, N, d.nq, device=device)
        qvel = torch.empty(B, N, d.nv, device=device)
        for c in range(3):
            m = clip == c
            qpos[m] = tabs_q[c][frame[m]]
            qvel[m] = tabs_v[c][frame[m]]
        idx = nsteps.to(torch.int32)
    else:
        ncl = len(tabs_q)
        clip = (torch.arange(N, device=device) % ncl).to(torch.int32).expand(B, N).contiguous() if clip_ids is None else clip_ids
        lens = torch.tensor(hp.clip_lens, device=device)
        frame = (torch.rand(B, N, generator=g, device=device) * lens[clip.long()]).long()
        qpos = torch.empty(B, N, d.nq, device=device)
        qvel = torch.empty(B, N, d.nv, device=device)
        for c in range(ncl):
            m = clip == c
            qpos[m] = tabs_q[c][frame[m]]
            qvel[m] = tabs_v[c][frame[m]]
        idx = frame.to(torch.int32)
    qpos[..., 7:] += 0.05 * torch.randn(B, N, d.nq - 7, generator=g, device=device)
    qpos[..., :7] += 0.02 * torch.randn(B, N, 7, generator=g, device=device)
    qvel += 0.5 * torch.randn(B, N, d.nv, generator=g, device=device)
    mc = hp.model_constants()
    lo = torch.tensor(mc["ctrl_lo"][:d.na], device=device)
    hi = torch.tensor(mc["ctrl_hi"][:d.na], device=device)
    if hp.combined:
        lo, hi = lo / 20, hi / 20
    action = lo + (hi - lo) * torch.rand(B, N, d.na, generator=g, device=device)
    contact = torch.randint(0, 4 if not hp.combined else 1024, (B, N), generator=g, device=device, dtype=torch.int32)
    if hp.combined:
        contact = contact << 2
    ep_len = torch.randint(0, 900, (B, N), generator=g, device=device, dtype=torch.int32)
    return dict(qpos=qpos.contiguous(), qvel=qvel.contiguous(), action=action.contiguous(), contact=contact, idx=idx,
                ep_len=ep_len, clip=clip.contiguous())


def recorded_states(hp, n_envs, n_batches, path, device):
    """Env e of batch b replays step (e + b) mod T of the recorded episode (qpos / qvel / action as logged, frame index
    idx_init + t as src/check_debug_log.py replays it)."""
    import torch
    from deepmimic_mujoco_b200.debug_log import read_debug_log
    log = read_debug_log(path)
    a = log.arrays()
    T = len(log)
    t = (torch.arange(n_envs, device=device)[None, :] + torch.arange(n_batches, device=device)[:, None]) % T
    dev = lambda x, dt: torch.tensor(x, dtype=dt, device=device)
    frames = dev(log.frame_indices(hp.clip_lens[0]), torch.int32)
    return dict(qpos=dev(a["qpos"], torch.float32)[t].contiguous(), qvel=dev(a["qvel"], torch.float32)[t].contiguous(),
                action=dev(a["action"], torch.float32)[t].contiguous(), contact=torch.zeros(n_batches, n_envs, dtype=torch.int32, device=device),
                idx=frames[t].contiguous(), ep_len=t.to(torch.int32).contiguous(),
                clip=torch.zeros(n_batches, n_envs, dtype=torch.int32, device=device))


class Workload:
    """Resident buffers + one-step launcher for the GPU arm."""

    def __init__(self, args, device, n_envs, seed, steps_hint):
        import torch
        from deepmimic_mujoco_b200.hotpath import HotPath
        self.torch, self.args, self.device, self.n = torch, args, device, n_envs
        self.hp = HotPath(args.robot if not args.combined else "unitree_g1", combined=args.combined, device=device.index)
        if args.combined:
            self.hp.load_combined()
        else:
            self.hp.load_motions(args.motion.split(","))
        self.hp.set_kernel_variant(args.variant)
        d = self.hp.dims
        self.kind = 1 if args.combined else 0
        self.bpe = BYTES_PER_ENV_STEP[(self.hp.robot, self.kind)]
        per_batch = n_envs * self.bpe
        self.B = max(2, min(int(np.ceil(1.3 * L2_BYTES / per_batch)), 256))
        if args.states == "recorded":
            path = args.recording or os.path.join(REPO, "tests", "golden", "recorded_%s_%s.json" % (self.hp.robot, args.motion.split(",")[0]))
            s = recorded_states(self.hp, n_envs, self.B, path, device)
        else:
            s = synth_states(self.hp, n_envs, self.B, seed, device)
        self.soa = args.layout == "soa"
        if self.soa:
            for k in ("qpos", "qvel", "action"):
                s[k] = s[k].transpose(1, 2).contiguous()  # [B, dim, N]
        self.s = s
        self.idx0, self.ep0 = s["idx"].clone(), s["ep_len"].clone()
        shape = (lambda dim: (self.B, dim, n_envs)) if self.soa else (lambda dim: (self.B, n_envs, dim))
        self.obs = torch.zeros(shape(d.obs_dim), device=device)
        self.torque = torch.zeros(shape(d.nu), device=device)
        self.reward = torch.zeros(self.B, n_envs, device=device)
        self.done = torch.zeros(self.B, n_envs, dtype=torch.uint8, device=device)
        self.reason = torch.zeros(self.B, n_envs, dtype=torch.uint8, device=device)
        self.ep_return = torch.zeros(self.B, n_envs, device=device)
        self.stats = not args.no_stats
        self.i = 0
        self._prepared = None

    def restore(self):
        self.s["idx"].copy_(self.idx0)
        self.s["ep_len"].copy_(self.ep0)

    def step(self):
        b = self.i % self.B
        self.i += 1
        if self._prepared is None:
            self._prepared = [self._prepare(k) for k in range(self.B)]
        self.hp.step_prepared(self._prepared[b])

    def _prepare(self, b):
        s = self.s
        return self.hp.prepare_step(self.n, s["qpos"][b], s["qvel"][b], s["action"][b], s["contact"][b], s["idx"][b], s["ep_len"][b],
                     s["clip"][b], self.obs[b], self.reward[b], self.done[b], self.reason[b], self.torque[b], None,
                     self.ep_return[b], soa_ld=(self.n if self.soa else 0), stats=self.stats)


def timed_region(wl, steps, warmup, world, allreduce_every, stats_t):
    """W untimed + exactly K timed steps, barrier + synchronize on both sides, CUDA events, max over ranks."""
    import torch
    import torch.distributed as dist
    from deepmimic_mujoco_b200 import dist as dmdist

    def run(k):
        for i in range(k):
            wl.step()
            if world > 1 and stats_t is not None and (i + 1) % allreduce_every == 0:
                dmdist.allreduce_(stats_t)

    run(warmup)
    if world > 1 and stats_t is not None:  # the collective is part of the warm-up too (NCCL connects lazily on first use)
        dmdist.allreduce_(stats_t)
    wl.restore()
    if stats_t is not None:
        wl.hp.stats_zero(torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = wl.hp.launch_count()
    e0.record()
    run(steps)
    if world > 1 and stats_t is not None:
        dmdist.allreduce_(stats_t)
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    launches = wl.hp.launch_count() - l0
    if world > 1:
        t = torch.tensor([ms], device=wl.device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms, launches


def kernel_time(wl, steps):
    """Average device time of one step's kernel work: `steps` launches queued back to back on torch's current stream,
    bracketed by CUDA events on that stream (the library launches on the stream it is handed).  Per-launch event
    pairs would fold host launch latency into short kernels, so the bracket spans the whole queue."""
    import torch
    wl.restore()
    for _ in range(3):
        wl.step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        wl.step()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps, steps


def policy_time(hp, device, n):
    """The consumer of the observation, reported next to the step: fused normaliser + 256/128 tanh MLP
    (dm_policy_act) on n resident observations, both kernels, CUDA events on the launch stream.  Not part of `value`."""
    import torch
    d = hp.dims
    rng = np.random.default_rng(5)
    out_dim = min(int(d.na), 32)
    W = [rng.normal(0, 0.1, sh).astype(np.float32) for sh in ((d.obs_dim, 256), (256,), (256, 128), (128,), (128, out_dim), (out_dim,))]
    hp.policy_upload(*W)
    hp.policy_set_normalizer(np.zeros(d.obs_dim, np.float32), np.ones(d.obs_dim, np.float32))
    obs, act = torch.randn(n, d.obs_dim, device=device), torch.empty(n, out_dim, device=device)
    res = {"envs": n, "net": "%d-256-128-%d tanh, normaliser on" % (d.obs_dim, out_dim)}
    for variant, key in ((1, "tensor_core_ms"), (0, "cuda_core_ms")):
        hp.set_policy_variant(variant)
        for _ in range(3):
            hp.policy_act(n, obs, act, normalize=True)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            hp.policy_act(n, obs, act, normalize=True)
        e1.record()
        torch.cuda.synchronize()
        res[key] = e0.elapsed_time(e1) / 10
    res["env_per_s"] = n / (res["tensor_core_ms"] * 1e-3)
    hp.set_policy_variant(1)
    return res


def e2e_region(args, steps, warmup, seed, local_rank=0, world=1):
    """Same metric through the public API a user calls -- DPVecEnv.step(actions, force_state=(qpos, qvel, contact)) with
    HOST numpy buffers: H2D of the step's inputs, kernel, D2H of obs / reward / done / torque every step."""
    import torch
    from deepmimic_mujoco_b200.vec_env import DPVecEnv
    env = DPVecEnv(args.envs, motion=args.motion.split(","), robot=args.robot, combined=args.combined, seed=seed, device=local_rank)
    dev = torch.device("cuda", local_rank)
    B = 8
    s = synth_states(env.hp, args.envs, B, seed + 1, dev)
    host = {k: v.cpu().numpy() for k, v in s.items()}
    host["contact"] = host["contact"].astype(np.uint32)
    env.reset()

    def one(i):
        b = i % B
        return env.step(host["action"][b], force_state=(host["qpos"][b], host["qvel"][b], host["contact"][b]))

    def over_ranks(dt):  # every rank steps its own environments at the same time; the job's time is the slowest rank's
        if world == 1:
            return dt
        import torch.distributed as dist
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            import torch.distributed as dist
            dist.barrier()

    for i in range(warmup):
        one(i)
    sync()
    t0 = time.perf_counter()
    for i in range(steps):
        obs, rew, done, infos = one(i)
    torch.cuda.synchronize()
    dt = over_ranks(time.perf_counter() - t0)
    assert obs.shape == (args.envs, env.hp.obs_dim) and len(infos) == args.envs
    out = dict(value=world * args.envs * steps / dt, unit="env-steps/s", h2d_bytes_per_step=int(env.h2d_bytes) * world,
               d2h_bytes_per_step=int(env.d2h_bytes) * world, ms_per_step=1e3 * dt / steps,
               api="DPVecEnv.step(actions, force_state=(qpos, qvel, contact)) with ordinary (pageable) numpy inputs and caller-owned "
                   "outputs: dm_step_host copies the inputs into the handle's pinned block in chunks of whole tiles (copies shared "
                   "with 3 helper threads); the step kernel of each chunk reads and writes the pinned block itself -- its TMA bulk "
                   "loads / stores cross PCIe, both directions at once -- while the host copies the next chunk in; the observations "
                   "land directly in the (pinned, recycled) array the caller receives")
    # the same call when the simulator keeps its state in the handle's pinned arrays (DPVecEnv.input_buffers()): no
    # host-side staging at all, one kernel launch on mapped memory per step
    qb, vb, cb, ab = env.input_buffers()
    qb[...] = host["qpos"][0]; vb[...] = host["qvel"][0]; cb[...] = host["contact"][0]; ab[...] = host["action"][0]
    env.copy_outputs = False
    for i in range(warmup):
        env.step(ab, force_state=(qb, vb, cb))
    sync()
    t0 = time.perf_counter()
    for i in range(steps):
        obs, rew, done, infos = env.step(ab, force_state=(qb, vb, cb))
    dt = over_ranks(time.perf_counter() - t0)
    out["pinned"] = dict(value=world * args.envs * steps / dt, ms_per_step=1e3 * dt / steps,
                         note="inputs written in place into DPVecEnv.input_buffers(), outputs read as views of the pinned block")
    env.close()
    return out


# ------------------------------------------------------------------------------------------- CPU baseline
def _cpu_worker(conn, robot, motion, combined, n_envs, seed):
    """One SubprocVecEnv-style worker: owns n_envs oracle envs, one (recv, step, send) per env per step
    (message pattern of reference src/profile_subproc_dpenv.py:10-25)."""
    from oracle.envs import MultiClipEnv, SingleClipEnv
    import random
    random.seed(seed)
    envs = [MultiClipEnv() if combined else SingleClipEnv(motion, robot) for _ in range(n_envs)]
    for e in envs:
        e.reset()
    conn.send("ready")
    while True:
        msg = conn.recv()
        if msg is None:
            break
        i, action, qpos, qvel, contact = msg
        obs, rew, done, info = envs[i].step(action, (qpos, qvel), contact)
        if done:
            info["terminal_observation"] = obs
            obs = envs[i].reset()
        conn.send((obs, rew, done, info))


def cpu_baseline(args, seconds, procs=None):
    """The oracle (float64 numpy restatement of the reference, one env at a time in Python) run as the reference runs:
    P worker processes connected by pipes, one message round trip per env per step."""
    import multiprocessing as mp
    from oracle.envs import SingleClipEnv, MultiClipEnv, tables_for
    P = procs or os.cpu_count() or 1
    per = 4
    ctx = mp.get_context("fork")
    robot = "unitree_g1" if args.combined else args.robot
    motion = args.motion.split(",")[0]
    if args.combined:
        MultiClipEnv()
    else:
        tables_for(robot, motion)  # build tables once before forking
    probe = MultiClipEnv() if args.combined else SingleClipEnv(motion, robot)
    rng = np.random.default_rng(0)
    nq, nv, na = probe.model.nq, probe.model.nv, probe.action_dim
    base_q = np.asarray(probe.mocap.data_config[0] if not args.combined else probe.walk_mocap.data_config[0])
    base_v = np.asarray(probe.mocap.data_vel[0] if not args.combined else probe.walk_mocap.data_vel[0])
    pipes, ps = [], []
    for p in range(P):
        a, b = ctx.Pipe()
        pr = ctx.Process(target=_cpu_worker, args=(b, robot, motion, args.combined, per, p), daemon=True)
        pr.start()
        pipes.append(a)
        ps.append(pr)
    for a in pipes:
        a.recv()
    steps_done, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < seconds:
        for i in range(per):  # vec step: send to every worker, then gather (step_async / step_wait)
            for a in pipes:
                q = base_q + np.concatenate([rng.normal(0, 0.02, 7), rng.normal(0, 0.05, nq - 7)])
                v = base_v + rng.normal(0, 0.5, nv)
                a.send((i, rng.uniform(-1, 1, na), q, v, int(rng.integers(0, 4))))
            for a in pipes:
                a.recv()
        steps_done += per * P
    dt = time.perf_counter() - t0
    for a in pipes:
        a.send(None)
    for pr in ps:
        pr.join(timeout=2)
    return dict(value=steps_done / dt, unit="env-steps/s", cores=P, kind="port",
                sample="%d env-steps of the same workload (%s %s, %d envs in %d worker processes over pipes, %.1f s)" %
                       (steps_done, robot, "combined" if args.combined else motion, per * P, P, dt))


def reference_arm(args, rank):
    """--impl reference: the reference's CPU implementation of the path (oracle port: MuJoCo / SB3 are not installable
    here) on all host cores; each "step" is a bounded sample so the run ends within minutes."""
    if rank != 0:
        return
    P = os.cpu_count() or 1
    budget = min(90.0, max(6.0, 0.15 * (args.steps + args.warmup)))
    cb = cpu_baseline(args, budget, P)
    line = {"impl": "reference", "metric": "env-steps/s (ref-lookup+FK+reward+obs+PD)", "value": cb["value"], "unit": "env-steps/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * args.envs / cb["value"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args), "envs_per_gpu": args.envs,
                       "note": "ms_per_step = time the CPU arm needs for one control step of all envs_per_gpu environments"},
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    RESULT.append(line)


def workload_name(args):
    if args.combined:
        return "unitree_g1 DPCombinedEnv (walk/run/getup_facedown_towalk), %d envs per GPU" % args.envs
    return "%s DPEnv '%s', %d envs per GPU (BASELINE.json configs[1])" % (args.robot, args.motion, args.envs)


class QuietStdout:
    """Everything but the final JSON line goes to stderr (NCCL prints its version banner on stdout)."""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        os.dup2(2, 1)
        return self

    def __exit__(self, *exc):
        sys.stdout.flush()
        os.dup2(self.saved, 1)
        os.close(self.saved)


def emit(line):
    print(json.dumps(line), flush=True)


def main():
    args = parse()
    from deepmimic_mujoco_b200 import dist as dmdist
    rank, local_rank, world = dmdist.env_info()
    if args.impl == "reference":
        reference_arm(args, rank)
        return
    import torch
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    rank, local_rank, world = dmdist.init("nccl")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    peak, peak_src = peaks()

    wl = Workload(args, device, args.envs, 1234 + rank, args.steps)
    stats_t = dmdist.stats_tensor(wl.hp) if wl.stats else None
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    ms, launches = timed_region(wl, args.steps, max(args.warmup, 3), world, args.allreduce_every, stats_t)
    clocks = sampler.stop() if sampler else None
    value = world * args.envs * args.steps / (ms * 1e-3)
    stats_host = wl.hp.stats_read() if wl.stats else None

    # roofline of the dominant kernel at this workload
    k_ms, k_n = kernel_time(wl, min(args.steps, 200))
    bytes_per_launch = wl.bpe * args.envs
    ach = bytes_per_launch / (k_ms * 1e-3) / 1e9 if k_ms > 0 else 0.0
    roof = {"bound": "hbm", "kernel": "step kernel (K1)", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
            "traffic": NCU_TRAFFIC.get((wl.hp.robot, wl.kind, args.envs), (None, None))[0],
            "traffic_source": NCU_TRAFFIC.get((wl.hp.robot, wl.kind, args.envs), (None, None))[1], "peak_source": peak_src, "bytes_per_env_step": wl.bpe, "envs_per_launch": args.envs,
            "kernel_ms": k_ms, "kernel_launches_timed": k_n}

    line = {"metric": "env-steps/s (ref-lookup+FK+reward+obs+PD)", "value": value, "unit": "env-steps/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic" if args.states == "synthetic" else "recorded (synthetic-recorded episode_debug_log, tools/make_recording.py)",
            "config": {"workload": workload_name(args), "layout": args.layout, "state_batches": wl.B,
                       "l2_policy": "inputs larger than L2: %d distinct resident batches (%.0f MB) cycled" %
                                    (wl.B, wl.B * bytes_per_launch / 1e6),
                       "stats": wl.stats, "allreduce_every": args.allreduce_every if world > 1 else None},
            "roofline": roof, "gpu_launches": int(launches), "clocks": clocks}
    if stats_host is not None:
        m = dmdist.moments(stats_host, wl.hp.obs_dim)
        line["stats"] = {"count": m["count"], "reward_mean": m["reward_mean"], "done_rate": m["done_rate"]}

    if rank == 0 and not args.no_sweep:
        sweep = []
        for n in [int(x) for x in args.sweep_envs.split(",") if x]:
            w2 = Workload(args, device, n, 99, 20)
            for _ in range(5):
                w2.step()
            km, kn = kernel_time(w2, 20)
            a = w2.bpe * n / (km * 1e-3) / 1e9
            sweep.append({"envs": n, "kernel_ms": km, "env_steps_per_s": n / (km * 1e-3), "GBps": a, "frac": a / peak})
            del w2
            torch.cuda.empty_cache()
        line["sweep"] = sweep
        line["policy"] = policy_time(wl.hp, device, 1048576)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    if not args.no_e2e:  # every rank steps its own environments through the host API at the same time
        del wl
        torch.cuda.empty_cache()
        e2e = e2e_region(args, min(args.steps, 200), 20, 7, local_rank, world)
        if rank == 0:
            line["e2e"] = e2e
    if rank == 0:
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline(args, args.cpu_seconds)
        RESULT.append(line)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


RESULT = []

if __name__ == "__main__":
    with QuietStdout():
        main()
    for line in RESULT:
        emit(line)
