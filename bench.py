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
STATE_SEED = 1234  # rank r's headline batches: synth_states(..., seed = STATE_SEED + r) on the CPU generator
# dram__bytes_read.sum + dram__bytes_write.sum of ONE step-kernel launch from `ncu --set full` captures of this very
# shape (tools/gpu_r2_ncu_all.sh -> profiles/r02_*_ncu_full_summary.txt), keyed by (robot, env kind, envs per GPU, layout).
# Small batches: the outputs are still dirty in L2 when the kernel ends, so only the reads reach DRAM inside the launch.
def _t(name, rd, wr):
    return (rd + wr, "profiles/r02_%s_ncu_full_summary.txt" % name)


NCU_TRAFFIC = {("humanoid3d", 0, 4096, "aos"): _t("h4096", 1.769728e6, 0.000000e6),
               ("humanoid3d", 0, 8192, "aos"): _t("c3_8192", 3.488512e6, 0.000000e6),
               ("humanoid3d", 0, 65536, "aos"): _t("c3_65536", 26.884864e6, 0.221696e6),
               ("humanoid3d", 0, 1048576, "aos"): _t("h1M", 444.581888e6, 362.651392e6),
               ("humanoid3d", 0, 2097152, "aos"): _t("h2M", 887.202560e6, 779.940352e6),
               ("unitree_g1", 1, 32768, "aos"): _t("g1c_32768", 15.305216e6, 0.000000e6),
               ("unitree_g1", 1, 262144, "aos"): _t("g1c_262144", 120.935168e6, 95.656704e6),
               ("unitree_g1", 1, 1048576, "aos"): _t("g1c_1M", 535.478528e6, 536.336384e6),
               ("humanoid3d", 0, 1048576, "soa"): _t("soa_h1M", 447.850752e6, 366.636800e6),
               ("unitree_g1", 1, 1048576, "soa"): _t("soa_g1c_1M", 524.789504e6, 535.347200e6)}


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
    ap.add_argument("--allreduce-every", type=int, default=1, help="steps between two statistics collectives (N > 1); BASELINE "
                    "configs[2] asks for one EVERY step")
    ap.add_argument("--collective", default="auto", choices=["auto", "reduce", "nccl", "off"],
                    help="auto: dm_step issues the device-side lagged exchange itself (side stream or exchange CTA, dm_peer.cuh); reduce: "
                         "the same exchange launched explicitly between the steps; nccl: snapshot + ncclAllReduce + commit")
    ap.add_argument("--collective-at-1", action="store_true", help="run the collective's local form at N = 1 too (cost probe)")
    ap.add_argument("--no-configs", action="store_true", help="skip the BASELINE configs block (configs 3, 4, 5, 1 M row)")
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
        elif getattr(args, "host_seeded", False):  # the headline workload: the CPU arm draws the SAME numbers (cpu_states)
            s = {k: v.to(device) for k, v in synth_states(self.hp, n_envs, self.B, seed, torch.device("cpu")).items()}
        else:
            s = synth_states(self.hp, n_envs, self.B, seed, device)
        self.soa = args.layout == "soa"
        if self.soa:
            for k in ("qpos", "qvel", "action"):
                s[k] = s[k].transpose(1, 2).contiguous()  # [B, dim, N]
        self.s = s
        self.idx0, self.ep0, self.clip0 = s["idx"].clone(), s["ep_len"].clone(), s["clip"].clone()
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
        self.s["clip"].copy_(self.clip0)  # the combined env's motion state is in-out too: every region starts from the seeded batch

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


def launch_gate_us(steps):
    """Length of the device-side gate in front of a timed region: ~10 us per step queued, 100 us .. 2 ms."""
    if "DM_BENCH_GATE_US" in os.environ:
        return float(os.environ["DM_BENCH_GATE_US"])
    return float(min(max(10 * steps, 100), 2000))


def timed_region(wl, steps, warmup, world, every, coll, mode):
    """W untimed + exactly K timed steps, barrier + synchronize on both sides, CUDA events, max over ranks.
    coll: the statistics collective (deepmimic_mujoco_b200.dist.StatsCollective) or None; with `mode` "auto" dm_step issues
    the exchange itself behind every `every`-th step, otherwise it is launched here between the steps."""
    import torch
    import torch.distributed as dist
    auto = coll is not None and mode == "auto" and coll.auto(every)

    def run(k):
        for i in range(k):
            wl.step()
            if coll is not None and not auto and (i + 1) % every == 0:
                coll.reduce()

    run(warmup)  # the collective is part of the warm-up too (peer mappings and NCCL connect lazily on first use)
    if coll is not None:
        if auto:
            coll.reduce()  # nothing of the warm-up may still be in flight when the statistics are zeroed
        coll.join()
    wl.restore()
    torch.cuda.synchronize()
    if wl.stats:
        wl.hp.stats_zero(torch.cuda.current_stream().cuda_stream)
        if auto:
            coll.auto(every)  # restart the step counter of the auto mode
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    # A short region (the driver's 20 steps of ~9 us) would otherwise start on an EMPTY launch queue and measure the host's
    # launch ramp -- on every rank, coupled through the exchange -- rather than the device (N = 2, 20 steps: 232 us without,
    # 194 us with the gate; 200 steps: 9.13 vs 9.11 us per step).  A device-side gate holds the stream while the host
    # queues the region's launches; the events sit behind the gate.  DM_BENCH_GATE_US=0 switches it off.
    gate_us = launch_gate_us(steps)
    if gate_us > 0:
        torch.cuda._sleep(int(gate_us * 2000))
    if coll is not None and world > 1 and not os.environ.get("DM_BENCH_NO_ALIGN"):
        # the host barrier releases the ranks tens of microseconds apart -- a tenth of a 20-step region at 4096 environments.
        # One exchange of (zeroed) statistics lines the GPUs up to within the NVLink round trip before the clock starts.
        coll.reduce()
        coll.join()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = wl.hp.launch_count()
    e0.record()
    run(steps)
    probe_no_final = bool(os.environ.get("DM_BENCH_NO_FINAL"))  # probe only: the flush after the clock stops
    if coll is not None and not probe_no_final:
        if auto or steps % every:
            coll.reduce()  # auto: collect the exchange still in flight and exchange what the last step added; else the remainder
        coll.join()        # the timed region ends when the last exchange has ended
    e1.record()
    if coll is not None and probe_no_final:
        coll.reduce()
        coll.join()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    launches = wl.hp.launch_count() - l0
    if auto:
        coll.auto(0)
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
    gate_us = launch_gate_us(steps)  # as in timed_region: the queue is filled before the first event
    if gate_us > 0:
        torch.cuda._sleep(int(gate_us * 2000))
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
    env = DPVecEnv(args.envs, motion=args.motion.split(","), robot=args.robot, combined=args.combined, seed=seed, device=local_rank,
                   pin_cpus="auto" if world > 1 else None)
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
    out = dict(value=world * args.envs * steps / dt, unit="env-steps/s", cpus_per_rank=None if env.cpus is None else len(env.cpus),
               h2d_bytes_per_step=int(env.h2d_bytes) * world,
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
class _OracleHandle:
    """What synth_states needs of a HotPath, built from the oracle's tables: the CPU arm never touches libdmhot."""

    def __init__(self, robot, motions, combined):
        from types import SimpleNamespace
        from oracle.assets import load_model
        from oracle.envs import tables_for
        self.combined = combined
        self.robot = "unitree_g1" if combined else robot
        m = load_model(self.robot)
        names = ["walk", "run", "getup_facedown_towalk"] if combined else list(motions)
        self.mocaps = [tables_for(self.robot, n) for n in names]
        self.clip_lens = [t.get_length() for t in self.mocaps] + ([180] if combined else [])
        n_hand = 14 if self.robot == "unitree_g1" else 0
        self.dims = SimpleNamespace(nq=m.nq, nv=m.nv, na=m.nu - n_hand, nu=m.nu)
        self._ctrl = m.actuator_ctrlrange

    def model_constants(self):
        return dict(ctrl_lo=self._ctrl[:, 0].astype(np.float32), ctrl_hi=self._ctrl[:, 1].astype(np.float32))


def cpu_states(args, n_envs, n_batches, seed):
    """The GPU arm's headline state batches (same generator, same seed, CPU device) as float64 numpy for the oracle."""
    import torch
    oh = _OracleHandle(args.robot, args.motion.split(","), args.combined)
    s = synth_states(oh, n_envs, n_batches, seed, torch.device("cpu"))
    return {k: v.numpy() for k, v in s.items()}


def _cpu_worker(conn, robot, motions, combined, n_envs, first_env, st):
    """One SubprocVecEnv-style worker: owns n_envs oracle envs, one (recv, step, send) per env per step (message pattern
    of reference src/profile_subproc_dpenv.py:10-25).  Env i of this worker is global env first_env + i: it starts from the
    counters of the seeded batch and is stepped with that env's states."""
    from oracle.envs import MultiClipEnv, SingleClipEnv
    envs = []
    for i in range(n_envs):
        g = first_env + i
        if combined:
            e = MultiClipEnv()
            e.reset()
        else:
            e = SingleClipEnv(motions[int(st["clip"][0, g])], robot)
            e.reset(idx_init=int(st["idx"][0, g]))
        envs.append(e)
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


def cpu_single_process(args, steps=10000, seconds=None):
    """BASELINE configs[0]: the reference's own single-process loop (src/profile_dpenv.py:6-10: 10 000 random-action steps,
    reset on done) on the oracle with DPEnv(motion, robot); states = force_state from the seeded generator."""
    from oracle.envs import MultiClipEnv, SingleClipEnv
    motions = args.motion.split(",")
    B = 64
    st = cpu_states(args, 1, B, STATE_SEED)
    env = MultiClipEnv() if args.combined else SingleClipEnv(motions[0], "unitree_g1" if args.combined else args.robot)
    env.reset()
    done_steps, t0 = 0, time.perf_counter()
    for i in range(steps):
        b = i % B
        obs, rew, done, info = env.step(st["action"][b, 0].astype(np.float64), (st["qpos"][b, 0].astype(np.float64), st["qvel"][b, 0].astype(np.float64)),
                                        int(st["contact"][b, 0]))
        if done:
            env.reset()
        done_steps += 1
        if seconds and time.perf_counter() - t0 > seconds:
            break
    dt = time.perf_counter() - t0
    return dict(value=done_steps / dt, unit="env-steps/s", cores=1, ms_per_step=1e3 * dt / done_steps, steps=done_steps,
                loop="reference src/profile_dpenv.py:6-10 (one process, one env, reset on done), %d steps" % done_steps)


def cpu_baseline(args, seconds, procs=None, per=None):
    """The oracle (float64 numpy restatement of the reference, one env at a time in Python) run as the reference runs:
    P worker processes connected by pipes, one message round trip per env per step.  The N_cpu = P x per environments are
    environments [0, N_cpu) of the GPU arm's headline batches (same seeded generator); the rate is what is reported, not
    an extrapolated time."""
    import multiprocessing as mp
    from oracle.envs import MultiClipEnv, tables_for
    P = procs or os.cpu_count() or 1
    per = per or max(1, 64 // P) * (2 if P <= 16 else 1)
    n_cpu = P * per
    while n_cpu % 32:  # N_cpu = 32 k
        per += 1
        n_cpu = P * per
    ctx = mp.get_context("fork")
    robot = "unitree_g1" if args.combined else args.robot
    motions = args.motion.split(",")
    B = 16
    st = cpu_states(args, n_cpu, B, STATE_SEED)
    if args.combined:
        MultiClipEnv()
    else:
        for m in motions:
            tables_for(robot, m)  # build tables once before forking
    f64 = {k: st[k].astype(np.float64) for k in ("qpos", "qvel", "action")}
    pipes, ps = [], []
    for p in range(P):
        a, b = ctx.Pipe()
        pr = ctx.Process(target=_cpu_worker, args=(b, robot, motions, args.combined, per, p * per, st), daemon=True)
        pr.start()
        pipes.append(a)
        ps.append(pr)
    for a in pipes:
        a.recv()
    steps_done, vec_steps, t0 = 0, 0, time.perf_counter()
    while time.perf_counter() - t0 < seconds:
        b = vec_steps % B
        for i in range(per):  # vec step: send to every worker, then gather (step_async / step_wait)
            for p, a in enumerate(pipes):
                g = p * per + i
                a.send((i, f64["action"][b, g], f64["qpos"][b, g], f64["qvel"][b, g], int(st["contact"][b, g])))
            for a in pipes:
                a.recv()
        steps_done += n_cpu
        vec_steps += 1
    dt = time.perf_counter() - t0
    for a in pipes:
        a.send(None)
    for pr in ps:
        pr.join(timeout=2)
    return dict(value=steps_done / dt, unit="env-steps/s", cores=P, kind="port", n_cpu_envs=n_cpu, vec_steps=vec_steps,
                sample="%d vec steps of environments [0, %d) of the GPU arm's seeded batches (%s %s; %d worker processes x %d envs "
                       "over pipes, %.1f s); a RATE measured on that sample, not extrapolated from a time" %
                       (vec_steps, n_cpu, robot, "combined" if args.combined else args.motion, P, per, dt))


def reference_arm(args, rank):
    """--impl reference: the reference's CPU implementation of the path (oracle port: MuJoCo / SB3 are not installable
    here) on all host cores; each "step" is a bounded sample so the run ends within minutes."""
    if rank != 0:
        return
    P = os.cpu_count() or 1
    budget = min(90.0, max(6.0, 0.15 * (args.steps + args.warmup)))
    cb = cpu_baseline(args, budget, P)
    cb["single_process"] = cpu_single_process(args, 10000, seconds=min(60.0, budget))
    line = {"impl": "reference", "metric": "env-steps/s (ref-lookup+FK+reward+obs+PD)", "value": cb["value"], "unit": "env-steps/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * args.envs / cb["value"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args), "envs_per_gpu": args.envs,
                       "note": "value = measured rate on the N_cpu-environment sample (cpu_baseline.sample); ms_per_step = what one "
                               "control step of all envs_per_gpu environments would take at that rate"},
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


def _variant_args(args, **kw):
    d = dict(vars(args))
    d.update(kw)
    return argparse.Namespace(**d)


def run_config(args, device, rank, world, envs_per_rank, steps, warmup, seed, with_coll, peak):
    """One workload on every rank at once (weak or sharded: the caller picks envs_per_rank): device-timed steps with the
    statistics collective every step when N > 1, plus the step kernel's own time and roofline fraction."""
    import torch
    from deepmimic_mujoco_b200 import dist as dmdist
    wl = Workload(args, device, envs_per_rank, seed + rank, steps)
    coll = None
    if wl.stats and with_coll and (world > 1 or args.collective_at_1) and args.collective != "off":
        coll = dmdist.StatsCollective(wl.hp, force_nccl=(args.collective == "nccl"))
    ms, launches = timed_region(wl, steps, warmup, world, 1, coll, args.collective)
    out = {"envs_total": envs_per_rank * world, "envs_per_gpu": envs_per_rank, "ms_per_step": ms / steps,
           "env_steps_per_s": world * envs_per_rank * steps / (ms * 1e-3), "gpu_launches": int(launches)}
    if coll is not None:
        tot = coll.totals()
        out["collective"] = {"mode": coll.mode, "issue": args.collective, "every": 1, "count": tot[0],
                             "count_expected": float(world * envs_per_rank * steps)}
        assert tot[0] == world * envs_per_rank * steps, "statistics count %r != N * E * steps = %r" % (tot[0], world * envs_per_rank * steps)
    k_ms, _ = kernel_time(wl, min(steps, 50))
    gbs = wl.bpe * envs_per_rank / (k_ms * 1e-3) / 1e9
    out.update(kernel_ms=k_ms, GBps=gbs, frac=gbs / peak, bytes_per_env_step=wl.bpe)
    tr = NCU_TRAFFIC.get((wl.hp.robot, wl.kind, envs_per_rank, args.layout))
    if tr:
        out["traffic"], out["traffic_source"] = tr
    wl.hp.close()
    del wl
    torch.cuda.empty_cache()
    return out


def configs_block(args, device, rank, world, peak):
    """BASELINE.json configs 3, 4, 5 and the 1 M-env row of the north-star target, measured in this very run."""
    steps, warmup = 20, 5
    blk = {}
    c3 = _variant_args(args, robot="humanoid3d", motion="spinkick,dance_b", combined=False, layout="aos", states="synthetic")
    blk["config3_spinkick_dance_b_65536_sharded"] = run_config(c3, device, rank, world, 65536 // world, steps, warmup, 300, True, peak)
    c4 = _variant_args(args, robot="unitree_g1", combined=True, layout="aos", states="synthetic")
    blk["config4_g1_combined_262144_sharded"] = run_config(c4, device, rank, world, 262144 // world, steps, warmup, 400, True, peak)
    if world > 1:
        blk["config4_g1_combined_262144_per_gpu"] = run_config(c4, device, rank, world, 262144, steps, warmup, 410, True, peak)
    h1 = _variant_args(args, robot="humanoid3d", motion="walk", combined=False, layout="aos", states="synthetic")
    blk["target_humanoid3d_walk_1M_per_gpu"] = run_config(h1, device, rank, world, 1 << 20, steps, warmup, 500, True, peak)
    blk["target_g1_combined_1M_per_gpu"] = run_config(c4, device, rank, world, 1 << 20, steps, warmup, 510, True, peak)
    sweep = []
    for e in range(10, 23):  # config 5: 2^10 .. 2^22 environments per GPU, both robots, every rank at once
        row = {"envs_per_gpu": 1 << e}
        for name, a in (("humanoid3d_walk", h1), ("g1_combined", c4)):
            r = run_config(a, device, rank, world, 1 << e, 10, 3, 600 + e, False, peak)
            row[name] = {k: r[k] for k in ("ms_per_step", "env_steps_per_s", "kernel_ms", "GBps", "frac")}
        sweep.append(row)
    blk["config5_sweep"] = sweep
    if args.layout == "aos":  # the SoA layout at 1 M environments next to the row-major one
        blk["layout_soa_humanoid3d_walk_1M"] = run_config(_variant_args(h1, layout="soa"), device, rank, world, 1 << 20, 10, 3, 700, False, peak)
        blk["layout_soa_g1_combined_1M"] = run_config(_variant_args(c4, layout="soa"), device, rank, world, 1 << 20, 10, 3, 710, False, peak)
    return blk


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

    wl = Workload(_variant_args(args, host_seeded=args.envs <= 65536), device, args.envs, STATE_SEED + rank, args.steps)
    coll = None
    if wl.stats and (world > 1 or args.collective_at_1) and args.collective != "off":
        coll = dmdist.StatsCollective(wl.hp, force_nccl=(args.collective == "nccl"))
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    ms, launches = timed_region(wl, args.steps, max(args.warmup, 3), world, args.allreduce_every, coll, args.collective)
    clocks = sampler.stop() if sampler else None
    value = world * args.envs * args.steps / (ms * 1e-3)
    stats_host = None
    if wl.stats:
        stats_host = coll.totals() if coll is not None else wl.hp.stats_read()
        want = (world if coll is not None else 1) * args.envs * args.steps
        assert stats_host[0] == want, "statistics count %r != %r" % (stats_host[0], want)

    # roofline of the dominant kernel at this workload
    k_ms, k_n = kernel_time(wl, min(args.steps, 200))
    bytes_per_launch = wl.bpe * args.envs
    ach = bytes_per_launch / (k_ms * 1e-3) / 1e9 if k_ms > 0 else 0.0
    roof = {"bound": "hbm", "kernel": "step kernel (K1)", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
            "traffic": NCU_TRAFFIC.get((wl.hp.robot, wl.kind, args.envs, args.layout), (None, None))[0],
            "traffic_source": NCU_TRAFFIC.get((wl.hp.robot, wl.kind, args.envs, args.layout), (None, None))[1], "peak_source": peak_src, "bytes_per_env_step": wl.bpe, "envs_per_launch": args.envs,
            "kernel_ms": k_ms, "kernel_launches_timed": k_n}

    line = {"metric": "env-steps/s (ref-lookup+FK+reward+obs+PD)", "value": value, "unit": "env-steps/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic" if args.states == "synthetic" else "recorded (synthetic-recorded episode_debug_log, tools/make_recording.py)",
            "config": {"workload": workload_name(args), "layout": args.layout, "state_batches": wl.B,
                       "l2_policy": "inputs larger than L2: %d distinct resident batches (%.0f MB) cycled" %
                                    (wl.B, wl.B * bytes_per_launch / 1e6),
                       "stats": wl.stats,
                       "launch_queue": "the K step launches are queued behind a device-side gate of %.0f us (torch.cuda._sleep) in front of "
                                       "the first event, so that the region times the device and not the host's launch ramp" % launch_gate_us(args.steps),
                       "collective": None if coll is None else {"mode": coll.mode, "issue": args.collective, "every": args.allreduce_every,
                                                                "why_not_peer": coll.why_not_peer}},
            "roofline": roof, "gpu_launches": int(launches), "clocks": clocks}
    if stats_host is not None:
        m = dmdist.moments(stats_host, wl.hp.obs_dim)
        line["stats"] = {"count": m["count"], "count_expected": float((world if coll is not None else 1) * args.envs * args.steps),
                         "scope": "all ranks" if coll is not None else "this rank", "reward_mean": m["reward_mean"], "done_rate": m["done_rate"]}
    if rank == 0 and not args.no_sweep:
        line["policy"] = policy_time(wl.hp, device, 1048576)
    wl.hp.close()
    del wl
    torch.cuda.empty_cache()
    if not args.no_configs and not args.no_sweep:  # every rank takes part (sharded workloads, the collective)
        blk = configs_block(args, device, rank, world, peak)
        if rank == 0:
            line["configs"] = blk
            # kept for readers of earlier rounds' lines: the two sizes the round-1 `sweep` key held
            line["sweep"] = [dict(envs=r["envs_per_gpu"], **r["humanoid3d_walk"]) for r in blk["config5_sweep"]
                             if r["envs_per_gpu"] in (65536, 1048576)]
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    if not args.no_e2e:  # every rank steps its own environments through the host API at the same time
        e2e = e2e_region(args, min(args.steps, 200), 20, 7, local_rank, world)
        if rank == 0:
            line["e2e"] = e2e
        if rank == 0 and world == 1 and "configs" in line:  # BASELINE configs[3] through the same host API
            c4 = _variant_args(args, robot="unitree_g1", combined=True, envs=262144)
            r = e2e_region(c4, 8, 3, 9, local_rank, world)
            line["configs"]["config4_g1_combined_262144_sharded"]["e2e"] = {k: r[k] for k in ("value", "unit", "ms_per_step", "h2d_bytes_per_step", "d2h_bytes_per_step")}
            line["configs"]["config4_g1_combined_262144_sharded"]["e2e"]["pinned"] = r["pinned"]
    if rank == 0:
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline(args, args.cpu_seconds)
            line["cpu_baseline"]["single_process"] = cpu_single_process(args, 10000, seconds=args.cpu_seconds)
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
