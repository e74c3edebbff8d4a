"""Edge cases of the C ABI on the GPU: empty and ragged batches, NaN / Inf / huge inputs (numpy's NaN-propagating
max / min in the observation-bounds guard, deepmimic_env.py:465-476), misaligned buffers, argument errors, and the
device RNG of reference-state initialisation."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-5, 5e-7


def _step_arrays(hp, n, seed=0):
    rng = np.random.default_rng(seed)
    m = hp.mocaps[0]
    L = m.get_length()
    f = rng.integers(0, L, n)
    q = (np.asarray(m.data_config)[f] + rng.normal(0, 0.03, (n, hp.nq))).astype(np.float32)
    v = (np.asarray(m.data_vel)[f] + rng.normal(0, 0.3, (n, hp.nv))).astype(np.float32)
    a = rng.uniform(-1, 1, (n, hp.na)).astype(np.float32)
    return f.astype(np.int32), q, v, a


def _host_step(hp, idx, q, v, a, contact=None):
    n = len(idx)
    out = dict(idx=idx.copy(), ep_len=np.zeros(n, np.int32), obs=np.zeros((n, hp.obs_dim), np.float32), reward=np.zeros(n, np.float32),
               done=np.zeros(n, np.uint8), reason=np.zeros(n, np.uint8), torque=np.zeros((n, hp.nu), np.float32))
    hp.step_host(n, q, v, a, contact, out["idx"], out["ep_len"], None, out["obs"], out["reward"], out["done"], out["reason"], out["torque"])
    return out


@pytest.fixture(scope="module")
def hp():
    from deepmimic_mujoco_b200.hotpath import HotPath
    h = HotPath("humanoid3d")
    h.load_motions(["walk"])
    return h


def test_empty_batch_is_a_no_op(hp):
    import torch
    z = lambda *s, dt=torch.float32: torch.zeros(*s, dtype=dt, device="cuda")
    k0 = hp.launch_count()
    hp.step(0, z(0, 35), z(0, 34), z(0, 28), None, z(0, dt=torch.int32), z(0, dt=torch.int32), None, z(0, 67), z(0), z(0, dt=torch.uint8))
    assert hp.launch_count() == k0
    idx, q, v, a = _step_arrays(hp, 0)
    _host_step(hp, idx, q, v, a)


@pytest.mark.parametrize("n", [1, 31, 127, 128, 129, 257, 4096 + 5])
def test_ragged_sizes_agree_with_the_oracle_path(hp, n):
    """Full 128-env tiles go to K1b, the tail to K1a: any n gives the same rows as n = 1 calls would."""
    idx, q, v, a = _step_arrays(hp, n, seed=n)
    big = _host_step(hp, idx, q, v, a)
    pick = sorted(set([0, n - 1, n // 2, min(n - 1, 127), min(n - 1, 128)]))
    for i in pick:
        one = _host_step(hp, idx[i:i + 1], q[i:i + 1], v[i:i + 1], a[i:i + 1])
        for k in big:
            np.testing.assert_array_equal(big[k][i], one[k][0], err_msg="%s row %d of %d" % (k, i, n))


def test_nan_inf_and_huge_inputs_follow_the_reference(hp):
    from oracle.envs import SingleClipEnv
    n = 256
    idx, q, v, a = _step_arrays(hp, n, seed=3)
    q[3, 10] = np.nan            # NaN joint angle: np.max(obs) > 100 is False -> no guard, NaN observation
    v[5, 20] = np.inf            # inf velocity -> obs 0.1 * inf > 100 -> guard
    v[7, 9] = 2000.0             # finite but out of bounds -> guard (zero obs, zero reward, done)
    q[9, 2] = 50.0               # root far above the floor -> high_z termination, no guard
    a[11, 4] = np.nan            # NaN action passes through mju_clip
    a[13, :] = 1e9               # clipped to ctrlrange
    got = _host_step(hp, idx, q, v, a)
    env = SingleClipEnv("walk", "humanoid3d")
    for i in (0, 3, 5, 7, 9, 11, 13, 200):
        env.idx_curr, env.episode_length = int(idx[i]), 0
        with np.errstate(all="ignore"):
            obs, rew, done, info = env.step(a[i].astype(np.float64), (q[i].astype(np.float64), v[i].astype(np.float64)), 0)
        assert bool(got["done"][i]) == bool(done), i
        np.testing.assert_allclose(got["obs"][i], obs, rtol=RTOL, atol=ATOL, equal_nan=True, err_msg=str(i))
        np.testing.assert_allclose(got["reward"][i], rew, rtol=RTOL, atol=ATOL, equal_nan=True, err_msg=str(i))
        np.testing.assert_allclose(got["torque"][i], env.torque, rtol=RTOL, atol=ATOL, equal_nan=True, err_msg=str(i))
    assert got["reason"][5] == 8 and got["reason"][7] == 8 and not got["obs"][7].any()
    assert got["reason"][9] == 2 and np.isnan(got["obs"][3]).any() and np.isnan(got["torque"][11, 4])


def test_misaligned_device_buffers_fall_back_to_the_strided_kernel(hp):
    import torch
    n = 256
    idx, q, v, a = _step_arrays(hp, n, seed=8)
    dev = torch.device("cuda")
    ref = _host_step(hp, idx, q, v, a)
    pad = torch.zeros(n * 35 + 1, device=dev)
    qd = pad[1:].view(n, 35)  # 4-byte aligned only: no TMA
    qd.copy_(torch.from_numpy(q))
    t = lambda x: torch.from_numpy(x).to(dev)
    obs, rew, done = torch.zeros(n, 67, device=dev), torch.zeros(n, device=dev), torch.zeros(n, dtype=torch.uint8, device=dev)
    i2, e2 = t(idx.copy()), torch.zeros(n, dtype=torch.int32, device=dev)
    hp.step(n, qd, t(v), t(a), None, i2, e2, None, obs, rew, done)
    torch.cuda.synchronize()
    np.testing.assert_array_equal(obs.cpu().numpy(), ref["obs"])
    np.testing.assert_array_equal(rew.cpu().numpy(), ref["reward"])


def test_argument_errors_are_reported_not_thrown(hp):
    from deepmimic_mujoco_b200 import _native as N
    from deepmimic_mujoco_b200.hotpath import HotPath
    L = N.lib()
    a = N.StepArgs()
    a.n = 4
    assert L.dm_step(hp.h, C.byref(a)) == -1 and b"NULL" in L.dm_last_error(hp.h)
    a.n = -1
    assert L.dm_step(hp.h, C.byref(a)) == -1
    assert L.dm_step(None, C.byref(a)) == -1
    h = C.c_void_p()
    assert L.dm_create(0, 7, 0, C.byref(h)) == -1 and b"unknown robot" in L.dm_last_error(None)
    assert L.dm_create(0, 0, 1, C.byref(h)) == -1  # DPCombinedEnv exists for unitree_g1 only
    assert L.dm_create(99, 0, 0, C.byref(h)) == -1
    fresh = HotPath("humanoid3d")
    idx, q, v, x = _step_arrays(hp, 4)
    with pytest.raises(N.DmError, match="no clips uploaded"):
        fresh.step_host(4, q, v, x, None, idx, np.zeros(4, np.int32), None, np.zeros((4, 67), np.float32), np.zeros(4, np.float32),
                        np.zeros(4, np.uint8))
    assert L.dm_upload_clips(fresh.h, 0, None) == -1
    fresh.close()
    with pytest.raises(N.DmError):
        HotPath("unitree_g1", combined=True).step_host(4, q, v, x, None, idx, idx, None, q, q, q)  # no clip buffer / no clips


def test_reset_rng_is_uniform_reproducible_and_counter_based():
    from deepmimic_mujoco_b200.vec_env import DPVecEnv
    n = 76 * 400
    env = DPVecEnv(n, seed=123)
    env.reset()
    a = env.idx.copy()
    assert a.min() == 0 and a.max() == 75
    cnt = np.bincount(a, minlength=76)
    assert abs(cnt - 400).max() < 6 * np.sqrt(400)  # uniform over the clip (random.randint(0, len - 1), deepmimic_env.py:313)
    np.testing.assert_array_equal(env.qpos, np.asarray(env.hp.mocaps[0].data_config, dtype=np.float32)[a])  # set_state(frame)
    env.reset()
    b = env.idx.copy()
    assert (a != b).mean() > 0.9  # the counter advanced
    env2 = DPVecEnv(n, seed=123)
    env2.reset()
    np.testing.assert_array_equal(env2.idx, a)  # same seed, same counter -> same draw
    env3 = DPVecEnv(n, seed=124)
    env3.reset()
    assert (env3.idx != a).mean() > 0.9
    for e in (env, env2, env3):
        e.close()
