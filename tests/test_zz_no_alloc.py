"""dm_batch_stats_noalloc (include/dmhot.h): the statistics reduction whose scratch is a static device array -- no
cudaMalloc / cudaFree inside the call.  No counterpart in the reference.

Called through the C ABI (ctypes) and compared with the EXISTING entry points for the same sums
(dm_step(accumulate_stats = 1) + dm_stats_read) at the tolerance tests/test_gpu_batch.py uses for the statistics
(rtol 2e-6, atol 1e-6), on two calls in a row that share the static scratch: a large batch first, then a smaller one of
another robot (another observation width), so stale rows and stale columns of the first call would show.
Collected last on purpose (test_zz_*)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL, ATOL = 2e-6, 1e-6


def _existing(robot, motion, combined, n, seed):
    """One batched step with the fused statistics through the existing entry points; returns the handle's sums and the
    device tensors of the step's outputs (the inputs of the new entry point)."""
    import torch
    import bench
    from deepmimic_mujoco_b200.hotpath import HotPath
    hp = HotPath(robot, combined=combined)
    hp.load_combined() if combined else hp.load_motions([motion])
    dev = torch.device("cuda", 0)
    s = {k: v[0].contiguous() for k, v in bench.synth_states(hp, n, 1, seed, dev).items()}
    s["qpos"][::9, 2] -= 0.7  # some environments fall: finished episodes feed the n_done / ep_return / ep_len sums
    d = hp.dims
    out = dict(obs=torch.zeros(n, d.obs_dim, device=dev), reward=torch.zeros(n, device=dev),
               done=torch.zeros(n, dtype=torch.uint8, device=dev), ep_return=torch.rand(n, device=dev),
               idx=s["idx"].clone(), ep_len=s["ep_len"].clone(), clip=s["clip"].clone())
    hp.stats_zero()
    hp.step(n, s["qpos"], s["qvel"], s["action"], s["contact"], out["idx"], out["ep_len"], out["clip"], out["obs"],
            out["reward"], out["done"], None, None, None, out["ep_return"], stats=True)
    torch.cuda.synchronize()
    want = hp.stats_read()
    assert want[0] == n and (combined or want[3] > 0)
    return hp, out, want


def _no_alloc(out, n, obs_dim, result):
    from deepmimic_mujoco_b200 import _native as N
    L = N.lib()
    rc = L.dm_batch_stats_noalloc(0, n, obs_dim, out["obs"].data_ptr(), out["reward"].data_ptr(), out["done"].data_ptr(),
                                  out["ep_return"].data_ptr(), out["ep_len"].data_ptr(), result.data_ptr(), None)
    assert rc == 0, L.dm_last_error(None)


def test_no_alloc_statistics_match_the_existing_entry_point_twice_in_a_row():
    import torch
    dev = torch.device("cuda", 0)
    # every buffer exists BEFORE the calls under test; `result` starts as garbage that must be fully overwritten
    cases = [("unitree_g1", None, True, 2048 * 5 + 77, 11), ("humanoid3d", "walk", False, 128 * 3 + 5, 12)]
    prepared = [_existing(*c) for c in cases]
    results = [torch.full((6 + 2 * hp.obs_dim,), float("nan"), dtype=torch.float64, device=dev) for hp, _, _ in prepared]
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated(dev)
    for (hp, out, _), c, res in zip(prepared, cases, results):  # two calls in a row on the default stream
        _no_alloc(out, c[3], hp.obs_dim, res)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated(dev) == before
    for (hp, out, want), res in zip(prepared, results):
        got = res.cpu().numpy()
        assert got.shape == want.shape and np.isfinite(got).all()
        np.testing.assert_array_equal(got[[0, 3]], want[[0, 3]])  # counts are exact
        np.testing.assert_allclose(got, want, rtol=RTOL, atol=ATOL)
    # the same call again gives the same bits (fixed summation order, scratch fully rewritten)
    hp, out, want = prepared[0]
    again = torch.zeros_like(results[0])
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(dev)[0]  # kernels are loaded by now: any cudaMalloc inside the call would show here
    _no_alloc(out, cases[0][3], hp.obs_dim, again)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(dev)[0] == free0
    np.testing.assert_array_equal(again.cpu().numpy(), results[0].cpu().numpy())
    for hp, _, _ in prepared:
        hp.close()


def test_no_alloc_rejects_sizes_beyond_the_static_scratch():
    import torch
    from deepmimic_mujoco_b200 import _native as N
    L = N.lib()
    dev = torch.device("cuda", 0)
    x = torch.zeros(256, device=dev)
    d8 = torch.zeros(256, dtype=torch.uint8, device=dev)
    o = torch.zeros(512, dtype=torch.float64, device=dev)
    call = lambda n, dim: L.dm_batch_stats_noalloc(0, n, dim, x.data_ptr(), x.data_ptr(), d8.data_ptr(), None, None, o.data_ptr(), None)
    assert call(1, N.NOALLOC_MAX_OBS_DIM + 1) != 0
    assert b"DM_NOALLOC_MAX_OBS_DIM" in L.dm_last_error(None)
    assert call(N.NOALLOC_MAX_ENVS + 1, 67) != 0
    assert b"DM_NOALLOC_MAX_ENVS" in L.dm_last_error(None)
    assert call(-1, 67) != 0 and call(4, 0) != 0
    assert L.dm_batch_stats_noalloc(0, 4, 8, None, x.data_ptr(), d8.data_ptr(), None, None, o.data_ptr(), None) != 0
    # the largest accepted width works, and an empty batch gives zeros
    n, dim = 2, N.NOALLOC_MAX_OBS_DIM
    obs = torch.arange(n * dim, dtype=torch.float32, device=dev).reshape(n, dim)
    assert L.dm_batch_stats_noalloc(0, n, dim, obs.data_ptr(), x.data_ptr(), d8.data_ptr(), None, None, o.data_ptr(), None) == 0
    torch.cuda.synchronize()
    got = o.cpu().numpy()[:6 + 2 * dim]
    ob = obs.cpu().numpy().astype(np.float64)
    np.testing.assert_array_equal(got, np.concatenate([[n, 0, 0, 0, 0, 0], ob.sum(0), (ob * ob).sum(0)]))
    assert call(0, 67) == 0
    torch.cuda.synchronize()
    assert not o.cpu().numpy()[:6 + 2 * 67].any()
