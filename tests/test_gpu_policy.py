"""Fused normaliser + policy MLP kernel (dm_policy_act) against the reference's ExtractedPolicy: its own
known-answer test (extracted_policy.py:480-485), its float64 outputs on 512 seeded observations
(tests/golden/extracted_policy.npz, written by tools/gen_policy_golden.py from the reference's Python), and the
oracle restatement for other shapes and for the normaliser.  Tolerance 1e-5 relative in fp32.  An action is a sum
of 128 products of O(1) magnitude on top of two more such layers, so "relative" is taken against the action scale:
|a - b| <= 1e-5 * max(|b|, 1) (fp32 rounding of the 66- and 256-term dot products alone reaches 1e-6 for the
unnormalised N(0, 2) half of the seeded observations); the reference's own np.allclose criterion is asserted as well
on its known-answer input."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "extracted_policy.npz")


def _close(got, want, rel=1e-5, floor=1.0):
    err = np.abs(got - want)
    tol = rel * np.maximum(np.abs(want), floor)
    assert np.all(err <= tol), "max err %.3e (tol %.3e) at %s" % (err.max(), tol.flat[err.argmax()], np.unravel_index(err.argmax(), err.shape))


VARIANTS = [0, 1]  # 0 = fp32 CUDA cores (K6a), 1 = tcgen05 tensor cores with fp16 hi/lo operand splitting (K6b)


@pytest.mark.parametrize("variant", VARIANTS)
def test_reference_known_answer_and_seeded_outputs(variant):
    from deepmimic_mujoco_b200.policy import ExtractedPolicy
    z = np.load(GOLD)
    pol = ExtractedPolicy.from_npz(GOLD)
    pol.hp.set_policy_variant(variant)
    assert (pol.obs_shape, pol.act_shape) == (66, 28)
    _close(pol.act(z["kat_obs"]), z["kat_expected"])
    assert np.allclose(pol.act(z["kat_obs"]), z["kat_expected"], rtol=1e-5, atol=1e-6)  # the reference's own criterion
    _close(pol.act(z["obs"]), z["act"])
    pol.close()


@pytest.mark.parametrize("variant", VARIANTS)
@pytest.mark.parametrize("dims,n", [((67, 256, 128, 28), 128 * 5 + 37), ((85, 256, 128, 23), 300), ((98, 200, 96, 23), 1000),
                                    ((3, 16, 8, 1), 129), ((128, 256, 128, 32), 128 * 150)])
def test_other_shapes_match_oracle(dims, n, variant):
    import torch
    from deepmimic_mujoco_b200.hotpath import HotPath
    from oracle.policy import mlp_act
    rng = np.random.default_rng(sum(dims))
    i, h0, h1, o = dims
    W = [rng.normal(0, 1.0 / np.sqrt(a), (a, b)).astype(np.float32) for a, b in ((i, h0), (h0, h1), (h1, o))]
    B = [rng.normal(0, 0.2, b).astype(np.float32) for b in (h0, h1, o)]
    hp = HotPath("humanoid3d")
    hp.set_policy_variant(variant)
    hp.policy_upload(W[0], B[0], W[1], B[1], W[2], B[2])
    obs = rng.normal(0, 1.5, (n, i)).astype(np.float32)
    want = mlp_act(obs.astype(np.float64), *[x.astype(np.float64) for x in (W[0], B[0], W[1], B[1], W[2], B[2])])
    for off in (0, 1):  # 16-byte aligned base (TMA tiles) and a base that is not (cooperative loads)
        buf = torch.zeros(n * i + 4, dtype=torch.float32, device="cuda")
        x = buf[off:off + n * i].view(n, i)
        x.copy_(torch.from_numpy(obs))
        act = torch.full((n, o), np.nan, dtype=torch.float32, device="cuda")
        hp.policy_act(n, x, act)
        _close(act.cpu().numpy(), want)
    hp.close()


@pytest.mark.parametrize("variant", VARIANTS)
def test_normaliser_reward_scaling_and_action_clip(variant):
    import torch
    from deepmimic_mujoco_b200.hotpath import HotPath
    from oracle import policy as OP
    z = np.load(GOLD)
    W = [z[k] for k in ("W0", "B0", "W2", "B2", "WA", "BA")]
    rng = np.random.default_rng(3)
    n = 128 * 3 + 5
    obs = (rng.normal(0, 3.0, (n, 66)) + rng.normal(0, 2, 66)).astype(np.float32)
    obs[7, 3] = 1e6  # far outside +-clip_obs
    rew = rng.normal(0.5, 0.3, n).astype(np.float32)
    mean, var = obs.mean(0).astype(np.float32), obs.var(0).astype(np.float32)
    hp = HotPath("humanoid3d")
    hp.set_policy_variant(variant)
    hp.policy_upload(*W)
    hp.policy_set_normalizer(mean, var, eps=1e-8, clip_obs=5.0, ret_var=0.09, clip_reward=2.0)
    d = torch.device("cuda")
    x, r = torch.from_numpy(obs).to(d), torch.from_numpy(rew).to(d)
    act, xn, rn = torch.empty(n, 28, device=d), torch.empty(n, 66, device=d), torch.empty(n, device=d)
    hp.policy_act(n, x, act, normalize=True, obs_norm=xn, reward=r, reward_norm=rn, act_lo=-1.0, act_hi=1.0)
    wn = OP.normalize_obs(obs.astype(np.float64), mean.astype(np.float64), var.astype(np.float64), 1e-8, 5.0)
    _close(xn.cpu().numpy(), wn, floor=0.5)
    _close(rn.cpu().numpy(), OP.normalize_reward(rew.astype(np.float64), 0.09, 1e-8, 2.0), floor=0.05)
    # the network sees the kernel's own fp32 normalised observations
    want = np.clip(OP.mlp_act(xn.cpu().numpy().astype(np.float64), *[w.astype(np.float64) for w in W]), -1.0, 1.0)
    _close(act.cpu().numpy(), want)
    assert np.abs(xn.cpu().numpy()).max() == 5.0 and np.abs(act.cpu().numpy()).max() == 1.0
    hp.close()


def test_normaliser_from_running_statistics():
    """The step kernel accumulates the moments, the normaliser kernel turns them into mean / inv_std on the device, the
    policy consumes the step's observation buffer in place: nothing visits the host."""
    import torch
    import bench
    from deepmimic_mujoco_b200.hotpath import HotPath
    from oracle import policy as OP
    hp = HotPath("humanoid3d")
    hp.load_motions(["walk"])
    n = 128 * 8
    s = bench.synth_states(hp, n, 1, 4, torch.device("cuda", 0))
    s = {k: v[0].contiguous() for k, v in s.items()}
    d = torch.device("cuda")
    obs, rew, done = torch.zeros(n, 67, device=d), torch.zeros(n, device=d), torch.zeros(n, dtype=torch.uint8, device=d)
    hp.stats_zero()
    hp.step(n, s["qpos"], s["qvel"], s["action"], s["contact"], s["idx"], s["ep_len"], s["clip"], obs, rew, done, stats=True)
    rng = np.random.default_rng(0)
    W = [rng.normal(0, 0.1, sh).astype(np.float32) for sh in ((67, 256), (256,), (256, 128), (128,), (128, 28), (28,))]
    hp.policy_upload(*W)
    hp.policy_normalizer_from_stats(eps=1e-4, clip_obs=10.0, clip_reward=10.0)
    act, xn, rn = torch.empty(n, 28, device=d), torch.empty(n, 67, device=d), torch.empty(n, device=d)
    hp.policy_act(n, obs, act, normalize=True, obs_norm=xn, reward=rew, reward_norm=rn)
    torch.cuda.synchronize()
    mean, var, _, r_var = OP.moments_from_stats(hp.stats_read(), 67)
    o64 = obs.cpu().numpy().astype(np.float64)
    np.testing.assert_allclose(mean, o64.mean(0), rtol=1e-6, atol=1e-6)
    _close(xn.cpu().numpy(), OP.normalize_obs(o64, mean, var, 1e-4, 10.0), rel=2e-5, floor=0.5)
    _close(rn.cpu().numpy(), OP.normalize_reward(rew.cpu().numpy().astype(np.float64), r_var, 1e-4, 10.0), rel=2e-5, floor=0.05)
    want = OP.mlp_act(xn.cpu().numpy().astype(np.float64), *[w.astype(np.float64) for w in W])
    _close(act.cpu().numpy(), want)
    hp.close()


def test_policy_argument_errors():
    import torch
    from deepmimic_mujoco_b200 import _native as N
    from deepmimic_mujoco_b200.hotpath import HotPath
    hp = HotPath("humanoid3d")
    x = torch.zeros(4, 66, device="cuda")
    a = torch.zeros(4, 28, device="cuda")
    with pytest.raises(N.DmError, match="no policy uploaded"):
        hp.policy_act(4, x, a)
    with pytest.raises(N.DmError, match="supported sizes"):
        hp.policy_upload(np.zeros((66, 512)), np.zeros(512), np.zeros((512, 256)), np.zeros(256), np.zeros((256, 28)), np.zeros(28))
    hp.policy_upload(np.zeros((66, 256)), np.zeros(256), np.zeros((256, 128)), np.zeros(128), np.zeros((128, 28)), np.zeros(28))
    hp.policy_act(0, None, None)  # empty batch: no-op
    with pytest.raises(N.DmError, match="go together"):
        hp.policy_act(4, x, a, reward=torch.zeros(4, device="cuda"))
    hp.close()


def test_policy_in_the_resident_loop():
    """DPVecEnv.step_tensors -> ExtractedPolicy.act_device on the env's own device buffers and handle, against the
    oracle MLP on the observations the env produced."""
    import torch
    from deepmimic_mujoco_b200.policy import ExtractedPolicy
    from deepmimic_mujoco_b200.vec_env import DPVecEnv
    from oracle.policy import mlp_act
    n = 128 * 4 + 9
    env = DPVecEnv(n, motion="walk", robot="humanoid3d", seed=3)
    rng = np.random.default_rng(1)
    W = [rng.normal(0, 0.1, sh).astype(np.float32) for sh in ((67, 256), (256,), (256, 128), (128,), (128, 28), (28,))]
    pol = ExtractedPolicy(*W, hotpath=env.hp)
    obs = env.reset_tensors()
    b = env.device_buffers()
    for _ in range(3):
        act = pol.act_device(obs, act_lo=-2.0, act_hi=2.0)
        want = np.clip(mlp_act(obs.cpu().numpy().astype(np.float64), *[w.astype(np.float64) for w in W]), -2.0, 2.0)
        _close(act.cpu().numpy(), want)
        obs, rew, done = env.step_tensors(act, b["qpos"], b["qvel"])  # kinematic replay of the reference states
        assert torch.isfinite(obs).all() and torch.isfinite(rew).all()
    env.close()
