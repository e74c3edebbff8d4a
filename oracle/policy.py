"""Policy inference and the observation / reward normaliser restated on the CPU (oracle; test infrastructure).

`mlp_act` follows the reference's src/extracted_policy.py:471-478 (ExtractedPolicy.act): two tanh layers and a
linear head, float64 numpy.  Pinned by tests/golden/extracted_policy.npz (the reference's own known-answer test and
its outputs on seeded observations, written by tools/gen_policy_golden.py).

The normaliser has no counterpart in the reference (its sb3_ppo.py trains on raw observations); it restates the
Stable-Baselines3 VecNormalize arithmetic the SURVEY names as the next consumer:
    obs'    = clip((obs - mean) / sqrt(var + eps), -clip_obs, clip_obs)
    reward' = clip(reward / sqrt(ret_var + eps), -clip_reward, clip_reward)
PARITY UNPINNED for the normaliser (SB3 is not installed here)."""
import numpy as np


def mlp_act(obs, W0, B0, W2, B2, WA, BA):
    f1 = np.tanh(np.matmul(obs, W0) + B0)
    f = np.tanh(np.matmul(f1, W2) + B2)
    return np.matmul(f, WA) + BA


def normalize_obs(obs, mean, var, eps=1e-8, clip_obs=10.0):
    return np.clip((obs - mean) / np.sqrt(var + eps), -clip_obs, clip_obs)


def normalize_reward(reward, ret_var, eps=1e-8, clip_reward=10.0):
    return np.clip(reward / np.sqrt(ret_var + eps), -clip_reward, clip_reward)


def moments_from_stats(stats, obs_dim):
    """mean / variance of the observation columns and of the reward from the step kernels' running sums
    [count, sum r, sum r^2, n_done, sum ep_return, sum ep_len, sum obs[D], sum obs^2[D]] (population variance, as
    SB3's RunningMeanStd keeps it)."""
    s = np.asarray(stats, dtype=np.float64)
    n = max(s[0], 1.0)
    mean = s[6:6 + obs_dim] / n
    var = np.maximum(s[6 + obs_dim:6 + 2 * obs_dim] / n - mean * mean, 0.0)
    r_mean = s[1] / n
    r_var = max(s[2] / n - r_mean * r_mean, 0.0)
    return mean, var, r_mean, r_var
