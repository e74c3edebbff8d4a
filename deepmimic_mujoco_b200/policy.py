"""ExtractedPolicy: the reference's numpy policy (src/extracted_policy.py:3-478: `act(obs)` -> action mean of a
66 -> 256 -> 128 -> 28 tanh MLP) evaluated by the fused device kernel (dm_policy_act) instead of three np.matmul calls,
optionally behind the observation normaliser.  `act` keeps the reference's signature: obs [N, in] -> a_mean [N, out].

`act_device` is the resident form: a CUDA tensor of observations (e.g. DPVecEnv's device obs buffer) in, a CUDA
tensor of actions out, nothing crosses PCIe."""
import numpy as np

from .hotpath import HotPath


class ExtractedPolicy(object):
    def __init__(self, W0, B0, W2, B2, WA, BA, hotpath=None, robot="humanoid3d", device=0):
        """Weights named as in the reference (W0/B0, W2/B2, WA/BA).  `hotpath`: share the handle of a DPVecEnv so the
        normaliser can read that env's running statistics; otherwise a private handle is created."""
        self.hp = hotpath if hotpath is not None else HotPath(robot, device=device)
        self._own = hotpath is None
        self.hp.policy_upload(W0, B0, W2, B2, WA, BA)
        self.obs_shape, self.act_shape = int(np.shape(W0)[0]), int(np.shape(WA)[1])
        self.normalize = False

    @classmethod
    def from_npz(cls, path, **kw):
        z = np.load(path)
        return cls(z["W0"], z["B0"], z["W2"], z["B2"], z["WA"], z["BA"], **kw)

    def set_normalizer(self, mean, var, eps=1e-8, clip_obs=10.0, ret_var=1.0, clip_reward=10.0):
        self.hp.policy_set_normalizer(mean, var, eps, clip_obs, ret_var, clip_reward)
        self.normalize = mean is not None

    def normalizer_from_stats(self, eps=1e-8, clip_obs=10.0, clip_reward=10.0):
        self.hp.policy_normalizer_from_stats(eps, clip_obs, clip_reward)
        self.normalize = True

    def act_device(self, obs, out=None, act_lo=-np.inf, act_hi=np.inf, obs_norm=None, reward=None, reward_norm=None):
        import torch
        n = obs.shape[0]
        assert obs.is_cuda and obs.dtype == torch.float32 and obs.is_contiguous() and obs.shape[1] == self.obs_shape
        if out is None:
            out = torch.empty((n, self.act_shape), dtype=torch.float32, device=obs.device)
        self.hp.policy_act(n, obs, out, normalize=self.normalize, obs_norm=obs_norm, reward=reward, reward_norm=reward_norm,
                           act_lo=act_lo, act_hi=act_hi)
        return out

    def act(self, obs):
        """obs [N, obs_shape] (numpy) -> a_mean [N, act_shape] (numpy float32)."""
        import torch
        x = torch.as_tensor(np.ascontiguousarray(obs, dtype=np.float32)).reshape(-1, self.obs_shape)
        dev = torch.device("cuda", self.hp.device)
        return self.act_device(x.to(dev)).cpu().numpy()

    def close(self):
        if self._own:
            self.hp.close()
