"""Host-side logic of the VecEnv mirror that needs no GPU: the lazily built `infos` sequence, the ring of caller-owned
output arrays (deepmimic_mujoco_b200/vec_env.py) and the conditional stable_baselines3 base class."""
import importlib
import sys
import types

import numpy as np

from deepmimic_mujoco_b200.vec_env import LazyInfos, _OutputRing


def test_lazy_infos_is_a_sequence_of_dicts():
    x = LazyInfos(6)
    assert len(x) == 6
    x[2]["a"] = 1
    x[np.int64(4)]["b"] = 2  # indices arrive as numpy integers from np.nonzero
    assert x[2] == {"a": 1} and x[4] == {"b": 2} and x[0] == {} and x[-2] == {"b": 2}
    assert [d for d in x] == [{}, {}, {"a": 1}, {}, {"b": 2}, {}]
    assert x[1:3] == [{}, {"a": 1}]
    assert x == [{}, {}, {"a": 1}, {}, {"b": 2}, {}] and list(x) == [{}, {}, {"a": 1}, {}, {"b": 2}, {}]
    assert x.touched() == [0, 1, 2, 3, 4, 5]  # the comparisons above looked at every entry
    y = LazyInfos(1 << 20)  # O(1): nothing is built until it is looked at
    y[12345]["k"] = 1
    assert y.touched() == [12345]
    try:
        x[6]
        assert False
    except IndexError:
        pass
    x[0] = {"replaced": True}
    assert x[0] == {"replaced": True}


def test_output_ring_reuses_a_slot_only_after_every_view_of_it_is_gone():
    r = _OutputRing((8, 3), np.float32, 4, pinned=False)
    a = r.get()
    a[:] = 1.0
    b = r.get()
    assert a.ctypes.data != b.ctypes.data  # still held by the caller: never handed out twice
    pa = a.ctypes.data
    view = a[1:]          # a derived view ...
    flat = a.reshape(-1)  # ... and a reshape keep the slot the caller's
    del a
    c = r.get()
    assert c.ctypes.data not in (pa, b.ctypes.data)
    del view
    d = r.get()
    assert d.ctypes.data != pa
    del flat  # nothing of the first hand-out is left
    e = r.get()
    assert e.ctypes.data == pa
    held = [r.get() for _ in range(6)]  # more live arrays than slots: plain fresh arrays, never a live slot
    ptrs = {h.ctypes.data for h in held} | {b.ctypes.data, c.ctypes.data, d.ctypes.data, e.ctypes.data}
    assert len(ptrs) == 10


def test_output_ring_sees_a_torch_alias():
    import torch
    r = _OutputRing((4, 2), np.float32, 1, pinned=False)
    a = r.get()
    pa = a.ctypes.data
    t = torch.from_numpy(a)  # SB3's obs_as_tensor
    del a
    assert r.get().ctypes.data != pa
    del t
    assert r.get().ctypes.data == pa


def test_dpvecenv_is_an_sb3_vecenv_when_sb3_is_importable():
    """SB3 wraps whatever is not an `isinstance(env, VecEnv)` in a DummyVecEnv (reference src/sb3_ppo.py:275-278,307 hands
    PPO a SubprocVecEnv).  stable_baselines3 is not installed here, so a stub package exposing the ABC stands in."""
    import abc

    class VecEnv(abc.ABC):  # the constructor signature and abstract surface of stable_baselines3 1.x / 2.x
        def __init__(self, num_envs, observation_space, action_space):
            self.num_envs, self.observation_space, self.action_space = num_envs, observation_space, action_space
            self.render_mode = None

        for _m in ("reset", "step_async", "step_wait", "close", "get_attr", "set_attr", "env_method", "env_is_wrapped"):
            locals()[_m] = abc.abstractmethod(lambda self, *a, **k: None)
        del _m

    names = ["stable_baselines3", "stable_baselines3.common", "stable_baselines3.common.vec_env"]
    saved = {k: sys.modules.get(k) for k in names}
    try:
        for k in names:
            sys.modules[k] = types.ModuleType(k)
        sys.modules[names[2]].VecEnv = VecEnv
        import deepmimic_mujoco_b200.vec_env as ve
        ve2 = importlib.reload(ve)
        assert issubclass(ve2.DPVecEnv, VecEnv)
        assert not getattr(ve2.DPVecEnv, "__abstractmethods__", None)  # every abstract method is implemented
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
        import deepmimic_mujoco_b200.vec_env as ve
        importlib.reload(ve)
    assert ve.DPVecEnv.__mro__[1] is object
