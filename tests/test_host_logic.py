"""Host-side logic of the VecEnv mirror that needs no GPU: the lazily built `infos` list and the recycling of
objects a step hands out (deepmimic_mujoco_b200/vec_env.py)."""
import sys

import numpy as np

from deepmimic_mujoco_b200.vec_env import DPVecEnv, LazyInfos


def test_lazy_infos_is_a_list_of_dicts_and_recycles():
    x = LazyInfos(6)
    assert isinstance(x, list) and len(x) == 6
    x[2]["a"] = 1
    x[np.int64(4)]["b"] = 2  # indices arrive as numpy integers from np.nonzero
    assert x[2] == {"a": 1} and x[4] == {"b": 2} and x[0] == {}
    assert [d for d in x] == [{}, {}, {"a": 1}, {}, {"b": 2}, {}]
    assert x[1:3] == [{}, {"a": 1}]
    kept = x[2]
    y = x._recycle()
    assert y is x and len(x) == 6 and all(d == {} for d in x)
    assert kept == {"a": 1}  # a dict the caller still holds is the caller's


class _Ring:
    """The part of DPVecEnv that hands out step outputs, without a device handle."""
    _fresh = DPVecEnv._fresh

    def __init__(self, n):
        self._rings = {"obs": [], "infos": []}
        self._makers = {"obs": lambda: np.empty((n, 3), np.float32), "infos": lambda: LazyInfos(n)}


def test_step_outputs_are_reused_only_after_the_caller_dropped_them():
    r = _Ring(8)
    a = r._fresh("obs")
    b = r._fresh("obs")
    assert a is not b  # still held by the caller: never handed out twice
    ida = id(a)
    view = a[1:]
    del a
    c = r._fresh("obs")
    assert c is not b and id(c) != ida  # a view keeps the array alive: still the caller's
    del view, c
    held = [r._fresh("obs") for _ in range(6)]  # more live arrays than ring slots: plain fresh arrays
    assert len({id(h) for h in held}) == 6 and len(r._rings["obs"]) == 4
    i1 = r._fresh("infos")
    i1[3]["terminal_observation"] = 1
    del i1
    i2 = r._fresh("infos")
    assert len(i2) == 8 and i2[3] == {} and sys.getrefcount(i2) >= 2
