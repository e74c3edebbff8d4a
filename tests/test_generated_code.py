"""The generated headers under deepmimic_mujoco_b200/csrc/generated/ are what the committed generators emit from the
committed model constants (assets/models.json): nobody edited them by hand and nobody forgot to regenerate them.
CPU only."""
import os
import sys

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GEN = os.path.join(REPO, "deepmimic_mujoco_b200", "csrc", "generated")
sys.path.insert(0, os.path.join(REPO, "tools"))


def _models():
    from deepmimic_mujoco_b200.mjcf import load_models
    return load_models(os.path.join(REPO, "deepmimic_mujoco_b200", "assets", "models.json"))


@pytest.mark.parametrize("robot", ["humanoid3d", "unitree_g1"])
def test_robot_header_is_what_gen_model_code_emits(robot):
    import gen_model_code as g
    src = g.emit_robot(robot, _models()[robot], g.ROBOT_CFG[robot])
    assert src == open(os.path.join(GEN, "robot_%s.cuh" % robot)).read()


@pytest.mark.parametrize("robot", ["humanoid3d", "unitree_g1"])
def test_parts_header_is_what_gen_parts_code_emits(robot):
    import gen_model_code as g
    import gen_parts_code as gp
    src = gp.emit_parts(robot, _models()[robot], g.ROBOT_CFG[robot])
    assert src == open(os.path.join(GEN, "parts_%s.cuh" % robot)).read()
