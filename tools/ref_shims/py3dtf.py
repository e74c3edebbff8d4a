"""py3dtf 0.2 stand-in: only Quaternion(x, y, z, w).to_rpy() (reference deepmimic_env.py:56,163,217,219)."""
from oracle.rotations import to_rpy


class Quaternion(object):
    def __init__(self, x, y, z, w):
        self.x, self.y, self.z, self.w = x, y, z, w

    def to_rpy(self):
        return to_rpy(self.x, self.y, self.z, self.w)
