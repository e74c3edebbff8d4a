"""pyquaternion 0.9.9 stand-in: the operations the reference's loader uses
(mocap_util.py:31-40, mocap_v2.py:350-362)."""
import numpy as np

from oracle import rotations as R


class Quaternion(object):
    def __init__(self, *args, **kw):
        if "matrix" in kw:
            self.q = R.pq_from_matrix(kw["matrix"])
        else:
            self.q = np.array([float(a) for a in args], dtype=np.float64)

    def __mul__(self, other):
        out = Quaternion(1, 0, 0, 0)
        out.q = R.pq_mul(self.q, other.q)
        return out

    @property
    def conjugate(self):
        out = Quaternion(1, 0, 0, 0)
        out.q = R.pq_conj(self.q)
        return out

    @property
    def elements(self):
        return self.q

    @property
    def axis(self):
        return R.pq_axis(self.q)

    @property
    def angle(self):
        return R.pq_angle(self.q)
