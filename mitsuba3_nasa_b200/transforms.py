"""ScalarTransform4f mirror (include/mitsuba/core/transform.h:173-330 of the reference).

Host-side scene set-up only; matrices are built in float64 and handed to the
C ABI as float32 row-major 4x4.
"""
import math
import numpy as np


class ScalarTransform4f:
    def __init__(self, matrix=None):
        if matrix is None:
            self.matrix = np.eye(4)
        elif isinstance(matrix, ScalarTransform4f):
            self.matrix = matrix.matrix.copy()
        else:
            m = np.asarray(matrix, dtype=np.float64)
            if m.shape == (3,):  # ScalarTransform4f([sx, sy, sz]) idiom used in the docs = scale
                self.matrix = np.diag([m[0], m[1], m[2], 1.0])
            else:
                self.matrix = m.reshape(4, 4).copy()

    # -- constructors (transform.h:173-330) --
    @staticmethod
    def translate(v):
        m = np.eye(4)
        m[:3, 3] = np.asarray(v, dtype=np.float64)
        return ScalarTransform4f(m)

    @staticmethod
    def scale(v):
        v = np.asarray(v, dtype=np.float64)
        if v.ndim == 0:
            v = np.array([v, v, v])
        return ScalarTransform4f(np.diag([v[0], v[1], v[2], 1.0]))

    @staticmethod
    def rotate(axis, angle):
        axis = np.asarray(axis, dtype=np.float64)
        axis = axis / np.linalg.norm(axis)
        a = math.radians(angle)
        s, c = math.sin(a), math.cos(a)
        x, y, z = axis
        m = np.eye(4)
        m[:3, :3] = np.array([
            [c + (1 - c) * x * x, (1 - c) * x * y - s * z, (1 - c) * x * z + s * y],
            [(1 - c) * y * x + s * z, c + (1 - c) * y * y, (1 - c) * y * z - s * x],
            [(1 - c) * z * x - s * y, (1 - c) * z * y + s * x, c + (1 - c) * z * z]])
        return ScalarTransform4f(m)

    @staticmethod
    def look_at(origin, target, up):
        origin, target, up = (np.asarray(a, dtype=np.float64) for a in (origin, target, up))
        d = target - origin
        d = d / np.linalg.norm(d)
        left = np.cross(up, d)
        left = left / np.linalg.norm(left)
        new_up = np.cross(d, left)
        m = np.eye(4)
        m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = left, new_up, d, origin
        return ScalarTransform4f(m)

    @staticmethod
    def perspective(fov, near, far):
        recip = 1.0 / (far - near)
        cot = 1.0 / math.tan(math.radians(fov * 0.5))
        m = np.zeros((4, 4))
        m[0, 0] = cot
        m[1, 1] = cot
        m[2, 2] = far * recip
        m[2, 3] = -near * far * recip
        m[3, 2] = 1.0
        return ScalarTransform4f(m)

    # -- algebra --
    def __matmul__(self, other):
        if isinstance(other, ScalarTransform4f):
            return ScalarTransform4f(self.matrix @ other.matrix)
        v = np.asarray(other, dtype=np.float64)
        r = self.matrix @ np.append(v, 1.0)
        return r[:3] / r[3]

    __mul__ = __matmul__

    def inverse(self):
        return ScalarTransform4f(np.linalg.inv(self.matrix))

    def transform_affine(self, p):
        return self.matrix[:3, :3] @ np.asarray(p, dtype=np.float64) + self.matrix[:3, 3]

    def transform_vector(self, v):
        return self.matrix[:3, :3] @ np.asarray(v, dtype=np.float64)

    def translation(self):
        return self.matrix[:3, 3].copy()

    def f32(self):
        return self.matrix.astype(np.float32)

    def __repr__(self):
        return "ScalarTransform4f(%s)" % np.array2string(self.matrix)


Transform4f = ScalarTransform4f


def coordinate_system(n):
    """include/mitsuba/core/vector.h:116-136 (host copy for set-up code)."""
    n = np.asarray(n, dtype=np.float64)
    sign = 1.0 if n[2] >= 0 else -1.0
    a = -1.0 / (sign + n[2])
    b = n[0] * n[1] * a
    s = np.array([sign * n[0] * n[0] * a + 1.0, sign * b, -sign * n[0]])
    t = np.array([b, n[1] * n[1] * a + sign, -n[1]])
    return s, t
