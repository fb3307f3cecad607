"""ScalarTransform4f look-alike (include/mitsuba/core/transform.h:173-260): translate / rotate / scale /
look_at with the reference's chaining semantics (`T.translate(a).rotate(ax, deg).scale(s)` multiplies on the
right). Host-side scene description only; matrices are float32 like ScalarTransform4f."""
from __future__ import annotations

import numpy as np

f32 = np.float32


class Transform4f:
    def __init__(self, matrix=None):
        self.matrix = np.eye(4, dtype=f32) if matrix is None else np.array(getattr(matrix, "matrix", matrix), dtype=f32).reshape(4, 4)

    def __matmul__(self, other):
        if isinstance(other, Transform4f):
            return Transform4f((self.matrix.astype(np.float64) @ other.matrix.astype(np.float64)).astype(f32))
        v = np.asarray(other, dtype=f32)
        return self.transform_affine(v)

    __mul__ = __matmul__

    def transform_affine(self, p):
        """transform.h:104-111: start from the translation column, then one multiply-add per matrix column in float32 —
        product and sum rounded separately, as the reference's portable scalar build evaluates dr::fmadd (the build the
        oracle is pinned to; a build with hardware FMA differs in the last bit). Mesh loaders (ply.cpp:295, obj.cpp:237,
        cube.cpp:148) apply exactly this to every vertex, so vertices come out bit-equal."""
        p = np.asarray(p, f32)
        m = self.matrix
        r = np.broadcast_to(m[:3, 3], p.shape).astype(f32)
        for i in range(3):
            r = (m[:3, i] * p[..., i:i + 1]).astype(f32) + r
        return r.astype(f32)

    def transform_vector(self, v):
        """transform.h:128-137: column 0 times x, then multiply-adds of the other two columns."""
        v = np.asarray(v, f32)
        m = self.matrix
        r = (m[:3, 0] * v[..., 0:1]).astype(f32)
        for i in (1, 2):
            r = (m[:3, i] * v[..., i:i + 1]).astype(f32) + r
        return r.astype(f32)

    def inverse(self):
        return Transform4f(np.linalg.inv(self.matrix.astype(np.float64)).astype(f32))

    def has_scale(self):
        m = self.matrix[:3, :3].astype(np.float64)
        return not np.allclose(m @ m.T, np.eye(3), atol=1e-3)

    # ---- static constructors double as chaining methods (class access -> static, instance access -> chained)
    class _Chain:
        def __init__(self, fn):
            self.fn = fn

        def __get__(self, obj, cls):
            if obj is None:
                return lambda *a, **k: Transform4f(self.fn(*a, **k))
            return lambda *a, **k: obj @ Transform4f(self.fn(*a, **k))

    @staticmethod
    def _translate(v):
        m = np.eye(4, dtype=np.float64); m[:3, 3] = np.asarray(v, np.float64); return m

    @staticmethod
    def _scale(v):
        v = np.asarray(v, np.float64)
        if v.ndim == 0:
            v = np.repeat(v, 3)
        m = np.eye(4, dtype=np.float64); m[0, 0], m[1, 1], m[2, 2] = v; return m

    @staticmethod
    def _rotate(axis, angle):
        a = np.asarray(axis, np.float64); a = a / np.linalg.norm(a)
        t = np.radians(angle); c, s = np.cos(t), np.sin(t)
        x, y, z = a
        m = np.eye(4, dtype=np.float64)
        m[:3, :3] = [[c + x * x * (1 - c), x * y * (1 - c) - z * s, x * z * (1 - c) + y * s],
                     [y * x * (1 - c) + z * s, c + y * y * (1 - c), y * z * (1 - c) - x * s],
                     [z * x * (1 - c) - y * s, z * y * (1 - c) + x * s, c + z * z * (1 - c)]]
        return m

    @staticmethod
    def _look_at(origin, target, up):
        o, t, u = (np.asarray(x, np.float64) for x in (origin, target, up))
        d = t - o; d /= np.linalg.norm(d)
        left = np.cross(u, d); left /= np.linalg.norm(left)
        new_up = np.cross(d, left)
        m = np.eye(4, dtype=np.float64)
        m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = left, new_up, d, o
        return m

    translate = _Chain(_translate.__func__)
    scale = _Chain(_scale.__func__)
    rotate = _Chain(_rotate.__func__)
    look_at = _Chain(_look_at.__func__)

    def __repr__(self):
        return f"Transform4f(\n{self.matrix})"


ScalarTransform4f = Transform4f
