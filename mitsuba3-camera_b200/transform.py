"""ScalarTransform4f look-alike (include/mitsuba/core/transform.h:173-260): translate / rotate / scale /
look_at with the reference's chaining semantics (`T.translate(a).rotate(ax, deg).scale(s)` multiplies on the
right). Host-side scene description only; matrices are float32 like ScalarTransform4f."""
from __future__ import annotations

import numpy as np

f32 = np.float32


def mm32(a, b):
    """4x4 product as Dr.Jit's scalar Matrix * Matrix evaluates it in float32 (ext/drjit/include/drjit/matrix.h:102-117): entry
    (r, j) = ((a[r,0] b[0,j] + a[r,1] b[1,j]) + a[r,2] b[2,j]) + a[r,3] b[3,j], every product and sum rounded (the reference's
    portable build does not fuse them)."""
    a = np.asarray(a, f32); b = np.asarray(b, f32)
    out = (a[:, 0:1] * b[0:1, :]).astype(f32)
    for i in range(1, 4):
        out = ((a[:, i:i + 1] * b[i:i + 1, :]).astype(f32) + out).astype(f32)
    return out


def inverse_f32(m):
    """Inverse of a float32 4x4 matrix with the operation order of the reference's own (Transform(matrix) ->
    dr::inverse_transpose, ext/drjit/include/drjit/matrix.h:384-443: 2x2 sub-determinants of column pairs, cofactor rows, one
    reciprocal of the determinant; every product and sum rounded to float32). Bit-exact against the pinned build on 201 matrices."""
    m = np.asarray(m, f32).reshape(4, 4)
    swap, half = [1, 0, 3, 2], [2, 3, 0, 1]
    a, b, c, d = m[:, 0].copy(), m[half, 1], m[:, 2].copy(), m[half, 3]
    mul = lambda x, y: (x * y).astype(f32)                                           # noqa: E731
    t = mul(c, d)[swap]; r0 = mul(b, t); r1 = mul(a, t)
    t = t[half]; r0 = mul(b, t) - r0; r1 = (mul(a, t) - r1)[half]
    t = mul(b, c)[swap]; r0 = mul(d, t) + r0; r3 = mul(a, t)
    t = t[half]; r0 = r0 - mul(d, t); r3 = (mul(a, t) - r3)[half]
    t = mul(b[half], d)[swap]; c = c[half]; r0 = mul(c, t) + r0; r2 = mul(a, t)
    t = t[half]; r0 = r0 - mul(c, t); r2 = (mul(a, t) - r2)[half]
    t = mul(a, b)[swap]; r2 = mul(d, t) + r2; r3 = mul(c, t) - r3
    t = t[half]; r2 = mul(d, t) - r2; r3 = r3 - mul(c, t)
    t = mul(a, d)[swap]; r1 = r1 - mul(c, t); r2 = mul(b, t) + r2
    t = t[half]; r1 = mul(c, t) + r1; r2 = r2 - mul(b, t)
    t = mul(a, c)[swap]; r1 = mul(d, t) + r1; r3 = r3 - mul(b, t)
    t = t[half]; r1 = r1 - mul(d, t); r3 = mul(b, t) + r3
    p = mul(a, r0)
    rdet = f32(1) / f32(f32(f32(p[0] + p[1]) + p[2]) + p[3])
    return np.stack([r0 * rdet, r1 * rdet, r2 * rdet, r3 * rdet], 0).astype(f32)


class Transform4f:
    """Carries the matrix AND its inverse, the latter built the way the reference builds it (transform.h:173-282): every
    constructor knows its exact inverse (translate(-v), scale(1/v), the transpose of a rotation, look_at's analytic inverse)
    and products multiply both sides. Scene descriptions are assembled in DOUBLE precision — the XML parser and `Properties`
    work with `Transform<Point<double, 4>>` (src/core/xml.cpp, include/mitsuba/core/properties.h) — and a plugin reads the
    result rounded once to float32, inverse included: `.matrix` / `.inverse_matrix` are those float32 arrays. A transform made
    from a bare matrix gets the float32 cofactor inverse the reference computes for it (`inverse_f32`)."""

    def __init__(self, matrix=None, inverse=None):
        if isinstance(matrix, Transform4f):
            matrix, inverse = matrix.m64, (matrix.inv64 if inverse is None else inverse)
        self.m64 = np.eye(4) if matrix is None else np.array(getattr(matrix, "matrix", matrix), dtype=np.float64).reshape(4, 4)
        self.inv64 = (inverse_f32(self.m64.astype(f32)).astype(np.float64) if inverse is None
                      else np.array(inverse, dtype=np.float64).reshape(4, 4))
        self.matrix = self.m64.astype(f32)
        self.inverse_matrix = self.inv64.astype(f32)

    def __matmul__(self, other):
        if isinstance(other, Transform4f):
            return Transform4f(self.m64 @ other.m64, other.inv64 @ self.inv64)
        v = np.asarray(other, dtype=f32)
        return self.transform_affine(v)

    __mul__ = __matmul__

    def matmul32(self, other):
        """The product as a PLUGIN forms it: ScalarTransform4f operands, float32 arithmetic (e.g. sphere.cpp:128-133
        `to_world * translate(center) * scale(radius)`), both for the matrix and for the inverse."""
        return Transform4f(mm32(self.matrix, other.matrix), mm32(other.inverse_matrix, self.inverse_matrix))

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
        return Transform4f(self.inv64, self.m64)

    def has_scale(self):
        m = self.matrix[:3, :3].astype(np.float64)
        return not np.allclose(m @ m.T, np.eye(3), atol=1e-3)

    # ---- static constructors double as chaining methods (class access -> static, instance access -> chained)
    class _Chain:
        def __init__(self, fn):
            self.fn = fn

        def __get__(self, obj, cls):
            if obj is None:
                return lambda *a, **k: Transform4f(*self.fn(*a, **k))
            return lambda *a, **k: obj @ Transform4f(*self.fn(*a, **k))

    # each constructor returns (matrix, inverse) as the reference's static constructors do, in double precision
    @staticmethod
    def _translate(v):
        v = np.asarray(v, np.float64).reshape(3)
        m = np.eye(4); m[:3, 3] = v
        inv = np.eye(4); inv[:3, 3] = -v
        return m, inv

    @staticmethod
    def _scale(v):
        v = np.asarray(v, np.float64)
        if v.ndim == 0:
            v = np.repeat(v, 3)
        m = np.eye(4); m[0, 0], m[1, 1], m[2, 2] = v
        inv = np.eye(4); inv[0, 0], inv[1, 1], inv[2, 2] = 1.0 / v
        return m, inv

    @staticmethod
    def _rotate(axis, angle):
        a = np.asarray(axis, np.float64); a = a / np.linalg.norm(a)
        t = np.radians(angle); c, s = np.cos(t), np.sin(t)
        x, y, z = a
        m = np.eye(4, dtype=np.float64)
        m[:3, :3] = [[c + x * x * (1 - c), x * y * (1 - c) - z * s, x * z * (1 - c) + y * s],
                     [y * x * (1 - c) + z * s, c + y * y * (1 - c), y * z * (1 - c) - x * s],
                     [z * x * (1 - c) - y * s, z * y * (1 - c) + x * s, c + z * z * (1 - c)]]
        return m, m.T.copy()                       # Transform(matrix, matrix): the inverse transpose IS the matrix

    @staticmethod
    def _look_at(origin, target, up):
        o, t, u = (np.asarray(x, np.float64) for x in (origin, target, up))
        d = t - o; d /= np.linalg.norm(d)
        left = np.cross(u, d); left /= np.linalg.norm(left)
        new_up = np.cross(d, left)
        m = np.eye(4, dtype=np.float64)
        m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = left, new_up, d, o
        # transform.h:272-279: inverse = transpose(rotation), its translation = inverse * (-origin, 1)
        inv = np.eye(4); inv[:3, :3] = m[:3, :3].T
        no = -o
        inv[:3, 3] = (inv[:3, 0] * no[0] + inv[:3, 1] * no[1]) + inv[:3, 2] * no[2]      # Matrix * Vector: column by column
        return m, inv

    translate = _Chain(_translate.__func__)
    scale = _Chain(_scale.__func__)
    rotate = _Chain(_rotate.__func__)
    look_at = _Chain(_look_at.__func__)

    def __repr__(self):
        return f"Transform4f(\n{self.matrix})"


ScalarTransform4f = Transform4f
