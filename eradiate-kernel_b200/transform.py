"""Transform4f as the reference keeps it: a matrix plus its inverse transpose, both in
the variant's scalar type (include/mitsuba/core/transform.h:20-330).  Compositions
multiply both members, so inverses of translate/scale/rotate/look_at chains are
analytic, as in the reference."""
import numpy as np


class Transform:
    def __init__(self, matrix=None, inverse_transpose=None, dtype=np.float32):
        self.dtype = np.dtype(dtype)
        if matrix is None:
            matrix = np.eye(4)
        self.matrix = np.asarray(matrix, dtype=self.dtype).reshape(4, 4).copy()
        if inverse_transpose is None:
            inverse_transpose = np.linalg.inv(self.matrix.astype(np.float64)).T
        self.inverse_transpose = np.asarray(inverse_transpose, dtype=self.dtype).reshape(4, 4).copy()

    # -- constructors (transform.h:150-330)
    @classmethod
    def translate(cls, v, dtype=np.float32):
        v = np.asarray(v, dtype=dtype)
        m = np.eye(4, dtype=dtype); m[:3, 3] = v
        it = np.eye(4, dtype=dtype); it[3, :3] = -v
        return cls(m, it, dtype)

    @classmethod
    def scale(cls, v, dtype=np.float32):
        v = np.asarray(v, dtype=dtype) * np.ones(3, dtype=dtype)
        m = np.diag(np.concatenate([v, np.ones(1, dtype=dtype)])).astype(dtype)
        it = np.diag(np.concatenate([(np.ones(3, dtype=dtype) / v), np.ones(1, dtype=dtype)])).astype(dtype)
        return cls(m, it, dtype)

    @classmethod
    def rotate(cls, axis, angle_deg, dtype=np.float32):
        dt = np.dtype(dtype).type
        a = np.asarray(axis, dtype=dtype)
        a = a / np.sqrt(np.dot(a, a))
        ang = dt(angle_deg) * dt(np.pi / 180.0)
        s, c = np.sin(ang), np.cos(ang)
        x, y, z = a
        t = dt(1) - c
        r = np.array([[t * x * x + c, t * x * y - s * z, t * x * z + s * y, 0],
                      [t * x * y + s * z, t * y * y + c, t * y * z - s * x, 0],
                      [t * x * z - s * y, t * y * z + s * x, t * z * z + c, 0],
                      [0, 0, 0, 1]], dtype=dtype)
        return cls(r, r.copy(), dtype)  # inverse = transpose, so inverse_transpose = r

    @classmethod
    def look_at(cls, origin, target, up, dtype=np.float32):
        """transform.h:241-270"""
        o = np.asarray(origin, dtype=dtype); t = np.asarray(target, dtype=dtype); up = np.asarray(up, dtype=dtype)

        def nrm(v):
            return (v / np.sqrt(np.dot(v, v))).astype(dtype)
        d = nrm(nrm(t - o))
        left = nrm(np.cross(up, d).astype(dtype))
        new_up = np.cross(d, left).astype(dtype)
        m = np.eye(4, dtype=dtype)
        m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = left, new_up, d, o
        inv = np.eye(4, dtype=dtype)
        inv[0, :3], inv[1, :3], inv[2, :3] = left, new_up, d
        inv[:, 3] = inv @ np.concatenate([-o, np.ones(1, dtype=dtype)]).astype(dtype)
        return cls(m, inv.T.copy(), dtype)

    # -- algebra
    def __matmul__(self, other):
        return Transform(self.matrix @ other.matrix, self.inverse_transpose @ other.inverse_transpose, self.dtype)

    def inverse(self):
        return Transform(self.inverse_transpose.T.copy(), self.matrix.T.copy(), self.dtype)

    def astype(self, dtype):
        return Transform(self.matrix.astype(dtype), self.inverse_transpose.astype(dtype), dtype)

    def point(self, p):
        p = np.asarray(p, dtype=self.dtype)
        r = self.matrix @ np.concatenate([p, np.ones(1, dtype=self.dtype)])
        return (r[:3] / r[3]).astype(self.dtype)

    def _fma(self, a, b, c):
        """enoki::fmadd on the variant's scalar type.  float32: the product is exact in float64 and the sum is
        rounded once more to float32; float64: numpy has no fused operation, a * b + c is used."""
        if self.dtype == np.float32:
            return (np.asarray(a, np.float64) * np.asarray(b, np.float64) + np.asarray(c, np.float64)).astype(np.float32)
        return a * b + c

    def point_affine(self, p):
        """transform.h:90-98: result = column 3; result = fmadd(column i, p[i], result) for i = 0, 1, 2.
        `p` may be one point or an array of points (n, 3)."""
        p = np.asarray(p, dtype=self.dtype)
        m = self.matrix
        r = np.broadcast_to(m[:3, 3], p.shape).astype(self.dtype)
        for i in range(3):
            r = self._fma(m[:3, i], p[..., i:i + 1], r)
        return r.astype(self.dtype)

    def vector(self, v):
        """transform.h:116-126: result = column 0 * v[0]; result = fmadd(column i, v[i], result) for i = 1, 2"""
        v = np.asarray(v, dtype=self.dtype)
        m = self.matrix
        r = (m[:3, 0] * v[..., 0:1]).astype(self.dtype)
        for i in (1, 2):
            r = self._fma(m[:3, i], v[..., i:i + 1], r)
        return r.astype(self.dtype)

    def normal(self, n):
        """transform.h:128-138: the same chain with the columns of the inverse transpose"""
        n = np.asarray(n, dtype=self.dtype)
        m = self.inverse_transpose
        r = (m[:3, 0] * n[..., 0:1]).astype(self.dtype)
        for i in (1, 2):
            r = self._fma(m[:3, i], n[..., i:i + 1], r)
        return r.astype(self.dtype)

    def colmajor(self):
        """16 doubles, m[c*4+r] (enoki::Matrix column storage)."""
        return [float(x) for x in self.matrix.T.reshape(-1)]


def coordinate_system(n, dtype=np.float32):
    """include/mitsuba/core/vector.h:116-136"""
    dt = np.dtype(dtype).type
    n = np.asarray(n, dtype=dtype)
    sign = dt(-1) if np.signbit(n[2]) else dt(1)
    a = -dt(1) / (sign + n[2])
    b = n[0] * n[1] * a
    ms = (lambda x, y: -x if np.signbit(y) else x)
    s = np.array([ms(n[0] * n[0] * a, n[2]) + dt(1), ms(b, n[2]), ms(-n[0], n[2])], dtype=dtype)
    t = np.array([b, sign + n[1] * n[1] * a, -n[1]], dtype=dtype)
    return s, t
