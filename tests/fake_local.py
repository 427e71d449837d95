"""A numpy stand-in for linear_algebra_b200's Vector/Matrix, used ONLY by tests/test_dist_gloo.py
to drive the sharding bookkeeping of linear_algebra_b200/dist.py on CPU under gloo, together with a host model
(`ShardedEngine`) of what liblinalg_b200's la_*_sharded entry points do across ranks: partials exchanged as bytes and
combined in rank order with the element type's own arithmetic (csrc/comm.cu: combine_host and the fused kernels).
It is test scaffolding, not a fallback: the product never imports it."""
import numpy as np

from linear_algebra_b200.traits import parameter


def _wrap(x, dt):
    with np.errstate(over="ignore"):
        return np.asarray(x).astype(dt.np)


class Vector:
    def __init__(self, data, dtype):
        self.dtype = parameter(dtype)
        self.data = np.ascontiguousarray(data, dtype=self.dtype.np).reshape(-1)

    @staticmethod
    def from_vec(v, dtype=None):
        v = np.asarray(v)
        return Vector(v, dtype if dtype is not None else v.dtype)

    def to_vec(self):
        return self.data.copy()

    def len(self):
        return self.data.size

    def _b(self, o):
        return o.data if isinstance(o, Vector) else self.dtype.np.type(o)

    def __add__(self, o):
        with np.errstate(over="ignore"):
            return Vector(_wrap(self.data + self._b(o), self.dtype), self.dtype)

    def __sub__(self, o):
        with np.errstate(over="ignore"):
            return Vector(_wrap(self.data - self._b(o), self.dtype), self.dtype)

    def __mul__(self, o):
        if isinstance(o, Matrix):
            with np.errstate(over="ignore"):
                acc = np.zeros(o.col_count, self.dtype.np)
                for r in range(o.row_count):         # row order == the device's strip order for tiny inputs
                    acc = _wrap(acc + self.data[r] * o.data.data[r * o.col_count:(r + 1) * o.col_count], self.dtype)
            return Vector(acc, self.dtype)
        with np.errstate(over="ignore"):
            return Vector(_wrap(self.data * self._b(o), self.dtype), self.dtype)

    def __truediv__(self, o):
        b = self._b(o)
        r = self.data / b if self.dtype.np.kind == "f" else np.trunc(self.data.astype(np.float64) / np.asarray(b, np.float64))
        return Vector(_wrap(r, self.dtype), self.dtype)

    def __iadd__(self, o):
        self.data = (self + o).data
        return self

    def __isub__(self, o):
        self.data = (self - o).data
        return self

    def __imul__(self, o):
        self.data = (self * o).data
        return self

    def __eq__(self, o):
        return self.data.size == o.data.size and bool(np.all(self.data == o.data))

    __hash__ = None

    def sum(self):
        with np.errstate(over="ignore"):
            return self.data.sum(dtype=self.dtype.np)


def dot(a, b):
    with np.errstate(over="ignore"):
        return (a.data * b.data).sum(dtype=a.dtype.np)


class Matrix:
    def __init__(self, data, rows, cols):
        self.data, self.row_count, self.col_count = data, rows, cols
        self.dtype = data.dtype

    @staticmethod
    def from_vec(v, rows, cols, dtype=None):
        return Matrix(Vector.from_vec(v, dtype), rows, cols)

    def get_row_count(self):
        return self.row_count

    def get_col_count(self):
        return self.col_count

    def to_vec(self):
        return self.data.to_vec()

    def _m(self):
        return self.data.data.reshape(self.row_count, self.col_count)

    def __add__(self, o):
        return Matrix(self.data + o.data, self.row_count, self.col_count)

    def __sub__(self, o):
        return Matrix(self.data - o.data, self.row_count, self.col_count)

    def __mul__(self, o):
        if isinstance(o, Matrix):
            with np.errstate(over="ignore"):
                c = (self._m().astype(np.int64) @ o._m().astype(np.int64)) if self.dtype.np.kind != "f" else self._m() @ o._m()
            return Matrix(Vector(_wrap(c, self.dtype), self.dtype), self.row_count, o.col_count)
        return Matrix(self.data * o, self.row_count, self.col_count)


def mul_transpose_mat(v, m):
    with np.errstate(over="ignore"):
        r = (m._m().astype(np.int64) @ v.data.astype(np.int64)) if v.dtype.np.kind != "f" else m._m() @ v.data
    return Vector(_wrap(r, v.dtype), v.dtype)


def synchronize():
    pass


def combine_scalars(parts, np_dtype):
    """Rank-ordered sum of per-rank partials with the element type's own arithmetic (integers wrap, f32 rounds after every
    add): what la_sum_vec_sharded / la_dot_vec_vec_sharded deliver on every rank."""
    dt = np.dtype(np_dtype)
    acc = dt.type(0)
    with np.errstate(over="ignore"):
        for p in parts:
            acc = dt.type(acc + np.frombuffer(p, dtype=dt, count=1)[0])
    return acc


class ShardedEngine:
    """Host model of the cross-rank steps (linear_algebra_b200/dist.py asks the local module for it when that module is not
    the real one)."""

    def __init__(self, comm):
        self.comm = comm

    def scalar(self, a, b):
        part = a.sum() if b is None else dot(a, b)
        return combine_scalars(self.comm.all_gather_host(np.asarray(part).tobytes()), a.dtype.np)

    def eq(self, a, b):
        mine = bool(a == b)
        return all(p == b"\x01" for p in self.comm.all_gather_host(b"\x01" if mine else b"\x00"))

    def vec_mat(self, v, mat, cols):
        partial = v * mat
        parts = self.comm.all_gather_host(partial.to_vec().tobytes())
        dt = partial.dtype.np
        acc = np.frombuffer(parts[0], dtype=dt).copy()
        with np.errstate(over="ignore"):
            for p in parts[1:]:
                acc = (acc + np.frombuffer(p, dtype=dt)).astype(dt)
        return Vector.from_vec(acc, partial.dtype)
