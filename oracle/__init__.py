"""oracle -- TEST INFRASTRUCTURE ONLY.  ctypes bindings for the two CPU checkers.

* ``oracle.port``  : oracle/liboracle.so, the plain-C restatement (oracle/la_oracle.c).
* ``oracle.ref``   : oracle/_ref/libref_kernels.so, the reference's own kernels.cl compiled
                     verbatim as C++ (oracle/build_ref.py); ``ref_fma`` is the same with
                     -ffp-contract=fast.  Present only where it was built (/root/reference).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  linear_algebra_b200 never does.
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

# (name, LA code, numpy dtype, OpenCL type name in kernels.cl / lib.rs:84-92)
DTYPES = [
    ("u8", 0, np.uint8, "uchar"),
    ("i8", 1, np.int8, "char"),
    ("u16", 2, np.uint16, "ushort"),
    ("i16", 3, np.int16, "short"),
    ("u32", 4, np.uint32, "uint"),
    ("i32", 5, np.int32, "int"),
    ("u64", 6, np.uint64, "ulong"),
    ("i64", 7, np.int64, "long"),
    ("f32", 8, np.float32, "float"),
    ("f64", 9, np.float64, "double"),
]
CODE = {n: c for n, c, _, _ in DTYPES}
NP = {n: d for n, _, d, _ in DTYPES}
CL = {n: t for n, _, _, t in DTYPES}
NAME_OF_NP = {np.dtype(d): n for n, _, d, _ in DTYPES}
ADD, SUB, MUL, DIV = 0, 1, 2, 3
WORK_GROUP_COUNT = 64  # lib.rs:114


def name_of(arr):
    return NAME_OF_NP[arr.dtype]


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def build(force=False):
    """Compile liboracle.so (and oracle/_ref when /root/reference is present)."""
    so = os.path.join(HERE, "liboracle.so")
    src = os.path.join(HERE, "la_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.run(["make", "-C", HERE, "liboracle.so"], check=True, stdout=subprocess.DEVNULL)
    if os.path.isdir("/root/reference"):
        ref = os.path.join(HERE, "_ref", "libref_kernels.so")
        if force or not os.path.exists(ref):
            subprocess.run(["python3", os.path.join(HERE, "build_ref.py")], check=True, stdout=subprocess.DEVNULL)


class Port:
    """The plain-C restatement (kind == "port")."""

    kind = "port"

    def __init__(self):
        build()
        self.lib = ctypes.CDLL(os.path.join(HERE, "liboracle.so"))
        L = ctypes.c_long
        V = ctypes.c_void_p
        I = ctypes.c_int
        self.lib.oracle_ew_vec_vec.argtypes = [I, I, V, V, V, L]
        self.lib.oracle_ew_assign_vec_vec.argtypes = [I, I, V, V, L]
        self.lib.oracle_ew_vec_scl.argtypes = [I, I, V, V, V, L]
        self.lib.oracle_ew_assign_vec_scl.argtypes = [I, I, V, V, L]
        self.lib.oracle_eq_vec.argtypes = [I, V, V, L]
        self.lib.oracle_reduce_groups.argtypes = [I, V, V, V, L, I, I]
        self.lib.oracle_reduce.argtypes = [I, V, V, L, I, I, V]
        self.lib.oracle_mul_vec_mat.argtypes = [I, V, V, V, L, L, I]
        self.lib.oracle_mul_vec_transpose_mat.argtypes = [I, V, V, V, L, L, I]
        self.lib.oracle_mul_mat_mat.argtypes = [I, V, V, V, L, L, L, I]
        self.lib.oracle_transpose.argtypes = [I, V, V, L, L]
        self.lib.oracle_f32_dot_truth.argtypes = [V, L, V, L, L, V, V]
        self.lib.oracle_f32_mul_mat_mat_truth.argtypes = [V, V, V, V, L, L, L]

    # ---- elementwise (kernels.cl:15-67)
    def ew_vec_vec(self, op, a, b):
        a = np.ascontiguousarray(a); b = np.ascontiguousarray(b, dtype=a.dtype)
        assert a.shape == b.shape
        c = np.empty_like(a)
        assert self.lib.oracle_ew_vec_vec(CODE[name_of(a)], op, _p(c), _p(a), _p(b), a.size) == 0
        return c

    def ew_assign_vec_vec(self, op, c, b):
        c = np.array(c, copy=True); b = np.ascontiguousarray(b, dtype=c.dtype)
        assert self.lib.oracle_ew_assign_vec_vec(CODE[name_of(c)], op, _p(c), _p(b), c.size) == 0
        return c

    def ew_vec_scl(self, op, a, s):
        a = np.ascontiguousarray(a); s = np.array([s], dtype=a.dtype)
        c = np.empty_like(a)
        assert self.lib.oracle_ew_vec_scl(CODE[name_of(a)], op, _p(c), _p(a), _p(s), a.size) == 0
        return c

    def ew_assign_vec_scl(self, op, c, s):
        c = np.array(c, copy=True); s = np.array([s], dtype=c.dtype)
        assert self.lib.oracle_ew_assign_vec_scl(CODE[name_of(c)], op, _p(c), _p(s), c.size) == 0
        return c

    # ---- eq (kernels.cl:71-74 + vector.rs:645-675: length mismatch is decided on the host)
    def eq_vec(self, a, b):
        a = np.ascontiguousarray(a); b = np.ascontiguousarray(b, dtype=a.dtype)
        if a.size != b.size:
            return False
        r = self.lib.oracle_eq_vec(CODE[name_of(a)], _p(a), _p(b), a.size)
        assert r >= 0
        return bool(r)

    # ---- sum / dot (kernels.cl:83-124 + vector.rs:256-293, 326-363)
    def reduce_groups(self, a, b=None, W=32, fma=False):
        a = np.ascontiguousarray(a)
        bp = None if b is None else _p(np.ascontiguousarray(b, dtype=a.dtype))
        res = np.empty(WORK_GROUP_COUNT, dtype=a.dtype)
        assert self.lib.oracle_reduce_groups(CODE[name_of(a)], _p(a), bp, _p(res), a.size, W, int(fma)) == 0
        return res

    def sum_vec(self, a, W=32):
        a = np.ascontiguousarray(a)
        out = np.zeros(1, dtype=a.dtype)
        assert self.lib.oracle_reduce(CODE[name_of(a)], _p(a), None, a.size, W, 0, _p(out)) == 0
        return out[0]

    def dot_vec_vec(self, a, b, W=32, fma=False):
        a = np.ascontiguousarray(a); b = np.ascontiguousarray(b, dtype=a.dtype)
        assert a.size == b.size
        out = np.zeros(1, dtype=a.dtype)
        assert self.lib.oracle_reduce(CODE[name_of(a)], _p(a), _p(b), a.size, W, int(fma), _p(out)) == 0
        return out[0]

    # ---- matrix-vector (kernels.cl:143-152), matrix-matrix (kernels.cl:161-169), transpose
    def mul_vec_mat(self, v, m, rows, cols, fma=False):
        v = np.ascontiguousarray(v); m = np.ascontiguousarray(m, dtype=v.dtype)
        assert v.size == rows and m.size == rows * cols
        c = np.empty(cols, dtype=v.dtype)
        assert self.lib.oracle_mul_vec_mat(CODE[name_of(v)], _p(c), _p(v), _p(m), rows, cols, int(fma)) == 0
        return c

    def mul_vec_transpose_mat(self, v, m, rows, cols, fma=False):
        v = np.ascontiguousarray(v); m = np.ascontiguousarray(m, dtype=v.dtype)
        assert v.size == cols and m.size == rows * cols
        c = np.empty(rows, dtype=v.dtype)
        assert self.lib.oracle_mul_vec_transpose_mat(CODE[name_of(v)], _p(c), _p(v), _p(m), rows, cols, int(fma)) == 0
        return c

    def mul_mat_mat(self, a, b, M, K, N, fma=False):
        a = np.ascontiguousarray(a); b = np.ascontiguousarray(b, dtype=a.dtype)
        assert a.size == M * K and b.size == K * N
        c = np.empty(M * N, dtype=a.dtype)
        assert self.lib.oracle_mul_mat_mat(CODE[name_of(a)], _p(c), _p(a), _p(b), M, K, N, int(fma)) == 0
        return c

    def transpose(self, a, R, C):
        a = np.ascontiguousarray(a)
        assert a.size == R * C
        out = np.empty(R * C, dtype=a.dtype)
        assert self.lib.oracle_transpose(CODE[name_of(a)], _p(out), _p(a), R, C) == 0
        return out

    # ---- f64 truth for f32 error bounds (not reference behaviour)
    def f32_dot_truth(self, a, b):
        a = np.ascontiguousarray(a, dtype=np.float32); b = np.ascontiguousarray(b, dtype=np.float32)
        d = ctypes.c_double(); t = ctypes.c_double()
        self.lib.oracle_f32_dot_truth(_p(a), 1, _p(b), 1, a.size, ctypes.byref(d), ctypes.byref(t))
        return d.value, t.value

    def f32_mul_mat_mat_truth(self, a, b, M, K, N):
        a = np.ascontiguousarray(a, dtype=np.float32); b = np.ascontiguousarray(b, dtype=np.float32)
        c = np.empty(M * N, dtype=np.float64); t = np.empty(M * N, dtype=np.float64)
        self.lib.oracle_f32_mul_mat_mat_truth(_p(c), _p(t), _p(a), _p(b), M, K, N)
        return c, t


class Ref:
    """The reference's kernels.cl compiled as C++ (kind == "reference").  Same method names as
    Port; the host-side halves of sum/dot/eq (vector.rs) are restated here in Python."""

    kind = "reference"

    def __init__(self, fma=False):
        name = "libref_kernels_fma.so" if fma else "libref_kernels.so"
        path = os.path.join(HERE, "_ref", name)
        if not os.path.exists(path):
            build()
        if not os.path.exists(path):
            raise FileNotFoundError(path + " (oracle/_ref is built only where /root/reference exists)")
        self.lib = ctypes.CDLL(path)

    def _f(self, name, arr):
        return getattr(self.lib, "ref_%s_%s" % (CL[name_of(arr)], name))

    def ew_vec_vec(self, op, a, b):
        a = np.ascontiguousarray(a); b = np.ascontiguousarray(b, dtype=a.dtype)
        c = np.empty_like(a)
        self._f("ew_vec_vec", a)(ctypes.c_int(op), _p(c), _p(a), _p(b), ctypes.c_long(a.size))
        return c

    def ew_assign_vec_vec(self, op, c, b):
        c = np.array(c, copy=True); b = np.ascontiguousarray(b, dtype=c.dtype)
        self._f("ew_assign_vec_vec", c)(ctypes.c_int(op), _p(c), _p(b), ctypes.c_long(c.size))
        return c

    def ew_vec_scl(self, op, a, s):
        a = np.ascontiguousarray(a); s = np.array([s], dtype=a.dtype)
        c = np.empty_like(a)
        self._f("ew_vec_scl", a)(ctypes.c_int(op), _p(c), _p(a), _p(s), ctypes.c_long(a.size))
        return c

    def ew_assign_vec_scl(self, op, c, s):
        c = np.array(c, copy=True); s = np.array([s], dtype=c.dtype)
        self._f("ew_assign_vec_scl", c)(ctypes.c_int(op), _p(c), _p(s), ctypes.c_long(c.size))
        return c

    def eq_vec(self, a, b):
        a = np.ascontiguousarray(a); b = np.ascontiguousarray(b, dtype=a.dtype)
        if a.size != b.size:          # vector.rs:649-651
            return False
        flag = np.array([1], dtype=np.uint8)   # vector.rs:658-661
        self._f("eq_vec", a)(_p(flag), _p(a), _p(b), ctypes.c_long(a.size))
        return bool(flag[0] != 0)     # vector.rs:673

    def reduce_groups(self, a, b=None, W=32, fma=False):
        a = np.ascontiguousarray(a)
        res = np.empty(WORK_GROUP_COUNT, dtype=a.dtype)
        if b is None:
            self._f("sum_vec", a)(_p(a), _p(res), ctypes.c_int(a.size), ctypes.c_int(WORK_GROUP_COUNT), ctypes.c_int(W))
        else:
            b = np.ascontiguousarray(b, dtype=a.dtype)
            self._f("dot_vec_vec", a)(_p(a), _p(b), _p(res), ctypes.c_int(a.size), ctypes.c_int(WORK_GROUP_COUNT), ctypes.c_int(W))
        return res

    @staticmethod
    def _host_sum(vals):
        # Rust `iter::sum()` in index order (vector.rs:269, 290, 339, 361), wrapping for integers.
        acc = vals.dtype.type(0)
        with np.errstate(over="ignore"):
            for x in vals:
                acc = vals.dtype.type(acc + x)
        return acc

    def sum_vec(self, a, W=32):
        a = np.ascontiguousarray(a)
        if a.size <= WORK_GROUP_COUNT:
            return self._host_sum(a)
        return self._host_sum(self.reduce_groups(a, None, W))

    def dot_vec_vec(self, a, b, W=32, fma=False):
        a = np.ascontiguousarray(a); b = np.ascontiguousarray(b, dtype=a.dtype)
        if a.size <= WORK_GROUP_COUNT:
            with np.errstate(over="ignore"):
                return self._host_sum(a * b)
        return self._host_sum(self.reduce_groups(a, b, W))

    def mul_vec_mat(self, v, m, rows, cols, fma=False):
        v = np.ascontiguousarray(v); m = np.ascontiguousarray(m, dtype=v.dtype)
        c = np.empty(cols, dtype=v.dtype)
        self._f("mul_vec_mat", v)(_p(c), _p(v), _p(m), ctypes.c_int(cols), ctypes.c_int(rows))
        return c

    def mul_vec_transpose_mat(self, v, m, rows, cols, fma=False):
        v = np.ascontiguousarray(v); m = np.ascontiguousarray(m, dtype=v.dtype)
        c = np.empty(rows, dtype=v.dtype)
        self._f("mul_vec_transpose_mat", v)(_p(c), _p(v), _p(m), ctypes.c_int(cols), ctypes.c_int(cols), ctypes.c_long(rows))
        return c

    def mul_mat_mat(self, a, b, M, K, N, fma=False):
        a = np.ascontiguousarray(a); b = np.ascontiguousarray(b, dtype=a.dtype)
        c = np.empty(M * N, dtype=a.dtype)
        self._f("mul_mat_mat", a)(_p(c), _p(a), _p(b), ctypes.c_int(N), ctypes.c_int(K), ctypes.c_long(M))
        return c


_port = None
_ref = {}


def port():
    global _port
    if _port is None:
        _port = Port()
    return _port


def ref(fma=False):
    if fma not in _ref:
        _ref[fma] = Ref(fma)
    return _ref[fma]


def ref_available():
    return os.path.exists(os.path.join(HERE, "_ref", "libref_kernels.so")) or os.path.isdir("/root/reference")
