"""Vector: host-side mirror of the reference's `Vector<T>` (src/vector.rs) over the C ABI.

Same names, argument meaning and error behaviour as the reference: shape mismatches raise
AssertionError where the reference's `assert_eq!` panics (vector.rs:97,115,327,513,545), every
operator blocks until the device has finished (vector.rs:108 `event.wait_for()`), out-of-place
operators allocate a fresh result (vector.rs:99).  Rust's by-value / by-reference operator pairs
(`&a + &b` vs `a + &b`) collapse to `a + b` (new vector) and `a += b` (in place)."""
import ctypes
import math

import numpy as np

from . import _lib
from ._lib import ADD, SUB, MUL, DIV, call
from .traits import Parameter, parameter

_SYNC = True      # the reference waits for every kernel; set_sync(False) makes ops stream-ordered
_STREAM = None    # None = the legacy default stream


def set_sync(flag):
    global _SYNC
    _SYNC = bool(flag)


def synchronize():
    call("la_stream_sync", _STREAM)


def _done():
    if _SYNC:
        call("la_stream_sync", _STREAM)


class _PinnedBlock:
    """Owner of one la_host_alloc block; numpy views keep it alive through `.base`."""

    def __init__(self, n, dt):
        p = ctypes.c_void_p()
        call("la_host_alloc", ctypes.byref(p), max(1, int(n)) * dt.size)
        self._p = p.value
        self.__array_interface__ = {"shape": (int(n),), "typestr": dt.np.str, "data": (self._p, False), "version": 3}

    def __del__(self):
        try:
            _lib.lib.la_host_free(ctypes.c_void_p(self._p))
        except Exception:
            pass


def pinned(n, dtype):
    """A numpy array of `n` elements in page-locked host memory (la_host_alloc): uploads and
    downloads from/to it run at full PCIe rate.  Freed when the last view is garbage collected."""
    return np.asarray(_PinnedBlock(n, parameter(dtype)))


class Event:
    """CUDA event on the library's stream (device-side timing)."""

    def __init__(self):
        p = ctypes.c_void_p()
        call("la_event_create", ctypes.byref(p))
        self._e = p.value

    def record(self):
        call("la_event_record", self._e, _STREAM)
        return self

    def elapsed_ms(self, stop):
        ms = ctypes.c_float()
        call("la_event_elapsed_ms", self._e, stop._e, ctypes.byref(ms))
        return ms.value

    def __del__(self):
        try:
            _lib.lib.la_event_destroy(ctypes.c_void_p(self._e))
        except Exception:
            pass


class Vector:
    """Owns one device buffer of `len` elements of `dtype` (vector.rs:23-27)."""

    __slots__ = ("dtype", "_ptr", "_len", "_owner")
    __hash__ = None

    def __init__(self, dtype, ptr, length, owner=True):
        self.dtype, self._ptr, self._len, self._owner = parameter(dtype), ptr, int(length), owner

    # ---- construction / transfer (vector.rs:30-94)
    @staticmethod
    def uninitialized(size, dtype):
        dt = parameter(dtype)
        p = ctypes.c_void_p()
        call("la_malloc", ctypes.byref(p), int(size) * dt.size)
        return Vector(dt, p.value, size)

    @staticmethod
    def new(default_val, size, dtype):
        res = Vector.uninitialized(size, dtype)
        val = np.array([default_val], dtype=res.dtype.np)
        call("la_fill", res.dtype.code, res._ptr, val.ctypes.data, res._len, _STREAM)
        _done()
        return res

    @staticmethod
    def synth(seed, size, dtype):
        """Device-generated synthetic data (la_synth); see linear_algebra_b200.synth for the host mirror."""
        res = Vector.uninitialized(size, dtype)
        call("la_synth", res.dtype.code, res._ptr, int(seed), res._len, _STREAM)
        _done()
        return res

    @staticmethod
    def from_vec(v, dtype=None):
        host = np.ascontiguousarray(v, dtype=None if dtype is None else parameter(dtype).np).reshape(-1)
        dt = parameter(dtype if dtype is not None else host.dtype)
        res = Vector.uninitialized(host.size, dt)
        call("la_upload", res._ptr, host.ctypes.data, host.nbytes, _STREAM)
        call("la_stream_sync", _STREAM)          # host memory may be pageable / temporary
        return res

    def to_vec(self, out=None):
        """Blocking download (vector.rs:71-78).  `out`: optional preallocated (e.g. pinned) array."""
        if out is None:
            out = np.empty(self._len, dtype=self.dtype.np)
        assert out.size == self._len and out.dtype == self.dtype.np and out.flags.c_contiguous
        call("la_download", out.ctypes.data, self._ptr, out.nbytes, _STREAM)
        return out

    def slice_to_vec(self, start, count):
        """Download elements [start, start+count) only (spot checks of operands too big to bring back whole)."""
        assert 0 <= start and start + count <= self._len
        out = np.empty(count, dtype=self.dtype.np)
        call("la_download", out.ctypes.data, self._ptr + start * self.dtype.size, out.nbytes, _STREAM)
        return out

    def len(self):
        return self._len

    __len__ = len

    @property
    def ptr(self):
        """Raw device pointer (replaces get_buffer / get_buf, vector.rs:153-159, 247-253)."""
        return self._ptr

    def clone(self):                              # vector.rs:700-708
        res = Vector.uninitialized(self._len, self.dtype)
        call("la_copy", res._ptr, self._ptr, self._len * self.dtype.size, _STREAM)
        _done()
        return res

    def __del__(self):
        if getattr(self, "_owner", False) and getattr(self, "_ptr", None):
            try:
                _lib.lib.la_free(ctypes.c_void_p(self._ptr))
            except Exception:
                pass
            self._ptr = None

    # ---- user kernels (vector.rs:80-151); `kernel` is a linear_algebra_b200.kernel.Kernel whose argument slots
    #      are named "C" (result / self), "A", "B" as the reference's helpers expect
    @staticmethod
    def generate(kernel, size, dtype=None):                       # vector.rs:80-90
        res = Vector.uninitialized(size, dtype if dtype is not None else kernel.dtype)
        kernel.set_arg_buf_named("C", res)
        kernel.enq(res._len, _STREAM)
        call("la_stream_sync", _STREAM)
        return res

    @staticmethod
    def from_for_each2(a, b, kernel):                             # vector.rs:96-112 (the `_lock_free` body)
        assert a._len == b._len, "assertion failed: `(left == right)` left: `%d`, right: `%d`" % (a._len, b._len)
        res = Vector.uninitialized(a._len, a.dtype)
        kernel.set_arg_buf_named("C", res)
        kernel.set_arg_buf_named("A", a)
        kernel.set_arg_buf_named("B", b)
        kernel.enq(res._len, _STREAM)
        call("la_stream_sync", _STREAM)
        return res

    def for_each_mut(self, other, kernel):                        # vector.rs:114-125
        assert self._len == other._len, "assertion failed: `(left == right)` left: `%d`, right: `%d`" % (self._len, other._len)
        kernel.set_arg_buf_named("C", self)
        kernel.set_arg_buf_named("B", other)
        kernel.enq(self._len, _STREAM)
        call("la_stream_sync", _STREAM)
        return self

    def map_mut(self, kernel):                                    # vector.rs:127-136
        kernel.set_arg_buf_named("C", self)
        kernel.enq(self._len, _STREAM)
        call("la_stream_sync", _STREAM)
        return self

    def map(self, kernel):                                        # vector.rs:138-151
        res = Vector.uninitialized(self._len, self.dtype)
        kernel.set_arg_buf_named("C", res)
        kernel.set_arg_buf_named("B", self)
        kernel.enq(res._len, _STREAM)
        call("la_stream_sync", _STREAM)
        return res

    # ---- helpers (vector.rs:96-125)
    def _check_same(self, other):
        assert isinstance(other, Vector), "expected a Vector"
        assert self.dtype.code == other.dtype.code, "element types differ"
        assert self._len == other._len, "assertion failed: `(left == right)` left: `%d`, right: `%d`" % (self._len, other._len)

    def _from_for_each2(self, other, op):
        self._check_same(other)
        res = Vector.uninitialized(self._len, self.dtype)
        call("la_ew_vec_vec", self.dtype.code, op, res._ptr, self._ptr, other._ptr, self._len, _STREAM)
        _done()
        return res

    def _for_each_mut(self, other, op):
        self._check_same(other)
        call("la_ew_assign_vec_vec", self.dtype.code, op, self._ptr, other._ptr, self._len, _STREAM)
        _done()
        return self

    def _scalar(self, s):
        return np.array([s], dtype=self.dtype.np)

    def _vec_scl(self, op, s):
        res = Vector.uninitialized(self._len, self.dtype)
        sv = self._scalar(s)
        call("la_ew_vec_scl", self.dtype.code, op, res._ptr, self._ptr, sv.ctypes.data, self._len, _STREAM)
        _done()
        return res

    def _assign_vec_scl(self, op, s):
        sv = self._scalar(s)
        call("la_ew_assign_vec_scl", self.dtype.code, op, self._ptr, sv.ctypes.data, self._len, _STREAM)
        _done()
        return self

    # ---- operators (vector.rs:366-643)
    def __add__(self, other):
        return self._from_for_each2(other, ADD)

    def __iadd__(self, other):
        return self._for_each_mut(other, ADD)

    def __sub__(self, other):
        return self._from_for_each2(other, SUB)

    def __isub__(self, other):
        return self._for_each_mut(other, SUB)

    def __mul__(self, other):
        from .matrix import Matrix
        if isinstance(other, Matrix):             # Vector * Matrix, vector.rs:508-539
            return _mul_vec_mat(self, other)
        if isinstance(other, Vector):             # Hadamard, vector.rs:482-496
            return self._from_for_each2(other, MUL)
        return self._vec_scl(MUL, other)          # scale, vector.rs:430-456

    def __imul__(self, other):
        if isinstance(other, Vector):
            return self._for_each_mut(other, MUL)
        return self._assign_vec_scl(MUL, other)

    def __truediv__(self, other):
        if isinstance(other, Vector):
            return self._from_for_each2(other, DIV)
        return self._vec_scl(DIV, other)

    def __itruediv__(self, other):
        if isinstance(other, Vector):
            return self._for_each_mut(other, DIV)
        return self._assign_vec_scl(DIV, other)

    def __eq__(self, other):                      # vector.rs:645-675
        if not isinstance(other, Vector):
            return NotImplemented
        if self._len != other._len or self.dtype.code != other.dtype.code:
            return False
        eq = ctypes.c_int(0)
        call("la_eq_vec", self.dtype.code, self._ptr, other._ptr, self._len, ctypes.byref(eq), _STREAM)
        return eq.value != 0

    def __ne__(self, other):
        r = self.__eq__(other)
        return r if r is NotImplemented else not r

    # ---- reductions (vector.rs:256-323)
    def sum(self):
        out = np.zeros(1, dtype=self.dtype.np)
        call("la_sum_vec", self.dtype.code, self._ptr, self._len, out.ctypes.data, _STREAM)
        return out[0]

    def length(self):                             # sqrt(dot(self, self)), vector.rs:303-305
        assert self.dtype.real, "length() needs a Real element type (traits.rs:76-78)"
        return self.dtype.np.type(math.sqrt(dot(self, self))) if self.dtype.np == np.float64 else np.sqrt(dot(self, self))

    def normalize(self):                          # consumes self in Rust; here: in place, returns self
        assert self.dtype.real, "normalize() needs a Real element type (traits.rs:76-78)"
        call("la_normalize_vec", self.dtype.code, self._ptr, self._ptr, self._len, _STREAM)     # dot, sqrt, 1/len, scale on the device
        _done()
        return self

    def normalized(self):
        assert self.dtype.real, "normalized() needs a Real element type (traits.rs:76-78)"
        res = Vector.uninitialized(self._len, self.dtype)
        call("la_normalize_vec", self.dtype.code, res._ptr, self._ptr, self._len, _STREAM)
        _done()
        return res

    # ---- persistence (vector.rs:183-245, util.rs:6-21): native-endian u64 element size, u64 element
    # count, raw data.  Files written by the reference load here and vice versa.
    def write_to_file(self, f):
        f.write(np.array([self.dtype.size, self._len], dtype=np.uint64).tobytes())
        f.write(self.to_vec().tobytes())

    @staticmethod
    def read_from_file(f, dtype):
        dt = parameter(dtype)
        elem_size, count = (int(x) for x in np.frombuffer(_read_exact(f, 16), dtype=np.uint64))
        if elem_size != dt.size:       # vector.rs:215-220
            raise AssertionError("Elem size from buffer does not seem to match, what was expected!Missmatch in endiannes?")
        return Vector.from_vec(np.frombuffer(_read_exact(f, count * dt.size), dtype=dt.np), dt)

    def save(self, path):
        with open(path, "wb") as f:
            self.write_to_file(f)

    @staticmethod
    def open(path, dtype):
        with open(path, "rb") as f:
            return Vector.read_from_file(f, dtype)

    # ---- formatting (vector.rs:678-698)
    def __str__(self):                            # Display, vector.rs:678-687: every element
        return "{ " + ", ".join(repr(x.item()) for x in self.to_vec()) + " }"

    def __repr__(self):
        """Debug form.  Long vectors are abbreviated (a failing `assert a == b` in a test must not format 2^28 elements)."""
        if self._len <= 64:
            return self.__str__()
        head = ", ".join(repr(x.item()) for x in self.slice_to_vec(0, 8))
        tail = ", ".join(repr(x.item()) for x in self.slice_to_vec(self._len - 4, 4))
        return "{ %s, ... (%d elements of %s) ..., %s }" % (head, self._len, self.dtype.name, tail)


def _read_exact(f, n):
    data = f.read(n)
    if len(data) != n:
        raise EOFError("Failed to read all data from file")       # vector.rs:244
    return data


def dot(a, b):                                    # vector.rs:326-363
    a._check_same(b)
    out = np.zeros(1, dtype=a.dtype.np)
    call("la_dot_vec_vec", a.dtype.code, a._ptr, b._ptr, a._len, out.ctypes.data, _STREAM)
    return out[0]


def _mul_vec_mat(vector, other_m):                # vector.rs:508-539
    assert vector._len == other_m.get_row_count(), "assertion failed: `(left == right)` left: `%d`, right: `%d`" % (vector._len, other_m.get_row_count())
    assert vector.dtype.code == other_m.dtype.code, "element types differ"
    res = Vector.uninitialized(other_m.get_col_count(), vector.dtype)
    call("la_mul_vec_mat", vector.dtype.code, res._ptr, vector._ptr, other_m.data._ptr,
         other_m.get_row_count(), other_m.get_col_count(), _STREAM)
    _done()
    return res


def mul_transpose_mat(vector, other_m):           # vector.rs:542-569
    """vector * transpose(other_m)  ==  other_m x vector (as a column)."""
    assert vector._len == other_m.get_col_count(), "assertion failed: `(left == right)` left: `%d`, right: `%d`" % (vector._len, other_m.get_col_count())
    assert vector.dtype.code == other_m.dtype.code, "element types differ"
    res = Vector.uninitialized(other_m.get_row_count(), vector.dtype)
    call("la_mul_vec_transpose_mat", vector.dtype.code, res._ptr, vector._ptr, other_m.data._ptr,
         other_m.get_row_count(), other_m.get_col_count(), _STREAM)
    _done()
    return res
