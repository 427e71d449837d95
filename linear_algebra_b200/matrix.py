"""Matrix: host-side mirror of the reference's `Matrix<T>` (src/matrix.rs): a Vector of
rows*cols elements plus the two counts, dense row-major (matrix.rs:9-13)."""
from . import vector as _v
from ._lib import call
from .vector import Vector


class Matrix:
    __slots__ = ("data", "row_count", "col_count")
    __hash__ = None

    def __init__(self, data, row_count, col_count):
        self.data, self.row_count, self.col_count = data, int(row_count), int(col_count)

    # ---- construction (matrix.rs:16-66)
    @staticmethod
    def new(default_value, row_count, col_count, dtype):
        return Matrix(Vector.new(default_value, row_count * col_count, dtype), row_count, col_count)

    @staticmethod
    def uninitialized(row_count, col_count, dtype):
        return Matrix(Vector.uninitialized(row_count * col_count, dtype), row_count, col_count)

    @staticmethod
    def from_vec(v, row_count, col_count, dtype=None):
        data = Vector.from_vec(v, dtype)
        assert data.len() == row_count * col_count, "assertion failed: `(left == right)` left: `%d`, right: `%d`" % (data.len(), row_count * col_count)
        return Matrix(data, row_count, col_count)

    @property
    def dtype(self):
        return self.data.dtype

    def len(self):
        return (self.row_count, self.col_count)

    def get_row_count(self):
        return self.row_count

    def get_col_count(self):
        return self.col_count

    def to_vec(self, out=None):
        return self.data.to_vec(out)

    def clone(self):
        return Matrix(self.data.clone(), self.row_count, self.col_count)

    def _check_shape(self, other):
        assert isinstance(other, Matrix), "expected a Matrix"
        assert self.row_count == other.row_count, "assertion failed: `(left == right)` left: `%d`, right: `%d`" % (self.row_count, other.row_count)
        assert self.col_count == other.col_count, "assertion failed: `(left == right)` left: `%d`, right: `%d`" % (self.col_count, other.col_count)

    # ---- elementwise operators delegate to Vector (matrix.rs:134-218, 268-277)
    def __add__(self, other):
        self._check_shape(other)
        return Matrix(self.data + other.data, self.row_count, self.col_count)

    def __iadd__(self, other):
        self._check_shape(other)
        self.data += other.data
        return self

    def __sub__(self, other):
        self._check_shape(other)
        return Matrix(self.data - other.data, self.row_count, self.col_count)

    def __isub__(self, other):
        self._check_shape(other)
        self.data -= other.data
        return self

    def __mul__(self, other):
        if isinstance(other, Matrix):             # matrix.rs:221-232
            assert self.col_count == other.row_count, "assertion failed: `(left == right)` left: `%d`, right: `%d`" % (self.col_count, other.row_count)
            return _mul_helper(self.data, other.data, self.row_count, self.col_count, other.col_count)
        return Matrix(self.data * other, self.row_count, self.col_count)     # scale, matrix.rs:195-204

    def __imul__(self, scalar):
        assert not isinstance(scalar, (Matrix, Vector)), "`*=` takes a scalar (matrix.rs:206-210)"
        self.data *= scalar
        return self

    def __truediv__(self, scalar):                # matrix.rs:268-277
        assert not isinstance(scalar, (Matrix, Vector)), "`/` takes a scalar"
        return Matrix(self.data / scalar, self.row_count, self.col_count)

    def __eq__(self, other):                      # matrix.rs:279-287
        if not isinstance(other, Matrix):
            return NotImplemented
        return self.row_count == other.row_count and self.col_count == other.col_count and self.data == other.data

    def __ne__(self, other):
        r = self.__eq__(other)
        return r if r is NotImplemented else not r

    def prepare(self):
        """This matrix as the right-hand operand of many products (la_gemm_prepare_b): the work a product does on B
        alone is done once.  `a.mul_prepared(p)` then equals `a * b` bit for bit.  No counterpart in the reference."""
        return PreparedMatrix(self)

    def mul_prepared(self, prepared):
        assert isinstance(prepared, PreparedMatrix), "expected a PreparedMatrix (Matrix.prepare())"
        assert self.col_count == prepared.matrix.row_count, "assertion failed: `(left == right)` left: `%d`, right: `%d`" % (self.col_count, prepared.matrix.row_count)
        assert self.dtype.code == prepared.matrix.dtype.code, "element types differ"
        res = Matrix.uninitialized(self.row_count, prepared.matrix.col_count, self.dtype)
        call("la_mul_mat_mat_prepared", res.data.ptr, self.data.ptr, prepared.handle, self.row_count, _v._STREAM)
        _v._done()
        return res

    def transpose(self):
        """Not in the reference; named by the north star: out[c*rows + r] = in[r*cols + c]."""
        res = Matrix.uninitialized(self.col_count, self.row_count, self.dtype)
        call("la_transpose", self.dtype.code, res.data.ptr, self.data.ptr, self.row_count, self.col_count, _v._STREAM)
        _v._done()
        return res

    # ---- persistence (matrix.rs:76-130): u64 element size, u64 rows, u64 cols, raw row-major data
    def write_to_file(self, f):
        import numpy as np
        f.write(np.array([self.dtype.size, self.row_count, self.col_count], dtype=np.uint64).tobytes())
        f.write(self.to_vec().tobytes())

    @staticmethod
    def read_from_file(f, dtype):
        import numpy as np
        from .traits import parameter
        from .vector import _read_exact
        dt = parameter(dtype)
        elem_size, rows, cols = (int(x) for x in np.frombuffer(_read_exact(f, 24), dtype=np.uint64))
        if elem_size != dt.size:       # matrix.rs:97-100
            raise AssertionError("Elem size from buffer does not seem to match, what was expected!Missmatch in endiannes?")
        return Matrix.from_vec(np.frombuffer(_read_exact(f, rows * cols * dt.size), dtype=dt.np), rows, cols, dt)

    def save(self, path):
        with open(path, "wb") as f:
            self.write_to_file(f)

    @staticmethod
    def open(path, dtype):
        with open(path, "rb") as f:
            return Matrix.read_from_file(f, dtype)

    def __repr__(self):
        """Debug form; large matrices are abbreviated (see Vector.__repr__)."""
        if self.row_count * self.col_count <= 256:
            return self.__str__()
        return "Mat%dx%d %r" % (self.row_count, self.col_count, self.data)

    def __str__(self):                            # matrix.rs:289-317: every element
        v = self.to_vec()
        rows = ["\t " + ", ".join(repr(x.item()) for x in v[r * self.col_count:(r + 1) * self.col_count]) for r in range(self.row_count)]
        return "Mat%dx%d {\n%s\n}\n" % (self.row_count, self.col_count, "\n".join(rows))


class PreparedMatrix:
    """Handle of la_gemm_prepare_b.  Keeps the Matrix alive (the library keeps B's pointer, not a copy)."""
    __slots__ = ("matrix", "handle")

    def __init__(self, matrix):
        import ctypes
        self.matrix = matrix
        self.handle = ctypes.c_void_p()
        call("la_gemm_prepare_b", matrix.dtype.code, matrix.data.ptr, matrix.row_count, matrix.col_count, _v._STREAM, ctypes.byref(self.handle))

    def free(self):
        if self.handle is not None and self.handle.value:
            call("la_gemm_prepared_free", self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def _mul_helper(a, b, a_row_count, a_col_count, b_col_count):     # matrix.rs:240-265
    assert a.dtype.code == b.dtype.code, "element types differ"
    res = Matrix.uninitialized(a_row_count, b_col_count, a.dtype)
    call("la_mul_mat_mat", a.dtype.code, res.data.ptr, a.ptr, b.ptr, a_row_count, a_col_count, b_col_count, _v._STREAM)
    _v._done()
    return res


def mul_column_row(column, row):                  # matrix.rs:236-238
    """Outer product: column (as an n x 1 matrix) times row (as a 1 x m matrix)."""
    return _mul_helper(column, row, column.len(), 1, row.len())


def mul_host(a_host, b_host, a_row_count, a_col_count, b_col_count, out=None):
    """C = A x B on HOST arrays in one streamed call (la_mul_mat_mat_host): what
    `Matrix.from_vec(a) * Matrix.from_vec(b)` followed by `to_vec()` computes, with the uploads,
    the multiply and the download overlapped.  Pass page-locked arrays (linear_algebra_b200.pinned)
    for full PCIe rate.  Returns the host array of a_row_count * b_col_count elements."""
    import numpy as np
    from .traits import parameter
    a = np.ascontiguousarray(a_host).reshape(-1); b = np.ascontiguousarray(b_host, dtype=a.dtype).reshape(-1)
    assert a.size == a_row_count * a_col_count, "assertion failed: `(left == right)` left: `%d`, right: `%d`" % (a.size, a_row_count * a_col_count)
    assert b.size == a_col_count * b_col_count, "assertion failed: `(left == right)` left: `%d`, right: `%d`" % (b.size, a_col_count * b_col_count)
    dt = parameter(a.dtype)
    if out is None:
        out = np.empty(a_row_count * b_col_count, dtype=a.dtype)
    assert out.size == a_row_count * b_col_count and out.dtype == a.dtype and out.flags.c_contiguous
    call("la_mul_mat_mat_host", dt.code, out.ctypes.data, a.ctypes.data, b.ctypes.data, a_row_count, a_col_count, b_col_count)
    return out
