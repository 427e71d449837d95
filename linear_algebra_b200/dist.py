"""Row-block sharding of Vector / Matrix over one process per GPU.

The reference is single-device (src/lib.rs:310-335: one context, one in-order queue); this is the Python face of the
multi-GPU layer the north star adds.  On GPUs it is a THIN caller: the group, the NVLink peer-memory exchanges and every
rank-ordered combination live below the C ABI (csrc/comm.cu: la_comm_create, la_*_sharded), and `torch.distributed` is
used for one thing only -- the byte all-gather that la_comm_create takes as its `exchange` callback.

  op                         partition                              exchange (all inside liblinalg_b200)
  elementwise / scale        co-sharded ranges                      none
  sum / dot                  la_sum_vec_sharded / la_dot_vec_vec_sharded: one launch per rank, partials delivered to the
                             peers' mailboxes over NVLink, added in rank order, scalar back through a zero-copy word
  ==                         la_eq_vec_sharded (one byte per rank through the callback)
  M x v (mul_transpose_mat)  M by rows, v replicated -> rows shard  none
  v x M                      la_mul_vec_mat_sharded: partial into an exported buffer, one kernel adds the peers' partials
  M x M                      A, C by rows, B replicated             none
  M x M on host buffers      la_mul_mat_mat_host_sharded: 1/world of B per PCIe link, slices over NVLink
  transpose                  la_transpose_sharded: pieces stored straight into the owners' memory

`Comm` wraps the process group; `local` is the module providing Vector/Matrix.  This file holds bookkeeping only (who owns which
range, which entry point to call): it contains no arithmetic and no exchange of its own.  The CPU tests (tests/test_dist_gloo.py)
drive the same bookkeeping under gloo with a numpy stand-in module that brings its own `ShardedEngine` (tests/fake_local.py: a host
model of the rank-ordered combinations); with the real module every cross-rank step is a la_*_sharded call."""
import ctypes

import numpy as np


def partition(n, world, rank):
    """la_partition (csrc/comm.cu): contiguous blocks, the first n % world ranks get one extra element."""
    from . import _lib
    lo, hi = ctypes.c_size_t(), ctypes.c_size_t()
    _lib.lib.la_partition(int(n), int(world), int(rank), ctypes.byref(lo), ctypes.byref(hi))
    return lo.value, hi.value


class Comm:
    """Byte-level collectives over a torch.distributed process group."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.backend = dist.get_backend(group)

    def all_gather_host(self, payload: bytes):
        """All ranks contribute `payload` (same length everywhere); returns the list in rank order."""
        import torch
        device = "cuda" if self.backend == "nccl" else "cpu"
        mine = torch.frombuffer(bytearray(payload), dtype=torch.uint8).to(device)
        out = [torch.empty_like(mine) for _ in range(self.world)]
        self.dist.all_gather(out, mine, group=self.group)
        return [bytes(t.cpu().numpy().tobytes()) for t in out]

    def barrier(self):
        self.dist.barrier(group=self.group)

    # ---- the communicator below the C ABI (csrc/comm.cu)
    def engine(self, m):
        """Who performs the cross-rank steps for local module `m`: liblinalg_b200 for the real module, the stand-in's own
        ShardedEngine otherwise (CPU tests)."""
        cache = self.__dict__.setdefault("_engines", {})
        if id(m) not in cache:
            cache[id(m)] = _DeviceEngine(self, m) if hasattr(m, "_lib") else m.ShardedEngine(self)
        return cache[id(m)]

    def lacomm(self, m):
        """Lazily: la_comm_create with a callback that all-gathers bytes over the process group."""
        h = getattr(self, "_la", None)
        if h is None:
            def xchg(_user, send, recv, nbytes):
                try:
                    parts = self.all_gather_host(ctypes.string_at(send, nbytes))
                    ctypes.memmove(recv, b"".join(parts), nbytes * self.world)
                    return 0
                except Exception:  # noqa: BLE001  (reported by the library as a failed exchange)
                    return 1
            self._cb = m._lib.EXCHANGE_FN(xchg)              # must outlive the communicator
            h = ctypes.c_void_p()
            m._lib.call("la_comm_create", self.world, self.rank, ctypes.cast(self._cb, ctypes.c_void_p), None, ctypes.byref(h))
            self._la = h
            w, r, pm = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
            m._lib.call("la_comm_info", h, ctypes.byref(w), ctypes.byref(r), ctypes.byref(pm))
            self._p2p = {"ok": bool(pm.value), "error": None if pm.value else m._lib.lib.la_last_error().decode("utf-8", "replace")}
            self._m = m
        return h

    def close(self):
        h = getattr(self, "_la", None)
        if h is not None:
            self._m._lib.call("la_comm_destroy", h)
            self._la = None


class _DeviceEngine:
    """Cross-rank steps with the real module: one la_*_sharded call each (csrc/comm.cu).  The few microseconds of Python around a
    call are what the ranks drift apart by between rounds, so the scalar path is kept short."""

    def __init__(self, comm, m):
        self.comm, self.m, self.h = comm, m, comm.lacomm(m)
        self.out = np.zeros(1, dtype=np.uint64)
        self.out_ptr = self.out.ctypes.data

    def scalar(self, a, b):
        """sum of a (b is None) or dot(a, b) over all ranks."""
        lib, dt = self.m._lib.lib, a.dtype
        if b is None:
            st = lib.la_sum_vec_sharded(self.h, dt.code, a.ptr, a.len(), self.out_ptr, None)
        else:
            st = lib.la_dot_vec_vec_sharded(self.h, dt.code, a.ptr, b.ptr, a.len(), self.out_ptr, None)
        if st:
            self.m._lib.check(st)
        return self.out.view(dt.np)[0].copy()        # the value sits in the low bytes of the 8-byte word

    def eq(self, a, b):
        eq = ctypes.c_int(0)
        self.m._lib.call("la_eq_vec_sharded", self.h, a.dtype.code, a.ptr, b.ptr, a.len(), ctypes.byref(eq), None)
        return bool(eq.value)

    def vec_mat(self, v, mat, cols):
        out = self.m.Vector.uninitialized(cols, v.dtype)
        self.m._lib.call("la_mul_vec_mat_sharded", self.h, v.dtype.code, out.ptr, v.ptr, mat.data.ptr, mat.get_row_count(), cols, None)
        self.m.synchronize()
        return out


class ShardedVector:
    """This rank's contiguous range [lo, hi) of a global vector of `global_len` elements."""

    def __init__(self, comm, local_vec, global_len, local_mod=None):
        self.comm, self.local, self.global_len = comm, local_vec, int(global_len)
        self.lo, self.hi = partition(global_len, comm.world, comm.rank)
        assert self.local.len() == self.hi - self.lo, "local shard has the wrong length"
        self._m = local_mod

    @staticmethod
    def from_global(comm, host_array, dtype=None, local_mod=None):
        """Every rank passes the same full host array and keeps only its range."""
        m = local_mod or _default_local()
        host = np.ascontiguousarray(host_array).reshape(-1)
        lo, hi = partition(host.size, comm.world, comm.rank)
        return ShardedVector(comm, m.Vector.from_vec(host[lo:hi], dtype), host.size, m)

    def _mod(self):
        return self._m or _default_local()

    def _wrap(self, v):
        return ShardedVector(self.comm, v, self.global_len, self._m)

    def _co(self, other):
        assert isinstance(other, ShardedVector) and other.global_len == self.global_len, \
            "assertion failed: `(left == right)` left: `%d`, right: `%d`" % (self.global_len, getattr(other, "global_len", -1))
        return other.local

    def len(self):
        return self.global_len

    # elementwise: purely local
    def __add__(self, o): return self._wrap(self.local + self._co(o))
    def __sub__(self, o): return self._wrap(self.local - self._co(o))
    def __iadd__(self, o): self.local += self._co(o); return self
    def __isub__(self, o): self.local -= self._co(o); return self

    def __mul__(self, o):
        if isinstance(o, ShardedMatrix):
            return _sharded_vec_mat(self, o)
        return self._wrap(self.local * (self._co(o) if isinstance(o, ShardedVector) else o))

    def __imul__(self, o):
        self.local *= (self._co(o) if isinstance(o, ShardedVector) else o)
        return self

    def __truediv__(self, o):
        return self._wrap(self.local / (self._co(o) if isinstance(o, ShardedVector) else o))

    # one exchange: a single element per rank, combined in rank order
    def sum(self):
        return self.comm.engine(self._mod()).scalar(self.local, None)

    def __eq__(self, other):
        if not isinstance(other, ShardedVector):
            return NotImplemented
        if self.global_len != other.global_len or self.local.dtype.code != other.local.dtype.code:
            return False                    # global properties: every rank decides the same without an exchange
        return self.comm.engine(self._mod()).eq(self.local, other.local)

    __hash__ = None

    def gather(self):
        """The full vector on every rank (host array)."""
        parts = self.comm.all_gather_host(_pad(self.local.to_vec().tobytes(), self._max_shard_bytes()))
        dt = self.local.dtype.np
        out = []
        for r, p in enumerate(parts):
            lo, hi = partition(self.global_len, self.comm.world, r)
            out.append(np.frombuffer(p, dtype=dt, count=hi - lo))
        return np.concatenate(out) if out else np.empty(0, dt)

    def _max_shard_bytes(self):
        lo, hi = partition(self.global_len, self.comm.world, 0)
        return (hi - lo) * self.local.dtype.size


def _pad(b, n):
    return b + b"\x00" * (n - len(b))


def dot(a, b):
    a._co(b)
    return a.comm.engine(a._mod()).scalar(a.local, b.local)


class ShardedMatrix:
    """Rows [lo, hi) of a global rows x cols matrix (all columns)."""

    def __init__(self, comm, local_mat, global_rows, local_mod=None):
        self.comm, self.local, self.global_rows = comm, local_mat, int(global_rows)
        self.lo, self.hi = partition(global_rows, comm.world, comm.rank)
        assert self.local.get_row_count() == self.hi - self.lo
        self.col_count = self.local.get_col_count()
        self._m = local_mod

    @staticmethod
    def from_global(comm, host_array, rows, cols, dtype=None, local_mod=None):
        m = local_mod or _default_local()
        host = np.ascontiguousarray(host_array).reshape(rows, cols)
        lo, hi = partition(rows, comm.world, comm.rank)
        return ShardedMatrix(comm, m.Matrix.from_vec(host[lo:hi].reshape(-1), hi - lo, cols, dtype), rows, m)

    def _mod(self):
        return self._m or _default_local()

    def len(self):
        return (self.global_rows, self.col_count)

    def _co(self, o):
        assert isinstance(o, ShardedMatrix) and o.len() == self.len(), "matrix shapes differ"
        return o.local

    def __add__(self, o): return ShardedMatrix(self.comm, self.local + self._co(o), self.global_rows, self._m)
    def __sub__(self, o): return ShardedMatrix(self.comm, self.local - self._co(o), self.global_rows, self._m)

    def __mul__(self, o):
        m = self._mod()
        if isinstance(o, m.Matrix):            # A (row-sharded) x B (replicated): no exchange
            return ShardedMatrix(self.comm, self.local * o, self.global_rows, self._m)
        assert not isinstance(o, (ShardedMatrix, ShardedVector)), "right operand of a sharded GEMM must be replicated"
        return ShardedMatrix(self.comm, self.local * o, self.global_rows, self._m)

    def gather(self):
        rows_max = partition(self.global_rows, self.comm.world, 0)
        nbytes = (rows_max[1] - rows_max[0]) * self.col_count * self.local.dtype.size
        parts = self.comm.all_gather_host(_pad(self.local.to_vec().tobytes(), nbytes))
        out = []
        for r, p in enumerate(parts):
            lo, hi = partition(self.global_rows, self.comm.world, r)
            out.append(np.frombuffer(p, dtype=self.local.dtype.np, count=(hi - lo) * self.col_count))
        return np.concatenate(out)


def transpose(sm):
    """Transpose of a row-sharded matrix, row-sharded again (rank p ends up with rows [p_lo, p_hi) of the transpose =
    columns of the input): la_transpose_sharded."""
    comm, m = sm.comm, sm._mod()
    rows_g, cols_g = sm.global_rows, sm.col_count
    if comm.world == 1:
        return ShardedMatrix(comm, sm.local.transpose(), cols_g, sm._m)
    assert hasattr(m, "_lib"), "the sharded transpose runs on GPUs (la_transpose_sharded)"
    lo, hi = partition(cols_g, comm.world, comm.rank)
    out = m.Matrix.uninitialized(hi - lo, rows_g, sm.local.dtype)
    m._lib.call("la_transpose_sharded", comm.lacomm(m), sm.local.dtype.code, out.data.ptr, sm.local.data.ptr, rows_g, cols_g, None)
    m.synchronize()
    return ShardedMatrix(comm, out, cols_g, sm._m)


def mul_host(comm, a_shard_host, b_host, a_rows_global, k, n, out=None, local_mod=None):
    """C = A x B on HOST buffers over all ranks: A and C are row-block sharded (each rank passes its
    own rows of A and gets its rows of C), every rank holds all of B on the host.  Each rank uploads
    only its 1/world slice of B over its own PCIe link and the slices are all_gathered over NVLink;
    then the streamed host-buffer GEMM (la_mul_mat_mat_host) runs on the local rows with the gathered,
    device-resident B."""
    m = local_mod or _default_local()
    from .traits import parameter
    a = np.ascontiguousarray(a_shard_host).reshape(-1); b = np.ascontiguousarray(b_host, dtype=a.dtype).reshape(-1)
    dt = parameter(a.dtype)
    lo, hi = partition(a_rows_global, comm.world, comm.rank)
    rows = hi - lo
    assert a.size == rows * k and b.size == k * n
    if out is None:
        out = np.empty(rows * n, dtype=a.dtype)
    if comm.world == 1:
        m._lib.call("la_mul_mat_mat_host", dt.code, out.ctypes.data, a.ctypes.data, b.ctypes.data, rows, k, n)
        return out
    m._lib.call("la_mul_mat_mat_host_sharded", comm.lacomm(m), dt.code, out.ctypes.data, a.ctypes.data, b.ctypes.data, a_rows_global, k, n)
    return out


def mul_transpose_mat(vector, sharded_m):
    """M x v with M row-sharded and v replicated (a plain Vector): each rank owns a slice of the
    result; no exchange."""
    m = sharded_m._mod()
    return ShardedVector(sharded_m.comm, m.mul_transpose_mat(vector, sharded_m.local), sharded_m.global_rows, sharded_m._m)


def _sharded_vec_mat(sv, sm):
    """v x M with v and M co-sharded by rows: every rank's full-length partial, added in rank order (la_mul_vec_mat_sharded).
    Returns a replicated plain Vector of `cols` elements."""
    assert sv.global_len == sm.global_rows, "assertion failed: `(left == right)` left: `%d`, right: `%d`" % (sv.global_len, sm.global_rows)
    if sv.comm.world == 1:
        return sv.local * sm.local
    return sv.comm.engine(sm._mod()).vec_mat(sv.local, sm.local, sm.col_count)


def _default_local():
    import linear_algebra_b200 as la
    return la
