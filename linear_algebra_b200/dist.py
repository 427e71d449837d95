"""Row-block sharding of Vector / Matrix over one process per GPU.

The reference is single-device (src/lib.rs:310-335: one context, one in-order queue); this is
the multi-GPU layer the north star adds.  `torch.distributed` (NCCL on GPUs, gloo in the CPU
tests) is only plumbing: ranks exchange raw bytes with all_gather and every rank finishes the
combination locally IN RANK ORDER, so results are deterministic, identical on all ranks, and
need no NCCL support for the element type (NCCL has no 16-bit integers).

  op                         partition                              exchange
  elementwise / scale        co-sharded ranges                      none
  sum / dot                  fused: the reduction kernel stores its partial into every peer's mailbox over NVLink
                             peer memory and adds the partials in rank order (la_*_allreduce, one launch per rank);
                             fallback / == : local partial + all_gather of one element
  M x v (mul_transpose_mat)  M by rows, v replicated -> rows shard  none
  v x M                      v and M by rows -> full-length partial fused: partial into an exported buffer, one kernel
                                                                    adds the peers' partials over NVLink in rank order
                                                                    (la_vec_allreduce); fallback: all_gather + adds
  M x M                      A, C by rows, B replicated             none
  M x M on host buffers      A, C by rows; every rank uploads 1/world all_gather of B over NVLink (mul_host)
                             of B over its own PCIe link
  transpose                  rows in -> rows of the transpose out   all_to_all of (rows_q x cols_p) blocks

`Comm` wraps the process group; `local` is the module providing Vector/Matrix (the real one by
default; the CPU tests inject a numpy stand-in to exercise this file's logic under gloo)."""
import numpy as np


def partition(n, world, rank):
    """Contiguous block partition of range(n): the first n % world ranks get one extra element."""
    base, extra = divmod(int(n), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


class Comm:
    """Byte-level collectives over a torch.distributed process group."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.backend = dist.get_backend(group)

    def _tensor(self, ptr, nbytes):
        import torch
        if self.backend == "gloo":
            raise RuntimeError("device pointers need the nccl backend")

        class _View:      # zero-copy view of library-owned device memory
            __cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 2, "strides": None}
        return torch.as_tensor(_View(), device="cuda")

    def all_gather_host(self, payload: bytes):
        """All ranks contribute `payload` (same length everywhere); returns the list in rank order."""
        import torch
        device = "cuda" if self.backend == "nccl" else "cpu"
        mine = torch.frombuffer(bytearray(payload), dtype=torch.uint8).to(device)
        out = [torch.empty_like(mine) for _ in range(self.world)]
        self.dist.all_gather(out, mine, group=self.group)
        return [bytes(t.cpu().numpy().tobytes()) for t in out]

    def all_gather_device(self, dst_ptr, src_ptr, nbytes):
        """dst (world * nbytes device bytes, rank order) <- every rank's src (nbytes device bytes)."""
        out = self._tensor(dst_ptr, nbytes * self.world)
        self.dist.all_gather_into_tensor(out, self._tensor(src_ptr, nbytes), group=self.group)

    def all_to_all_device(self, dst_ptr, dst_counts, src_ptr, src_counts):
        """Byte counts per rank; chunks are contiguous and in rank order on both sides."""
        out = self._tensor(dst_ptr, sum(dst_counts)); inp = self._tensor(src_ptr, sum(src_counts))
        self.dist.all_to_all_single(out, inp, output_split_sizes=list(dst_counts), input_split_sizes=list(src_counts), group=self.group)

    def barrier(self):
        self.dist.barrier(group=self.group)

    # ---- NVLink peer-memory mailboxes for the fused reductions (la_*_allreduce)
    def p2p(self, m):
        """Lazily: create this rank's mailbox, exchange the IPC handles, map the peers' mailboxes.  Returns
        (ctypes array of `world` device pointers, next sequence number) or None when unavailable."""
        import ctypes
        st = getattr(self, "_p2p", None)
        if st is None:
            st = self._p2p = {"ok": False, "seq": 0}
            if self.backend == "nccl" and self.world <= 16:
                try:
                    hb = m._lib.lib.la_p2p_handle_bytes()

                    def exported(nbytes):
                        mine = ctypes.c_void_p(); handle = ctypes.create_string_buffer(hb)
                        m._lib.call("la_p2p_buffer_create", nbytes, ctypes.byref(mine), handle)
                        ptrs = (ctypes.c_void_p * self.world)()
                        for r, h in enumerate(self.all_gather_host(handle.raw)):
                            if r == self.rank:
                                ptrs[r] = mine.value
                            else:
                                p = ctypes.c_void_p()
                                m._lib.call("la_p2p_buffer_open", ctypes.create_string_buffer(h, hb), ctypes.byref(p))
                                ptrs[r] = p.value
                        return ptrs
                    st.update(ok=True, ptrs=exported(m._lib.lib.la_p2p_mailbox_bytes()), vec_half=4 << 20,
                              vecs=exported(2 * (4 << 20)), out=m.Vector.uninitialized(8, "u8"))
                except Exception as e:  # noqa: BLE001  (no peer access, IPC refused, ...): keep the NCCL path
                    st["error"] = str(e)
                flags = self.all_gather_host(b"\x01" if st["ok"] else b"\x00")       # all ranks or none
                st["ok"] = all(f == b"\x01" for f in flags)
        if not st["ok"]:
            return None
        st["seq"] += 1
        return st["ptrs"], st["seq"], st["out"]


def combine_scalars(parts, np_dtype):
    """Rank-ordered sum of per-rank partials with the element type's own arithmetic (integers
    wrap, f32 rounds after every add): what a single device would produce from those partials."""
    dt = np.dtype(np_dtype)
    acc = dt.type(0)
    with np.errstate(over="ignore"):
        for p in parts:
            acc = dt.type(acc + np.frombuffer(p, dtype=dt, count=1)[0])
    return acc


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
        if self.comm.backend == "nccl":
            return _device_scalar_exchange(self, None)
        part = self.local.sum()
        return combine_scalars(self.comm.all_gather_host(np.asarray(part).tobytes()), self.local.dtype.np)

    def __eq__(self, other):
        if not isinstance(other, ShardedVector):
            return NotImplemented
        same_shape = self.global_len == other.global_len and self.local.dtype.code == other.local.dtype.code
        mine = bool(same_shape and self.local == other.local)
        return all(p == b"\x01" for p in self.comm.all_gather_host(b"\x01" if mine else b"\x00"))

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


def _device_scalar_exchange(a, b):
    """NCCL path of sum (b is None) / dot: the local partial stays on the device (la_*_dev writes it into
    this rank's 8-byte slot of a small gather buffer), one in-place all_gather over NVLink, one download
    of `world` slots, rank-ordered combination on the host."""
    m, comm, dt = a._mod(), a.comm, a.local.dtype
    fused = comm.p2p(m)
    if fused is not None:
        # ONE launch per rank: reduce, store the partial into every peer's mailbox over NVLink, add in rank order
        ptrs, seq, out = fused
        if b is None:
            m._lib.call("la_sum_vec_allreduce", dt.code, a.local.ptr, a.local.len(), out.ptr, ptrs, comm.world, comm.rank, seq, None)
        else:
            m._lib.call("la_dot_vec_vec_allreduce", dt.code, a.local.ptr, b.local.ptr, a.local.len(), out.ptr, ptrs, comm.world, comm.rank, seq, None)
        return np.frombuffer(out.to_vec().tobytes()[:dt.size], dtype=dt.np, count=1)[0]
    slots = getattr(comm, "_scalar_slots", None)
    if slots is None:
        slots = comm._scalar_slots = m.Vector.uninitialized(8 * comm.world, "u8")
    mine = slots.ptr + 8 * comm.rank
    if b is None:
        m._lib.call("la_sum_vec_dev", dt.code, a.local.ptr, a.local.len(), mine, None)
    else:
        m._lib.call("la_dot_vec_vec_dev", dt.code, a.local.ptr, b.local.ptr, a.local.len(), mine, None)
    comm.all_gather_device(slots.ptr, mine, 8)           # in place; ordered after the kernel on the default stream
    raw = slots.to_vec().tobytes()
    return combine_scalars([raw[8 * r:8 * r + dt.size] for r in range(comm.world)], dt.np)


def dot(a, b):
    a._co(b)
    if a.comm.backend == "nccl":
        return _device_scalar_exchange(a, b)
    part = a._mod().dot(a.local, b.local)
    return combine_scalars(a.comm.all_gather_host(np.asarray(part).tobytes()), a.local.dtype.np)


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
    """Transpose of a row-sharded matrix, row-sharded again (rank p ends up with rows [p_lo, p_hi) of
    the transpose = columns of the input).  Local transpose, one all_to_all of contiguous
    (cols_p x rows_q) blocks over NVLink, then strided placement of the received blocks."""
    comm, m = sm.comm, sm._mod()
    rows_g, cols_g = sm.global_rows, sm.col_count
    if comm.world == 1:
        return ShardedMatrix(comm, sm.local.transpose(), cols_g, sm._m)
    assert comm.backend == "nccl", "the sharded transpose exchanges device buffers (NCCL)"
    dt = sm.local.dtype
    my_rows = sm.hi - sm.lo
    local_t = sm.local.transpose()                          # cols_g x my_rows: rows [p_lo, p_hi) are rank p's block
    spans = [partition(cols_g, comm.world, p) for p in range(comm.world)]
    row_spans = [partition(rows_g, comm.world, q) for q in range(comm.world)]
    my_lo, my_hi = spans[comm.rank]
    my_cols = my_hi - my_lo
    send_counts = [(hi - lo) * my_rows * dt.size for lo, hi in spans]
    recv_counts = [my_cols * (hi - lo) * dt.size for lo, hi in row_spans]
    recv = m.Vector.uninitialized(sum(recv_counts) // dt.size, dt)
    m.synchronize()
    comm.all_to_all_device(recv.ptr, recv_counts, local_t.data.ptr, send_counts)
    import torch
    torch.cuda.current_stream().synchronize()
    out = m.Matrix.uninitialized(my_cols, rows_g, dt)
    off = 0
    for q, (lo, hi) in enumerate(row_spans):                # block from rank q: my_cols x (hi - lo), goes to columns [lo, hi)
        m._lib.call("la_copy_2d", out.data.ptr + lo * dt.size, rows_g * dt.size, recv.ptr + off, (hi - lo) * dt.size, (hi - lo) * dt.size, my_cols, None)
        off += recv_counts[q]
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
    if comm.world == 1 or comm.backend != "nccl" or (k % comm.world) != 0:
        m._lib.call("la_mul_mat_mat_host", dt.code, out.ctypes.data, a.ctypes.data, b.ctypes.data, rows, k, n)
        return out
    import torch
    bdev = m.Vector.uninitialized(k * n, dt)
    slice_elems = (k // comm.world) * n
    m._lib.call("la_upload", bdev.ptr + comm.rank * slice_elems * dt.size, b.ctypes.data + comm.rank * slice_elems * dt.size, slice_elems * dt.size, None)
    m.synchronize()
    comm.all_gather_device(bdev.ptr, bdev.ptr + comm.rank * slice_elems * dt.size, slice_elems * dt.size)     # in place
    torch.cuda.current_stream().synchronize()
    m._lib.call("la_mul_mat_mat_host", dt.code, out.ctypes.data, a.ctypes.data, bdev.ptr, rows, k, n)
    return out


def mul_transpose_mat(vector, sharded_m):
    """M x v with M row-sharded and v replicated (a plain Vector): each rank owns a slice of the
    result; no exchange."""
    m = sharded_m._mod()
    return ShardedVector(sharded_m.comm, m.mul_transpose_mat(vector, sharded_m.local), sharded_m.global_rows, sharded_m._m)


def _sharded_vec_mat(sv, sm):
    """v x M with v and M co-sharded by rows: local full-length partial, then one all_gather and
    rank-ordered adds.  Returns a replicated plain Vector of `cols` elements."""
    assert sv.global_len == sm.global_rows, "assertion failed: `(left == right)` left: `%d`, right: `%d`" % (sv.global_len, sm.global_rows)
    m = sm._mod()
    comm = sv.comm
    if comm.world == 1:
        return sv.local * sm.local
    fused = comm.p2p(m) if comm.backend == "nccl" else None
    dtp = sv.local.dtype
    if fused is not None and sm.col_count * dtp.size <= comm._p2p["vec_half"]:
        # local partial straight into this rank's exported buffer, then ONE kernel reads every peer's partial over
        # NVLink and adds them in rank order
        ptrs, seq, _ = fused
        st = comm._p2p
        mine = st["vecs"][comm.rank] + (seq & 1) * st["vec_half"]
        m._lib.call("la_mul_vec_mat", dtp.code, mine, sv.local.ptr, sm.local.data.ptr, sm.local.get_row_count(), sm.col_count, None)
        out = m.Vector.uninitialized(sm.col_count, dtp)
        m._lib.call("la_vec_allreduce", dtp.code, out.ptr, sm.col_count, st["vecs"], st["vec_half"], ptrs, comm.world, comm.rank, seq, None)
        m.synchronize()
        return out
    partial = sv.local * sm.local                  # la_mul_vec_mat on the local rows
    if comm.backend == "nccl":
        dt = partial.dtype
        nbytes = partial.len() * dt.size
        gathered = m.Vector.uninitialized(partial.len() * comm.world, dt)
        m.synchronize()
        comm.all_gather_device(gathered.ptr, partial.ptr, nbytes)
        import torch
        torch.cuda.current_stream().synchronize()
        acc = m.Vector(dt, gathered.ptr, partial.len(), owner=False)
        for r in range(1, comm.world):             # ascending rank order: deterministic
            acc += m.Vector(dt, gathered.ptr + r * nbytes, partial.len(), owner=False)
        out = acc.clone()
        del acc
        return out
    # host path (gloo / CPU tests)
    parts = comm.all_gather_host(partial.to_vec().tobytes())
    dt = partial.dtype.np
    acc = np.frombuffer(parts[0], dtype=dt).copy()
    with np.errstate(over="ignore"):
        for p in parts[1:]:
            acc = (acc + np.frombuffer(p, dtype=dt)).astype(dt)
    return m.Vector.from_vec(acc, partial.dtype)


def _default_local():
    import linear_algebra_b200 as la
    return la
