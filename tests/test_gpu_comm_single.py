"""The multi-GPU entry points with a group of ONE rank: everything the communicator does except the exchange itself.  A one-GPU
box (where tests/test_gpu_dist.py and tests/test_gpu_comm_cpp.py skip) still drives la_comm_create / la_*_sharded through the
C ABI here and checks them against the oracle; with world = 1 they must equal the plain entry points bit for bit."""
import ctypes

import numpy as np
import pytest

import gen
import oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def la():
    import linear_algebra_b200 as la
    la.init(0)
    return la


@pytest.fixture()
def comm(la):
    h = ctypes.c_void_p()
    la._lib.call("la_comm_create", 1, 0, None, None, ctypes.byref(h))
    yield h
    la._lib.call("la_comm_destroy", h)


def test_partition_matches_the_python_mirror(la):
    from linear_algebra_b200.dist import partition
    lo, hi = ctypes.c_size_t(), ctypes.c_size_t()
    for n in (0, 1, 7, 1000, 8192, (1 << 32) + 5):
        for world in (1, 2, 3, 8, 16):
            for rank in range(world):
                la._lib.lib.la_partition(n, world, rank, ctypes.byref(lo), ctypes.byref(hi))
                assert (lo.value, hi.value) == partition(n, world, rank)


def test_comm_info_and_argument_checks(la, comm):
    w, r, pm = ctypes.c_int(-1), ctypes.c_int(-1), ctypes.c_int(-1)
    la._lib.call("la_comm_info", comm, ctypes.byref(w), ctypes.byref(r), ctypes.byref(pm))
    assert (w.value, r.value, pm.value) == (1, 0, 0)
    h = ctypes.c_void_p()
    assert la._lib.lib.la_comm_create(2, 0, None, None, ctypes.byref(h)) == la._lib.LA_ERR_ARG      # more than one rank needs the callback
    assert la._lib.lib.la_comm_create(17, 0, None, None, ctypes.byref(h)) == la._lib.LA_ERR_ARG
    assert la._lib.lib.la_comm_create(1, 1, None, None, ctypes.byref(h)) == la._lib.LA_ERR_ARG
    la._lib.call("la_comm_barrier", comm)


@pytest.mark.parametrize("name", ["i32", "u16", "f32"])
def test_sharded_reductions_with_one_rank(la, comm, name):
    P = oracle.port()
    dt = la.parameter(name)
    n = 300007
    a = gen.values(name, n, 110); b = gen.values(name, n, 111)
    A = la.Vector.from_vec(a, dt); B = la.Vector.from_vec(b, dt)
    s = np.zeros(1, dt.np); d = np.zeros(1, dt.np); eq = ctypes.c_int(-1)
    la._lib.call("la_sum_vec_sharded", comm, dt.code, A.ptr, n, s.ctypes.data, None)
    la._lib.call("la_dot_vec_vec_sharded", comm, dt.code, A.ptr, B.ptr, n, d.ctypes.data, None)
    assert s[0].tobytes() == np.asarray(A.sum()).tobytes() and d[0].tobytes() == np.asarray(la.dot(A, B)).tobytes()
    if name != "f32":
        assert s[0] == P.sum_vec(a) and d[0] == P.dot_vec_vec(a, b)
    la._lib.call("la_eq_vec_sharded", comm, dt.code, A.ptr, A.ptr, n, ctypes.byref(eq), None); assert eq.value == 1
    la._lib.call("la_eq_vec_sharded", comm, dt.code, A.ptr, B.ptr, n, ctypes.byref(eq), None); assert eq.value == 0


def test_sharded_products_and_transpose_with_one_rank(la, comm):
    P = oracle.port()
    rows, cols = 1037, 2052
    m = gen.values("i32", rows * cols, 112); v = gen.values("i32", rows, 113)
    M = la.Vector.from_vec(m); V = la.Vector.from_vec(v)
    C = la.Vector.uninitialized(cols, "i32")
    la._lib.call("la_mul_vec_mat_sharded", comm, 5, C.ptr, V.ptr, M.ptr, rows, cols, None)
    la.synchronize()
    assert C.to_vec().tobytes() == P.mul_vec_mat(v, m, rows, cols).tobytes()
    T = la.Vector.uninitialized(rows * cols, "i32")
    la._lib.call("la_transpose_sharded", comm, 5, T.ptr, M.ptr, rows, cols, None)
    la.synchronize()
    assert T.to_vec().tobytes() == P.transpose(m, rows, cols).tobytes()
    Mg, K, N = 300, 200, 100
    a = gen.values("i32", Mg * K, 114); b = gen.values("i32", K * N, 115)
    c = np.empty(Mg * N, np.int32)
    la._lib.call("la_mul_mat_mat_host_sharded", comm, 5, c.ctypes.data, a.ctypes.data, b.ctypes.data, Mg, K, N)
    assert c.tobytes() == P.mul_mat_mat(a, b, Mg, K, N).tobytes()
