"""la_mul_vec_mat_noalloc: the v x M entry point for callers that cannot allocate device memory
(a host program whose own allocator already owns the GPU's memory, or a latency-critical section
that cannot take the device-wide synchronisation of a cudaMalloc).  It keeps its per-strip
partials in a fixed 8 MiB block inside the library instead of grown scratch.

Checked here, through the C ABI: on two calls in a row (different shapes and data, so stale
contents of the shared block would show) its output equals la_mul_vec_mat's bit for bit -- for
floats too, because both entry points cut the matrix into the same strips and add them in the
same order -- and agrees with the CPU oracle directly (integers bit-exact, f32 within the bar of
tests/test_gpu_parity.py); the device's free-memory figure does not move across repeated calls
(it does move when the allocating entry point has to grow its scratch, which shows the probe
sees such things); shapes whose partials exceed the block are refused with a clear message.

This file sorts last on purpose: a run that stops at the first failure has then already run
every other test.  It also passes when run alone in a fresh process."""
import ctypes

import numpy as np
import pytest

import gen
import oracle

pytestmark = pytest.mark.gpu

NORMWISE_TOL = 1e-5          # the f32 bar of tests/test_gpu_parity.py


@pytest.fixture(scope="module")
def la():
    import linear_algebra_b200 as la
    la.init(0)
    return la


def both(la, name, rows, cols, seed):
    dt = la.I8_CHAR if name == "i8" else la.parameter(name)
    m = gen.values(name, rows * cols, seed); v = gen.values(name, rows, seed + 1)
    M = la.Vector.from_vec(m, dt); V = la.Vector.from_vec(v, dt)
    expect = la.Vector.uninitialized(cols, dt); got = la.Vector.uninitialized(cols, dt)
    la._lib.call("la_mul_vec_mat", dt.code, expect.ptr, V.ptr, M.ptr, rows, cols, None)
    la._lib.call("la_mul_vec_mat_noalloc", dt.code, got.ptr, V.ptr, M.ptr, rows, cols, None)
    la.synchronize()
    return got.to_vec(), expect.to_vec()


@pytest.mark.parametrize("name", ["f32", "i32", "u8", "f64"])
def test_two_calls_in_a_row_match_the_allocating_entry_point(la, name):
    # call 1: many strips, 16-byte packets; call 2: other shape, ragged columns (element kernel),
    # fewer strips -> reads a different part of the block that call 1 left dirty
    for rows, cols, seed in ((3000, 4096, 40), (1500, 1003, 50), (3000, 4096, 60)):
        got, expect = both(la, name, rows, cols, seed)
        assert got.tobytes() == expect.tobytes(), (name, rows, cols)      # same strips, same order: identical bits, floats included


@pytest.mark.parametrize("name", ["i32", "f32"])
def test_against_the_oracle_directly(la, name):
    """Not only equal to the sibling entry point: equal to the reference's kernel (src/kernels.cl:143-147 as restated in
    oracle/la_oracle.c) -- integers bit for bit, f32 inside the gate every f32 reduction here passes."""
    P = oracle.port()
    for rows, cols, seed in ((3000, 4096, 41), (1500, 1003, 51)):
        got, _ = both(la, name, rows, cols, seed)
        m = gen.values(name, rows * cols, seed); v = gen.values(name, rows, seed + 1)
        want = P.mul_vec_mat(v, m, rows, cols)
        if name == "i32":
            assert got.tobytes() == want.tobytes(), (rows, cols)
        else:
            m64 = m.astype(np.float64).reshape(rows, cols); v64 = v.astype(np.float64)
            truth = v64 @ m64; scale = np.abs(v64) @ np.abs(m64)
            e_ours = (np.abs(got - truth) / scale).max()
            e_ref = max((np.abs(want - truth) / scale).max(), (np.abs(P.mul_vec_mat(v, m, rows, cols, fma=True) - truth) / scale).max())
            assert e_ours <= max(e_ref, 2.0 ** -22), (rows, cols, e_ours, e_ref)
            assert np.linalg.norm(got - want) <= NORMWISE_TOL * np.linalg.norm(want)


def free_device_bytes():
    """Free memory of the current device as the driver reports it (cuMemGetInfo on the context la.init made current)."""
    cu = ctypes.CDLL("libcuda.so.1")
    free = ctypes.c_size_t(0); total = ctypes.c_size_t(0)
    assert cu.cuMemGetInfo_v2(ctypes.byref(free), ctypes.byref(total)) == 0
    return free.value


def test_repeated_calls_leave_free_device_memory_unchanged(la):
    """"Allocates nothing", measured: after one warm-up call (which loads the kernels), further calls of several shapes
    leave the driver's free-memory figure exactly where it was.  The same probe does see la_mul_vec_mat allocate its scratch
    on a stream it has not served yet, so an allocation inside the no-alloc path would not go unnoticed."""
    dt = la.parameter("f32")
    shapes = ((3000, 4096), (1500, 1003), (8192, 8192))
    bufs = []
    for rows, cols in shapes:                                     # operands first: the caller's own allocations
        M = la.Vector.from_vec(gen.values("f32", rows * cols, 80), dt); V = la.Vector.from_vec(gen.values("f32", rows, 81), dt)
        bufs.append((M, V, la.Vector.uninitialized(cols, dt), rows, cols))
    for M, V, C, rows, cols in bufs:                              # warm-up: module load, symbol look-up, every kernel variant used below
        la._lib.call("la_mul_vec_mat_noalloc", dt.code, C.ptr, V.ptr, M.ptr, rows, cols, None)
    la.synchronize()
    before = free_device_bytes()
    for _ in range(3):
        for M, V, C, rows, cols in bufs:
            la._lib.call("la_mul_vec_mat_noalloc", dt.code, C.ptr, V.ptr, M.ptr, rows, cols, None)
    la.synchronize()
    assert free_device_bytes() == before, "la_mul_vec_mat_noalloc changed the device's free memory"
    # the probe is live: on a stream of its own (scratch is kept per stream, so whatever earlier tests made the library
    # allocate for the default stream does not matter) la_mul_vec_mat has to allocate room for its partials
    rows, cols = 512, 1 << 18                                      # 5 strips x 2^18 x 4 B = 5 MiB of partials
    M = la.Vector.uninitialized(rows * cols, dt); V = la.Vector.uninitialized(rows, dt); C = la.Vector.uninitialized(cols, dt)
    stream = ctypes.c_void_p()
    la._lib.call("la_stream_create", ctypes.byref(stream))
    try:
        la._lib.call("la_mul_vec_mat_noalloc", dt.code, C.ptr, V.ptr, M.ptr, rows, cols, stream)     # warm: same kernels
        la._lib.call("la_stream_sync", stream)
        before = free_device_bytes()
        la._lib.call("la_mul_vec_mat_noalloc", dt.code, C.ptr, V.ptr, M.ptr, rows, cols, stream)
        la._lib.call("la_stream_sync", stream)
        assert free_device_bytes() == before
        la._lib.call("la_mul_vec_mat", dt.code, C.ptr, V.ptr, M.ptr, rows, cols, stream)
        la._lib.call("la_stream_sync", stream)
        assert free_device_bytes() < before, "probe did not see la_mul_vec_mat allocate its scratch"
    finally:
        la._lib.call("la_stream_destroy", stream)


def test_single_strip_and_empty_shapes(la):
    got, expect = both(la, "i32", 5, 70000, 70)        # one strip: no partials at all
    assert got.tobytes() == expect.tobytes()
    out = la.Vector.new(7, 16, "i32")
    la._lib.call("la_mul_vec_mat_noalloc", 5, out.ptr, out.ptr, out.ptr, 0, 16, None)   # rows == 0 -> zeros
    assert not out.to_vec().any()


def test_shape_beyond_the_fixed_block_is_refused(la):
    rows, cols = 64, 1 << 21                            # u8: 2 strips x 2^21 x 4 B = 16 MiB of partials
    M = la.Vector.uninitialized(rows * cols, "u8"); V = la.Vector.uninitialized(rows, "u8"); C = la.Vector.uninitialized(cols, "u8")
    st = la._lib.lib.la_mul_vec_mat_noalloc(0, ctypes.c_void_p(C.ptr), ctypes.c_void_p(V.ptr), ctypes.c_void_p(M.ptr), rows, cols, None)
    assert st == la._lib.LA_ERR_UNSUPPORTED
    msg = la._lib.lib.la_last_error().decode()
    assert "fixed block" in msg and "la_mul_vec_mat" in msg
    # the allocating entry point serves the same shape
    la._lib.call("la_mul_vec_mat", 0, C.ptr, V.ptr, M.ptr, rows, cols, None)
    la.synchronize()
