"""la_mul_vec_mat_noalloc: the v x M entry point for callers that cannot allocate device memory
(a host program whose own allocator already owns the GPU's memory, or a latency-critical section
that cannot take the device-wide synchronisation of a cudaMalloc).  It keeps its per-strip
partials in a fixed 8 MiB block inside the library instead of grown scratch.

Checked here, through the C ABI: on two calls in a row (different shapes and data, so stale
contents of the shared block would show) its output equals la_mul_vec_mat's -- bit-for-bit for
integers, and within the f32 bar of tests/test_gpu_parity.py (normwise 1e-5) for floats; shapes
whose partials exceed the block are refused with a clear message.

This file sorts last on purpose: a run that stops at the first failure has then already run
every other test.  It also passes when run alone in a fresh process."""
import ctypes

import numpy as np
import pytest

import gen

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
        if np.dtype(got.dtype).kind == "f":
            assert np.linalg.norm(got.astype(np.float64) - expect.astype(np.float64)) <= NORMWISE_TOL * np.linalg.norm(expect.astype(np.float64)), (name, rows, cols)
        else:
            assert got.tobytes() == expect.tobytes(), (name, rows, cols)


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
