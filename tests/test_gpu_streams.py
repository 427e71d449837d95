"""Buffers and streams through the C ABI: la_malloc / la_free are pool operations ordered on the legacy default stream, while
la_stream_create hands out non-blocking streams the legacy stream does not order itself against.  The library bridges the
two (csrc/runtime.cu): a block freed while a kernel on a created stream still uses it must not be handed out and overwritten
before that kernel has run, and a fresh block must not be used on a created stream before legacy-stream work on its previous
owner has drained.  (The reference has one in-order queue, src/lib.rs:310-335, so the question does not arise there.)"""
import ctypes

import numpy as np
import pytest

import gen

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def la():
    import linear_algebra_b200 as la
    la.init(0)
    return la


def dev_alloc(la, nbytes):
    p = ctypes.c_void_p()
    la._lib.call("la_malloc", ctypes.byref(p), nbytes)
    return p


def test_free_right_after_a_launch_on_a_created_stream(la):
    call = la._lib.call
    n = 1 << 26                                   # 256 MiB per operand: the add takes ~0.1 ms, long enough to still be running at la_free
    a = gen.values("f32", n, 90); b = gen.values("f32", n, 91)
    A = la.Vector.from_vec(a); B = la.Vector.from_vec(b)
    want = a + b
    zero = np.zeros(1, np.float32)
    stream = ctypes.c_void_p()
    call("la_stream_create", ctypes.byref(stream))
    try:
        for _ in range(4):
            C = dev_alloc(la, n * 4); keep = dev_alloc(la, n * 4)
            call("la_ew_vec_vec", 8, 0, C, A.ptr, B.ptr, n, stream)           # C = A + B on the created stream
            call("la_copy", keep, C, n * 4, stream)                           # ... and a copy of it, also on that stream
            call("la_free", C)                                                # freed while both may still be pending
            E = dev_alloc(la, n * 4)                                          # the pool may hand the same block straight back
            call("la_fill", 8, E, zero.ctypes.data, n, None)                  # and the legacy stream scribbles over it
            call("la_stream_sync", stream)
            got = np.empty(n, np.float32)
            call("la_download", got.ctypes.data, keep, n * 4, stream)
            assert got.tobytes() == want.tobytes(), "a block was reused while a kernel on a created stream still needed it"
            call("la_stream_sync", None)
            call("la_free", keep); call("la_free", E)
    finally:
        call("la_stream_destroy", stream)


def test_fresh_block_on_a_created_stream_waits_for_the_legacy_stream(la):
    call = la._lib.call
    n = 1 << 26
    a = gen.values("i32", n, 92); A = la.Vector.from_vec(a)
    seven = np.array([7], np.int32); three = np.array([3], np.int32)
    stream = ctypes.c_void_p()
    call("la_stream_create", ctypes.byref(stream))
    try:
        for _ in range(4):
            X = dev_alloc(la, n * 4)
            call("la_fill", 5, X, seven.ctypes.data, n, None)                 # legacy-stream work on X ...
            call("la_ew_assign_vec_vec", 5, 0, X, A.ptr, n, None)             # X = 7 + A
            call("la_free", X)                                                # ... which is then freed
            Y = dev_alloc(la, n * 4)                                          # probably the same block
            call("la_fill", 5, Y, three.ctypes.data, n, stream)               # its new owner works on a created stream
            call("la_stream_sync", stream)
            call("la_stream_sync", None)
            got = np.empty(n, np.int32)
            call("la_download", got.ctypes.data, Y, n * 4, None)
            assert (got == 3).all(), "work of the block's previous owner landed after the new owner's"
            call("la_free", Y)
    finally:
        call("la_stream_destroy", stream)


def test_reductions_and_normalize_on_a_created_stream(la):
    """Scratch is kept per (device, stream): the first reduction on a new stream allocates that stream's block, and the result
    slot / ticket in its header must stay valid across the calls la_normalize_vec chains."""
    call = la._lib.call
    n = 300001
    a = gen.values("f32", n, 93); A = la.Vector.from_vec(a)
    stream = ctypes.c_void_p()
    call("la_stream_create", ctypes.byref(stream))
    try:
        C = dev_alloc(la, n * 4)
        call("la_normalize_vec", 8, C, A.ptr, n, stream)                      # first use of this stream: scratch is made inside
        out = np.zeros(1, np.float32)
        call("la_dot_vec_vec", 8, C, C, n, out.ctypes.data, stream)
        assert abs(float(out[0]) - 1.0) < 1e-5
        ref = la.Vector.uninitialized(n, "f32")
        call("la_normalize_vec", 8, ref.ptr, A.ptr, n, None)
        la.synchronize()
        got = np.empty(n, np.float32)
        call("la_download", got.ctypes.data, C, n * 4, stream)
        assert got.tobytes() == ref.to_vec().tobytes()
        call("la_free", C)
    finally:
        call("la_stream_destroy", stream)


def test_two_host_threads_on_their_own_streams(la):
    """include/linalg_b200.h: "any thread may call any function; ... concurrent calls must use distinct streams".  Two host threads,
    each with its own la_stream_create stream, run elementwise ops, reductions and a tensor-engine product at the same
    time (ctypes releases the GIL during the calls); each must get exactly the result a lone caller gets."""
    import threading
    call = la._lib.call
    n = 1 << 22
    M = K = N = 512
    want = {}
    data = {}
    for t in range(2):
        a = gen.values("f32", n, 120 + t); b = gen.values("f32", n, 130 + t)
        ga = gen.values("f32", M * K, 140 + t); gb = gen.values("f32", K * N, 150 + t)
        A = la.Vector.from_vec(a); B = la.Vector.from_vec(b)
        GA = la.Matrix.from_vec(ga, M, K); GB = la.Matrix.from_vec(gb, K, N)
        data[t] = (A, B, GA, GB)
        want[t] = ((A + B).to_vec().tobytes(), np.asarray(la.dot(A, B)).tobytes(), (GA * GB).to_vec().tobytes())
    errors = []

    def worker(t):
        try:
            A, B, GA, GB = data[t]
            stream = ctypes.c_void_p()
            call("la_stream_create", ctypes.byref(stream))
            try:
                for _ in range(20):
                    C = dev_alloc(la, n * 4); G = dev_alloc(la, M * N * 4)
                    call("la_ew_vec_vec", 8, 0, C, A.ptr, B.ptr, n, stream)
                    d = np.zeros(1, np.float32)
                    call("la_dot_vec_vec", 8, A.ptr, B.ptr, n, d.ctypes.data, stream)
                    call("la_mul_mat_mat", 8, G, GA.data.ptr, GB.data.ptr, M, K, N, stream)
                    got_c = np.empty(n, np.float32); got_g = np.empty(M * N, np.float32)
                    call("la_download", got_c.ctypes.data, C, n * 4, stream)
                    call("la_download", got_g.ctypes.data, G, M * N * 4, stream)
                    assert got_c.tobytes() == want[t][0], "elementwise result differs under concurrency"
                    assert d.tobytes() == want[t][1], "dot differs under concurrency"
                    assert got_g.tobytes() == want[t][2], "GEMM differs under concurrency"
                    call("la_free", C); call("la_free", G)
            finally:
                call("la_stream_destroy", stream)
        except Exception as e:  # noqa: BLE001
            errors.append((t, repr(e)))

    threads = [threading.Thread(target=worker, args=(t,)) for t in range(2)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors
