#!/usr/bin/env python3
"""GB/s of the HBM-bound ops at their BASELINE.json sizes (quick tuning loop; GPU only)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import linear_algebra_b200 as la  # noqa: E402

la.init(0)
lib = la._lib


def timed(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    la.synchronize()
    e0, e1 = la.Event(), la.Event()
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    return e0.elapsed_ms(e1) / iters


n = 1 << 28
A = la.Vector.synth(1, n, "f32"); B = la.Vector.synth(2, n, "f32"); C = la.Vector.uninitialized(n, "f32")
res = la.Vector.uninitialized(4, "f32"); s = np.array([1.5], np.float32)
for name, fn, nbytes in (("add", lambda: lib.call("la_ew_vec_vec", 8, 0, C.ptr, A.ptr, B.ptr, n, None), 12 * n),
                         ("scale", lambda: lib.call("la_ew_vec_scl", 8, 2, C.ptr, A.ptr, s.ctypes.data, n, None), 8 * n),
                         ("dot", lambda: lib.call("la_dot_vec_vec_dev", 8, A.ptr, B.ptr, n, res.ptr, None), 8 * n),
                         ("sum", lambda: lib.call("la_sum_vec_dev", 8, A.ptr, n, res.ptr, None), 4 * n)):
    ms = timed(fn)
    print("%-8s %.4f ms  %.0f GB/s" % (name, ms, nbytes / ms / 1e6), flush=True)
del A, B, C
r = 32768
M = la.Vector.synth(3, r * r, "f32"); v = la.Vector.synth(4, r, "f32"); y = la.Vector.uninitialized(r, "f32")
for name, fn in (("M*v", lambda: lib.call("la_mul_vec_transpose_mat", 8, y.ptr, v.ptr, M.ptr, r, r, None)),
                 ("v*M", lambda: lib.call("la_mul_vec_mat", 8, y.ptr, v.ptr, M.ptr, r, r, None))):
    ms = timed(fn, 10, 2)
    print("%-8s %.4f ms  %.0f GB/s" % (name, ms, 4 * (r * r + 2 * r) / ms / 1e6), flush=True)
for rr in (4096, 8192, 1024, 256):
    Ms = la.Vector(M.dtype, M.ptr, rr * r, owner=False)
    ms = timed(lambda: lib.call("la_mul_vec_transpose_mat", 8, y.ptr, v.ptr, Ms.ptr, rr, r, None), 10, 2)
    print("M*v %dx%d  %.4f ms  %.0f GB/s" % (rr, r, ms, 4 * (rr * r + rr + r) / ms / 1e6), flush=True)
