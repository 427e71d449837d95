#!/usr/bin/env python3
"""GB/s of elementwise add/mul/div, sum, dot and M x v for every element type (1 GiB operands).  GPU only."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import linear_algebra_b200 as la
la.init(0); lib = la._lib

def timed(fn, iters=10, warm=2):
    for _ in range(warm): fn()
    la.synchronize()
    e0, e1 = la.Event(), la.Event(); e0.record()
    for _ in range(iters): fn()
    e1.record()
    return e0.elapsed_ms(e1) / iters

for name in ("u8", "i8c", "u16", "i16", "u32", "i32", "u64", "i64", "f32", "f64"):
    dt = la.parameter(name); n = (1 << 30) // dt.size
    A = la.Vector.synth(1, n, dt); B = la.Vector.synth(2, n, dt); C = la.Vector.uninitialized(n, dt); res = la.Vector.uninitialized(2, "u64")
    one = np.array([3], dt.np)
    lib.call("la_ew_vec_scl", dt.code, 2, B.ptr, B.ptr, one.ctypes.data, n, None)   # avoid trivial divisors; zeros possible but harmless on GPU
    row = [name]
    for op, nm in ((0, "add"), (2, "mul"), (3, "div")):
        ms = timed(lambda: lib.call("la_ew_vec_vec", dt.code, op, C.ptr, A.ptr, B.ptr, n, None))
        row.append("%s %.0f" % (nm, 3 * n * dt.size / ms / 1e6))
    ms = timed(lambda: lib.call("la_sum_vec_dev", dt.code, A.ptr, n, res.ptr, None)); row.append("sum %.0f" % (n * dt.size / ms / 1e6))
    ms = timed(lambda: lib.call("la_dot_vec_vec_dev", dt.code, A.ptr, B.ptr, n, res.ptr, None)); row.append("dot %.0f" % (2 * n * dt.size / ms / 1e6))
    r = 16384; c = n // r
    y = la.Vector.uninitialized(max(r, c), dt); v = la.Vector.synth(3, max(r, c), dt)
    ms = timed(lambda: lib.call("la_mul_vec_transpose_mat", dt.code, y.ptr, v.ptr, A.ptr, r, c, None)); row.append("M*v %.0f" % (n * dt.size / ms / 1e6))
    ms = timed(lambda: lib.call("la_mul_vec_mat", dt.code, y.ptr, v.ptr, A.ptr, r, c, None)); row.append("v*M %.0f" % (n * dt.size / ms / 1e6))
    ms = timed(lambda: lib.call("la_transpose", dt.code, C.ptr, A.ptr, r, c, None)); row.append("T %.0f" % (2 * n * dt.size / ms / 1e6))
    print("  ".join(row), "GB/s", flush=True)
    del A, B, C
