#!/usr/bin/env python3
"""Runs each HBM-bound op of the hot path once or twice at its BASELINE.json size, for ncu.
GPU only.  Usage: tools/profile_ops.py [ops...]   (default: all)"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import linear_algebra_b200 as la  # noqa: E402


def main():
    la.init(0)
    lib = la._lib
    want = set(sys.argv[1:])
    one = np.array([1.0], np.float32); s = np.array([1.5], np.float32); three = np.array([3], np.int32)

    def on(name):
        return not want or name in want

    reps = 2
    if on("ew") or on("reduce"):
        n = 1 << 28
        A = la.Vector.uninitialized(n, "f32"); B = la.Vector.uninitialized(n, "f32"); C = la.Vector.uninitialized(n, "f32")
        lib.call("la_fill", 8, A.ptr, one.ctypes.data, n, None); lib.call("la_fill", 8, B.ptr, s.ctypes.data, n, None)
        res = la.Vector.uninitialized(4, "f32")
        for _ in range(reps):
            if on("ew"):
                lib.call("la_ew_vec_vec", 8, 0, C.ptr, A.ptr, B.ptr, n, None)
                lib.call("la_ew_vec_scl", 8, 2, C.ptr, A.ptr, s.ctypes.data, n, None)
            if on("reduce"):
                lib.call("la_dot_vec_vec_dev", 8, A.ptr, B.ptr, n, res.ptr, None)
                lib.call("la_sum_vec_dev", 8, A.ptr, n, res.ptr, None)
        la.synchronize()
        del A, B, C
    if on("gemv"):
        r = 32768
        M = la.Vector.uninitialized(r * r, "f32"); v = la.Vector.uninitialized(r, "f32"); y = la.Vector.uninitialized(r, "f32")
        lib.call("la_fill", 8, M.ptr, one.ctypes.data, r * r, None); lib.call("la_fill", 8, v.ptr, s.ctypes.data, r, None)
        for _ in range(reps):
            lib.call("la_mul_vec_transpose_mat", 8, y.ptr, v.ptr, M.ptr, r, r, None)
            lib.call("la_mul_vec_mat", 8, y.ptr, v.ptr, M.ptr, r, r, None)
        la.synchronize()
        del M, v, y
    if on("transpose"):
        k = 8192
        X = la.Vector.uninitialized(k * k, "i32"); Z = la.Vector.uninitialized(k * k, "i32")
        lib.call("la_fill", 5, X.ptr, three.ctypes.data, k * k, None)
        for _ in range(reps):
            lib.call("la_transpose", 5, Z.ptr, X.ptr, k, k, None)
        la.synchronize()
    print("done")


if __name__ == "__main__":
    main()
