#!/usr/bin/env python3
"""Runs the 3xTF32 tensor-core GEMM engine a few times at n^3 (default 8192) for ncu."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import linear_algebra_b200 as la  # noqa: E402

la.init(0)
lib = la._lib
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
A = la.Vector.synth(0x5EED0003, n * n, "f32"); B = la.Vector.synth(0x5EED1003, n * n, "f32"); C = la.Vector.uninitialized(n * n, "f32")   # U[-1,1)
lib.call("la_set_gemm_mode", 1)
for _ in range(3):
    lib.call("la_mul_mat_mat", 8, C.ptr, A.ptr, B.ptr, n, n, n, None)
la.synchronize()
e0, e1 = la.Event(), la.Event()
e0.record()
for _ in range(5):
    lib.call("la_mul_mat_mat", 8, C.ptr, A.ptr, B.ptr, n, n, n, None)
e1.record()
ms = e0.elapsed_ms(e1) / 5
print("3xTF32 n=%d: %.3f ms %.1f TFLOP/s engine %d" % (n, ms, 2.0 * n ** 3 / ms / 1e9, lib.lib.la_gemm_last_engine()))
