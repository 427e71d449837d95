"""Two cold f32 tensor-engine products at 8192^3 and one la_gemm_prepare_b (run under ncu to capture the split pre-pass kernels)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import linear_algebra_b200 as la
la.init(0)
L = la._lib
n = 8192
A = la.Vector.synth(1, n * n, "f32"); B = la.Vector.synth(2, n * n, "f32"); C = la.Vector.uninitialized(n * n, "f32")
for _ in range(2):
    L.call("la_mul_mat_mat", 8, C.ptr, A.ptr, B.ptr, n, n, n, None)
la.synchronize()
P = la.Matrix(B, n, n).prepare()
la.synchronize()
