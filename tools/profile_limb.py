"""A few i32 / u16 / u8 8192^3 products on the INT8-limb tensor engine (run under ncu for the launch list)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import linear_algebra_b200 as la
la.init(0)
L = la._lib
n = 8192
for name, code in (("i32", 5), ("u16", 2), ("u8", 0)):
    A = la.Vector.synth(1, n * n, name); B = la.Vector.synth(2, n * n, name); C = la.Vector.uninitialized(n * n, name)
    for _ in range(2):
        L.call("la_mul_mat_mat", code, C.ptr, A.ptr, B.ptr, n, n, n, None)
    la.synchronize()
    e0, e1 = la.Event(), la.Event(); e0.record()
    for _ in range(5):
        L.call("la_mul_mat_mat", code, C.ptr, A.ptr, B.ptr, n, n, n, None)
    e1.record()
    print(name, "%.3f ms" % (e0.elapsed_ms(e1) / 5), flush=True)
    del A, B, C
