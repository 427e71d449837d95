"""One f32 product of the size given on the command line (for ncu launch lists)."""
import sys
sys.path.insert(0, ".")
import linear_algebra_b200 as la
la.init(0); lib = la._lib
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
dt = sys.argv[2] if len(sys.argv) > 2 else "f32"
import os
if os.environ.get('LA_RASTER'): lib.lib.la_gemm_set_raster(int(os.environ['LA_RASTER']))
A = la.Vector.synth(1, n * n, dt); B = la.Vector.synth(2, n * n, dt); C = la.Vector.uninitialized(n * n, dt)
for _ in range(3):
    lib.call("la_mul_mat_mat", la.parameter(dt).code, C.ptr, A.ptr, B.ptr, n, n, n, None)
la.synchronize()
print("engine", lib.lib.la_gemm_last_engine())
