import sys, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
import linear_algebra_b200 as la
la.init(0); lib = la._lib.lib
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
t0 = time.time()
def p(*a): print("%.2fs" % (time.time() - t0), *a, flush=True)
A = la.Matrix(la.Vector.synth(0x5EED0003, n * n, "f32"), n, n); B = la.Matrix(la.Vector.synth(0x5EED0004, n * n, "f32"), n, n)
la.synchronize(); p("synth done")
C = A * B
la.synchronize(); p("gemm done engine", lib.la_gemm_last_engine())
Bt = B.transpose(); la.synchronize(); p("transpose done")
x = la.Vector.from_vec((np.random.default_rng(1).integers(0, 2, n) * 2 - 1).astype(np.float32))
lhs = la.mul_transpose_mat(x, C).to_vec(); p("Mv done")
lib.la_gemm_set_raster(0)
D = A * B; la.synchronize(); p("gemm m-major done")
p("equal", D == C)
lib.la_gemm_set_stream_tail(0)
E = A * B; la.synchronize(); p("gemm no split done")
e = E.to_vec(); c = C.to_vec(); p("max rel diff split vs not", float(np.max(np.abs(e - c)) / np.max(np.abs(e))))
