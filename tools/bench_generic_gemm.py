import sys, numpy as np
sys.path.insert(0, ".")
import linear_algebra_b200 as la
la.init(0); lib = la._lib
n = 4096
variants = [int(v) for v in sys.argv[1:]] or [0]
for name, code, variant in [(nm, c, v) for nm, c in (("f64", 9), ("u64", 6)) for v in variants + [-1]] + [("i16", 3, 0)]:
    lib.lib.la_gemm_set_variant(variant)
    A = la.Vector.synth(1, n*n, name); B = la.Vector.synth(2, n*n, name); C = la.Vector.uninitialized(n*n, name)
    lib.call("la_set_gemm_mode", 0 if name == "i16" else 2)
    f = lambda: lib.call("la_mul_mat_mat", code, C.ptr, A.ptr, B.ptr, n, n, n, None)
    f(); la.synchronize()
    e0, e1 = la.Event(), la.Event(); e0.record()
    for _ in range(3): f()
    e1.record(); ms = e0.elapsed_ms(e1) / 3
    print("%s %d^3 variant %d engine %d: %.3f ms %.2f T(FL)OP/s" % (name, n, variant, lib.lib.la_gemm_last_engine(), ms, 2.0*n**3/ms/1e9), flush=True)
