"""f32 GEMM at small and mid sizes, per engine (SIMT / tensor / what AUTO picks): is the AUTO threshold where it should be?"""
import sys
sys.path.insert(0, ".")
import linear_algebra_b200 as la
la.init(0); lib = la._lib
import os
DT = os.environ.get("DTYPE", "f32"); CODE = la.parameter(DT).code
for n in (128, 192, 256, 384, 512, 640, 768, 1024, 1536, 2048, 3072):
    A = la.Vector.synth(1, n * n, DT); B = la.Vector.synth(2, n * n, DT); C = la.Vector.uninitialized(n * n, DT)
    out = []
    for mode, label in ((0, "simt"), (1, "3xtf32"), (2, "auto")):
        lib.call("la_set_gemm_mode", mode)
        f = lambda: lib.call("la_mul_mat_mat", CODE, C.ptr, A.ptr, B.ptr, n, n, n, None)
        for _ in range(5): f()
        la.synchronize()
        best = 1e9
        for _ in range(5):
            e0, e1 = la.Event(), la.Event(); e0.record()
            for _ in range(10): f()
            e1.record(); best = min(best, e0.elapsed_ms(e1) / 10)
        out.append("%s(engine %d) %8.1f us %6.1f TF/s" % (label, lib.lib.la_gemm_last_engine(), best * 1e3, 2.0 * n ** 3 / best / 1e9))
    print("%5d^3  " % n + "   ".join(out), flush=True)
lib.call("la_set_gemm_mode", 2)
