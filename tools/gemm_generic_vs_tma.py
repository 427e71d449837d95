"""SIMT engines at small and mid sizes: the generic 64x64 kernel (variant -1) against the TMA-fed kernel (variant 0), LA_GEMM_SIMT."""
import os, sys
sys.path.insert(0, ".")
import linear_algebra_b200 as la
la.init(0); lib = la._lib
lib.call("la_set_gemm_mode", 0)
for DT in os.environ.get("DTYPES", "f32,i32,f64").split(","):
    CODE = la.parameter(DT).code
    for n in (128, 256, 384, 512, 768, 1024, 1280, 1536, 2048):
        A = la.Vector.synth(1, n * n, DT); B = la.Vector.synth(2, n * n, DT); C = la.Vector.uninitialized(n * n, DT)
        out = []
        for variant in (-1, 0):
            lib.lib.la_gemm_set_variant(variant)
            f = lambda: lib.call("la_mul_mat_mat", CODE, C.ptr, A.ptr, B.ptr, n, n, n, None)
            for _ in range(5): f()
            la.synchronize()
            best = 1e9
            for _ in range(5):
                e0, e1 = la.Event(), la.Event(); e0.record()
                for _ in range(10): f()
                e1.record(); best = min(best, e0.elapsed_ms(e1) / 10)
            out.append("variant %2d (engine %d) %8.1f us %6.1f T/s" % (variant, lib.lib.la_gemm_last_engine(), best * 1e3, 2.0 * n ** 3 / best / 1e9))
        print("%s %5d^3  " % (DT, n) + "   ".join(out), flush=True)
lib.lib.la_gemm_set_variant(0)
