"""The generic 64 x 64 x 16 GEMM kernel (small / ragged products, every dtype): time per product and rate."""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import linear_algebra_b200 as la
la.init(0); L = la._lib
L.lib.la_gemm_set_variant(-1)            # force the generic engine
L.call("la_set_gemm_mode", 0)
for name, code in (("f32", 8), ("i32", 5), ("f64", 9), ("u16", 2), ("u8", 0)):
    for n in (256, 512, 1024, 2048):
        A = la.Vector.synth(1, n * n, name); B = la.Vector.synth(2, n * n, name); C = la.Vector.uninitialized(n * n, name)
        f = lambda: L.call("la_mul_mat_mat", code, C.ptr, A.ptr, B.ptr, n, n, n, None)
        for _ in range(3): f()
        la.synchronize()
        batch = max(1, int(2e8 / n ** 3) * 4)
        ev = [la.Event() for _ in range(11)]; ev[0].record()
        for i in range(10):
            for _ in range(batch): f()
            ev[i + 1].record()
        ms = statistics.median(ev[i].elapsed_ms(ev[i + 1]) / batch for i in range(10))
        print("%s %4d^3 generic: %.4f ms %.2f T(FL)OP/s" % (name, n, ms, 2.0 * n ** 3 / ms / 1e9), flush=True)
L.lib.la_gemm_set_variant(0); L.call("la_set_gemm_mode", 2)
