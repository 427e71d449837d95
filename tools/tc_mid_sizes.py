"""f32 tensor engine at mid sizes: default launch against the persistent grid (la_gemm_tc_set_persistent) and tile widths,
cold and with B prepared.  Median of 10 batches of back-to-back launches."""
import ctypes, os, statistics, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import linear_algebra_b200 as la
la.init(0)
L = la._lib; lib = L.lib

def med(fn, batch, iters=10):
    for _ in range(3): fn()
    la.synchronize()
    ev = [la.Event() for _ in range(iters + 1)]
    ev[0].record()
    for i in range(iters):
        for _ in range(batch): fn()
        ev[i + 1].record()
    return statistics.median(ev[i].elapsed_ms(ev[i + 1]) / batch for i in range(iters))

L.call("la_set_gemm_mode", 1)
for n in [int(x) for x in sys.argv[1:]] or [1024, 2048, 4096, 8192]:
    A = la.Vector.synth(1, n * n, "f32"); B = la.Vector.synth(2, n * n, "f32"); C = la.Vector.uninitialized(n * n, "f32")
    h = ctypes.c_void_p(); L.call("la_gemm_prepare_b", 8, B.ptr, n, n, None, ctypes.byref(h))
    batch = max(1, int(2e9 / n ** 3))
    out = []
    for persist in (0, 1):
        lib.la_gemm_tc_set_persistent(persist)
        for bn in (0, 256, 224, 192):
            lib.la_gemm_tc_set_tile_n(bn)
            cold = med(lambda: L.call("la_mul_mat_mat", 8, C.ptr, A.ptr, B.ptr, n, n, n, None), batch)
            prep = med(lambda: L.call("la_mul_mat_mat_prepared", C.ptr, A.ptr, h, n, None), batch)
            out.append("persist %d bn %3d: cold %.4f ms %.0f TF/s | prepared %.4f ms %.0f TF/s" % (persist, bn, cold, 2e-9 * n ** 3 / cold, prep, 2e-9 * n ** 3 / prep))
    lib.la_gemm_tc_set_persistent(0); lib.la_gemm_tc_set_tile_n(0)
    print("n = %d\n  " % n + "\n  ".join(out), flush=True)
    L.call("la_gemm_prepared_free", h)
