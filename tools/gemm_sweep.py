#!/usr/bin/env python3
"""Times every tile-shape variant of the TMA SIMT GEMM engine (la_gemm_set_variant) on square
f32 / i32 problems and spot-checks each result.  GPU only.  Usage: tools/gemm_sweep.py [n ...]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import linear_algebra_b200 as la  # noqa: E402

VARIANTS = {0: "16x8 256x128 BK32 ffma2 (default)", 12: "16x8 256x128 BK16 ffma2", 1: "16x8 256x128 ffma", 2: "8x8 128x128 x2/SM ffma2", 3: "8x8 128x128 x2/SM ffma",
            4: "8x16 128x256 ffma2", 5: "16x8 128x256 ffma2", 6: "16x8 256x128 BK8 ffma2", 7: "8x8 128x128 x1/SM ffma2", 8: "8x16 128x256 ffma",
            9: "8x8 192x128 12 warps ffma2", 10: "8x8 128x192 12 warps ffma2"}


def main():
    la.init(0)
    lib = la._lib
    sizes = [int(x) for x in sys.argv[1:]] or [8192]
    for n in sizes:
        rng = np.random.default_rng(5)
        a = rng.random(n * n, dtype=np.float32) * 2 - 1; b = rng.random(n * n, dtype=np.float32) * 2 - 1
        A = la.Matrix.from_vec(a, n, n); B = la.Matrix.from_vec(b, n, n); C = la.Matrix.uninitialized(n, n, "f32")
        rows = rng.integers(0, n, 16); cols = rng.integers(0, n, 16)
        truth = np.array([a[r * n:(r + 1) * n].astype(np.float64) @ b[c::n].astype(np.float64) for r, c in zip(rows, cols)])
        lib.call("la_set_gemm_mode", 0)      # SIMT engine variants first
        for dtype_code, label in ((8, "f32"), (5, "i32")):
            for v, desc in VARIANTS.items():
                if label == "i32" and v not in (0, 4, 7):
                    continue
                lib.lib.la_gemm_set_variant(v)
                try:
                    def f():
                        lib.call("la_mul_mat_mat", dtype_code, C.data.ptr, A.data.ptr, B.data.ptr, n, n, n, None)
                    for _ in range(2):
                        f()
                    la.synchronize()
                    e0, e1 = la.Event(), la.Event()
                    iters = 5
                    e0.record()
                    for _ in range(iters):
                        f()
                    e1.record()
                    ms = e0.elapsed_ms(e1) / iters
                    ok = ""
                    if label == "f32":
                        c = C.to_vec()
                        got = np.array([c[r * n + cc] for r, cc in zip(rows, cols)])
                        ok = "maxerr %.2e" % float(np.max(np.abs(got - truth)))
                    print("n=%d %s variant %d (%s): %.3f ms  %.2f T%s/s  %s" % (n, label, v, desc, ms, 2.0 * n ** 3 / ms / 1e9, "FLOP" if label == "f32" else "OP", ok), flush=True)
                except Exception as e:  # noqa: BLE001
                    print("n=%d %s variant %d FAILED: %s" % (n, label, v, e), flush=True)
        lib.lib.la_gemm_set_variant(0)
        # tensor-core engine
        lib.call("la_set_gemm_mode", 1)
        for tv, desc in ((0, "3xTF32 chunked 2-CTA 256x256 cta_group::2"), (5, "3xTF32 chunked 128x256 cluster2 B-multicast"), (3, "3xTF32 chunked 128x256"), (1, "3xTF32 single-chain 128x256")):
            lib.lib.la_gemm_tc_set_variant(tv)
            try:
                def f():
                    lib.call("la_mul_mat_mat", 8, C.data.ptr, A.data.ptr, B.data.ptr, n, n, n, None)
                for _ in range(2):
                    f()
                la.synchronize()
                e0, e1 = la.Event(), la.Event()
                e0.record()
                for _ in range(5):
                    f()
                e1.record()
                ms = e0.elapsed_ms(e1) / 5
                c = C.to_vec()
                got = np.array([c[r * n + cc] for r, cc in zip(rows, cols)])
                print("n=%d f32 tc variant %d (%s): %.3f ms  %.2f TFLOP/s  maxerr %.2e  engine %d" % (n, tv, desc, ms, 2.0 * n ** 3 / ms / 1e9, float(np.max(np.abs(got - truth))), lib.lib.la_gemm_last_engine()), flush=True)
            except Exception as e:  # noqa: BLE001
                print("n=%d tc variant %d FAILED: %s" % (n, tv, e), flush=True)
        lib.call("la_set_gemm_mode", 2)
        lib.lib.la_gemm_tc_set_variant(0)


if __name__ == "__main__":
    main()
