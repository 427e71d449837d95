"""Tile-order sweep of the tensor-core GEMM engines: la_gemm_set_raster(group).  Groups are interleaved over several
rounds (the GPU's power state drifts over a run, so back-to-back blocks per setting mislead)."""
import os
import sys
sys.path.insert(0, ".")
import linear_algebra_b200 as la
la.init(0); lib = la._lib
groups = [int(g) for g in sys.argv[1:]] or [0, 4, 8, 16]
hook = {"stream": lib.lib.la_gemm_set_stream_tail, "bn": lib.lib.la_gemm_tc_set_tile_n}.get(os.environ.get("KNOB"), lib.lib.la_gemm_set_raster)
def shape(spec):
    d = [int(x) for x in spec.split("x")]
    return (d[0], d[0], d[0]) if len(d) == 1 else tuple(d)
cases = [(c.split(":")[0], 8 if c.startswith("f32") else 5, shape(c.split(":")[1])) for c in os.environ.get("CASES", "f32:8192,f32:16384,i32:8192,i32:16384").split(",")]
rounds = int(os.environ.get("ROUNDS", "4"))
for name, code, (M, K, N) in cases:
    n = max(M, K, N)
    A = la.Vector.synth(1, M * K, name); B = la.Vector.synth(2, K * N, name); C = la.Vector.uninitialized(M * N, name)
    f = lambda: lib.call("la_mul_mat_mat", code, C.ptr, A.ptr, B.ptr, M, K, N, None)
    for _ in range(6): f()
    la.synchronize()
    res = {g: [] for g in groups}
    for r in range(rounds):
        for g in (groups if r % 2 == 0 else groups[::-1]):
            hook(g)
            f(); la.synchronize()
            reps = 4 if n <= 8192 else 2
            e0, e1 = la.Event(), la.Event(); e0.record()
            for _ in range(reps): f()
            e1.record(); res[g].append(e0.elapsed_ms(e1) / reps)
    for g in groups:
        ms = sorted(res[g])[len(res[g]) // 2]
        print("%s %dx%dx%d knob %2d engine %d: median %8.3f ms %7.2f T(FL)OP/s   all: %s" % (name, M, K, N, g, lib.lib.la_gemm_last_engine(), ms, 2.0 * M * K * N / ms / 1e9, " ".join("%.3f" % x for x in res[g])), flush=True)
    del A, B, C
