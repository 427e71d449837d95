"""One GEMM configuration in a fresh process, timed the way bench.py times the headline (3 warm-ups, 5 timed steps, CUDA events):
   python tools/fresh_gemm.py <dtype> <n | MxKxN> [env LA_GEMM_RASTER=...]"""
import os, sys
sys.path.insert(0, ".")
import linear_algebra_b200 as la
la.init(0); lib = la._lib
if os.environ.get("LA_CHUNK"): lib.lib.la_gemm_tc_set_chunk(int(os.environ["LA_CHUNK"]))
if os.environ.get("LA_PERSIST"): lib.lib.la_gemm_tc_set_persistent(int(os.environ["LA_PERSIST"]))
dt = sys.argv[1]; dims = [int(x) for x in sys.argv[2].split("x")]
M, K, N = dims if len(dims) == 3 else (dims[0],) * 3
A = la.Vector.synth(1, M * K, dt); B = la.Vector.synth(2, K * N, dt); C = la.Vector.uninitialized(M * N, dt)
f = lambda: lib.call("la_mul_mat_mat", la.parameter(dt).code, C.ptr, A.ptr, B.ptr, M, K, N, None)
for _ in range(3): f()
la.synchronize()
e0, e1 = la.Event(), la.Event(); e0.record()
for _ in range(5): f()
e1.record(); ms = e0.elapsed_ms(e1) / 5
print("%s %dx%dx%d persist=%s chunk=%s raster=%s engine %d: %.4f ms  %.1f T(FL)OP/s" % (dt, M, K, N, os.environ.get("LA_PERSIST", "default"), os.environ.get("LA_CHUNK", "auto"), os.environ.get("LA_GEMM_RASTER", "auto"), lib.lib.la_gemm_last_engine(), ms, 2.0 * M * K * N / ms / 1e9), flush=True)
