"""Where may LA_GEMM_AUTO start using the f32 tensor engine?  For M = N = 4096 (16M outputs) and a range of K, the worst error of
the tensor engine (LA_GEMM_3XTF32) and of the SIMT engine in units of sum|a||b|, beside the reference order's own worst error in
both contraction modes -- the gate is e(ours) <= max(E_oracle, 2^-22).
LA_CHUNK / LA_CROSS in the environment force the chain length (k-blocks of 32) / the cross-term type (0 tf32, 1 bf16)."""
import os, sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, gen, oracle
import linear_algebra_b200 as la
la.init(0); lib = la._lib.lib
P = oracle.port()
if os.environ.get("LA_CHUNK"): lib.la_gemm_tc_set_chunk(int(os.environ["LA_CHUNK"]))
if os.environ.get("LA_ADIRECT"): lib.la_gemm_tc_set_a_direct(int(os.environ["LA_ADIRECT"]))
if os.environ.get("LA_CROSS"): lib.la_gemm_tc_set_cross(int(os.environ["LA_CROSS"]))
Ks = [int(x) for x in sys.argv[1:]] or [64, 96, 128, 160, 192, 224, 256, 320, 384, 512, 1024]
M = N = int(os.environ.get("LA_MN", "4096"))
for K in Ks:
    for seed in (65, 165):
        a = gen.values("f32", M * K, seed); b = gen.values("f32", K * N, seed + 1)
        truth, absab = P.f32_mul_mat_mat_truth(a, b, M, K, N)
        e0 = (np.abs(P.mul_mat_mat(a, b, M, K, N).astype(np.float64) - truth) / absab).max()
        e1 = (np.abs(P.mul_mat_mat(a, b, M, K, N, fma=True).astype(np.float64) - truth) / absab).max()
        A = la.Matrix.from_vec(a, M, K); B = la.Matrix.from_vec(b, K, N)
        out = []
        for mode in (1, 0):
            la._lib.call("la_set_gemm_mode", mode)
            C = (A * B).to_vec().astype(np.float64)
            out.append("%s(eng %d) %.3e" % ("tensor" if mode else "simt", lib.la_gemm_last_engine(), (np.abs(C - truth) / absab).max()))
        la._lib.call("la_set_gemm_mode", 2)
        bar = max(e0, e1, 2.0 ** -22)
        print("K=%d seed %d: %s | E_oracle %.3e (fma %.3e) bar %.3e" % (K, seed, "  ".join(out), e0, e1, bar), flush=True)
