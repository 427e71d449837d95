import sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, gen, oracle
import linear_algebra_b200 as la
la.init(0); lib = la._lib.lib
P = oracle.port()
for (M, K, N) in ((19200, 64, 256), (19200, 32, 256), (19200, 128, 256), (19200, 256, 256), (4096, 64, 4096)):
    a = gen.values("f32", M * K, 65); b = gen.values("f32", K * N, 66)
    truth, absab = P.f32_mul_mat_mat_truth(a, b, M, K, N)
    orc = P.mul_mat_mat(a, b, M, K, N).astype(np.float64)
    e_orc = (np.abs(orc - truth) / absab).max()
    for mode, st in ((0, 0), (1, 0), (1, 1)):
        la._lib.call("la_set_gemm_mode", mode); lib.la_gemm_set_stream_tail(st)
        C = (la.Matrix.from_vec(a, M, K) * la.Matrix.from_vec(b, K, N)).to_vec().astype(np.float64)
        e = (np.abs(C - truth) / absab).max()
        print((M, K, N), "mode", mode, "stream", st, "engine", lib.la_gemm_last_engine(), "e_ours %.3e  e_oracle %.3e  bound %.3e" % (e, e_orc, max(2 * e_orc, 2.0 ** -22)), flush=True)
