"""Tensor-engine error on structured (non-random) data per cross-term type and chain length; units of sum|a||b|."""
import sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, oracle
import linear_algebra_b200 as la
la.init(0); lib = la._lib.lib; P = oracle.port()
la._lib.call("la_set_gemm_mode", 1)
rng = np.random.default_rng(77)
for K in (2048, 8192):
    M, N = 256, 384
    cases = {
        "constant": (np.full(M * K, 1.0 / 3.0, np.float32), np.full(K * N, 0.7, np.float32)),
        "same sign": (rng.random(M * K, dtype=np.float32), rng.random(K * N, dtype=np.float32)),
        "rows of one value": (np.repeat(rng.random(M, dtype=np.float32), K), rng.random(K * N, dtype=np.float32)),
        "random +-": (rng.random(M * K, dtype=np.float32) * 2 - 1, rng.random(K * N, dtype=np.float32) * 2 - 1),
    }
    for what, (a, b) in cases.items():
        truth, absab = P.f32_mul_mat_mat_truth(a, b, M, K, N)
        e_orc = float((np.abs(P.mul_mat_mat(a, b, M, K, N).astype(np.float64) - truth) / absab).max())
        out = []
        for cross in (0, 1):
            for chunk in (2, 4, 8):
                lib.la_gemm_tc_set_cross(cross); lib.la_gemm_tc_set_chunk(chunk)
                C = (la.Matrix.from_vec(a, M, K) * la.Matrix.from_vec(b, K, N)).to_vec().astype(np.float64)
                out.append("%s/%d %.2e" % ("bf16" if cross else "tf32", chunk * 32, float((np.abs(C - truth) / absab).max())))
        print("K=%d %-18s oracle %.2e | %s" % (K, what, e_orc, "  ".join(out)), flush=True)
