"""Error of the f32 tensor engine against f64, in units of sum|a||b|, with tf32 or bf16 cross terms (la_gemm_tc_set_cross),
beside the reference order's own error (serial f32 sum) and the bar max(E_oracle, 2^-22) (see also tools/short_k_probe.py)."""
import sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, gen, oracle
import linear_algebra_b200 as la
la.init(0); lib = la._lib.lib
P = oracle.port()
shapes = [tuple(int(x) for x in s.split("x")) for s in sys.argv[1:]] or [(512, 256, 512), (1024, 1024, 1024), (2048, 2048, 1024), (1024, 4096, 1024), (768, 8192, 768)]
import os
if os.environ.get("LA_CHUNK"): lib.la_gemm_tc_set_chunk(int(os.environ["LA_CHUNK"]))
la._lib.call("la_set_gemm_mode", 1)
for (M, K, N) in shapes:
    a = gen.values("f32", M * K, 65); b = gen.values("f32", K * N, 66)
    truth, absab = P.f32_mul_mat_mat_truth(a, b, M, K, N)
    orc = P.mul_mat_mat(a, b, M, K, N).astype(np.float64)
    e_orc = (np.abs(orc - truth) / absab).max()
    out = []
    for cross in (0, 1):
        lib.la_gemm_tc_set_cross(cross)
        C = (la.Matrix.from_vec(a, M, K) * la.Matrix.from_vec(b, K, N)).to_vec().astype(np.float64)
        e = np.abs(C - truth) / absab
        out.append("%s cross: max %.3e rms %.3e" % ("bf16" if cross else "tf32", e.max(), np.sqrt((e ** 2).mean())))
    print((M, K, N), "engine", lib.la_gemm_last_engine(), " | ".join(out), "| E_oracle %.3e  bar %.3e" % (e_orc, max(e_orc, 2.0 ** -22)), flush=True)
lib.la_gemm_tc_set_cross(-1)
