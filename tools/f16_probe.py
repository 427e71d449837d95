"""The f16-split form of the f32 tensor engine (la_gemm_tc_set_f16) beside the tf32 form:
    python tools/f16_probe.py acc [K ...]     worst error / sum|a||b| of both forms against the reference order's own (the gate, factor 1)
    python tools/f16_probe.py time            event-timed steps of both forms, cold and with B prepared
LA_MN (default 2048) sets M = N for `acc`; LA_CHUNK forces the chain length (k-blocks of 32)."""
import ctypes, os, sys, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, gen, oracle
import torch
import linear_algebra_b200 as la

la.init(0); lib = la._lib.lib; call = la._lib.call
call("la_set_gemm_mode", 1)
if os.environ.get("LA_CHUNK"): lib.la_gemm_tc_set_chunk(int(os.environ["LA_CHUNK"]))


def acc(Ks):
    P = oracle.port()
    M = N = int(os.environ.get("LA_MN", "2048"))
    for K in Ks:
        for seed in (65, 165):
            a = gen.values("f32", M * K, seed); b = gen.values("f32", K * N, seed + 1)
            if seed == 165:          # rows / columns of very different scales, full-precision mantissas
                rng = np.random.default_rng(seed)
                a = (rng.standard_normal(M * K).astype(np.float32).reshape(M, K) * np.exp2(rng.integers(-20, 20, M)).astype(np.float32)[:, None]).reshape(-1)
                b = (rng.standard_normal(K * N).astype(np.float32).reshape(K, N) * np.exp2(rng.integers(-20, 20, N)).astype(np.float32)[None, :]).reshape(-1)
            truth, absab = P.f32_mul_mat_mat_truth(a, b, M, K, N)
            e0 = (np.abs(P.mul_mat_mat(a, b, M, K, N).astype(np.float64) - truth) / absab).max()
            e1 = (np.abs(P.mul_mat_mat(a, b, M, K, N, fma=True).astype(np.float64) - truth) / absab).max()
            A = la.Matrix.from_vec(a, M, K); B = la.Matrix.from_vec(b, K, N)
            out = []
            for f16 in (1, 0):
                lib.la_gemm_tc_set_f16(f16)
                C = (A * B).to_vec().astype(np.float64)
                eng = lib.la_gemm_last_engine()
                Cp = A.mul_prepared(B.prepare()).to_vec().astype(np.float64)
                out.append("%s %.3e (engine %d, prepared identical %s)" % ("f16" if f16 else "tf32", (np.abs(C - truth) / absab).max(), eng, np.array_equal(C, Cp)))
            bar = max(e0, e1, 2.0 ** -22)
            print("K=%d %s: %s | reference order %.3e (fma %.3e) bar %.3e" % (K, "uniform" if seed == 65 else "scaled normal", "  ".join(out), e0, e1, bar), flush=True)
    lib.la_gemm_tc_set_f16(-1)


def timed(fn, steps=10, warm=3):
    for _ in range(warm): fn()
    la.synchronize()
    lib.la_gemm_kernel_timing(1)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps): fn()
    e1.record(); torch.cuda.synchronize()
    ms = ctypes.c_float(); n = ctypes.c_int()
    lib.la_gemm_kernel_time_ms(ctypes.byref(ms), ctypes.byref(n)); lib.la_gemm_kernel_timing(0)
    return e0.elapsed_time(e1) / steps, ms.value / max(n.value, 1)


def bench(shapes):
    for (M, K, N) in shapes:
        A = la.Vector.synth(1, M * K, "f32"); B = la.Vector.synth(2, K * N, "f32"); C = la.Vector.uninitialized(M * N, "f32")
        flops = 2.0 * M * K * N
        for f16 in (0, 1, 0, 1):
            lib.la_gemm_tc_set_f16(f16)
            P = la.Matrix(B, K, N).prepare()
            time.sleep(0.5)
            cold = timed(lambda: call("la_mul_mat_mat", 8, C.ptr, A.ptr, B.ptr, M, K, N, None))
            time.sleep(0.5)
            steady = timed(lambda: call("la_mul_mat_mat_prepared", C.ptr, A.ptr, P.handle, M, None))
            print("%5d x %5d x %5d %-4s: cold %.4f ms (%.1f TFLOP/s, kernel %.4f) | B prepared %.4f ms (%.1f TFLOP/s, kernel %.4f)"
                  % (M, K, N, "f16" if f16 else "tf32", cold[0], flops / cold[0] / 1e9, cold[1], steady[0], flops / steady[0] / 1e9, steady[1]), flush=True)
            P.free()
    lib.la_gemm_tc_set_f16(-1)


def tiny():
    """Elements 2^-29 and more below their row's / column's maximum are left out of the planes and added afterwards in f32."""
    P = oracle.port()
    M, K, N = 384, 1024, 512
    rng = np.random.default_rng(5)
    a = rng.standard_normal(M * K).astype(np.float32); b = rng.standard_normal(K * N).astype(np.float32)
    for idx in rng.integers(0, M * K, 40): a[idx] = np.float32(1e-12) * np.float32(rng.standard_normal())
    for idx in rng.integers(0, K * N, 40): b[idx] = np.float32(1e-13) * np.float32(rng.standard_normal())
    a[5 * K + 7] = 1e-12; b[7 * N + 9] = 1e-12                  # a noted A element meeting a noted B element
    truth, absab = P.f32_mul_mat_mat_truth(a, b, M, K, N)
    e0 = (np.abs(P.mul_mat_mat(a, b, M, K, N).astype(np.float64) - truth) / absab).max()
    A = la.Matrix.from_vec(a, M, K); B = la.Matrix.from_vec(b, K, N)
    for f16 in (1, 0):
        lib.la_gemm_tc_set_f16(f16)
        C1 = (A * B).to_vec(); C2 = (A * B).to_vec(); Cp = A.mul_prepared(B.prepare()).to_vec()
        print("%s: worst error %.3e (reference order %.3e); reproducible %s; prepared identical %s" % ("f16" if f16 else "tf32", (np.abs(C1.astype(np.float64) - truth) / absab).max(), e0,
              C1.tobytes() == C2.tobytes(), C1.tobytes() == Cp.tobytes()), flush=True)
    # one column of B made of tiny values only next to one large one: exact zero planes there, everything through the fix
    b2 = b.copy().reshape(K, N); b2[:, 3] = 1e-20; b2[0, 3] = 1.0; b2 = b2.reshape(-1)
    truth, absab = P.f32_mul_mat_mat_truth(a, b2, M, K, N)
    B2 = la.Matrix.from_vec(b2, K, N)
    lib.la_gemm_tc_set_f16(1)
    C = (A * B2).to_vec()
    print("column of 1023 left-out elements: worst error %.3e" % (np.abs(C.astype(np.float64) - truth) / absab).max(), flush=True)
    b3 = b.copy().reshape(K, N); b3[:, 3:6] = 1e-20; b3[0, 3:6] = 1.0; b3 = b3.reshape(-1)          # 3069 > the list's capacity: the tf32 form takes over
    truth, absab = P.f32_mul_mat_mat_truth(a, b3, M, K, N)
    C = (A * la.Matrix.from_vec(b3, K, N)).to_vec()
    print("three such columns (list overflows): worst error %.3e" % (np.abs(C.astype(np.float64) - truth) / absab).max(), flush=True)
    lib.la_gemm_tc_set_f16(-1)
    # outputs that are essentially ONE product: B = identity / a permutation with signs / a diagonal (the reference is exact there)
    M = K = N = 1024
    a = rng.standard_normal(M * K).astype(np.float32)
    eye = np.eye(K, dtype=np.float32)
    perm = eye[rng.permutation(K)] * rng.choice(np.float32([-1.0, 1.0]), K)[:, None]
    diag = np.diag(rng.standard_normal(K).astype(np.float32))
    A = la.Matrix.from_vec(a, M, K)
    for name, bm in (("identity", eye), ("signed permutation", perm), ("diagonal", diag)):
        b = np.ascontiguousarray(bm).reshape(-1)
        truth, absab = P.f32_mul_mat_mat_truth(a, b, M, K, N)
        e0 = (np.abs(P.mul_mat_mat(a, b, M, K, N).astype(np.float64) - truth) / np.maximum(absab, 1e-300)).max()
        out = []
        for f16 in (1, 0):
            lib.la_gemm_tc_set_f16(f16)
            C = (A * la.Matrix.from_vec(b, K, N)).to_vec().astype(np.float64)
            out.append("%s %.3e" % ("f16" if f16 else "tf32", (np.abs(C - truth) / np.maximum(absab, 1e-300)).max()))
        print("B = %s: %s | reference order %.3e, bar %.3e" % (name, "  ".join(out), e0, max(e0, 2.0 ** -22)), flush=True)
    lib.la_gemm_tc_set_f16(-1)


if __name__ == "__main__":
    if sys.argv[1] == "tiny":
        tiny(); sys.exit(0)
    if sys.argv[1] == "acc":
        acc([int(x) for x in sys.argv[2:]] or [1024, 2048, 4096, 8192])
    else:
        args = [int(x) for x in sys.argv[2:]]
        bench([tuple(args[i:i + 3]) for i in range(0, len(args), 3)] or [(8192, 8192, 8192), (1024, 8192, 8192), (4096, 4096, 4096), (2048, 2048, 2048)])
