"""The f32 tensor engine's pre-pass BESIDE the product (la_gemm_tc_set_pipeline) against the serial form (pre-pass, then product):
bit-identity of the results over a set of shapes (ragged ones included, B prepared or not), then event-timed steps of both.
    python tools/pipe_probe.py check            # identity only
    python tools/pipe_probe.py time [M K N]...  # timing (default: the bench shapes)"""
import ctypes, sys, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, gen
import torch
import linear_algebra_b200 as la

la.init(0); lib = la._lib.lib; call = la._lib.call
call("la_set_gemm_mode", 1)                       # LA_GEMM_3XTF32: the tensor engine whatever the size


def product(A, B, pipe, prepared=None):
    lib.la_gemm_tc_set_pipeline(pipe)
    C = A.mul_prepared(prepared) if prepared is not None else A * B
    la.synchronize()
    return C


def check():
    shapes = [(256, 1024, 256), (2048, 1024, 2048), (1024, 8192, 2048), (1000, 1028, 516), (132, 64, 136), (388, 260, 772),
              (4096, 2048, 4096), (1024, 8192, 8192)]
    bad = 0
    for (M, K, N) in shapes:
        a = gen.values("f32", M * K, 7); b = gen.values("f32", K * N, 8)
        A = la.Matrix.from_vec(a, M, K); B = la.Matrix.from_vec(b, K, N)
        ref = product(A, B, 0).to_vec()
        eng0 = lib.la_gemm_last_engine()
        for rep in range(3):
            got = product(A, B, 1).to_vec()
            same = got.tobytes() == ref.tobytes()
            bad += not same
            print("%5d x %5d x %5d  pipelined == serial: %s (engine %d / %d)" % (M, K, N, same, eng0, lib.la_gemm_last_engine()), flush=True)
        P = B.prepare()
        got = product(A, B, 1, P).to_vec()
        same = got.tobytes() == ref.tobytes()
        bad += not same
        print("%5d x %5d x %5d  pipelined, B prepared == serial: %s" % (M, K, N, same), flush=True)
    # an operand outside the split's range must still end on the SIMT engine
    M = K = N = 1024
    a = gen.values("f32", M * K, 9); b = gen.values("f32", K * N, 10); a[12345] = np.inf
    A = la.Matrix.from_vec(a, M, K); B = la.Matrix.from_vec(b, K, N)
    r0 = product(A, B, 0).to_vec(); r1 = product(A, B, 1).to_vec()
    same = r0.tobytes() == r1.tobytes()
    bad += not same
    print("inf in A: same result %s, engine %d" % (same, lib.la_gemm_last_engine()), flush=True)
    print("MISMATCHES: %d" % bad)
    return bad


def timed(fn, steps=10, warm=3):
    for _ in range(warm): fn()
    la.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    ev[0].record()
    for i in range(steps):
        fn(); ev[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(steps))
    return ts[0], ts[len(ts) // 2]


def bench(shapes):
    for (M, K, N) in shapes:
        a = gen.values("f32", M * K, 7); b = gen.values("f32", K * N, 8)
        A = la.Matrix.from_vec(a, M, K); B = la.Matrix.from_vec(b, K, N)
        C = la.Matrix.uninitialized(M, N, "f32")
        P = B.prepare()
        flops = 2.0 * M * K * N
        for pipe in (0, 1, 0, 1):
            lib.la_gemm_tc_set_pipeline(pipe)
            time.sleep(0.5)
            cold = timed(lambda: call("la_mul_mat_mat", 8, C.data.ptr, A.data.ptr, B.data.ptr, M, K, N, None))
            time.sleep(0.5)
            steady = timed(lambda: call("la_mul_mat_mat_prepared", C.data.ptr, A.data.ptr, P.handle, M, None))
            print("%5d x %5d x %5d pipe %d: cold best %.4f median %.4f ms (%.1f TFLOP/s) | B prepared best %.4f median %.4f ms (%.1f TFLOP/s)"
                  % (M, K, N, pipe, cold[0], cold[1], flops / cold[1] / 1e9, steady[0], steady[1], flops / steady[1] / 1e9), flush=True)


def one(M, K, N, reps):
    a = gen.values("f32", M * K, 7); b = gen.values("f32", K * N, 8)
    A = la.Matrix.from_vec(a, M, K); B = la.Matrix.from_vec(b, K, N)
    ref = product(A, B, 0).to_vec()
    for rep in range(reps):
        t0 = time.time()
        got = product(A, B, 1)
        dt = time.time() - t0
        print("%d x %d x %d call %d: %.2f ms wall, identical %s" % (M, K, N, rep, dt * 1e3, got.to_vec().tobytes() == ref.tobytes()), flush=True)


def stall_note():
    out = (ctypes.c_uint * 7)()
    lib.la_gemm_tc_last_stall(None, out)
    tags = {0: "none", 1: "producer: A strips (a = block, b = segment)", 2: "producer: B tiles (a = group, b = segment)", 3: "producer: empty slot (a = k-block count)",
            4: "MMA: accumulator free (a = chunk count)", 5: "MMA: slot full (a = k-block count)", 6: "epilogue: accumulator full (a = chunk count)"}
    return "stall note: tag %d (%s), block %d, a %d, b %d, seen %d, wanted %d" % (out[1], tags.get(out[1], "?"), out[2], out[3], out[4], out[5], out[6])


def back_to_back(M, K, N, calls):
    a = gen.values("f32", M * K, 7); b = gen.values("f32", K * N, 8)
    A = la.Matrix.from_vec(a, M, K); B = la.Matrix.from_vec(b, K, N)
    C = la.Matrix.uninitialized(M, N, "f32")
    ref = product(A, B, 0).to_vec()
    lib.la_gemm_tc_set_pipeline(1)
    try:
        for i in range(calls):
            call("la_mul_mat_mat", 8, C.data.ptr, A.data.ptr, B.data.ptr, M, K, N, None)
        la.synchronize()
        print("%d x %d x %d: %d calls back to back: identical %s" % (M, K, N, calls, C.to_vec().tobytes() == ref.tobytes()), flush=True)
    except Exception as e:  # noqa: BLE001
        print("%d x %d x %d: FAILED: %s" % (M, K, N, str(e)[:160]), flush=True)
        print(stall_note(), flush=True)


def kernels(M, K, N):
    """B prepared; per setting: step time and the product kernel's own time (events around its launch)."""
    a = gen.values("f32", M * K, 7); b = gen.values("f32", K * N, 8)
    A = la.Matrix.from_vec(a, M, K); B = la.Matrix.from_vec(b, K, N)
    C = la.Matrix.uninitialized(M, N, "f32")
    P = B.prepare()
    flops = 2.0 * M * K * N
    for name, pipe, persist in (("serial, usual kernel", 0, 0), ("serial, usual kernel, persistent grid", 0, 1), ("serial, lean kernel", 2, 0),
                                ("serial, lean kernel, persistent grid", 2, 1), ("pipelined", 1, 0), ("serial, usual kernel", 0, 0)):
        lib.la_gemm_tc_set_pipeline(pipe); lib.la_gemm_tc_set_persistent(persist)
        time.sleep(0.5)
        fn = lambda: call("la_mul_mat_mat_prepared", C.data.ptr, A.data.ptr, P.handle, M, None)
        for _ in range(3): fn()
        la.synchronize()
        lib.la_gemm_kernel_timing(1)
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): fn()
        e1.record(); torch.cuda.synchronize()
        ms = ctypes.c_float(); n = ctypes.c_int()
        lib.la_gemm_kernel_time_ms(ctypes.byref(ms), ctypes.byref(n))
        lib.la_gemm_kernel_timing(0)
        step = e0.elapsed_time(e1) / 10
        print("%5d x %5d x %5d  %-40s step %.4f ms (%.1f TFLOP/s), product kernel %.4f ms x %d" % (M, K, N, name, step, flops / step / 1e9, ms.value / max(n.value, 1), n.value), flush=True)
    lib.la_gemm_tc_set_persistent(0)


if __name__ == "__main__":
    if sys.argv[1] == "kernels":
        args = [int(x) for x in sys.argv[2:]]
        for i in range(0, len(args), 3): kernels(*args[i:i + 3])
        sys.exit(0)
    if sys.argv[1] == "b2b":
        back_to_back(int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5]))
        sys.exit(0)
    if sys.argv[1] == "one":
        one(int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5]))
        sys.exit(0)
    if sys.argv[1] == "check":
        sys.exit(1 if check() else 0)
    args = [int(x) for x in sys.argv[2:]]
    shapes = [tuple(args[i:i + 3]) for i in range(0, len(args), 3)] or [(8192, 8192, 8192), (1024, 8192, 8192), (4096, 4096, 4096), (2048, 2048, 2048)]
    bench(shapes)
