"""oracle/cl_bench.py -- TEST INFRASTRUCTURE ONLY.  Times the reference's UNMODIFIED OpenCL kernels on the OpenCL
device the reference itself would pick (first GPU, lib.rs:356-373) -- on the GPU box that is the same B200 --
with the NDRanges of its Rust host code, kernel time from OpenCL event profiling.  Prints one JSON object.
Run as a subprocess by bench.py (`extra.reference_opencl_on_this_gpu`): a reported baseline, not a gate."""
import json
import sys
import os

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from oracle import cl_ref
    try:
        cl = cl_ref.cl_ref()
    except cl_ref.ClUnavailable as e:
        print(json.dumps({"unavailable": str(e)[:300]}))
        return
    out = {"device": cl.device_name, "opencl_library": cl.libname, "program": "kernels.cl templated as lib.rs:337-354 (9 default TYPES)"}
    rng = np.random.default_rng(3)

    def best(fn, reps=3):
        t = []
        for _ in range(reps):
            fn()
            t.append(cl.last_kernel_ns)
        return min(t) / 1e6        # ms

    # elementwise add / dot / sum over 2^28 f32 (gws = N; 64 groups x W for the reductions)
    n = 1 << 28
    a = rng.random(n, dtype=np.float32); b = rng.random(n, dtype=np.float32)
    A, B, C = cl._buf(a), cl._buf(b), cl._buf(nbytes=a.nbytes)
    ms = best(lambda: cl._launch(cl._kernel("float", "add_vec_vec"), [("buf", C), ("buf", A), ("buf", B)], n))
    out["add_vec_vec_f32_2^28"] = {"ms": round(ms, 4), "GB/s": round(12 * n / ms / 1e6, 1)}
    k = cl._kernel("float", "dot_vec_vec"); W = cl.preferred_multiple(k)
    T = cl._buf(nbytes=64 * 4)
    cnt = np.array([n], np.int32)
    ms = best(lambda: cl._launch(k, [("buf", A), ("buf", B), ("buf", T), ("val", cnt), ("local", W * 4)], 64 * W, W), 2)
    out["dot_vec_vec_f32_2^28"] = {"ms": round(ms, 3), "GB/s": round(8 * n / ms / 1e6, 1), "launch": "64 work-groups x %d (lib.rs:114)" % W}
    k = cl._kernel("float", "sum_vec")
    ms = best(lambda: cl._launch(k, [("buf", A), ("buf", T), ("val", cnt), ("local", W * 4)], 64 * W, W), 2)
    out["sum_vec_f32_2^28"] = {"ms": round(ms, 3), "GB/s": round(4 * n / ms / 1e6, 1)}
    cl._free(A, B, C, T)
    del a, b
    # matrix-vector at 32768^2 needs a 4 GiB cl_mem: beyond int indexing in kernels.cl (i * B_col_count is int): use 16384^2
    r = 16384
    m = rng.random(r * r, dtype=np.float32); v = rng.random(r, dtype=np.float32)
    Mx, V, Y = cl._buf(m), cl._buf(v), cl._buf(nbytes=r * 4)
    ri = np.array([r], np.int32)
    ms = best(lambda: cl._launch(cl._kernel("float", "mul_vec_transpose_mat"), [("buf", Y), ("buf", V), ("buf", Mx), ("val", ri), ("val", ri)], r), 2)
    out["mul_vec_transpose_mat_f32_16384^2"] = {"ms": round(ms, 3), "GB/s": round(4 * (r * r + 2 * r) / ms / 1e6, 1)}
    ms = best(lambda: cl._launch(cl._kernel("float", "mul_vec_mat"), [("buf", Y), ("buf", V), ("buf", Mx), ("val", ri), ("val", ri)], r), 2)
    out["mul_vec_mat_f32_16384^2"] = {"ms": round(ms, 3), "GB/s": round(4 * (r * r + 2 * r) / ms / 1e6, 1)}
    cl._free(Mx, V, Y)
    del m
    # GEMM: gws = M*N, one work-item per C element (matrix.rs:258-262)
    for g in (4096, 8192):
        a = rng.random(g * g, dtype=np.float32); b = rng.random(g * g, dtype=np.float32)
        A, B, C = cl._buf(a), cl._buf(b), cl._buf(nbytes=a.nbytes)
        gi = np.array([g], np.int32)
        ms = best(lambda: cl._launch(cl._kernel("float", "mul_mat_mat"), [("buf", C), ("buf", A), ("buf", B), ("val", gi), ("val", gi)], g * g), 1 if g > 4096 else 2)
        out["mul_mat_mat_f32_%d^2" % g] = {"ms": round(ms, 2), "TFLOP/s": round(2.0 * g ** 3 / ms / 1e9, 3)}
        ai = rng.integers(-2**31, 2**31 - 1, g * g, dtype=np.int32) if g == 4096 else None
        cl._free(A, B, C)
        if ai is not None:
            A, B, C = cl._buf(ai), cl._buf(ai), cl._buf(nbytes=ai.nbytes)
            ms = best(lambda: cl._launch(cl._kernel("int", "mul_mat_mat"), [("buf", C), ("buf", A), ("buf", B), ("val", gi), ("val", gi)], g * g), 1)
            out["mul_mat_mat_i32_%d^2" % g] = {"ms": round(ms, 2), "Top/s": round(2.0 * g ** 3 / ms / 1e9, 3)}
            cl._free(A, B, C)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
