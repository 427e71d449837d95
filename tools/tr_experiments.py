"""Transpose kernels side by side (GPU only): the LDG/STG tile kernel (default), the TMA load + TMA store kernel (variants 20..,
also its two halves alone) and the TMA load + st.global kernel (variants 30..), each with several band heights of the tile
walk.  Prints GB/s counted as 2 * R * C * s for every line (for the half-path modes that is NOT the traffic: halve it)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import linear_algebra_b200 as la
la.init(0)
L = la._lib; lib = L.lib

def run(R, C, code=5, name="i32", reps=20):
    sz = lib.la_dtype_size(code)
    X = la.Vector.uninitialized(R * C, name); Z = la.Vector.uninitialized(R * C, name)
    v = np.array([3], la.parameter(name).np)
    if os.environ.get("LA_RANDOM"):
        L.call("la_synth", code, X.ptr, 0x5EED, R * C, None)          # random words: HBM runs slower on them than on a constant fill
    else:
        L.call("la_fill", code, X.ptr, v.ctypes.data, R * C, None)
    for _ in range(3):
        L.call("la_transpose", code, Z.ptr, X.ptr, R, C, None)
    la.synchronize()
    e0, e1 = la.Event(), la.Event()
    e0.record()
    for _ in range(reps):
        L.call("la_transpose", code, Z.ptr, X.ptr, R, C, None)
    e1.record()
    ms = e0.elapsed_ms(e1) / reps
    return ms, 2.0 * R * C * sz / ms / 1e6

shapes = [(8192, 8192), (32768, 32768)]
def line(tag):
    print("%-44s" % tag, "  ".join("%dx%d %.4f ms %5.0f GB/s" % (R, C, *run(R, C, reps=20 if R < 20000 else 5)) for R, C in shapes), flush=True)

lib.la_transpose_set_variant(0); line("default (LDG/STG 64x64)")
variants = [int(x) for x in sys.argv[1:]] or [20, 30]
for v in variants:
    lib.la_transpose_set_variant(v)
    if 20 <= v < 30:
        for mode in (1, 2):
            lib.la_transpose_set_mode(mode)
            line("v%d %s" % (v, "LOADS ONLY" if mode == 1 else "STORES ONLY"))
        lib.la_transpose_set_mode(0)
        for hints in (1, 2, 3):
            lib.la_transpose_set_hints(hints); line("v%d evict-first hints %d" % (v, hints))
        lib.la_transpose_set_hints(0)
    for band in (0, 4, 16):
        lib.la_transpose_set_band(band)
        line("v%d band %d" % (v, band))
    lib.la_transpose_set_band(-1)
lib.la_transpose_set_variant(0)
