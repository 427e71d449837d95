#!/usr/bin/env python3
"""Generates tests/golden/ref_vectors.npz from the REFERENCE ITSELF: /root/reference/src/kernels.cl
compiled verbatim as C++ (oracle/_ref, built by oracle/build_ref.py) and driven with the NDRanges
the reference's Rust host code enqueues.  Run in the container that has /root/reference:

    python tests/golden/make_ref_vectors.py

Seeded inputs and the reference's outputs are stored for every element type and every kernel of
the hot path, including the ops the reference's own tests do not pin (sub, div, Hadamard, *_assign,
scalar forms, eq, the 64 per-group partials of sum/dot, both matrix-vector products, mat-mat)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE))); sys.path.insert(0, os.path.dirname(HERE))
import gen      # noqa: E402
import oracle   # noqa: E402
from oracle import ADD, SUB, MUL, DIV  # noqa: E402


def main():
    R = oracle.ref()
    out = {}
    for name in gen.ALL_NAMES:
        n = 257
        a = gen.values(name, n, 101); b = gen.values(name, n, 102); d = gen.divisors(name, n, 103)
        out[name + ".a"] = a; out[name + ".b"] = b; out[name + ".d"] = d
        for opname, op, rhs in (("add", ADD, b), ("sub", SUB, b), ("mul", MUL, b), ("div", DIV, d)):
            out["%s.%s_vec_vec" % (name, opname)] = R.ew_vec_vec(op, a, rhs)
            out["%s.%s_assign_vec_vec" % (name, opname)] = R.ew_assign_vec_vec(op, a, rhs)
        out[name + ".mul_vec_scl"] = R.ew_vec_scl(MUL, a, b[0]); out[name + ".div_vec_scl"] = R.ew_vec_scl(DIV, a, d[0])
        out[name + ".mul_assign_vec_scl"] = R.ew_assign_vec_scl(MUL, a, b[0]); out[name + ".div_assign_vec_scl"] = R.ew_assign_vec_scl(DIV, a, d[0])
        c = a.copy(); c[200] = c[200] + a.dtype.type(1)
        out[name + ".eq"] = np.array([R.eq_vec(a, a), R.eq_vec(a, c), R.eq_vec(a, a[:-1])], dtype=np.uint8)
        big_a = gen.values(name, 1500, 104); big_b = gen.values(name, 1500, 105)
        out[name + ".red_a"] = big_a; out[name + ".red_b"] = big_b
        out[name + ".sum_groups_W32"] = R.reduce_groups(big_a, None, 32)
        out[name + ".dot_groups_W32"] = R.reduce_groups(big_a, big_b, 32)
        out[name + ".sum"] = np.array([R.sum_vec(big_a, 32)]); out[name + ".dot"] = np.array([R.dot_vec_vec(big_a, big_b, 32)])
        rows, cols = 37, 53
        m = gen.values(name, rows * cols, 106); vr = gen.values(name, rows, 107); vc = gen.values(name, cols, 108)
        out[name + ".m"] = m; out[name + ".vr"] = vr; out[name + ".vc"] = vc
        out[name + ".mul_vec_mat"] = R.mul_vec_mat(vr, m, rows, cols)
        out[name + ".mul_vec_transpose_mat"] = R.mul_vec_transpose_mat(vc, m, rows, cols)
        M, K, N = 33, 17, 29
        ma = gen.values(name, M * K, 109); mb = gen.values(name, K * N, 110)
        out[name + ".ma"] = ma; out[name + ".mb"] = mb
        out[name + ".mul_mat_mat"] = R.mul_mat_mat(ma, mb, M, K, N)
    path = os.path.join(HERE, "ref_vectors.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
