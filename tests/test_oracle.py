"""Pins the oracle (CPU, no GPU needed).

1. oracle/la_oracle.c (the "port") reproduces every golden vector the reference's own
   src/tests.rs holds for the hot path (values transcribed in tests/golden/reference_tests.json
   with the tests.rs line each one comes from).
2. The port agrees BIT-EXACTLY -- floats included, same summation order -- with the reference's
   kernels.cl itself compiled as C++ (oracle/_ref), for every dtype and every kernel, on
   seeded random inputs.  Skipped only where oracle/_ref could not be built.
"""
import json
import os

import numpy as np
import pytest

import oracle
from oracle import ADD, SUB, MUL, DIV, NP
import gen

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "reference_tests.json")))
needs_ref = pytest.mark.skipif(not oracle.ref_available(), reason="oracle/_ref not built")


def impls():
    out = [oracle.port()]
    if oracle.ref_available():
        out.append(oracle.ref())
    return out


# ------------------------------------------------------------------ 1. reference golden vectors
@pytest.mark.parametrize("impl", impls(), ids=lambda i: i.kind)
def test_golden_vec_mat_mul(impl):           # tests.rs:5-21 (f32), :125-141 (u32)
    g = GOLD["vec_mat_mul"]
    for dt in (np.float32, np.uint32):
        p = impl.mul_vec_mat(np.array(g["a"], dt), np.array(g["m"], dt), g["rows"], g["cols"])
        assert p.tolist() == g["expect"]


@pytest.mark.parametrize("impl", impls(), ids=lambda i: i.kind)
def test_golden_vec_transpose_mat_mul(impl):  # tests.rs:166-181
    g = GOLD["vec_transpose_mat_mul"]
    p = impl.mul_vec_transpose_mat(np.array(g["a"], np.uint32), np.array(g["m"], np.uint32), g["rows"], g["cols"])
    assert p.tolist() == g["expect"]


@pytest.mark.parametrize("impl", impls(), ids=lambda i: i.kind)
def test_golden_mat_mat_mul(impl):            # tests.rs:143-164
    g = GOLD["mat_mat_mul"]
    c = impl.mul_mat_mat(np.array(g["a"], np.uint32), np.array(g["b"], np.uint32), g["M"], g["K"], g["N"])
    assert c.tolist() == g["expect"]


@pytest.mark.parametrize("impl", impls(), ids=lambda i: i.kind)
def test_golden_vec_sum(impl):                # tests.rs:46-57
    for size, expect in GOLD["vec_sum"]["cases"]:
        assert int(impl.sum_vec(np.arange(size, dtype=np.uint32))) == expect


@pytest.mark.parametrize("impl", impls(), ids=lambda i: i.kind)
def test_golden_vec_vec_dot(impl):            # tests.rs:59-79
    a = np.arange(1, 971, dtype=np.uint32)
    b = np.arange(3, 973, dtype=np.uint32)[::-1].copy()
    assert int(impl.dot_vec_vec(a, b)) == GOLD["vec_vec_dot"]["expect"]


@pytest.mark.parametrize("impl", impls(), ids=lambda i: i.kind)
def test_golden_vec_eq_add_scale(impl):       # tests.rs:23-44, 183-206
    a = np.array([1, 2, 3, 4], np.uint32); b = np.array([1, 2, 3, 5], np.uint32)
    c = np.array([1, 2, 3], np.uint32); d = np.array([1, 2, 3, 4], np.uint32)
    assert impl.eq_vec(a, a) and impl.eq_vec(b, b) and impl.eq_vec(a, d) and impl.eq_vec(d, a)
    assert not impl.eq_vec(a, c) and not impl.eq_vec(c, a)
    assert not impl.eq_vec(a, b) and not impl.eq_vec(b, a)
    assert impl.ew_vec_vec(ADD, a, a).tolist() == GOLD["vec_add"]["expect"]
    assert impl.ew_vec_scl(MUL, a, 2).tolist() == GOLD["vec_mul_scl"]["expect"]


@pytest.mark.parametrize("impl", impls(), ids=lambda i: i.kind)
def test_golden_length_and_normalize(impl):   # tests.rs:240-271
    for size in GOLD["length_vec"]["sizes"]:
        a = np.arange(size, dtype=np.float32)
        length = np.sqrt(impl.dot_vec_vec(a, a))
        expect = np.sqrt(np.float32(sum(float(x) * float(x) for x in range(size))))
        assert abs(float(length) - float(expect)) < 1e-4
        v = np.arange(1, size + 1, dtype=np.float32)
        inv = np.float32(1.0) / np.sqrt(impl.dot_vec_vec(v, v))
        n = impl.ew_vec_scl(MUL, v, inv)
        assert abs(float(np.sqrt(impl.dot_vec_vec(n, n))) - 1.0) < 0.01
    assert abs(float(np.sqrt(oracle.port().dot_vec_vec(np.arange(128, dtype=np.float32),
                                                      np.arange(128, dtype=np.float32)))) - 831.19192) < 1e-3


def test_nan_is_not_equal_to_itself():        # kernels.cl:72 uses !=
    a = np.array([1.0, np.nan], np.float32)
    for impl in impls():
        assert not impl.eq_vec(a, a)
        assert impl.eq_vec(np.array([0.0], np.float32), np.array([-0.0], np.float32))


# ------------------------------------------------------------------ 2. port == kernels.cl-as-C++
def same(x, y):
    x = np.asarray(x); y = np.asarray(y)
    return x.dtype == y.dtype and x.shape == y.shape and x.tobytes() == y.tobytes()


@needs_ref
@pytest.mark.parametrize("name", gen.ALL_NAMES)
def test_port_vs_ref_elementwise(name):
    P, R = oracle.port(), oracle.ref()
    for n in (1, 7, 1000, 4099):
        a = gen.values(name, n, 1); b = gen.values(name, n, 2); d = gen.divisors(name, n, 3)
        for op, rhs in ((ADD, b), (SUB, b), (MUL, b), (DIV, d)):
            assert same(P.ew_vec_vec(op, a, rhs), R.ew_vec_vec(op, a, rhs)), (name, op, n)
            assert same(P.ew_assign_vec_vec(op, a, rhs), R.ew_assign_vec_vec(op, a, rhs)), (name, op, n)
            assert same(P.ew_vec_vec(op, a, rhs), P.ew_assign_vec_vec(op, a, rhs))
        for op, s in ((MUL, b[0]), (DIV, d[0])):
            assert same(P.ew_vec_scl(op, a, s), R.ew_vec_scl(op, a, s)), (name, op, n)
            assert same(P.ew_assign_vec_scl(op, a, s), R.ew_assign_vec_scl(op, a, s)), (name, op, n)


@needs_ref
@pytest.mark.parametrize("name", gen.ALL_NAMES)
def test_port_vs_ref_eq(name):
    P, R = oracle.port(), oracle.ref()
    a = gen.values(name, 3001, 4)
    for pos in (None, 0, 1500, 3000):
        b = a.copy()
        if pos is not None:
            b[pos] = b[pos] + NP[name](1)
        assert P.eq_vec(a, b) == R.eq_vec(a, b) == (pos is None)
    assert not P.eq_vec(a, a[:-1]) and not R.eq_vec(a, a[:-1])


@needs_ref
@pytest.mark.parametrize("name", gen.ALL_NAMES)
@pytest.mark.parametrize("W", [8, 32, 64])
def test_port_vs_ref_sum_dot(name, W):
    P, R = oracle.port(), oracle.ref()
    for n in (1, 64, 65, 971, 5000, 70001):
        a = gen.values(name, n, 5); b = gen.values(name, n, 6)
        if n > 64:
            assert same(P.reduce_groups(a, None, W), R.reduce_groups(a, None, W)), (name, n, W)
            assert same(P.reduce_groups(a, b, W), R.reduce_groups(a, b, W)), (name, n, W)
        assert same(P.sum_vec(a, W), R.sum_vec(a, W)), (name, n, W)
        assert same(P.dot_vec_vec(a, b, W), R.dot_vec_vec(a, b, W)), (name, n, W)


@needs_ref
def test_port_vs_ref_dot_contracted():
    # -ffp-contract=fast build of kernels.cl vs the port's explicit fma() path.  The compiler is
    # free not to contract, so only require the result to lie within the off/fast bracket + 1 ulp.
    P, Rf = oracle.port(), oracle.ref(fma=True)
    a = gen.values("f32", 50000, 7); b = gen.values("f32", 50000, 8)
    lo, hi = sorted([float(P.dot_vec_vec(a, b, 32, False)), float(P.dot_vec_vec(a, b, 32, True))])
    r = float(Rf.dot_vec_vec(a, b, 32))
    tol = 4 * np.spacing(np.float32(max(abs(lo), abs(hi))))
    assert lo - tol <= r <= hi + tol


@needs_ref
@pytest.mark.parametrize("name", gen.ALL_NAMES)
def test_port_vs_ref_matvec_matmat(name):
    P, R = oracle.port(), oracle.ref()
    for rows, cols in ((1, 1), (4, 3), (3, 4), (37, 53), (128, 64)):
        m = gen.values(name, rows * cols, 9)
        vr = gen.values(name, rows, 10); vc = gen.values(name, cols, 11)
        assert same(P.mul_vec_mat(vr, m, rows, cols), R.mul_vec_mat(vr, m, rows, cols)), (name, rows, cols)
        assert same(P.mul_vec_transpose_mat(vc, m, rows, cols), R.mul_vec_transpose_mat(vc, m, rows, cols))
    for M, K, N in ((1, 1, 1), (2, 3, 2), (5, 1, 7), (33, 17, 29), (64, 64, 64)):
        a = gen.values(name, M * K, 12); b = gen.values(name, K * N, 13)
        assert same(P.mul_mat_mat(a, b, M, K, N), R.mul_mat_mat(a, b, M, K, N)), (name, M, K, N)


def test_outer_product_is_k1_gemm():          # matrix.rs:236-238 mul_column_row
    P = oracle.port()
    col = gen.values("i32", 5, 14); row = gen.values("i32", 7, 15)
    with np.errstate(over="ignore"):
        expect = (col[:, None] * row[None, :]).reshape(-1)
    assert same(P.mul_mat_mat(col, row, 5, 1, 7), expect)


@pytest.mark.parametrize("name", gen.ALL_NAMES)
def test_transpose_definition(name):
    P = oracle.port()
    for R_, C_ in ((1, 1), (1, 9), (9, 1), (33, 65), (64, 32)):
        a = gen.values(name, R_ * C_, 16)
        assert same(P.transpose(a, R_, C_), a.reshape(R_, C_).T.copy().reshape(-1))


def test_integer_wrap_is_order_independent():
    # the property that makes bit-exact integer parity possible whatever order the GPU sums in
    P = oracle.port()
    for name in gen.INT_NAMES:
        a = gen.values(name, 10007, 17); b = gen.values(name, 10007, 18)
        with np.errstate(over="ignore"):
            expect = (a.astype(np.uint64) * b.astype(np.uint64)).sum(dtype=np.uint64).astype(NP[name])
            if np.dtype(NP[name]).kind == "i":
                expect = (a.astype(np.int64).astype(np.uint64) * b.astype(np.int64).astype(np.uint64)).sum(dtype=np.uint64).astype(NP[name])
        assert P.dot_vec_vec(a, b) == expect
        assert P.sum_vec(a) == a.astype(np.uint64 if np.dtype(NP[name]).kind == "u" else np.int64).sum().astype(NP[name])
