"""GPU parity: the CUDA path (through the C ABI) against the oracle on the same seeded inputs.

Bars (DESIGN.md "Parity"):
  * integers, every op incl. sum/dot/GEMV/GEMM/transpose: BIT-EXACT (wrap-around arithmetic is
    order independent);
  * f32/f64 add/sub/mul/div/scale and transpose: BIT-EXACT (single IEEE operation);
  * f32 sum/dot/GEMV/GEMM (different, fixed summation order): with x64 the f64 truth and
    e(x) = |x - x64| / sum_k |a_k b_k|, require for every output (as written)
        e(ours) <= max(E_oracle, 2^-22),   E_oracle = max over the call's outputs of e(oracle),
    i.e. no output is worse than the worst error the reference's own summation order commits on
    that problem (rounding errors of two orders are independent draws, so an element-by-element
    comparison would fail by chance), and normwise ||ours - oracle|| / ||oracle|| <= 1e-5.
    "The reference's own order" is src/kernels.cl's k-ascending loop in the two forms OpenCL
    permits its compiler to emit: `res += a*b` as a multiply and an add, or contracted into one
    fma ("-ffp-contract=off AND =fast to bracket"); E_oracle is the larger
    of the two.  The only place a factor 2 survives is the explicit opt-in LA_GEMM_3XTF32 below
    K = LA_GEMM_AUTO_MIN_K (tc_factor): LA_GEMM_AUTO never sends such products to the tensor engine.
"""
import ctypes

import numpy as np
import pytest

import gen
import oracle
from oracle import ADD, SUB, MUL, DIV, NP

pytestmark = pytest.mark.gpu

F32_EPS_FLOOR = 2.0 ** -22
NORMWISE_TOL = 1e-5


@pytest.fixture(scope="module")
def la():
    import linear_algebra_b200 as la
    la.init(0)
    return la


def same(x, y):
    x = np.asarray(x); y = np.asarray(y)
    return x.dtype == y.dtype and x.shape == y.shape and x.tobytes() == y.tobytes()


def dt(la, name):
    # oracle name -> device Parameter ("i8" in the oracle is OpenCL char = I8_CHAR here)
    return la.I8_CHAR if name == "i8" else la.parameter(name)


# ------------------------------------------------------------------ elementwise
@pytest.mark.parametrize("name", gen.ALL_NAMES)
def test_elementwise_all_forms(la, name):
    P = oracle.port()
    for n in (1, 7, 1000, 4099, (1 << 20) + 3):
        a = gen.values(name, n, 1); b = gen.values(name, n, 2); d = gen.divisors(name, n, 3)
        A = la.Vector.from_vec(a, dt(la, name)); B = la.Vector.from_vec(b, dt(la, name)); D = la.Vector.from_vec(d, dt(la, name))
        assert same((A + B).to_vec(), P.ew_vec_vec(ADD, a, b)), (name, n, "add")
        assert same((A - B).to_vec(), P.ew_vec_vec(SUB, a, b)), (name, n, "sub")
        assert same((A * B).to_vec(), P.ew_vec_vec(MUL, a, b)), (name, n, "mul")
        assert same((A / D).to_vec(), P.ew_vec_vec(DIV, a, d)), (name, n, "div")
        assert same((A * b[0]).to_vec(), P.ew_vec_scl(MUL, a, b[0])), (name, n, "mul_scl")
        assert same((A / d[0]).to_vec(), P.ew_vec_scl(DIV, a, d[0])), (name, n, "div_scl")
        for op, rhs, RHS, f in ((ADD, b, B, "__iadd__"), (SUB, b, B, "__isub__"), (MUL, b, B, "__imul__"), (DIV, d, D, "__itruediv__")):
            C = A.clone(); getattr(C, f)(RHS)
            assert same(C.to_vec(), P.ew_assign_vec_vec(op, a, rhs)), (name, n, f)
        C = A.clone(); C *= b[0]
        assert same(C.to_vec(), P.ew_assign_vec_scl(MUL, a, b[0]))
        C = A.clone(); C /= d[0]
        assert same(C.to_vec(), P.ew_assign_vec_scl(DIV, a, d[0]))
        assert same(A.to_vec(), a)        # operands untouched


def test_rust_i8_maps_to_uchar_like_the_reference(la):
    # traits.rs:36: Vector<i8> runs the "uchar" kernels -> unsigned divide
    P = oracle.port()
    a = np.array([-128, -7, 100, -1], np.int8); d = np.array([3, 2, 7, 5], np.int8)
    got = (la.Vector.from_vec(a, la.I8) / la.Vector.from_vec(d, la.I8)).to_vec()
    assert same(got.view(np.uint8), P.ew_vec_vec(DIV, a.view(np.uint8), d.view(np.uint8)))
    got_signed = (la.Vector.from_vec(a, la.I8_CHAR) / la.Vector.from_vec(d, la.I8_CHAR)).to_vec()
    assert same(got_signed, P.ew_vec_vec(DIV, a, d))
    assert not same(got, got_signed)


@pytest.mark.parametrize("name", ["u8", "i16", "f32", "u64"])
def test_elementwise_unaligned_pointers(la, name):
    """Raw C-ABI call with operands that are not 16-byte aligned (element-granular path)."""
    P = oracle.port(); lib = la._lib.lib; code = dt(la, name).code; sz = np.dtype(NP[name]).itemsize
    n = 10001
    a = gen.values(name, n + 1, 4); b = gen.values(name, n + 1, 5)
    A = la.Vector.from_vec(a, dt(la, name)); B = la.Vector.from_vec(b, dt(la, name)); C = la.Vector.uninitialized(n + 1, dt(la, name))
    la._lib.call("la_ew_vec_vec", code, MUL, C.ptr + sz, A.ptr + sz, B.ptr + sz, n, None)
    la.synchronize()
    assert same(C.to_vec()[1:], P.ew_vec_vec(MUL, a[1:], b[1:]))
    out = np.zeros(1, NP[name])
    la._lib.call("la_dot_vec_vec", code, A.ptr + sz, B.ptr + sz, n, out.ctypes.data, None)
    if np.dtype(NP[name]).kind != "f":
        assert out[0] == P.dot_vec_vec(a[1:], b[1:])
    eq = ctypes.c_int(-1)
    la._lib.call("la_eq_vec", code, A.ptr + sz, A.ptr + sz, n, ctypes.byref(eq), None)
    assert eq.value == 1
    assert lib.la_last_error() is not None


# ------------------------------------------------------------------ eq
@pytest.mark.parametrize("name", gen.ALL_NAMES)
def test_eq(la, name):
    a = gen.values(name, 300001, 6)
    A = la.Vector.from_vec(a, dt(la, name))
    for pos in (None, 0, 150000, 300000):
        b = a.copy()
        if pos is not None:
            b[pos] = b[pos] + NP[name](1)
        B = la.Vector.from_vec(b, dt(la, name))
        assert (A == B) == (pos is None) == oracle.port().eq_vec(a, b)
        assert (A != B) == (pos is not None)
    assert A != la.Vector.from_vec(a[:-1], dt(la, name))      # vector.rs:649-651
    assert la.Vector.from_vec([], dt(la, name)) == la.Vector.from_vec([], dt(la, name))


def test_eq_float_semantics(la):
    nan = la.Vector.from_vec(np.array([1.0, np.nan, 2.0], np.float32))
    assert not (nan == nan)                   # kernels.cl:72 uses !=
    assert la.Vector.from_vec(np.array([0.0], np.float32)) == la.Vector.from_vec(np.array([-0.0], np.float32))


# ------------------------------------------------------------------ sum / dot
SIZES_RED = (1, 2, 64, 65, 971, 70001, (1 << 22) + 5)


@pytest.mark.parametrize("name", gen.INT_NAMES)
def test_sum_dot_integers_bit_exact(la, name):
    P = oracle.port()
    for n in SIZES_RED:
        a = gen.values(name, n, 7); b = gen.values(name, n, 8)
        A = la.Vector.from_vec(a, dt(la, name)); B = la.Vector.from_vec(b, dt(la, name))
        assert A.sum() == P.sum_vec(a), (name, n)
        assert la.dot(A, B) == P.dot_vec_vec(a, b), (name, n)
    assert la.Vector.from_vec([], dt(la, name)).sum() == 0


def check_float_dot_like(ours, orc, truth, absdot, what, orc_fma=None, factor=1.0):
    """`orc` = the reference's order compiled without contraction, `orc_fma` = with (both are the
    reference's own order); `factor` is 1 everywhere except the one documented opt-in case."""
    ours = np.atleast_1d(np.asarray(ours, np.float64)); orc = np.atleast_1d(np.asarray(orc, np.float64))
    truth = np.atleast_1d(truth); absdot = np.maximum(np.atleast_1d(absdot), np.finfo(np.float64).tiny)
    e_ours = np.abs(ours - truth) / absdot
    e_orc = float((np.abs(orc - truth) / absdot).max())
    if orc_fma is not None:
        e_orc = max(e_orc, float((np.abs(np.atleast_1d(np.asarray(orc_fma, np.float64)) - truth) / absdot).max()))
    bound = max(factor * e_orc, F32_EPS_FLOOR)
    assert float(e_ours.max()) <= bound, (what, float(e_ours.max()), e_orc)
    denom = np.linalg.norm(orc)
    if denom > 0:
        assert np.linalg.norm(ours - orc) / denom <= NORMWISE_TOL, what


def test_sum_dot_f32_tolerance(la):
    P = oracle.port()
    for n in SIZES_RED:
        a = gen.values("f32", n, 9); b = gen.values("f32", n, 10)
        A = la.Vector.from_vec(a); B = la.Vector.from_vec(b)
        d64, ad = P.f32_dot_truth(a, b)
        check_float_dot_like(la.dot(A, B), P.dot_vec_vec(a, b), d64, ad, ("dot", n), P.dot_vec_vec(a, b, fma=True))
        s64, asum = P.f32_dot_truth(a, np.ones_like(a))
        check_float_dot_like(A.sum(), P.sum_vec(a), s64, asum, ("sum", n))


def test_sum_dot_f64(la):
    a = gen.values("f64", 100003, 11); b = gen.values("f64", 100003, 12)
    A = la.Vector.from_vec(a); B = la.Vector.from_vec(b)
    assert abs(float(la.dot(A, B)) - float(np.dot(a, b))) <= 1e-12 * float(np.abs(a * b).sum())
    assert abs(float(A.sum()) - float(a.sum())) <= 1e-12 * float(np.abs(a).sum())


def test_sum_dot_are_reproducible(la):
    a = gen.values("f32", (1 << 22) + 5, 13); A = la.Vector.from_vec(a)
    first = la.dot(A, A)
    assert all(same(la.dot(A, A), first) for _ in range(5))


def test_dev_result_variants(la):
    a = gen.values("i32", 100000, 14); A = la.Vector.from_vec(a)
    out = la.Vector.uninitialized(2, "i32")
    la._lib.call("la_sum_vec_dev", A.dtype.code, A.ptr, A.len(), out.ptr, None)
    la._lib.call("la_dot_vec_vec_dev", A.dtype.code, A.ptr, A.ptr, A.len(), out.ptr + 4, None)
    got = out.to_vec()
    assert got[0] == oracle.port().sum_vec(a) and got[1] == oracle.port().dot_vec_vec(a, a)


# ------------------------------------------------------------------ matrix x vector
SHAPES_MV = ((1, 1), (4, 3), (3, 4), (37, 53), (128, 64), (1000, 4096), (4096, 1000), (513, 2052), (5000, 128), (2, 40000))


@pytest.mark.parametrize("name", gen.INT_NAMES)
def test_matvec_integers_bit_exact(la, name):
    P = oracle.port()
    for rows, cols in SHAPES_MV:
        m = gen.values(name, rows * cols, 15); vr = gen.values(name, rows, 16); vc = gen.values(name, cols, 17)
        M = la.Matrix.from_vec(m, rows, cols, dt(la, name))
        assert same((la.Vector.from_vec(vr, dt(la, name)) * M).to_vec(), P.mul_vec_mat(vr, m, rows, cols)), (name, rows, cols, "v*M")
        assert same(la.mul_transpose_mat(la.Vector.from_vec(vc, dt(la, name)), M).to_vec(), P.mul_vec_transpose_mat(vc, m, rows, cols)), (name, rows, cols, "M*v")


def test_matvec_f32_tolerance(la):
    P = oracle.port()
    for rows, cols in SHAPES_MV:
        m = gen.values("f32", rows * cols, 18); vr = gen.values("f32", rows, 19); vc = gen.values("f32", cols, 20)
        M = la.Matrix.from_vec(m, rows, cols)
        m64 = m.astype(np.float64).reshape(rows, cols)
        ours = (la.Vector.from_vec(vr) * M).to_vec()
        check_float_dot_like(ours, P.mul_vec_mat(vr, m, rows, cols), vr.astype(np.float64) @ m64, np.abs(vr.astype(np.float64)) @ np.abs(m64), ("v*M", rows, cols),
                             P.mul_vec_mat(vr, m, rows, cols, fma=True))
        ours = la.mul_transpose_mat(la.Vector.from_vec(vc), M).to_vec()
        check_float_dot_like(ours, P.mul_vec_transpose_mat(vc, m, rows, cols), m64 @ vc.astype(np.float64), np.abs(m64) @ np.abs(vc.astype(np.float64)), ("M*v", rows, cols),
                             P.mul_vec_transpose_mat(vc, m, rows, cols, fma=True))


# ------------------------------------------------------------------ matrix x matrix
SHAPES_MM = ((1, 1, 1), (2, 3, 2), (5, 1, 7), (33, 17, 29), (64, 64, 64), (256, 256, 256), (320, 100, 192), (512, 1000, 384), (130, 70, 258))


@pytest.mark.parametrize("name", gen.INT_NAMES)
def test_matmat_integers_bit_exact(la, name):
    P = oracle.port()
    for M, K, N in SHAPES_MM:
        a = gen.values(name, M * K, 21); b = gen.values(name, K * N, 22)
        C = la.Matrix.from_vec(a, M, K, dt(la, name)) * la.Matrix.from_vec(b, K, N, dt(la, name))
        assert C.len() == (M, N)
        assert same(C.to_vec(), P.mul_mat_mat(a, b, M, K, N)), (name, M, K, N)


AUTO_MIN_K = 256          # LA_GEMM_AUTO_MIN_K of include/linalg_b200.h


def tc_factor(K):
    """Bar factor for a product FORCED onto the tensor engine with la_set_gemm_mode(LA_GEMM_3XTF32).  From K = 256 (where
    LA_GEMM_AUTO starts using that engine) the gate is the one every engine meets: factor 1.  Below that a product is one
    to four accumulation chains inside the tensor cores, whose FP32 accumulation truncates, and can land a little ABOVE the
    reference order's own worst error (K = 64 over 16M outputs: 3.1e-7 against 3.0e-7): the explicit opt-in is gated at
    twice the reference order's error there, and the header says so."""
    return 1.0 if K >= AUTO_MIN_K else 2.0


def check_f32_gemm(la, a, b, M, K, N, what, factor=1.0):
    P = oracle.port()
    C = (la.Matrix.from_vec(a, M, K) * la.Matrix.from_vec(b, K, N)).to_vec()
    truth, absab = P.f32_mul_mat_mat_truth(a, b, M, K, N)
    check_float_dot_like(C, P.mul_mat_mat(a, b, M, K, N), truth, absab, what, P.mul_mat_mat(a, b, M, K, N, fma=True), factor)
    return C


def test_matmat_f32_tolerance(la):
    for M, K, N in SHAPES_MM:
        check_f32_gemm(la, gen.values("f32", M * K, 23), gen.values("f32", K * N, 24), M, K, N, (M, K, N))


@pytest.mark.parametrize("variant", list(range(0, 11)) + [12, 13, -1])
def test_matmat_every_engine_variant(la, variant):
    """All tile shapes of the TMA engine (and the generic engine, -1) on shapes with ragged
    edge tiles and a K tail.  Variant 0 is the default policy: the generic kernel while the TMA kernel's
    big tiles would leave most SMs idle (fewer than 96 of them), the TMA kernel from there on."""
    P = oracle.port(); lib = la._lib.lib
    lib.la_gemm_set_variant(variant)
    la._lib.call("la_set_gemm_mode", la._lib.GEMM_SIMT)
    try:
        for M, K, N in ((512, 256, 384), (388, 1001, 260), (2560, 64, 1280)):
            check_f32_gemm(la, gen.values("f32", M * K, 25), gen.values("f32", K * N, 26), M, K, N, (variant, M, K, N))
            want = 0 if variant < 0 else 1 if (variant > 0 or M == 2560) else 0
            assert lib.la_gemm_last_engine() == want, (variant, M, K, N)
            a = gen.values("i32", M * K, 27); b = gen.values("i32", K * N, 28)
            C = la.Matrix.from_vec(a, M, K) * la.Matrix.from_vec(b, K, N)
            assert same(C.to_vec(), P.mul_mat_mat(a, b, M, K, N)), (variant, M, K, N)
            assert lib.la_gemm_last_engine() == want
    finally:
        lib.la_gemm_set_variant(0)
        la._lib.call("la_set_gemm_mode", la._lib.GEMM_AUTO)


@pytest.mark.parametrize("tc_variant", [0, 3, 5])
def test_matmat_3xtf32_engine_meets_the_f32_bar(la, tc_variant):
    """LA_GEMM_3XTF32 (tcgen05 tensor cores, hi/lo split) must pass the SAME gate as the SIMT
    engine, and must really be the engine that ran."""
    lib = la._lib.lib
    la._lib.call("la_set_gemm_mode", la._lib.GEMM_3XTF32)
    lib.la_gemm_tc_set_variant(tc_variant)
    try:
        # (4096, 64, 4096): a lone accumulation chain over 16M outputs, where the tensor cores' truncating accumulation shows most
        for M, K, N in ((128, 32, 128), (256, 256, 256), (512, 1000, 384), (388, 1004, 260), (1024, 2048, 512), (4096, 64, 4096), (2048, 96, 2048), (2048, 160, 2048)):
            check_f32_gemm(la, gen.values("f32", M * K, 41), gen.values("f32", K * N, 42), M, K, N, ("3xtf32", tc_variant, M, K, N), tc_factor(K))
            assert lib.la_gemm_last_engine() == 2, "3xTF32 engine did not take %dx%dx%d" % (M, K, N)
        # shapes it does not cover fall through to the SIMT / generic engines
        check_f32_gemm(la, gen.values("f32", 33 * 17, 43), gen.values("f32", 17 * 29, 44), 33, 17, 29, "fallthrough")
        assert lib.la_gemm_last_engine() == 0
    finally:
        la._lib.call("la_set_gemm_mode", la._lib.GEMM_AUTO)
        lib.la_gemm_tc_set_variant(0)


def test_matmat_3xtf32_leaves_non_finite_and_extreme_operands_to_simt(la):
    """inf / NaN / magnitudes outside [2^-100, 2^100] cannot be split faithfully: the pre-pass
    spots them and the SIMT engine computes the product, so infinities propagate as in f32."""
    lib = la._lib.lib
    M = K = N = 512
    la._lib.call("la_set_gemm_mode", la._lib.GEMM_3XTF32)
    try:
        for poison, where in ((np.inf, "a"), (np.nan, "b"), (1e-35, "a"), (3e33, "b")):
            a = gen.values("f32", M * K, 45).copy(); b = gen.values("f32", K * N, 46).copy()
            (a if where == "a" else b)[777] = poison
            C = (la.Matrix.from_vec(a, M, K) * la.Matrix.from_vec(b, K, N)).to_vec()
            assert lib.la_gemm_last_engine() in (0, 1), (poison, "tensor engine should have declined")
            with np.errstate(invalid="ignore", over="ignore"):
                ref = (a.reshape(M, K).astype(np.float64) @ b.reshape(K, N).astype(np.float64)).reshape(-1)
            assert np.array_equal(np.isnan(C), np.isnan(ref)) and np.array_equal(np.isinf(C), np.isinf(ref)), poison
        a = gen.values("f32", M * K, 47); b = gen.values("f32", K * N, 48)      # clean data: tensor engine again
        la.Matrix.from_vec(a, M, K) * la.Matrix.from_vec(b, K, N)
        assert lib.la_gemm_last_engine() == 2
        # from K = 1024 the f16-split form runs first: it declines the same non-finite / huge values (and so does the tf32 form it
        # hands over to), but a lone 1e-35 is just an element far below its row's maximum, which it leaves out of its planes and
        # adds in plain f32 afterwards
        K = 1024
        for poison, where in ((np.inf, "a"), (np.nan, "b"), (3e33, "b"), (-np.inf, "b")):
            a = gen.values("f32", M * K, 45).copy(); b = gen.values("f32", K * N, 46).copy()
            (a if where == "a" else b)[777] = poison
            C = (la.Matrix.from_vec(a, M, K) * la.Matrix.from_vec(b, K, N)).to_vec()
            assert lib.la_gemm_last_engine() in (0, 1), (poison, "tensor engine should have declined")
            with np.errstate(invalid="ignore", over="ignore"):
                ref = (a.reshape(M, K).astype(np.float64) @ b.reshape(K, N).astype(np.float64)).reshape(-1)
            assert np.array_equal(np.isnan(C), np.isnan(ref)) and np.array_equal(np.isinf(C), np.isinf(ref)), poison
        a = gen.values("f32", M * K, 45).copy(); b = gen.values("f32", K * N, 46).copy()
        a[777] = 1e-35
        check_f32_gemm(la, a, b, M, K, N, ("1e-35 among ordinary values", M, K, N))
        assert lib.la_gemm_last_engine() == 2
    finally:
        la._lib.call("la_set_gemm_mode", la._lib.GEMM_AUTO)


def test_matmat_f16_split_form_handles_what_its_planes_leave_out(la):
    """The f16-split form (K >= 1024) scales every row of A / column of B to [2^14, 2^15) and keeps elements 2^-23 and more below
    their row's / column's maximum out of its fp16 planes; their products are added afterwards in plain f32, one owner per output
    and in a fixed order.  Inside the gate, reproducible, identical with a prepared B; a list that overflows (more than 1024
    such elements in an operand) sends the product to the tf32 form."""
    lib = la._lib.lib; P = oracle.port()
    la._lib.call("la_set_gemm_mode", la._lib.GEMM_3XTF32)
    try:
        M, K, N = 384, 1024, 512
        rng = np.random.default_rng(5)
        a = rng.standard_normal(M * K).astype(np.float32); b = rng.standard_normal(K * N).astype(np.float32)
        for idx in rng.integers(0, M * K, 40): a[idx] = np.float32(1e-12) * np.float32(rng.standard_normal())
        for idx in rng.integers(0, K * N, 40): b[idx] = np.float32(1e-13) * np.float32(rng.standard_normal())
        a[5 * K + 7] = 1e-12; b[7 * N + 9] = 1e-12                  # a left-out element of A meeting a left-out element of B
        a[9 * K: 9 * K + 200: 2] = 3e-11                            # a row with 100 of them (more than one round of the fix-up takes)
        C1 = check_f32_gemm(la, a, b, M, K, N, ("left-out elements", M, K, N))
        assert lib.la_gemm_last_engine() == 2
        A = la.Matrix.from_vec(a, M, K); B = la.Matrix.from_vec(b, K, N)
        assert same(C1, (A * B).to_vec()), "not reproducible"
        assert same(C1, A.mul_prepared(B.prepare()).to_vec()), "prepared B differs"
        # rows / columns of very different scales (2^-20 .. 2^20): scaled away exactly
        a2 = (rng.standard_normal(M * K).astype(np.float32).reshape(M, K) * np.exp2(rng.integers(-20, 20, M)).astype(np.float32)[:, None]).reshape(-1)
        b2 = (rng.standard_normal(K * N).astype(np.float32).reshape(K, N) * np.exp2(rng.integers(-20, 20, N)).astype(np.float32)[None, :]).reshape(-1)
        check_f32_gemm(la, a2, b2, M, K, N, ("scaled rows and columns", M, K, N))
        assert lib.la_gemm_last_engine() == 2
        # ragged sizes (multiples of 4 in M and N, of 8 in K)
        Mr, Kr, Nr = 388, 2056, 260
        ar = gen.values("f32", Mr * Kr, 87); br = gen.values("f32", Kr * Nr, 88)
        Cr = check_f32_gemm(la, ar, br, Mr, Kr, Nr, ("ragged", Mr, Kr, Nr))
        assert lib.la_gemm_last_engine() == 2 and lib.la_gemm_tc_last_form() == 1
        assert same(Cr, la.Matrix.from_vec(ar, Mr, Kr).mul_prepared(la.Matrix.from_vec(br, Kr, Nr).prepare()).to_vec())
        for Ms, Ks, Ns in ((132, 1024, 136), (256, 1032, 1028), (1000, 4104, 516)):
            check_f32_gemm(la, gen.values("f32", Ms * Ks, 89), gen.values("f32", Ks * Ns, 90), Ms, Ks, Ns, ("small / ragged", Ms, Ks, Ns))
            assert lib.la_gemm_last_engine() == 2 and lib.la_gemm_tc_last_form() == 1
        # an A outside the f16 form's range (a row maximum of 2^60) with a prepared B: the tf32 form, from B itself, as without a handle
        ah = a.copy(); ah[17 * K: 18 * K] *= np.float32(2.0 ** 60)
        Ah = la.Matrix.from_vec(ah, M, K)
        Ch = check_f32_gemm(la, ah, b, M, K, N, ("row maximum 2^60", M, K, N))
        assert lib.la_gemm_last_engine() == 2 and lib.la_gemm_tc_last_form() == 0
        assert same(Ch, Ah.mul_prepared(B.prepare()).to_vec())
        # 3000 left-out elements in B: the list overflows, the tf32 form takes over
        b3 = b.copy(); b3[rng.choice(K * N, 3000, replace=False)] = 1e-14
        check_f32_gemm(la, a, b3, M, K, N, ("list overflow", M, K, N))
        # all-zero rows and columns
        a4 = a.copy().reshape(M, K); a4[3] = 0; b4 = b.copy().reshape(K, N); b4[:, 11] = 0
        C4 = check_f32_gemm(la, a4.reshape(-1), b4.reshape(-1), M, K, N, ("zero row / column", M, K, N)).reshape(M, N)
        assert not C4[3].any() and not C4[:, 11].any()
    finally:
        la._lib.call("la_set_gemm_mode", la._lib.GEMM_AUTO)


def test_matmat_one_product_outputs_go_to_the_exact_engine(la):
    """A split product carries an error of up to ~2^-21 |a||b| per PRODUCT, which the gate only absorbs in a sum of several
    comparable products.  B = identity / permutation / diagonal, one-hot rows, a row with one outlier: every output is essentially
    one product, which the reference computes exactly -- and so must we (the f16-split form's pre-pass spots rows of A / columns
    of B whose sum of magnitudes is less than 4 times their largest magnitude and hands the product to the SIMT engine)."""
    lib = la._lib.lib; P = oracle.port()
    M = K = N = 1024
    rng = np.random.default_rng(9)
    a = rng.standard_normal(M * K).astype(np.float32)
    eye = np.eye(K, dtype=np.float32)
    cases = {"identity": eye, "signed permutation": eye[rng.permutation(K)] * rng.choice(np.float32([-1.0, 1.0]), K)[:, None],
             "diagonal": np.diag(rng.standard_normal(K).astype(np.float32))}
    for name, bm in cases.items():
        b = np.ascontiguousarray(bm, dtype=np.float32).reshape(-1)
        for prepared in (False, True):
            A = la.Matrix.from_vec(a, M, K); B = la.Matrix.from_vec(b, K, N)
            C = (A.mul_prepared(B.prepare()) if prepared else A * B).to_vec()
            assert lib.la_gemm_last_engine() in (0, 1), (name, prepared, "the tensor engine should have declined")
            assert same(C, P.mul_mat_mat(a, b, M, K, N, fma=True)), (name, prepared)
    # the same on the left: A = a matrix with one outlier per row
    a1 = (rng.standard_normal(M * K).astype(np.float32) * np.float32(1e-4)).reshape(M, K); a1[np.arange(M), rng.integers(0, K, M)] = 5.0
    b1 = rng.standard_normal(K * N).astype(np.float32)
    C = (la.Matrix.from_vec(a1.reshape(-1), M, K) * la.Matrix.from_vec(b1, K, N)).to_vec()
    assert lib.la_gemm_last_engine() in (0, 1)
    assert same(C, P.mul_mat_mat(a1.reshape(-1), b1, M, K, N, fma=True))
    # ... but outliers that do not outweigh the rest of their row (100 x the typical element, K = 1024) keep the tensor cores
    a2 = rng.standard_normal(M * K).astype(np.float32).reshape(M, K); a2[np.arange(M), rng.integers(0, K, M)] = 100.0
    check_f32_gemm(la, a2.reshape(-1), b1, M, K, N, ("outliers 100 x typical", M, K, N))
    assert lib.la_gemm_last_engine() == 2 and lib.la_gemm_tc_last_form() == 1


def test_matmat_tf32_form_still_meets_the_bar_from_k_1024(la):
    """la_gemm_tc_set_f16(0): the tf32 form alone (what K % 8 != 0, out-of-range operands and the host-buffer product use), tf32
    cross terms at every K; and the f16-split form forced below K = 1024 (a tuning setting: AUTO keeps the tf32 form there)."""
    lib = la._lib.lib
    la._lib.call("la_set_gemm_mode", la._lib.GEMM_3XTF32)
    try:
        lib.la_gemm_tc_set_f16(0)
        for M, K, N in ((512, 1024, 384), (388, 2052, 260), (256, 8192, 512)):
            check_f32_gemm(la, gen.values("f32", M * K, 81), gen.values("f32", K * N, 82), M, K, N, ("tf32 form", M, K, N))
            assert lib.la_gemm_last_engine() == 2
        lib.la_gemm_tc_set_f16(-1)
        M, K, N = 388, 1028, 260                                  # K % 8 != 0: the tf32 form by default
        check_f32_gemm(la, gen.values("f32", M * K, 83), gen.values("f32", K * N, 84), M, K, N, ("K % 8 != 0", M, K, N))
        lib.la_gemm_tc_set_f16(1)
        for M, K, N in ((512, 256, 384), (1024, 512, 1024), (388, 264, 260), (2048, 256, 2048)):
            check_f32_gemm(la, gen.values("f32", M * K, 85), gen.values("f32", K * N, 86), M, K, N, ("f16-split form, short K", M, K, N))
            assert lib.la_gemm_last_engine() == 2
    finally:
        lib.la_gemm_tc_set_f16(-1)
        la._lib.call("la_set_gemm_mode", la._lib.GEMM_AUTO)


def test_matmat_auto_mode_picks_the_engine_by_size(la):
    lib = la._lib.lib
    mode = __import__("ctypes").c_int(-1)
    la._lib.call("la_get_gemm_mode", __import__("ctypes").byref(mode))
    assert mode.value == la._lib.GEMM_AUTO
    check_f32_gemm(la, gen.values("f32", 256 * 256, 49), gen.values("f32", 256 * 256, 50), 256, 256, 256, "auto small")
    assert lib.la_gemm_last_engine() == 0          # small: the generic SIMT kernel (the TMA kernel's 256 x 128 tiles would use 2 SMs)
    check_f32_gemm(la, gen.values("f32", 512 * 1024, 51), gen.values("f32", 1024 * 512, 52), 512, 1024, 512, "auto large")
    assert lib.la_gemm_last_engine() == 2          # tensor cores
    a = gen.values("i32", 512 * 1024, 53); b = gen.values("i32", 1024 * 512, 54)
    C = la.Matrix.from_vec(a, 512, 1024) * la.Matrix.from_vec(b, 1024, 512)
    assert same(C.to_vec(), oracle.port().mul_mat_mat(a, b, 512, 1024, 512)) and lib.la_gemm_last_engine() == 3   # integers: INT8-limb tensor engine


@pytest.mark.parametrize("K", [64, 128, 192])
def test_matmat_auto_mode_keeps_short_products_on_the_simt_engine(la, K):
    """The case DESIGN.md calls worst for the tensor engine: 16M outputs of a short product (one to four accumulation chains).
    LA_GEMM_AUTO must give these to the FP32 SIMT engine -- the reference's own k-ascending order -- and they must pass the
    gate at factor 1.  With one fma per k in ascending order the result is, bit for bit, what src/kernels.cl:134-140 gives
    when its `res += a*b` is contracted (the form NVIDIA's OpenCL compiler emits)."""
    lib = la._lib.lib
    M = N = 4096
    a = gen.values("f32", M * K, 55); b = gen.values("f32", K * N, 56)
    C = check_f32_gemm(la, a, b, M, K, N, ("auto short K", M, K, N))
    assert lib.la_gemm_last_engine() == 1, "short products belong to the SIMT engine under LA_GEMM_AUTO"
    assert same(C, oracle.port().mul_mat_mat(a, b, M, K, N, fma=True))


def test_matmat_auto_mode_uses_the_tensor_engine_from_k_256(la):
    lib = la._lib.lib
    for M, K, N in ((4096, 256, 4096), (2048, 512, 2048), (1024, 384, 1024), (512, 256, 1024)):
        check_f32_gemm(la, gen.values("f32", M * K, 57), gen.values("f32", K * N, 58), M, K, N, ("auto", M, K, N))
        assert lib.la_gemm_last_engine() == 2


def test_matmat_f64(la):
    M, K, N = 70, 33, 65
    a = gen.values("f64", M * K, 29); b = gen.values("f64", K * N, 30)
    C = (la.Matrix.from_vec(a, M, K) * la.Matrix.from_vec(b, K, N)).to_vec()
    np.testing.assert_allclose(C, (a.reshape(M, K) @ b.reshape(K, N)).reshape(-1), rtol=1e-12, atol=1e-12)


def test_mul_column_row(la):                   # matrix.rs:236-238
    col = gen.values("i32", 50, 31); row = gen.values("i32", 70, 32)
    m = la.mul_column_row(la.Vector.from_vec(col), la.Vector.from_vec(row))
    assert m.len() == (50, 70)
    assert same(m.to_vec(), oracle.port().mul_mat_mat(col, row, 50, 1, 70))


def test_matrix_elementwise_delegates(la):     # matrix.rs:134-218, 268-277
    P = oracle.port()
    a = gen.values("i32", 6 * 7, 33); b = gen.values("i32", 6 * 7, 34)
    A = la.Matrix.from_vec(a, 6, 7); B = la.Matrix.from_vec(b, 6, 7)
    assert same((A + B).to_vec(), P.ew_vec_vec(ADD, a, b)) and same((A - B).to_vec(), P.ew_vec_vec(SUB, a, b))
    assert same((A * 3).to_vec(), P.ew_vec_scl(MUL, a, 3)) and same((A / 3).to_vec(), P.ew_vec_scl(DIV, a, 3))
    C = A.clone(); C += B; C -= B; C *= 5
    assert same(C.to_vec(), P.ew_vec_scl(MUL, a, 5))


# ------------------------------------------------------------------ transpose
@pytest.mark.parametrize("name", gen.ALL_NAMES)
def test_transpose(la, name):
    P = oracle.port()
    for R, C in ((1, 1), (1, 9), (9, 1), (33, 65), (64, 32), (128, 256), (1000, 516), (520, 1000), (2049, 31), (2064, 1056), (4112, 48), (16, 4096)):
        a = gen.values(name, R * C, 35)
        T = la.Matrix.from_vec(a, R, C, dt(la, name)).transpose()
        assert T.len() == (C, R)
        assert same(T.to_vec(), P.transpose(a, R, C)), (name, R, C)


@pytest.mark.parametrize("variant", [1, 2, 3, 4, 6, 20, 21, 22, 23, 30, 31, 32, 33, 34])
def test_transpose_tile_variants(la, variant):
    """la_transpose_set_variant: other tile shapes of the LDG/STG 4-byte kernel (1, 2), its persistent, software-pipelined tile
    walk (3..8 = blocks per SM), and the warp arrangements / ring depths of the two TMA kernels (20..23: TMA loads and TMA
    stores; 30..34: TMA loads and st.global stores; 4- and 8-byte elements) against the oracle, on shapes with partial edge
    boxes, with sub-tiles wholly outside the matrix (fewer rows than a block's warps cover), with more tiles than the
    persistent grid holds, and with a banded tile walk."""
    P = oracle.port(); lib = la._lib.lib
    lib.la_transpose_set_variant(variant)
    try:
        for band in ((-1, 3) if variant >= 20 else (-1,)):
            lib.la_transpose_set_band(band)
            for R, C in ((64, 32), (1000, 516), (2064, 1056), (4096, 8192), (16, 4096), (4, 20000), (20000, 4), (36, 100)):
                a = gen.values("i32", R * C, 36)
                assert same(la.Matrix.from_vec(a, R, C).transpose().to_vec(), P.transpose(a, R, C)), (variant, band, R, C)
            if variant >= 20:
                for R, C in ((66, 34), (1000, 518), (2050, 4098), (2, 30000)):
                    a = gen.values("u64", R * C, 37)
                    assert same(la.Matrix.from_vec(a, R, C, dt(la, "u64")).transpose().to_vec(), P.transpose(a, R, C)), (variant, band, "u64", R, C)
    finally:
        lib.la_transpose_set_variant(0); lib.la_transpose_set_band(-1)


# ------------------------------------------------------------------ device-side synthetic data
@pytest.mark.parametrize("name", gen.ALL_NAMES)
def test_synth_matches_host_mirror(la, name):
    from linear_algebra_b200 import synth
    n = 100003
    got = la.Vector.synth(0x5EED0001, n, dt(la, name)).to_vec()
    assert same(got, synth.values(dt(la, name), 0x5EED0001, 0, n)), name
    if np.dtype(got.dtype).kind == "f":
        assert got.min() >= -1.0 and got.max() < 1.0 and abs(float(got.mean())) < 0.02


# ------------------------------------------------------------------ one B, many products
@pytest.mark.parametrize("name,shapes", [
    ("f32", ((512, 1024, 512), (1024, 256, 768), (388, 1004, 260), (2560, 512, 2304), (64, 64, 64), (4096, 64, 4096), (130, 70, 258))),
    ("i32", ((512, 1024, 512), (700, 304, 260), (64, 64, 64), (130, 70, 258))),
    ("u16", ((512, 1024, 512),)), ("u8", ((512, 1024, 512),)), ("f64", ((256, 256, 256),)), ("i64", ((130, 70, 258),))])
def test_prepared_b_is_bit_identical_to_the_plain_product(la, name, shapes):
    """la_gemm_prepare_b / la_mul_mat_mat_prepared: the same engine selection and kernels as la_mul_mat_mat with the work on
    B done once -- so the SAME BITS, for every element type, on shapes of every engine (tensor, limb, TMA SIMT, generic,
    padded), for several A against one handle, and in every mode."""
    lib = la._lib.lib
    for M, K, N in shapes:
        b = gen.values(name, K * N, 95)
        B = la.Matrix.from_vec(b, K, N, dt(la, name))
        for mode in (la._lib.GEMM_AUTO, la._lib.GEMM_SIMT, la._lib.GEMM_3XTF32):
            la._lib.call("la_set_gemm_mode", mode)
            try:
                P = B.prepare()
                for seed in (96, 97):
                    A = la.Matrix.from_vec(gen.values(name, M * K, seed), M, K, dt(la, name))
                    plain = (A * B).to_vec(); engine = lib.la_gemm_last_engine()
                    prepared = A.mul_prepared(P).to_vec()
                    assert lib.la_gemm_last_engine() == engine, (name, M, K, N, mode)
                    assert same(plain, prepared), (name, M, K, N, mode, engine)
                P.free()
            finally:
                la._lib.call("la_set_gemm_mode", la._lib.GEMM_AUTO)


def test_prepared_b_with_non_finite_values_takes_the_simt_engine(la):
    lib = la._lib.lib
    M = K = N = 512
    a = gen.values("f32", M * K, 98); b = gen.values("f32", K * N, 99).copy()
    b[1234] = np.inf
    A = la.Matrix.from_vec(a, M, K); B = la.Matrix.from_vec(b, K, N)
    P = B.prepare()
    got = A.mul_prepared(P).to_vec()
    assert lib.la_gemm_last_engine() in (0, 1)
    assert same(got, (A * B).to_vec())
    with np.errstate(invalid="ignore", over="ignore"):
        ref = (a.reshape(M, K).astype(np.float64) @ b.reshape(K, N).astype(np.float64)).reshape(-1)
    assert np.array_equal(np.isinf(got), np.isinf(ref)) and np.array_equal(np.isnan(got), np.isnan(ref))
    # an A the split cannot represent, against a clean prepared B: same fallback
    b2 = gen.values("f32", K * N, 100); a2 = a.copy(); a2[77] = np.nan
    B2 = la.Matrix.from_vec(b2, K, N); P2 = B2.prepare(); A2 = la.Matrix.from_vec(a2, M, K)
    got2 = A2.mul_prepared(P2).to_vec()
    assert lib.la_gemm_last_engine() in (0, 1)
    assert same(got2, (A2 * B2).to_vec())


def test_mul_mat_mat_info_reports_the_engine(la):
    info = (ctypes.c_int * 2)(-1, -1)
    M = K = N = 512
    A = la.Matrix.from_vec(gen.values("f32", M * K, 101), M, K); B = la.Matrix.from_vec(gen.values("f32", K * N, 102), K, N)
    C = la.Matrix.uninitialized(M, N, "f32")
    la._lib.call("la_mul_mat_mat_info", 8, C.data.ptr, A.data.ptr, B.data.ptr, M, K, N, None, ctypes.cast(info, ctypes.c_void_p))
    la.synchronize()
    assert (info[0], info[1]) == (2, 0)
    assert same(C.to_vec(), (A * B).to_vec())
    A3 = la.Matrix.from_vec(gen.values("f32", 33 * 17, 103), 33, 17); B3 = la.Matrix.from_vec(gen.values("f32", 17 * 29, 104), 17, 29)
    C3 = la.Matrix.uninitialized(33, 29, "f32")
    la._lib.call("la_mul_mat_mat_info", 8, C3.data.ptr, A3.data.ptr, B3.data.ptr, 33, 17, 29, None, ctypes.cast(info, ctypes.c_void_p))
    assert info[0] == 0


# ------------------------------------------------------------------ streamed host-buffer GEMM
@pytest.mark.parametrize("name,shape", [("f32", (1000, 512, 768)), ("f32", (2304, 1024, 640)), ("i32", (700, 300, 260)), ("u8", (300, 200, 100)), ("f32", (5, 7, 3))])
def test_mul_host_equals_the_three_call_sequence(la, name, shape):
    M, K, N = shape
    a = gen.values(name, M * K, 60); b = gen.values(name, K * N, 61)
    expect = (la.Matrix.from_vec(a, M, K, dt(la, name)) * la.Matrix.from_vec(b, K, N, dt(la, name))).to_vec()
    ha = la.pinned(M * K, dt(la, name)); hb = la.pinned(K * N, dt(la, name)); hc = la.pinned(M * N, dt(la, name))
    ha[:] = a; hb[:] = b
    for _ in range(2):                         # twice: stream / event state must be reusable
        hc[:] = 0
        got = la.mul_host(ha, hb, M, K, N, out=hc)
        if np.dtype(NP[name]).kind == "f":
            P = oracle.port()
            truth, absab = P.f32_mul_mat_mat_truth(a, b, M, K, N)
            check_float_dot_like(got, P.mul_mat_mat(a, b, M, K, N), truth, absab, ("mul_host", shape), P.mul_mat_mat(a, b, M, K, N, fma=True))
            assert np.linalg.norm(got - expect) <= NORMWISE_TOL * np.linalg.norm(expect)
        else:
            assert same(got, expect), (name, shape)
    assert same(la.mul_host(a, b, M, K, N), got) or np.dtype(NP[name]).kind == "f"      # pageable host memory works too


def test_mul_host_redoes_the_product_in_f32_for_non_finite_operands(la):
    M = K = N = 1024
    a = gen.values("f32", M * K, 62).copy(); b = gen.values("f32", K * N, 63).copy()
    a[5 * K + 9] = np.inf
    got = la.mul_host(a, b, M, K, N)
    with np.errstate(invalid="ignore", over="ignore"):
        ref = (a.reshape(M, K).astype(np.float64) @ b.reshape(K, N).astype(np.float64)).reshape(-1)
    assert np.array_equal(np.isinf(got), np.isinf(ref)) and not np.isnan(got).any()
    finite = np.isfinite(ref)
    assert np.linalg.norm(got[finite] - ref[finite]) <= NORMWISE_TOL * np.linalg.norm(ref[finite])


# ------------------------------------------------------------------ integer GEMM on the INT8 tensor cores
@pytest.mark.parametrize("name", ["u8", "i8", "u16", "i16", "u32", "i32"])
def test_matmat_limb_engine_is_bit_exact(la, name):
    """The limb engine (byte-limb products on tcgen05 kind::i8, s32 accumulators that may wrap) must give
    the reference's bits: integers are exact mod 2^n whatever the order or the machinery."""
    P = oracle.port(); lib = la._lib.lib
    for M, K, N in ((512, 1024, 512), (388, 2048, 272), (1024, 4096, 128)):
        a = gen.values(name, M * K, 70); b = gen.values(name, K * N, 71)
        C = la.Matrix.from_vec(a, M, K, dt(la, name)) * la.Matrix.from_vec(b, K, N, dt(la, name))
        assert lib.la_gemm_last_engine() == 3, "limb engine did not take %s %dx%dx%d" % (name, M, K, N)
        assert same(C.to_vec(), P.mul_mat_mat(a, b, M, K, N)), (name, M, K, N)
    # all-ones bytes: every limb product is 255*255 and the s32 accumulators overflow for 32-bit types
    M, K, N = 256, 32768, 128
    a = np.full(M * K, -1, NP[name]) if np.dtype(NP[name]).kind == "i" else np.full(M * K, np.iinfo(NP[name]).max, NP[name])
    C = la.Matrix.from_vec(a, M, K, dt(la, name)) * la.Matrix.from_vec(a[:K * N], K, N, dt(la, name))
    assert lib.la_gemm_last_engine() == 3
    assert same(C.to_vec(), P.mul_mat_mat(a, a[:K * N], M, K, N)), (name, "wrap")
    # SIMT mode keeps the IMAD engines
    la._lib.call("la_set_gemm_mode", la._lib.GEMM_SIMT)
    try:
        a = gen.values(name, 512 * 1024, 72); b = gen.values(name, 1024 * 512, 73)
        C = la.Matrix.from_vec(a, 512, 1024, dt(la, name)) * la.Matrix.from_vec(b, 1024, 512, dt(la, name))
        assert lib.la_gemm_last_engine() in (0, 1) and same(C.to_vec(), P.mul_mat_mat(a, b, 512, 1024, 512))
    finally:
        la._lib.call("la_set_gemm_mode", la._lib.GEMM_AUTO)


# ------------------------------------------------------------------ tile order of the tensor engines
@pytest.mark.parametrize("group", [1, 3, 8])
def test_matmat_banded_tile_order_changes_nothing(la, group):
    """la_gemm_set_raster: bands of `group` tile rows instead of m-major order (the default above M = 8192).  Which CTA
    computes which tile cannot change a bit of the result: f32 tensor engine and i32 limb engine against their m-major
    runs, on shapes whose last band is ragged (8 and 5 tile rows of 256) and with partial edge tiles."""
    lib = la._lib.lib
    P = oracle.port()
    try:
        for M, K, N in ((1800, 160, 1100), (1280, 256, 640)):
            a = gen.values("f32", M * K, 61); b = gen.values("f32", K * N, 62)
            ai = gen.values("i32", M * K, 63); bi = gen.values("i32", K * N, 64)
            out = {}
            for g in (0, group):
                lib.la_gemm_set_raster(g)
                la._lib.call("la_set_gemm_mode", la._lib.GEMM_3XTF32)
                out["f", g] = (la.Matrix.from_vec(a, M, K) * la.Matrix.from_vec(b, K, N)).to_vec()
                assert lib.la_gemm_last_engine() == 2
                la._lib.call("la_set_gemm_mode", la._lib.GEMM_AUTO)
                out["i", g] = (la.Matrix.from_vec(ai, M, K) * la.Matrix.from_vec(bi, K, N)).to_vec()
                assert lib.la_gemm_last_engine() == 3
            assert same(out["f", 0], out["f", group]) and same(out["i", 0], out["i", group]), (group, M, K, N)
            assert same(out["i", group], P.mul_mat_mat(ai, bi, M, K, N))
    finally:
        lib.la_gemm_set_raster(-1)
        la._lib.call("la_set_gemm_mode", la._lib.GEMM_AUTO)


@pytest.mark.parametrize("cross", [0, 1])
@pytest.mark.parametrize("chunk", [0, 2, 8])
def test_matmat_tensor_engine_cross_terms_and_chain_lengths_meet_the_bar(la, cross, chunk):
    """The two cross terms hi*lo + lo*hi as tf32 MMAs (cross 0) or as bf16 MMAs on bf16(hi), bf16(x - hi) planes (cross 1,
    the default from K = 256), with accumulation chains of 64 / 256 k or the default by K: every combination must pass the
    f32 gate on short and long products."""
    lib = la._lib.lib
    la._lib.call("la_set_gemm_mode", la._lib.GEMM_3XTF32)
    lib.la_gemm_tc_set_cross(cross); lib.la_gemm_tc_set_f16(0)          # the tf32 form (from K = 1024 the f16-split form would run instead)
    try:
        for M, K, N in ((256, 64, 256), (512, 256, 384), (388, 1000, 260), (1024, 2048, 512), (512, 4096, 768)):
            # a forced chain length only where it leaves at least 8 chains: the truncating accumulation inside one chain is
            # what the FP32 registers are there to average out (the defaults: 32-chains up to K = 128, 64 below 2048,
            # 128 below 8192, 256 from there)
            lib.la_gemm_tc_set_chunk(chunk if K >= 8 * 32 * chunk else 0)
            # bf16 cross terms below K = 1024 are a tuning setting the defaults never choose (0.8 - 0.9 of the reference order's
            # error at K = 256: inside the gate on average, not with margin): gated like the other opt-in, at twice that error
            factor = tc_factor(K) if (cross == 0 or K >= 1024) else 2.0
            check_f32_gemm(la, gen.values("f32", M * K, 71), gen.values("f32", K * N, 72), M, K, N, ("cross", cross, "chunk", chunk, M, K, N), factor)
            assert lib.la_gemm_last_engine() == 2
    finally:
        lib.la_gemm_tc_set_cross(-1); lib.la_gemm_tc_set_chunk(0); lib.la_gemm_tc_set_f16(-1)
        la._lib.call("la_set_gemm_mode", la._lib.GEMM_AUTO)


def test_matmat_tensor_engine_persistent_grid_changes_nothing(la):
    """la_gemm_tc_set_persistent: 74 CTA pairs walk all work items instead of one pair per item (a tuning variant: +0.4 % at
    8192^3, -6 % at 16384^3, so off by default).  Same MMAs per output, so the bits must not change; with and without
    split last-wave tiles."""
    lib = la._lib.lib
    la._lib.call("la_set_gemm_mode", la._lib.GEMM_3XTF32)
    try:
        for M, K, N in ((2560, 512, 2304), (4096, 1024, 5000), (4800, 96, 2300)):
            a = gen.values("f32", M * K, 73); b = gen.values("f32", K * N, 74)
            A = la.Matrix.from_vec(a, M, K); B = la.Matrix.from_vec(b, K, N)
            lib.la_gemm_tc_set_persistent(0); C0 = (A * B).to_vec()
            lib.la_gemm_tc_set_persistent(1); C1 = (A * B).to_vec()
            assert lib.la_gemm_last_engine() == 2 and same(C0, C1), (M, K, N)
    finally:
        lib.la_gemm_tc_set_persistent(0)
        la._lib.call("la_set_gemm_mode", la._lib.GEMM_AUTO)


def test_matmat_tensor_engine_on_structured_data(la):
    """Data that defeats the 'rounding errors are random' argument: constant matrices (every bf16 rounding of a cross term errs
    the same way, so the errors add up linearly in K), same-sign data, and integer-valued f32 whose products and sums the
    reference computes exactly.  Long K, so the bf16 cross terms are in use."""
    lib = la._lib.lib; P = oracle.port()
    la._lib.call("la_set_gemm_mode", la._lib.GEMM_3XTF32)
    try:
        M, N = 256, 384
        rng = np.random.default_rng(77)
        cases = {}
        for K in (2048, 8192):                      # 128-long and 256-long accumulation chains
            cases["constant 1/3 x 0.7, K=%d" % K] = (K, np.full(M * K, 1.0 / 3.0, np.float32), np.full(K * N, 0.7, np.float32))
            cases["same sign, K=%d" % K] = (K, rng.random(M * K, dtype=np.float32), rng.random(K * N, dtype=np.float32))
            cases["rows of one value, K=%d" % K] = (K, np.repeat(rng.random(M, dtype=np.float32), K), rng.random(K * N, dtype=np.float32))
        for what, (K, a, b) in cases.items():
            # against the f64 truth only: on such data the reference order itself drifts (constant data: 1.2e-5 of the
            # result), so "close to the oracle" is no criterion here, "no further from the truth than the oracle" is
            C = (la.Matrix.from_vec(a, M, K) * la.Matrix.from_vec(b, K, N)).to_vec().astype(np.float64)
            assert lib.la_gemm_last_engine() == 2
            truth, absab = P.f32_mul_mat_mat_truth(a, b, M, K, N)
            e_ours = float((np.abs(C - truth) / absab).max())
            e_orc = float((np.abs(P.mul_mat_mat(a, b, M, K, N).astype(np.float64) - truth) / absab).max())
            print("%s: e_ours %.3e  e_oracle %.3e" % (what, e_ours, e_orc))
            # here even without the factor 2 of the random-data gate (measured: 0.05x - 0.8x of the oracle's error; the
            # worst, same-sign rows, is the truncation bias of one accumulation chain, which is why chains grow only with K)
            assert e_ours <= max(e_orc, F32_EPS_FLOOR), (what, e_ours, e_orc)
        K = 2048
        # integers small enough that every partial sum stays below 2^24: the reference is exact, so must we be
        a = rng.integers(-60, 61, M * K).astype(np.float32); b = rng.integers(-60, 61, K * N).astype(np.float32)
        C = (la.Matrix.from_vec(a, M, K) * la.Matrix.from_vec(b, K, N)).to_vec()
        exact = (a.reshape(M, K).astype(np.int64) @ b.reshape(K, N).astype(np.int64)).reshape(-1)
        assert np.abs(exact).max() < 2 ** 24 and np.array_equal(C.astype(np.int64), exact)
        # 13-bit integers times small ones: hi has more bits than bf16 keeps, but its partner's lo is zero
        a = rng.integers(4096, 8192, M * K).astype(np.float32); b = rng.integers(0, 2, K * N).astype(np.float32)
        C = (la.Matrix.from_vec(a, M, K) * la.Matrix.from_vec(b, K, N)).to_vec()
        exact = (a.reshape(M, K).astype(np.int64) @ b.reshape(K, N).astype(np.int64)).reshape(-1)
        assert np.abs(exact).max() < 2 ** 24 and np.array_equal(C.astype(np.int64), exact)
    finally:
        la._lib.call("la_set_gemm_mode", la._lib.GEMM_AUTO)


def test_matmat_tensor_engine_a_read_in_place_stays_inside_the_gate(la):
    """la_gemm_tc_set_a_direct (an experiment, off by default): A is read in place as its own hi plane -- the tensor
    cores truncate the f32 word to tf32 -- and its lo part is taken against the truncated value.  Less accurate than the rounded
    split (measured 1.2 - 1.4 x its error) but still inside the unrelaxed gate wherever the engine uses bf16 cross terms
    (K >= 1024); below that the flag has no effect.  The prepared-B path must agree with the direct call bit for bit."""
    lib = la._lib.lib
    la._lib.call("la_set_gemm_mode", la._lib.GEMM_3XTF32)
    lib.la_gemm_tc_set_a_direct(1); lib.la_gemm_tc_set_f16(0); lib.la_gemm_tc_set_cross(1)      # the tf32 form with bf16 cross terms: both tuning settings now
    try:
        for M, K, N in ((512, 1024, 384), (388, 2052, 260), (256, 4096, 512)):
            a = gen.values("f32", M * K, 73); b = gen.values("f32", K * N, 74)
            check_f32_gemm(la, a, b, M, K, N, ("A in place", M, K, N), 1.0)
            assert lib.la_gemm_last_engine() == 2
            A = la.Matrix.from_vec(a, M, K); B = la.Matrix.from_vec(b, K, N)
            assert same((A * B).to_vec(), A.mul_prepared(B.prepare()).to_vec()), (M, K, N)
        M, K, N = 512, 256, 384                       # tf32 cross terms: the flag changes nothing
        lib.la_gemm_tc_set_cross(-1)
        a = gen.values("f32", M * K, 75); b = gen.values("f32", K * N, 76)
        A = la.Matrix.from_vec(a, M, K); B = la.Matrix.from_vec(b, K, N)
        with_flag = (A * B).to_vec()
        lib.la_gemm_tc_set_a_direct(0)
        assert same(with_flag, (A * B).to_vec())
    finally:
        lib.la_gemm_tc_set_a_direct(0); lib.la_gemm_tc_set_f16(-1); lib.la_gemm_tc_set_cross(-1)
        la._lib.call("la_set_gemm_mode", la._lib.GEMM_AUTO)


@pytest.mark.parametrize("bn", [224, 192])
def test_matmat_3xtf32_narrower_tiles_change_nothing(la, bn):
    """la_gemm_tc_set_tile_n: pair tiles 256 x 224 / 256 x 192 (picked when they pack the waves better, e.g. the 8-GPU shard
    1024 x 8192 x 8192).  Every output is computed by the same chain of MMAs whatever the tile width, so with the K-split of
    the last wave off the results are bit-identical to the 256-wide tiles'; with it on they meet the same bar."""
    lib = la._lib.lib
    la._lib.call("la_set_gemm_mode", la._lib.GEMM_3XTF32)
    try:
        for M, K, N in ((512, 256, 1000), (1300, 1004, 2300), (256, 96, 224 * 3), (1024, 64, 8192)):
            a = gen.values("f32", M * K, 67); b = gen.values("f32", K * N, 68)
            A = la.Matrix.from_vec(a, M, K); B = la.Matrix.from_vec(b, K, N)
            lib.la_gemm_set_stream_tail(0)
            lib.la_gemm_tc_set_tile_n(256); C256 = (A * B).to_vec()
            lib.la_gemm_tc_set_tile_n(bn); Cbn = (A * B).to_vec()
            assert lib.la_gemm_last_engine() == 2
            assert same(C256, Cbn), (bn, M, K, N)
            lib.la_gemm_set_stream_tail(-1)
            check_f32_gemm(la, a, b, M, K, N, ("tile width", bn, M, K, N), tc_factor(K))
    finally:
        lib.la_gemm_tc_set_tile_n(0); lib.la_gemm_set_stream_tail(-1)
        la._lib.call("la_set_gemm_mode", la._lib.GEMM_AUTO)


@pytest.mark.parametrize("raster", [0, 3])
def test_matmat_3xtf32_stream_tail_meets_the_bar_and_is_reproducible(la, raster):
    """The ragged last wave of the tensor engine cut into equal spans (la_gemm_set_stream_tail): tiles then get up to two
    partial sums added with atomics onto zeroed memory.  Same f32 bar as every engine; bit-identical from run to run
    (two addends commute); shapes with one and with several waves, a K tail, ragged edges, both tile orders."""
    lib = la._lib.lib
    la._lib.call("la_set_gemm_mode", la._lib.GEMM_3XTF32)
    lib.la_gemm_set_stream_tail(1); lib.la_gemm_set_raster(raster)
    try:
        for M, K, N in ((2560, 512, 2304), (2500, 1000, 2300), (4096, 136, 5000), (256 * 75, 64, 256)):
            a = gen.values("f32", M * K, 65); b = gen.values("f32", K * N, 66)
            A = la.Matrix.from_vec(a, M, K); B = la.Matrix.from_vec(b, K, N)
            C1 = (A * B).to_vec()
            assert lib.la_gemm_last_engine() == 2
            if M * K * N <= 3e9:
                P = oracle.port()
                truth, absab = P.f32_mul_mat_mat_truth(a, b, M, K, N)
                check_float_dot_like(C1, P.mul_mat_mat(a, b, M, K, N), truth, absab, ("stream tail", raster, M, K, N), P.mul_mat_mat(a, b, M, K, N, fma=True), tc_factor(K))
            else:
                rows = [0, 255, 256, M // 2, M - 1]
                ref = a.reshape(M, K)[rows].astype(np.float64) @ b.reshape(K, N).astype(np.float64)
                assert np.linalg.norm(C1.reshape(M, N)[rows] - ref) / np.linalg.norm(ref) <= 1e-5
            for _ in range(3):
                assert same((A * B).to_vec(), C1), ("not reproducible", raster, M, K, N)
            lib.la_gemm_set_stream_tail(0)
            C0 = (A * B).to_vec()
            lib.la_gemm_set_stream_tail(1)
            assert np.linalg.norm(C0.astype(np.float64) - C1) / np.linalg.norm(C0.astype(np.float64)) <= 1e-6
    finally:
        lib.la_gemm_set_stream_tail(-1); lib.la_gemm_set_raster(-1)
        la._lib.call("la_set_gemm_mode", la._lib.GEMM_AUTO)


# ------------------------------------------------------------------ 8-byte element types on the TMA-fed SIMT engine
@pytest.mark.parametrize("variant", [0, 1, 2, 3, 4, 5, 13])
def test_matmat_64bit_types_on_the_tma_engine(la, variant):
    """f64 (DFMA, k ascending, one fused multiply-add per k) and u64/i64 (wrapping) through gemm_tma_kernel's 8-byte
    instantiation: interior and ragged edge tiles, a K tail, every tile shape.  Integers bit-exact; f64 against the oracle's
    own k-ascending f64 sum within 1e-13 of sum|a||b| (fused vs unfused rounding of each step)."""
    P = oracle.port(); lib = la._lib.lib
    lib.la_gemm_set_variant(variant)
    try:
        for M, K, N in ((512, 256, 384), (386, 1001, 262), (64, 16, 64), (1024, 32, 1024)):
            want = 1 if (variant > 0 or M == 1024) else 0          # variant 0: by size (48 tiles of 128 x 128 and up)
            for name in ("u64", "i64"):
                a = gen.values(name, M * K, 91); b = gen.values(name, K * N, 92)
                C = la.Matrix.from_vec(a, M, K, dt(la, name)) * la.Matrix.from_vec(b, K, N, dt(la, name))
                assert lib.la_gemm_last_engine() == want, (name, M, K, N)
                assert same(C.to_vec(), P.mul_mat_mat(a, b, M, K, N)), (variant, name, M, K, N)
            a = gen.values("f64", M * K, 93); b = gen.values("f64", K * N, 94)
            C = (la.Matrix.from_vec(a, M, K, "f64") * la.Matrix.from_vec(b, K, N, "f64")).to_vec()
            assert lib.la_gemm_last_engine() == want
            ref = P.mul_mat_mat(a, b, M, K, N)
            scale = (np.abs(a.reshape(M, K)) @ np.abs(b.reshape(K, N))).reshape(-1)
            assert np.max(np.abs(C - ref) / scale) < 1e-13, (variant, M, K, N)
        # odd M or N: the 16-byte chunks do not divide the rows -> generic engine (small) or padding (large)
        a = gen.values("i64", 65 * 32, 95); b = gen.values("i64", 32 * 67, 96)
        C = la.Matrix.from_vec(a, 65, 32, "i64") * la.Matrix.from_vec(b, 32, 67, "i64")
        assert lib.la_gemm_last_engine() == 0 and same(C.to_vec(), P.mul_mat_mat(a, b, 65, 32, 67))
        M, K, N = 515, 1031, 523
        a = gen.values("u64", M * K, 97); b = gen.values("u64", K * N, 98)
        C = la.Matrix.from_vec(a, M, K, "u64") * la.Matrix.from_vec(b, K, N, "u64")
        assert lib.la_gemm_last_padded() == 1 and lib.la_gemm_last_engine() == (1 if variant > 0 else 0)
        assert same(C.to_vec(), P.mul_mat_mat(a, b, M, K, N))
    finally:
        lib.la_gemm_set_variant(0)


# ------------------------------------------------------------------ ragged large products: padded onto the fast engines
def test_matmat_large_ragged_shapes_take_the_fast_engines_through_padding(la):
    P = oracle.port(); lib = la._lib.lib
    M, K, N = 517, 1031, 523                       # nothing is a multiple of 4
    a = gen.values("f32", M * K, 80); b = gen.values("f32", K * N, 81)
    check_f32_gemm(la, a, b, M, K, N, "padded f32")
    assert lib.la_gemm_last_padded() == 1 and lib.la_gemm_last_engine() == 2
    for name in ("i32", "u16", "u8"):
        ai = gen.values(name, M * K, 82); bi = gen.values(name, K * N, 83)
        C = la.Matrix.from_vec(ai, M, K, dt(la, name)) * la.Matrix.from_vec(bi, K, N, dt(la, name))
        assert lib.la_gemm_last_padded() == 1 and lib.la_gemm_last_engine() == 3, name
        assert same(C.to_vec(), P.mul_mat_mat(ai, bi, M, K, N)), name
    # unaligned base pointers
    a = gen.values("i32", 512 * 1024 + 1, 84); b = gen.values("i32", 1024 * 512 + 1, 85)
    A = la.Vector.from_vec(a); B = la.Vector.from_vec(b); Cv = la.Vector.uninitialized(512 * 512 + 1, "i32")
    la._lib.call("la_mul_mat_mat", 5, Cv.ptr + 4, A.ptr + 4, B.ptr + 4, 512, 1024, 512, None)
    la.synchronize()
    assert lib.la_gemm_last_padded() == 1
    assert same(Cv.to_vec()[1:], P.mul_mat_mat(a[1:], b[1:], 512, 1024, 512))
    # small ragged products stay on the generic kernel
    check_f32_gemm(la, gen.values("f32", 33 * 17, 86), gen.values("f32", 17 * 29, 87), 33, 17, 29, "small ragged")
    assert lib.la_gemm_last_padded() == 0 and lib.la_gemm_last_engine() == 0


# ------------------------------------------------------------------ fused normalize
@pytest.mark.parametrize("name", ["f32", "f64"])
def test_normalize_is_the_three_step_sequence(la, name):
    """la_normalize_vec (dot, sqrt, 1/length, scale on the device) gives the bits of the reference's sequence
    length() -> 1/length -> `* scalar` (vector.rs:303-323)."""
    for n in (1, 5, 1000, (1 << 20) + 3):
        a = gen.values(name, n, 90)
        A = la.Vector.from_vec(a, name)
        inv = NP[name](1.0) / A.length()
        expect = (A * inv).to_vec()
        assert same(A.normalized().to_vec(), expect), (name, n)
        B = A.clone().normalize()
        assert same(B.to_vec(), expect)
        assert abs(float(B.length()) - 1.0) < 1e-5
    with pytest.raises(AssertionError):
        la.Vector.from_vec([1, 2, 3], "i32").normalized()


def test_8bit_division_is_exact_for_every_operand_pair(la):
    """8-bit `/` runs through the FP32 pipe (common.cuh la_div_u8 / la_div_i8): all 256 x 255 (u8) and
    256 x 254 (i8, divisors 0 and -1 excluded as in OpenCL) operand pairs against the oracle."""
    P = oracle.port()
    a = np.repeat(np.arange(256, dtype=np.uint8), 255); b = np.tile(np.arange(1, 256, dtype=np.uint8), 256)
    assert same((la.Vector.from_vec(a, "u8") / la.Vector.from_vec(b, "u8")).to_vec(), P.ew_vec_vec(DIV, a, b))
    divs = np.array([d for d in range(-128, 128) if d not in (0, -1)], dtype=np.int8)
    a = np.repeat(np.arange(-128, 128, dtype=np.int8), divs.size); b = np.tile(divs, 256)
    assert same((la.Vector.from_vec(a, la.I8_CHAR) / la.Vector.from_vec(b, la.I8_CHAR)).to_vec(), P.ew_vec_vec(DIV, a, b))
    s = la.Vector.from_vec(np.arange(256, dtype=np.uint8), "u8")
    for d in (1, 2, 3, 7, 255):
        assert same((s / np.uint8(d)).to_vec(), P.ew_vec_scl(DIV, np.arange(256, dtype=np.uint8), np.uint8(d)))
