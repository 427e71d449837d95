"""Our CUDA results against the reference's UNMODIFIED OpenCL kernels running on the SAME B200 through a real
OpenCL driver (oracle/cl_ref.py; program assembled exactly as /root/reference/src/lib.rs:337-354 does, text
embedded in oracle/_ref/libref_clsrc.so).  Skipped where no OpenCL platform or no oracle/_ref exists.

Bars: integers bit-exact for every kernel; f32 add/sub/mul/scale bit-exact; f32 divide within OpenCL's own
2.5 ulp allowance (the driver's default divide is not correctly rounded; ours is); f32 sum/dot/GEMV/GEMM
within the stated f32 bar (normwise 1e-5, scalars 2^-20 of sum|a||b|)."""
import numpy as np
import pytest

import gen
import oracle
from oracle import ADD, SUB, MUL, DIV, NP

pytestmark = pytest.mark.gpu

NAMES = [n for n in gen.ALL_NAMES if n != "f64"]          # "double" is not in the reference's default TYPES (lib.rs:84-92)


@pytest.fixture(scope="module")
def cl():
    from oracle import cl_ref
    try:
        return cl_ref.cl_ref()
    except cl_ref.ClUnavailable as e:
        pytest.skip(str(e))


@pytest.fixture(scope="module")
def la():
    import linear_algebra_b200 as la
    la.init(0)
    return la


def dt(la, name):
    return la.I8_CHAR if name == "i8" else la.parameter(name)


def same(x, y):
    return x.dtype == y.dtype and x.shape == y.shape and x.tobytes() == y.tobytes()


def test_reference_program_builds_on_this_device(cl):
    assert "B200" in cl.device_name or cl.device_name      # whatever device the reference's own pick would choose
    assert cl.preferred_multiple(cl._kernel("float", "sum_vec")) in (8, 16, 32, 64)


@pytest.mark.parametrize("name", NAMES)
def test_elementwise_and_eq_against_opencl_reference(la, cl, name):
    n = 100003
    a = gen.values(name, n, 201); b = gen.values(name, n, 202); d = gen.divisors(name, n, 203)
    A = la.Vector.from_vec(a, dt(la, name)); B = la.Vector.from_vec(b, dt(la, name)); D = la.Vector.from_vec(d, dt(la, name))
    assert same((A + B).to_vec(), cl.ew_vec_vec(ADD, a, b))
    assert same((A - B).to_vec(), cl.ew_vec_vec(SUB, a, b))
    assert same((A * B).to_vec(), cl.ew_vec_vec(MUL, a, b))
    assert same((A * b[0]).to_vec(), cl.ew_vec_scl(MUL, a, b[0]))
    C = A.clone(); C -= B
    assert same(C.to_vec(), cl.ew_assign_vec_vec(SUB, a, b))
    ours, ref = (A / D).to_vec(), cl.ew_vec_vec(DIV, a, d)
    if name == "f32":
        ulp = np.spacing(np.abs(ref)).astype(np.float64)
        assert np.all(np.abs(ours.astype(np.float64) - ref.astype(np.float64)) <= 2.5 * ulp)
        assert same(ours, oracle.port().ew_vec_vec(DIV, a, d))        # ours is the correctly rounded quotient
    else:
        assert same(ours, ref)
    c = a.copy(); c[n // 2] = c[n // 2] + NP[name](1)
    assert (A == A) == cl.eq_vec(a, a) == True and (A == la.Vector.from_vec(c, dt(la, name))) == cl.eq_vec(a, c) == False  # noqa: E712


@pytest.mark.parametrize("name", NAMES)
def test_reductions_and_products_against_opencl_reference(la, cl, name):
    n = 70001
    a = gen.values(name, n, 204); b = gen.values(name, n, 205)
    A = la.Vector.from_vec(a, dt(la, name)); B = la.Vector.from_vec(b, dt(la, name))
    rows, cols = 129, 260
    m = gen.values(name, rows * cols, 206); vr = gen.values(name, rows, 207); vc = gen.values(name, cols, 208)
    Mx = la.Matrix.from_vec(m, rows, cols, dt(la, name))
    M_, K_, N_ = 130, 70, 92
    ma = gen.values(name, M_ * K_, 209); mb = gen.values(name, K_ * N_, 210)
    vm = (la.Vector.from_vec(vr, dt(la, name)) * Mx).to_vec()
    mv = la.mul_transpose_mat(la.Vector.from_vec(vc, dt(la, name)), Mx).to_vec()
    mm = (la.Matrix.from_vec(ma, M_, K_, dt(la, name)) * la.Matrix.from_vec(mb, K_, N_, dt(la, name))).to_vec()
    if name != "f32":
        assert A.sum() == cl.sum_vec(a) and la.dot(A, B) == cl.dot_vec_vec(a, b)
        assert same(vm, cl.mul_vec_mat(vr, m, rows, cols)) and same(mv, cl.mul_vec_transpose_mat(vc, m, rows, cols))
        assert same(mm, cl.mul_mat_mat(ma, mb, M_, K_, N_))
    else:
        scale = float(np.abs(a.astype(np.float64) * b.astype(np.float64)).sum())
        assert abs(float(la.dot(A, B)) - float(cl.dot_vec_vec(a, b))) <= 2.0 ** -20 * scale
        assert abs(float(A.sum()) - float(cl.sum_vec(a))) <= 2.0 ** -20 * float(np.abs(a.astype(np.float64)).sum())
        for got, ref in ((vm, cl.mul_vec_mat(vr, m, rows, cols)), (mv, cl.mul_vec_transpose_mat(vc, m, rows, cols)), (mm, cl.mul_mat_mat(ma, mb, M_, K_, N_))):
            assert np.linalg.norm(got.astype(np.float64) - ref) <= 1e-5 * np.linalg.norm(ref.astype(np.float64))


def test_port_matches_the_opencl_reference_order_for_f32(cl):
    """The plain-C restatement follows the reference's summation ORDER: with the device's own work-group size
    its f32 sum/dot must land inside the contracted / uncontracted bracket the OpenCL compiler may choose from."""
    P = oracle.port()
    a = gen.values("f32", 50000, 211); b = gen.values("f32", 50000, 212)
    W = cl.preferred_multiple(cl._kernel("float", "dot_vec_vec"))
    assert float(cl.sum_vec(a)) == float(P.sum_vec(a, W))            # no multiply: nothing to contract, bit-exact
    lo, hi = sorted([float(P.dot_vec_vec(a, b, W, False)), float(P.dot_vec_vec(a, b, W, True))])
    tol = 4 * float(np.spacing(np.float32(max(abs(lo), abs(hi)))))
    assert lo - tol <= float(cl.dot_vec_vec(a, b)) <= hi + tol
    for M_, K_, N_ in ((33, 17, 29),):
        ma = gen.values("f32", M_ * K_, 213); mb = gen.values("f32", K_ * N_, 214)
        ref = cl.mul_mat_mat(ma, mb, M_, K_, N_)
        assert same(ref, P.mul_mat_mat(ma, mb, M_, K_, N_, fma=False)) or same(ref, P.mul_mat_mat(ma, mb, M_, K_, N_, fma=True))
