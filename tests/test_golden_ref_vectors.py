"""Golden vectors produced by the reference's own kernels.cl (tests/golden/ref_vectors.npz, generated
by tests/golden/make_ref_vectors.py from oracle/_ref).  CPU: the plain-C restatement reproduces them
bit for bit.  GPU: the CUDA path, through the Vector/Matrix mirror, reproduces them -- bit-exact for
integers and single-operation float kernels, within the f32 bar for the summing kernels."""
import os

import numpy as np
import pytest

import gen
import oracle
from oracle import ADD, SUB, MUL, DIV

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_vectors.npz"))
OPS = (("add", ADD, "b"), ("sub", SUB, "b"), ("mul", MUL, "b"), ("div", DIV, "d"))


def same(x, y):
    x = np.asarray(x); y = np.asarray(y)
    return x.dtype == y.dtype and x.shape == y.shape and x.tobytes() == y.tobytes()


@pytest.mark.parametrize("name", gen.ALL_NAMES)
def test_port_reproduces_the_reference_outputs(name):
    P = oracle.port()
    a, b, d = G[name + ".a"], G[name + ".b"], G[name + ".d"]
    for opname, op, rhs in OPS:
        assert same(P.ew_vec_vec(op, a, G[name + "." + rhs]), G["%s.%s_vec_vec" % (name, opname)])
        assert same(P.ew_assign_vec_vec(op, a, G[name + "." + rhs]), G["%s.%s_assign_vec_vec" % (name, opname)])
    assert same(P.ew_vec_scl(MUL, a, b[0]), G[name + ".mul_vec_scl"]) and same(P.ew_vec_scl(DIV, a, d[0]), G[name + ".div_vec_scl"])
    ra, rb = G[name + ".red_a"], G[name + ".red_b"]
    assert same(P.reduce_groups(ra, None, 32), G[name + ".sum_groups_W32"]) and same(P.reduce_groups(ra, rb, 32), G[name + ".dot_groups_W32"])
    assert same(np.array([P.sum_vec(ra, 32)]), G[name + ".sum"]) and same(np.array([P.dot_vec_vec(ra, rb, 32)]), G[name + ".dot"])
    assert same(P.mul_vec_mat(G[name + ".vr"], G[name + ".m"], 37, 53), G[name + ".mul_vec_mat"])
    assert same(P.mul_vec_transpose_mat(G[name + ".vc"], G[name + ".m"], 37, 53), G[name + ".mul_vec_transpose_mat"])
    assert same(P.mul_mat_mat(G[name + ".ma"], G[name + ".mb"], 33, 17, 29), G[name + ".mul_mat_mat"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", gen.ALL_NAMES)
def test_cuda_reproduces_the_reference_outputs(name):
    import linear_algebra_b200 as la
    la.init(0)
    dt = la.I8_CHAR if name == "i8" else la.parameter(name)
    is_float = np.dtype(oracle.NP[name]).kind == "f"
    a, b, d = G[name + ".a"], G[name + ".b"], G[name + ".d"]
    A = la.Vector.from_vec(a, dt)
    rhs_dev = {"b": la.Vector.from_vec(b, dt), "d": la.Vector.from_vec(d, dt)}
    for (opname, op, rhs), fn in zip(OPS, (lambda x, y: x + y, lambda x, y: x - y, lambda x, y: x * y, lambda x, y: x / y)):
        assert same(fn(A, rhs_dev[rhs]).to_vec(), G["%s.%s_vec_vec" % (name, opname)]), (name, opname)
    C = A.clone(); C += rhs_dev["b"]
    assert same(C.to_vec(), G[name + ".add_assign_vec_vec"])
    C = A.clone(); C /= rhs_dev["d"]
    assert same(C.to_vec(), G[name + ".div_assign_vec_vec"])
    assert same((A * b[0]).to_vec(), G[name + ".mul_vec_scl"]) and same((A / d[0]).to_vec(), G[name + ".div_vec_scl"])
    c = a.copy(); c[200] = c[200] + a.dtype.type(1)
    assert [int(A == A), int(A == la.Vector.from_vec(c, dt)), int(A == la.Vector.from_vec(a[:-1], dt))] == G[name + ".eq"].tolist()
    ra, rb = G[name + ".red_a"], G[name + ".red_b"]
    RA = la.Vector.from_vec(ra, dt); RB = la.Vector.from_vec(rb, dt)
    M = la.Matrix.from_vec(G[name + ".m"], 37, 53, dt)
    vm = (la.Vector.from_vec(G[name + ".vr"], dt) * M).to_vec()
    mv = la.mul_transpose_mat(la.Vector.from_vec(G[name + ".vc"], dt), M).to_vec()
    mm = (la.Matrix.from_vec(G[name + ".ma"], 33, 17, dt) * la.Matrix.from_vec(G[name + ".mb"], 17, 29, dt)).to_vec()
    if not is_float:
        assert RA.sum() == G[name + ".sum"][0] and la.dot(RA, RB) == G[name + ".dot"][0]
        assert same(vm, G[name + ".mul_vec_mat"]) and same(mv, G[name + ".mul_vec_transpose_mat"]) and same(mm, G[name + ".mul_mat_mat"])
    else:
        tol = 1e-5 if name == "f32" else 1e-13
        scale = float(np.abs(ra.astype(np.float64) * rb.astype(np.float64)).sum())
        assert abs(float(la.dot(RA, RB)) - float(G[name + ".dot"][0])) <= tol * scale
        assert abs(float(RA.sum()) - float(G[name + ".sum"][0])) <= tol * float(np.abs(ra.astype(np.float64)).sum())
        for got, key in ((vm, ".mul_vec_mat"), (mv, ".mul_vec_transpose_mat"), (mm, ".mul_mat_mat")):
            ref = G[name + key].astype(np.float64)
            assert np.linalg.norm(got - ref) <= tol * np.linalg.norm(ref), key
