"""The reference's own unit tests (/root/reference/src/tests.rs), mirrored 1:1 against the
CUDA back end through the host-side Vector/Matrix mirror.  Same inputs, same expected values
(tests/golden/reference_tests.json), same assertions."""
import json
import math
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "reference_tests.json")))
TestType = "u32"      # tests.rs:3


@pytest.fixture(scope="module")
def la():
    import linear_algebra_b200 as la
    la.init(0)
    return la


def test_f32_vec_mat_mul(la):                 # tests.rs:5-21
    g = GOLD["vec_mat_mul"]
    a = la.Vector.from_vec(g["a"], "f32"); b = la.Vector.from_vec(g["expect"], "f32")
    m = la.Matrix.from_vec(g["m"], g["rows"], g["cols"], "f32")
    p = a * m
    assert p == b


def test_vec_eq(la):                          # tests.rs:23-44
    g = GOLD["vec_eq"]
    a, b, c, d = (la.Vector.from_vec(g[k], TestType) for k in "abcd")
    assert a == a and b == b
    assert a != c and c != a
    assert a != b and b != a
    assert a == d and d == a


def test_vec_sum(la):                         # tests.rs:46-57
    for size, expect in GOLD["vec_sum"]["cases"]:
        b = la.Vector.from_vec(np.arange(size, dtype=np.uint32)).sum()
        assert int(b) == expect == sum(range(size))


def test_vec_vec_dot(la):                     # tests.rs:59-79
    rng_a = np.arange(1, 971, dtype=np.uint32); rng_b = np.arange(3, 973, dtype=np.uint32)[::-1].copy()
    a = int((rng_a.astype(np.uint64) * rng_b).sum() & 0xFFFFFFFF)
    b = la.dot(la.Vector.from_vec(rng_a), la.Vector.from_vec(rng_b))
    assert a == int(b) == GOLD["vec_vec_dot"]["expect"]


def test_mat_eq(la):                          # tests.rs:81-123
    g = GOLD["mat_eq"]
    a = la.Matrix.from_vec(g["a"], *g["a_shape"], TestType); b = la.Matrix.from_vec(g["b"], *g["b_shape"], TestType)
    c = la.Matrix.from_vec(g["c"], *g["c_shape"], TestType); d = la.Matrix.from_vec(g["a"], *g["a_shape"], TestType)
    assert a == a and b == b
    assert a != c and c != a
    assert a != b and b != a
    assert a == d and d == a


def test_vec_mat_mul(la):                     # tests.rs:125-141
    g = GOLD["vec_mat_mul"]
    a = la.Vector.from_vec(g["a"], TestType); b = la.Vector.from_vec(g["expect"], TestType)
    m = la.Matrix.from_vec(g["m"], g["rows"], g["cols"], TestType)
    assert a * m == b


def test_mat_mat_mul(la):                     # tests.rs:143-164
    g = GOLD["mat_mat_mul"]
    a = la.Matrix.from_vec(g["a"], g["M"], g["K"], TestType); b = la.Matrix.from_vec(g["b"], g["K"], g["N"], TestType)
    c = la.Matrix.from_vec(g["expect"], g["M"], g["N"], TestType)
    assert c == a * b


def test_vec_transpose_mat_mul(la):           # tests.rs:166-181
    g = GOLD["vec_transpose_mat_mul"]
    a = la.Vector.from_vec(g["a"], TestType); b = la.Vector.from_vec(g["expect"], TestType)
    m = la.Matrix.from_vec(g["m"], g["rows"], g["cols"], TestType)
    assert la.mul_transpose_mat(a, m) == b


def test_vec_add(la):                         # tests.rs:183-193
    a = la.Vector.from_vec(GOLD["vec_add"]["a"], TestType); r = la.Vector.from_vec(GOLD["vec_add"]["expect"], TestType)
    assert r == a + a


def test_vec_mul_scl(la):                     # tests.rs:195-206
    a = la.Vector.from_vec(GOLD["vec_mul_scl"]["a"], TestType); r = la.Vector.from_vec(GOLD["vec_mul_scl"]["expect"], TestType)
    assert r == a * 2


def test_save_load_vec(la, tmp_path):         # tests.rs:208-221
    path = str(tmp_path / "save_load_vec.tmp")
    a = la.Vector.from_vec(GOLD["save_load_vec"]["a"], TestType)
    a.save(path)
    b = la.Vector.open(path, TestType)
    assert a == b
    raw = open(path, "rb").read()              # util.rs:6-21 / vector.rs:202-207: u64 elem size, u64 count, data
    assert np.frombuffer(raw[:16], np.uint64).tolist() == [4, 4]
    assert np.frombuffer(raw[16:], np.uint32).tolist() == GOLD["save_load_vec"]["a"]
    with pytest.raises(AssertionError):       # vector.rs:215-220
        la.Vector.open(path, "u64")


def test_save_load_mat(la, tmp_path):         # tests.rs:223-238
    path = str(tmp_path / "save_load_mat.tmp")
    g = GOLD["save_load_mat"]
    a = la.Matrix.from_vec(g["a"], *g["shape"], TestType)
    a.save(path)
    b = la.Matrix.open(path, TestType)
    assert a == b and b.len() == tuple(g["shape"])
    raw = open(path, "rb").read()
    assert np.frombuffer(raw[:24], np.uint64).tolist() == [4] + g["shape"]


def test_normalize_vec(la):                   # tests.rs:240-252
    for size in GOLD["normalize_vec"]["sizes"]:
        v = np.arange(1, size + 1, dtype=np.float32)
        a = la.Vector.from_vec(v).normalized()
        b = la.Vector.from_vec(v).normalize()
        assert abs(float(a.length()) - 1.0) < 0.01, a.length()
        assert abs(float(b.length()) - 1.0) < 0.01, b.length()


def test_length_vec(la):                      # tests.rs:254-271
    for size in GOLD["length_vec"]["sizes"]:
        a = la.Vector.from_vec(np.arange(size, dtype=np.float32)).length()
        b = np.sqrt(np.float32(sum(float(x * x) for x in range(size))))
        assert abs(float(a) - float(b)) < 0.0001, (a, b)


def test_vec_clone(la):                       # tests.rs:273-284
    g = GOLD["vec_clone"]
    a = la.Vector.from_vec(g["a"], TestType); b = la.Vector.from_vec(g["b"], TestType)
    c = a.clone()
    assert a == c and a != b


def test_shape_mismatch_panics_like_the_reference(la):
    a = la.Vector.from_vec([1, 2, 3], TestType); b = la.Vector.from_vec([1, 2], TestType)
    m = la.Matrix.from_vec([1, 2, 3, 4, 5, 6], 2, 3, TestType)
    with pytest.raises(AssertionError):       # vector.rs:97
        a + b
    with pytest.raises(AssertionError):       # vector.rs:115
        a += b
    with pytest.raises(AssertionError):       # vector.rs:327
        la.dot(a, b)
    with pytest.raises(AssertionError):       # vector.rs:513 (needs len == rows == 2)
        a * m
    with pytest.raises(AssertionError):       # vector.rs:545 (needs len == cols == 3)
        la.mul_transpose_mat(b, m)
    with pytest.raises(AssertionError):       # matrix.rs:229
        m * m
    with pytest.raises(AssertionError):       # matrix.rs:40
        la.Matrix.from_vec([1, 2, 3], 2, 2, TestType)
    with pytest.raises(AssertionError):       # matrix.rs:137-138
        m + la.Matrix.from_vec([1, 2, 3, 4, 5, 6], 3, 2, TestType)
