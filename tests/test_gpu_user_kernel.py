"""User kernels on the GPU (la_user_kernel_*; the reference's create_kernel + Vector::generate / from_for_each2 /
for_each_mut / map_mut / map, src/lib.rs:156-171, src/vector.rs:80-151).

The same kernel text runs three ways: through NVRTC on the B200 (the product), through the OpenCL-to-C++ shim on
the host (oracle/user_cl.py) and -- where an OpenCL driver exists -- appended to the reference's own program
exactly as its `get_src` does and run by that driver on the same GPU (oracle/cl_ref.py).  Integers must agree
bit for bit; f32 single-rounding expressions too; `exp` within 4 ulp (neither OpenCL's nor CUDA's is correctly
rounded); a*b+c within either contraction choice (OpenCL permits fusing)."""
import numpy as np
import pytest

import gen
from test_user_kernel import SAMPLE

pytestmark = pytest.mark.gpu

MORE = SAMPLE + """
kernel void {T}_blend(global {T}* C, global {T}* A, global {T}* B) { C[i] = A[i] * B[i] + A[i]; }
kernel void {T}_take(global {T}* C, global {T}* B) { C[i] = C[i] - B[i]; }
kernel void {T}_ramp(global {T}* C) { C[i] = ({T})(i * 3 + get_global_size(0)); }
kernel void {T}_shift(global {T}* C, {T} by) { C[i] = C[i] + by; }
"""
SIZES = [1, 255, 256, 257, 100003]           # below, at and around the 256-thread block; 100003 is prime


@pytest.fixture(scope="module")
def la():
    import linear_algebra_b200 as la
    la.init(0)
    return la


def dt(la, name):
    return la.I8_CHAR if name == "i8" else la.parameter(name)


_ORACLES = {}


def oracle_kernel(name, kernel, args, contract="off"):
    from oracle.user_cl import UserKernelOracle
    ty = dt(__import__("linear_algebra_b200"), name).cl_name
    key = (ty, kernel, contract)
    if key not in _ORACLES:
        _ORACLES[key] = UserKernelOracle(MORE, ty, ty + "_" + kernel, args, contract)
    return _ORACLES[key]


def same(x, y):
    return x.dtype == y.dtype and x.shape == y.shape and x.tobytes() == y.tobytes()


@pytest.mark.parametrize("name", gen.ALL_NAMES)
def test_map_doubles_every_element(la, name):
    d = dt(la, name)
    k = la.create_kernel(d, d.cl_name + "_twice", MORE).arg_buf_named("C").arg_buf_named("B")
    orc = oracle_kernel(name, "twice", ["buf", "buf"])
    for n in SIZES:
        b = gen.values(name, n, 900 + n)
        got = la.Vector.from_vec(b, d).map(k).to_vec()
        want = np.zeros_like(b)
        orc.run(n, want, b)
        assert same(got, want), (name, n)


@pytest.mark.parametrize("name", gen.ALL_NAMES)
def test_from_for_each2_and_for_each_mut(la, name):
    d = dt(la, name)
    blend = la.create_kernel(d, d.cl_name + "_blend", MORE).arg_buf_named("C").arg_buf_named("A").arg_buf_named("B")
    take = la.create_kernel(d, d.cl_name + "_take", MORE).arg_buf_named("C").arg_buf_named("B")
    n = 70001
    a = gen.values(name, n, 31); b = gen.values(name, n, 32)
    A = la.Vector.from_vec(a, d); B = la.Vector.from_vec(b, d)
    got = la.Vector.from_for_each2(A, B, blend).to_vec()
    wants = []
    for contract in ("off", "fast"):
        w = np.zeros_like(a)
        oracle_kernel(name, "blend", ["buf", "buf", "buf"], contract).run(n, w, a, b)
        wants.append(w)
    if name in ("f32", "f64"):
        assert all(g == w0 or g == w1 or (np.isnan(g) and np.isnan(w0)) for g, w0, w1 in zip(got, wants[0], wants[1]))
    else:
        assert same(got, wants[0])
    C = A.clone()
    C.for_each_mut(B, take)
    w = a.copy()
    oracle_kernel(name, "take", ["buf", "buf"]).run(n, w, b)
    assert same(C.to_vec(), w)
    with pytest.raises(AssertionError):
        la.Vector.from_for_each2(A, la.Vector.new(1, n - 1, d), blend)        # vector.rs:97 assert_eq!
    with pytest.raises(AssertionError):
        C.for_each_mut(la.Vector.new(1, n + 1, d), take)                     # vector.rs:115


@pytest.mark.parametrize("name", ["u8", "i16", "u32", "i64", "f32", "f64"])
def test_generate_and_map_mut_with_a_scalar(la, name):
    d = dt(la, name)
    ramp = la.create_kernel(d, d.cl_name + "_ramp", MORE).arg_buf_named("C")
    shift = la.create_kernel(d, d.cl_name + "_shift", MORE).arg_buf_named("C").arg_scl_named("by", 5)
    for n in (0, 1, 300, 4099):
        v = la.Vector.generate(ramp, n)
        with np.errstate(over="ignore"):
            want = (np.arange(n, dtype=np.uint64) * 3 + n).astype(d.np)
        assert same(v.to_vec(), want), (name, n)
        v.map_mut(shift)
        with np.errstate(over="ignore"):
            assert same(v.to_vec(), (want + d.np.type(5)).astype(d.np))
    shift.set_arg_scl_named("by", 9)
    v = la.Vector.new(1, 10, d).map_mut(shift)
    assert same(v.to_vec(), np.full(10, 10, d.np))


@pytest.mark.parametrize("name", ["i32", "u64", "f32"])
def test_helper_of_the_core_source_and_global_size(la, name):
    d = dt(la, name)
    rows, n = 1031, 77
    k = (la.create_kernel(d, d.cl_name + "_rows", MORE).arg_buf_named("C").arg_buf_named("A").arg_buf_named("B").arg_scl_named("n", n, "i32"))
    a = gen.values(name, rows * n, 41); b = gen.values(name, n, 42)
    A = la.Vector.from_vec(a, d); B = la.Vector.from_vec(b, d)
    # C, A and B have three different lengths here, so bind them by name and launch by hand (what a reference user
    # does with the ocl::Kernel directly) instead of going through from_for_each2
    C = la.Vector.uninitialized(rows, d)
    k.set_arg_buf_named("C", C).set_arg_buf_named("A", A).set_arg_buf_named("B", B)
    k.enq(rows)
    la.synchronize()
    want = np.zeros(rows, d.np)
    oracle_kernel(name, "rows", ["buf", "buf", "buf", "scl:int"]).run(rows, want, a, b, n)
    got = C.to_vec()
    if name == "f32":
        assert np.allclose(got, want, rtol=1e-4, atol=1e-4)                  # 77-term f32 sums, contraction allowed
    else:
        assert same(got, want)


def test_exp_helper_guarded_by_type(la):
    k = la.create_kernel(la.F32, "float_grow_kernel", MORE).arg_buf_named("C").arg_buf_named("B")
    x = np.linspace(-20, 20, 50001).astype(np.float32)
    got = la.Vector.from_vec(x).map(k).to_vec()
    want = np.exp(x.astype(np.float64))
    assert np.all(np.abs(got - want) <= 4 * np.spacing(want.astype(np.float32)))
    with pytest.raises(la.LaError):                                          # the guard leaves it out for integers
        la.create_kernel(la.I32, "int_grow_kernel", MORE)


def test_local_memory_and_barrier(la):
    k = la.create_kernel(la.I32, "int_flip", MORE).arg_buf_named("C").arg_buf_named("B")
    n = 256 * 37
    b = np.arange(n, dtype=np.int32)
    got = la.Vector.from_vec(b).map(k).to_vec()
    assert np.array_equal(got, b.reshape(-1, 256)[:, ::-1].reshape(-1))      # each work-group of 256 reversed


def test_cuda_dialect_kernel(la):
    k = la.Kernel(la.F32, "float_fill", 'extern "C" __global__ void float_fill(float* C, float v) { C[i] = v + (float)i; }')
    k.arg_buf_named("C").arg_scl_named("v", 0.5)
    v = la.Vector.generate(k, 1000)
    assert np.array_equal(v.to_vec(), np.arange(1000, dtype=np.float32) + 0.5)


def test_errors(la):
    with pytest.raises(la.LaError) as e:
        la.create_kernel(la.F32, "float_bad", "kernel void {T}_bad(global {T}* C) { C[i] = nothing; }")
    assert "nothing" in str(e.value)
    with pytest.raises(la.LaError) as e:
        la.create_kernel(la.F32, "float_absent", MORE)
    assert "float_absent" in str(e.value)
    k = la.create_kernel(la.F32, "float_twice", MORE).arg_buf_named("C").arg_buf_named("B")
    with pytest.raises(ValueError):
        k.enq(10)                                                            # arguments never set
    with pytest.raises(KeyError):
        k.set_arg_buf_named("Z", la.Vector.new(0, 1, la.F32))
    # the library still works afterwards
    assert (la.Vector.new(1.0, 8, la.F32) + la.Vector.new(2.0, 8, la.F32)).to_vec().tolist() == [3.0] * 8


def test_create_kernel_reads_src_extra_kernels_cl(la, tmp_path, monkeypatch):
    (tmp_path / "src").mkdir()
    (tmp_path / "src" / "extra_kernels.cl").write_text(MORE)
    monkeypatch.chdir(tmp_path)
    k = la.create_kernel(la.U16, "ushort_twice").arg_buf_named("C").arg_buf_named("B")      # lib.rs:165, :173-180
    assert la.Vector.from_vec(np.array([1, 2, 40000], np.uint16)).map(k).to_vec().tolist() == [2, 4, 80000 - 65536]


@pytest.mark.parametrize("name", [n for n in gen.ALL_NAMES if n != "f64"])
def test_same_text_through_the_reference_program_on_opencl(la, name):
    """The user's text appended to the reference's program as lib.rs:337-354 does, built and run by the OpenCL driver."""
    from oracle import cl_ref
    try:
        cl = _cl_with_extra(cl_ref)
    except cl_ref.ClUnavailable as e:
        pytest.skip(str(e))
    d = dt(la, name)
    n = 100003
    a = gen.values(name, n, 51); b = gen.values(name, n, 52)
    A = la.Vector.from_vec(a, d); B = la.Vector.from_vec(b, d)
    twice = la.create_kernel(d, d.cl_name + "_twice", MORE).arg_buf_named("C").arg_buf_named("B")
    ref = cl.run_user_kernel(d.cl_name + "_twice", [np.zeros_like(b), b], n)[0]
    assert same(B.map(twice).to_vec(), ref)
    take = la.create_kernel(d, d.cl_name + "_take", MORE).arg_buf_named("C").arg_buf_named("B")
    ref = cl.run_user_kernel(d.cl_name + "_take", [a.copy(), b], n)[0]
    C = A.clone()
    assert same(C.for_each_mut(B, take).to_vec(), ref)
    if name != "f32":                                                        # a*b+a: the two drivers may fuse differently
        blend = la.create_kernel(d, d.cl_name + "_blend", MORE).arg_buf_named("C").arg_buf_named("A").arg_buf_named("B")
        ref = cl.run_user_kernel(d.cl_name + "_blend", [np.zeros_like(a), a, b], n)[0]
        assert same(la.Vector.from_for_each2(A, B, blend).to_vec(), ref)


_CL = {}


def _cl_with_extra(cl_ref):
    if "x" not in _CL:
        try:
            _CL["x"] = cl_ref.ClRef(extra_source=MORE)
        except cl_ref.ClUnavailable as e:
            _CL["x"] = e
    if isinstance(_CL["x"], Exception):
        raise _CL["x"]
    return _CL["x"]
