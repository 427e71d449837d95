"""User kernels without a GPU: NVRTC compiles the reference-dialect text for every element type (la_user_kernel_compile
needs no device), compile errors come back with the build log, and the CPU oracle for user kernels
(oracle/user_cl.py: the same text through the OpenCL-to-C++ shim) agrees with plain numpy on known answers, so it
can be the checker in tests/test_gpu_user_kernel.py."""
import ctypes

import numpy as np
import pytest

import linear_algebra_b200 as la
from linear_algebra_b200 import _lib

# the sample that ships as the reference's src/extra_kernels.cl is of this shape: a helper only for the real types, a
# kernel for all of them (written here from its description in SURVEY.md, not copied)
SAMPLE = """
#if defined(IS_FLOAT) || defined(IS_DOUBLE)
{T} {T}_grow({T} x) {
    return exp(x);
}
kernel void {T}_grow_kernel(global {T}* C, global {T}* B) { C[i] = {T}_grow(B[i]); }
#endif

kernel void {T}_twice(global {T}* C, global {T}* B) {
    C[i] = B[i] * ({T})(2);
}
/* a comment that says global, local and kernel */
__kernel void {T}_rows(__global {T}* restrict C, __global {T}* A, __global {T}* B, int n) {
    // one work-item per row of A (n columns): dot with B, plus the NDRange size
    C[i] = {T}_dot_product(A + i * n, 1, B, 1, n) + ({T})get_global_size(0);
}
kernel void {T}_flip(global {T}* C, global {T}* B) {
    local {T} tile[256];
    tile[get_local_id(0)] = B[i];
    barrier(CLK_LOCAL_MEM_FENCE);
    C[i] = tile[get_local_size(0) - 1 - get_local_id(0)];
}
"""

ALL = [la.U8, la.I8_CHAR, la.U16, la.I16, la.U32, la.I32, la.U64, la.I64, la.F32, la.F64]


@pytest.mark.parametrize("dt", ALL, ids=lambda d: d.name)
def test_reference_dialect_compiles_for_every_type(dt):
    n = ctypes.c_size_t(0)
    _lib.call("la_user_kernel_compile", dt.code, SAMPLE.encode(), ctypes.byref(n))
    assert n.value > 4096                                   # a cubin with more in it than the launch-state kernel
    assert _lib.lib.la_user_kernel_type_name(dt.code).decode() == dt.cl_name
    assert la.check_source(dt, SAMPLE) == ""


def test_kernels_really_are_in_the_cubin():
    """`#define global` style mappings silently demote kernels to plain functions under NVRTC (the reason the
    qualifiers are rewritten as identifiers): the cubin must grow when a kernel is added."""
    a, b = ctypes.c_size_t(0), ctypes.c_size_t(0)
    _lib.call("la_user_kernel_compile", la.F32.code, b"", ctypes.byref(a))
    _lib.call("la_user_kernel_compile", la.F32.code, b"kernel void {T}_k(global {T}* C) { C[i] = C[i] * ({T})3; }", ctypes.byref(b))
    assert b.value > a.value


def test_cuda_dialect_is_accepted_too():
    assert la.check_source(la.F32, 'extern "C" __global__ void float_fill(float* C, float v) { C[i] = v; }') == ""


def test_compile_error_carries_the_build_log():
    with pytest.raises(la.LaError) as e:
        la.check_source(la.I32, "kernel void {T}_bad(global {T}* C) {\n  C[i] = no_such_name;\n}")
    assert e.value.status == _lib.LA_ERR_ARG
    assert "no_such_name" in str(e.value) and "extra_kernels.cl(2)" in str(e.value)      # the user's own line number
    assert "no_such_name" in _lib.lib.la_user_kernel_build_log().decode()


def test_type_guards_select_code_per_type():
    src = "#ifdef IS_UINT\nkernel void {T}_only(global {T}* C) { C[i] = 1; }\n#else\n#error not for this type\n#endif\n"
    assert la.check_source(la.U32, src) == ""
    with pytest.raises(la.LaError):
        la.check_source(la.I32, src)


def test_unknown_dtype_and_null_source_are_rejected():
    assert _lib.lib.la_user_kernel_compile(42, b"", None) == _lib.LA_ERR_ARG
    assert _lib.lib.la_user_kernel_compile(la.F32.code, None, None) == _lib.LA_ERR_ARG
    assert _lib.lib.la_user_kernel_type_name(42) is None


def test_missing_extra_kernels_file_is_an_empty_source(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    assert la.load_extra_src() == ""                         # lib.rs:340 unwrap_or_default
    (tmp_path / "src").mkdir()
    (tmp_path / "src" / "extra_kernels.cl").write_text("kernel void {T}_x(global {T}* C) { C[i] = 0; }")
    assert "{T}_x" in la.load_extra_src()


# ---------------------------------------------------------------- the CPU oracle for user kernels, against numpy
def test_user_kernel_oracle_known_answers():
    from oracle.user_cl import UserKernelOracle
    b = np.arange(1000, dtype=np.int32) * 7 - 300
    c = np.zeros_like(b)
    UserKernelOracle(SAMPLE, "int", "int_twice", ["buf", "buf"]).run(b.size, c, b)
    assert np.array_equal(c, b * 2)

    b8 = np.arange(256, dtype=np.uint8)
    c8 = np.zeros_like(b8)
    UserKernelOracle(SAMPLE, "uchar", "uchar_twice", ["buf", "buf"]).run(256, c8, b8)
    assert np.array_equal(c8, (b8.astype(np.uint32) * 2).astype(np.uint8))          # wraps mod 2^8

    rows, n = 37, 19
    rng = np.random.default_rng(5)
    A = rng.integers(-1000, 1000, rows * n).astype(np.int64); B = rng.integers(-1000, 1000, n).astype(np.int64)
    C = np.zeros(rows, np.int64)
    UserKernelOracle(SAMPLE, "long", "long_rows", ["buf", "buf", "buf", "scl:int"]).run(rows, C, A, B, n)
    assert np.array_equal(C, A.reshape(rows, n) @ B + rows)

    x = np.linspace(-3, 3, 101).astype(np.float32)
    y = np.zeros_like(x)
    UserKernelOracle(SAMPLE, "float", "float_grow_kernel", ["buf", "buf"]).run(x.size, y, x)
    assert np.allclose(y, np.exp(x.astype(np.float64)), rtol=3e-7)
