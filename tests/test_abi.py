"""CPU-only checks of the drop-in boundary: the library loads, exports exactly what
include/linalg_b200.h declares, and fails loudly (no CPU fallback) without a device."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "linalg_b200.h")
TUNING_HEADER = os.path.join(ROOT, "include", "linalg_b200_tuning.h")


def declared_functions(header=HEADER):
    src = open(header).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(la_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_reference_kernel_surface():
    names = declared_functions()
    # one entry point per kernel family of src/kernels.cl + buffers + transpose
    for required in ["la_ew_vec_vec", "la_ew_assign_vec_vec", "la_ew_vec_scl", "la_ew_assign_vec_scl", "la_eq_vec",
                     "la_sum_vec", "la_dot_vec_vec", "la_mul_vec_mat", "la_mul_vec_transpose_mat", "la_mul_mat_mat",
                     "la_transpose", "la_malloc", "la_free", "la_upload", "la_download", "la_copy", "la_last_error"]:
        assert required in names


def test_library_exports_every_declared_symbol():
    from linear_algebra_b200 import _lib
    for name in declared_functions():
        assert hasattr(_lib.lib, name), "liblinalg_b200.so does not export %s" % name
    # and the ctypes table covers the header (so the binding cannot drift silently)
    bound = set(_lib.SIGNATURES) | set(_lib.NON_STATUS)
    assert set(declared_functions()) == bound
    assert set(declared_functions(TUNING_HEADER)) == set(_lib.TUNING)


def test_library_exports_nothing_that_is_not_declared():
    """The converse: every dynamic symbol the library defines is declared in one of the two headers (the boundary in
    linalg_b200.h, the tuning / introspection hooks in linalg_b200_tuning.h).  C++ helpers shared between the .cu files and the
    statically linked CUDA runtime must not leak out (csrc/exports.map)."""
    import subprocess
    from linear_algebra_b200 import _lib
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], check=True, capture_output=True, text=True).stdout
    exported = {line.split()[-1] for line in out.splitlines() if len(line.split()) >= 3}
    declared = set(declared_functions()) | set(declared_functions(TUNING_HEADER))
    assert exported == declared, (sorted(exported - declared), sorted(declared - exported))


def test_tuning_header_is_plain_c(tmp_path):
    import subprocess
    c = tmp_path / "t.c"
    c.write_text('#include "linalg_b200_tuning.h"\nint main(void){ return 0; }\n')
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"), "-c", str(c), "-o", str(tmp_path / "t.o")], check=True)


def test_header_is_plain_c(tmp_path):
    import subprocess
    c = tmp_path / "t.c"
    c.write_text('#include "linalg_b200.h"\nint main(void){ return (int)sizeof(la_dtype) == 0; }\n')
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"), "-c", str(c), "-o", str(tmp_path / "t.o")], check=True)


def test_dtype_sizes_and_version():
    from linear_algebra_b200 import _lib, traits
    for p in traits.BY_NAME.values():
        assert _lib.lib.la_dtype_size(p.code) == (1 if p.name == "i8" else p.size) or p.size == _lib.lib.la_dtype_size(p.code)
    assert _lib.lib.la_dtype_size(99) == 0
    assert b"sm_100a" in _lib.lib.la_version()


def test_traits_mirror_reference_type_map():
    from linear_algebra_b200 import traits
    # src/traits.rs:30-57
    expect = {"u8": "uchar", "i8": "uchar", "u16": "ushort", "i16": "short", "u32": "uint", "i32": "int",
              "u64": "ulong", "i64": "long", "usize": "ulong", "isize": "long", "f32": "float", "f64": "double"}
    for name, cl in expect.items():
        assert traits.parameter(name).type_to_str() == cl
    assert traits.parameter(np.float32) is traits.F32 and traits.F32.real and not traits.U32.real
    with pytest.raises(TypeError):
        traits.parameter("f16")


def test_no_cpu_fallback():
    """Without a CUDA device every entry point must fail with LA_ERR_CUDA, never compute."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import linear_algebra_b200 as la
    with pytest.raises(la.LaError) as e:
        la.init(0)
    assert e.value.status == la._lib.LA_ERR_CUDA
    with pytest.raises(la.LaError):
        la.Vector.from_vec(np.arange(4, dtype=np.uint32))
    out = np.zeros(1, np.float32)
    st = la._lib.lib.la_sum_vec(8, ctypes.c_void_p(16), 4, out.ctypes.data, None)
    assert st == la._lib.LA_ERR_CUDA and b"CUDA" in la._lib.lib.la_last_error()


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "linear_algebra_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("no CPU fallback", ""), "%s mentions the oracle" % f
