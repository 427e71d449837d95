"""oracle/user_cl.py -- TEST INFRASTRUCTURE ONLY (the product never imports it).

CPU oracle for USER kernels (the reference's extension point, /root/reference/src/lib.rs:156-180,
src/vector.rs:80-151): the user's OpenCL-C text is templated exactly as `get_src` does it
(src/lib.rs:337-354: "#define IS_<TY>", every "{T}" replaced by the type name), compiled as C++ through
oracle/cl_shim.hpp with g++, and run one work-item per index in an OpenMP loop -- the same route the
reference's own kernels take in oracle/build_ref.py.  Standing in for the core source the reference pastes
in front of the user's text are the two things a user kernel can rely on from it: `i` = get_global_id(0)
(src/kernels.cl:14) and `{T}_dot_product` (src/kernels.cl:134-140), restated below.

Only barrier-free kernels (every launch in vector.rs:80-151 is `gws = len`, no local size)."""
import ctypes
import hashlib
import os
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

# result = sum over j < count of a[j*stride_a] * b[j*stride_b], accumulated in T, j ascending (kernels.cl:134-140)
_CORE = """
#define __kernel static
#define __global
#define __local
#define restrict __restrict__
#undef barrier
#define barrier(flags)          /* barrier kernels may be present in the text but are never run by this oracle */
#define i get_global_id(0)
static inline {T} {T}_dot_product(const {T}* a, int stride_a, const {T}* b, int stride_b, int count) {
    {T} acc = 0;
    for (int j = 0; j != count; j++) acc += a[j * stride_a] * b[j * stride_b];
    return acc;
}
"""

_C_SCALAR = {"int": ctypes.c_int, "uint": ctypes.c_uint, "long": ctypes.c_long, "ulong": ctypes.c_ulong, "float": ctypes.c_float,
             "double": ctypes.c_double, "short": ctypes.c_short, "ushort": ctypes.c_ushort, "char": ctypes.c_byte, "uchar": ctypes.c_ubyte}


class UserKernelOracle:
    """`args`: the kernel's parameter list in order, each "buf" (a `global {T}*`) or "scl:<cl type>"."""

    def __init__(self, source, ty, kernel_name, args, contract="off"):
        self.ty, self.args = ty, list(args)
        params, call_args = [], []
        for j, a in enumerate(self.args):
            if a == "buf":
                call_args.append("(%s*)argv[%d]" % (ty, j))
            else:
                call_args.append("*(%s*)argv[%d]" % (a.split(":", 1)[1], j))
        text = "#define IS_%s\n" % ty.upper() + (_CORE + source + "\n").replace("{T}", ty)
        math = "#include <cmath>\n" + "".join("using std::%s; " % f for f in ("exp", "log", "sqrt", "fabs", "sin", "cos", "pow", "fma", "floor", "ceil")) + "\n"
        tu = ('#include "cl_shim.hpp"\n' + math + text + '\n#include "cl_driver.hpp"\n'
              'extern "C" void run_user_kernel(long n, void** argv) {\n    CL_NDRANGE(n, %s(%s))\n}\n' % (kernel_name, ", ".join(call_args)))
        tag = hashlib.sha1((tu + contract).encode()).hexdigest()[:16]
        self._dir = tempfile.mkdtemp(prefix="la_user_oracle_")
        so = os.path.join(self._dir, "user_%s.so" % tag)
        cmd = ["g++", "-x", "c++", "-std=c++17", "-O2", "-march=x86-64-v3", "-fopenmp", "-fPIC", "-shared", "-fwrapv",
               "-ffp-contract=%s" % contract, "-Wno-unused-function", "-I", HERE, "-o", so, "-"]
        r = subprocess.run(cmd, input=tu.encode(), capture_output=True)
        if r.returncode != 0:
            raise RuntimeError("oracle build of the user kernel failed:\n" + r.stderr.decode("utf-8", "replace"))
        self._lib = ctypes.CDLL(so)
        self._lib.run_user_kernel.argtypes = [ctypes.c_long, ctypes.POINTER(ctypes.c_void_p)]
        self._lib.run_user_kernel.restype = None

    def run(self, n, *values):
        """Buffers are numpy arrays (modified in place where the kernel writes them); scalars are Python numbers."""
        assert len(values) == len(self.args)
        argv = (ctypes.c_void_p * max(1, len(values)))()
        keep = []
        for j, (a, v) in enumerate(zip(self.args, values)):
            if a == "buf":
                assert isinstance(v, np.ndarray) and v.flags.c_contiguous
                argv[j] = v.ctypes.data
            else:
                c = _C_SCALAR[a.split(":", 1)[1]](v)
                keep.append(c)
                argv[j] = ctypes.addressof(c)
        self._lib.run_user_kernel(int(n), argv)
