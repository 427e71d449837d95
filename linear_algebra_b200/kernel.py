"""User kernels: host-side mirror of the reference's extension point over the C ABI.

Reference: `create_kernel::<T>(name)` (src/lib.rs:156-171) returns an `ocl::Kernel` from the text of
src/extra_kernels.cl, which is compiled with the library's own kernels once per element type with
`{T}` replaced by the type's OpenCL name and `IS_<TYPE>` defined (src/lib.rs:173-180, 337-354).  The
user then declares the kernel's arguments by name, in signature order (`arg_buf_named("C", None)`,
as src/lib.rs:189-274 does for the built-in kernels), and hands the kernel to the Vector helpers
(src/vector.rs:80-151), which fill the slots named "C", "A", "B" and launch one work-item per element.

Here `create_kernel(dtype, name, source)` compiles the same text with NVRTC for sm_100a
(la_user_kernel_create); the text can stay in the reference's OpenCL-C dialect."""
import ctypes
import os

import numpy as np

from . import _lib
from ._lib import call
from .traits import parameter

EXTRA_KERNELS_PATH = os.path.join("src", "extra_kernels.cl")      # src/lib.rs:176, relative to the working directory


def load_extra_src(path=EXTRA_KERNELS_PATH):
    """src/lib.rs:173-180; a missing file is an empty string (`unwrap_or_default`, lib.rs:340)."""
    try:
        with open(path, "r") as f:
            return f.read()
    except OSError:
        return ""


def check_source(dtype, source):
    """Compile only (works without a GPU).  Returns NVRTC's log (warnings); raises LaError with the log on errors."""
    call("la_user_kernel_compile", parameter(dtype).code, source.encode(), None)
    return _lib.lib.la_user_kernel_build_log().decode("utf-8", "replace")


class Kernel:
    """One compiled kernel plus its named argument slots (the `ocl::Kernel` of the reference)."""

    def __init__(self, dtype, kernel_name, source):
        self.dtype = parameter(dtype)
        self.name = kernel_name
        self._handle = None
        self._names, self._values, self._keep = [], {}, {}
        h = ctypes.c_void_p()
        call("la_user_kernel_create", self.dtype.code, source.encode(), kernel_name.encode(), ctypes.byref(h))
        self._handle = h.value

    # ---- declaring arguments, in the order of the kernel's parameter list
    def arg_buf_named(self, name, vector=None):
        self._names.append(name)
        self._values[name] = None
        if vector is not None:
            self.set_arg_buf_named(name, vector)
        return self

    def arg_scl_named(self, name, value=None, dtype=None):
        self._names.append(name)
        self._values[name] = None
        self._keep[name] = parameter(dtype if dtype is not None else self.dtype)
        if value is not None:
            self.set_arg_scl_named(name, value)
        return self

    # ---- filling them
    def set_arg_buf_named(self, name, vector):
        if name not in self._values:
            raise KeyError("kernel %s has no argument named %r" % (self.name, name))     # ocl: Err -> unwrap() panics
        self._values[name] = ctypes.c_void_p(vector.ptr if hasattr(vector, "ptr") else int(vector))
        return self

    def set_arg_scl_named(self, name, value):
        if name not in self._values or name not in self._keep:
            raise KeyError("kernel %s has no scalar argument named %r" % (self.name, name))
        self._values[name] = np.array([value], dtype=self._keep[name].np)
        return self

    def enq(self, gws, stream=None):
        """Launch with global work size `gws` (`kernel.cmd().gws(n).enq()`); stream-ordered, does not wait."""
        argv = (ctypes.c_void_p * max(1, len(self._names)))()
        for j, name in enumerate(self._names):
            v = self._values[name]
            if v is None:
                raise ValueError("kernel %s: argument %r was never set" % (self.name, name))
            argv[j] = ctypes.addressof(v) if isinstance(v, ctypes.c_void_p) else v.ctypes.data
        call("la_user_kernel_launch", self._handle, int(gws), argv, stream)

    def __del__(self):
        if getattr(self, "_handle", None):
            try:
                _lib.lib.la_user_kernel_destroy(ctypes.c_void_p(self._handle))
            except Exception:
                pass
            self._handle = None


def create_kernel(dtype, kernel_name, source=None):
    """src/lib.rs:165-171.  `kernel_name` is the full name (`{T}` already substituted, e.g. "float_test_kernel").
    `source` defaults to the text of ./src/extra_kernels.cl, where the reference looks for it."""
    return Kernel(dtype, kernel_name, load_extra_src() if source is None else source)
