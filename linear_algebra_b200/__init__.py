"""linear_algebra_b200 -- B200-native back end for the hot path of usbalbin/linear_algebra.

`liblinalg_b200.so` (csrc/, sm_100a CUDA behind the C ABI of include/linalg_b200.h) plus a thin
host-side mirror of the reference's public surface: Vector, Matrix, dot, mul_transpose_mat,
mul_column_row, create_kernel (user kernels) and the Parameter/Real dtype registry."""
from . import _lib
from ._lib import LaError, lib
from .traits import Parameter, parameter, U8, I8, I8_CHAR, U16, I16, U32, I32, U64, I64, USIZE, ISIZE, F32, F64
from .vector import Vector, dot, mul_transpose_mat, set_sync, synchronize, pinned, Event
from .matrix import Matrix, mul_column_row, mul_host
from .kernel import Kernel, create_kernel, load_extra_src, check_source


def init(device=0):
    _lib.call("la_init", device)


def version():
    return lib.la_version().decode()
