"""Element-type registry: the host-side mirror of src/traits.rs.

`Parameter` (traits.rs:13-15) maps a Rust element type to its device type; `Real`
(traits.rs:18-28) marks the floating types that have sqrt/one/zero.  Here a dtype is a small
record instead of a trait impl, and the `ocl::OclPrm` supertrait is gone."""
import numpy as np


class Parameter:
    def __init__(self, name, code, np_dtype, cl_name, real=False):
        self.name, self.code, self.np, self.cl_name, self.real = name, code, np.dtype(np_dtype), cl_name, real
        self.size = self.np.itemsize

    def type_to_str(self):            # traits.rs:14
        return self.cl_name

    def __repr__(self):
        return "Parameter(%s)" % self.name


# codes = la_dtype in include/linalg_b200.h
U8 = Parameter("u8", 0, np.uint8, "uchar")
I8_CHAR = Parameter("i8c", 1, np.int8, "char")          # OpenCL "char": compiled by the reference, unreachable from Rust
U16 = Parameter("u16", 2, np.uint16, "ushort")
I16 = Parameter("i16", 3, np.int16, "short")
U32 = Parameter("u32", 4, np.uint32, "uint")
I32 = Parameter("i32", 5, np.int32, "int")
U64 = Parameter("u64", 6, np.uint64, "ulong")
I64 = Parameter("i64", 7, np.int64, "long")
F32 = Parameter("f32", 8, np.float32, "float", real=True)
F64 = Parameter("f64", 9, np.float64, "double", real=True)
# traits.rs:36 maps Rust i8 to "uchar" (not "char"): Vector<i8> computes with unsigned 8-bit
# device arithmetic.  Identical bits for + - * sum dot eq; `/` is an UNSIGNED divide.  Kept
# for drop-in parity; use I8_CHAR for a true signed 8-bit divide.
I8 = Parameter("i8", 0, np.int8, "uchar")
USIZE, ISIZE = U64, I64                                  # traits.rs:42-43 on a 64-bit host

BY_NAME = {p.name: p for p in (U8, I8, I8_CHAR, U16, I16, U32, I32, U64, I64, F32, F64)}
BY_NAME.update({"usize": U64, "isize": I64})
_BY_NP = {np.dtype(np.uint8): U8, np.dtype(np.int8): I8, np.dtype(np.uint16): U16, np.dtype(np.int16): I16,
          np.dtype(np.uint32): U32, np.dtype(np.int32): I32, np.dtype(np.uint64): U64, np.dtype(np.int64): I64,
          np.dtype(np.float32): F32, np.dtype(np.float64): F64}


def parameter(dtype):
    """Resolve a Parameter from a Parameter, a name ("u32"), or a numpy dtype."""
    if isinstance(dtype, Parameter):
        return dtype
    if isinstance(dtype, str) and dtype in BY_NAME:
        return BY_NAME[dtype]
    try:
        return _BY_NP[np.dtype(dtype)]
    except (TypeError, KeyError):
        raise TypeError("Invalid type %r" % (dtype,))    # traits.rs:52 panic!("Invalid type")
