//! Element types.  `Parameter` keeps `type_to_str()` (the {T} a user kernel sees) and gains the library's dtype code
//! in place of the reference's `ocl::OclPrm` supertrait (src/traits.rs:13-15 there); `Real` is unchanged in meaning.
use std::ops::{Add, AddAssign, Div, DivAssign, Mul, MulAssign, Neg, Sub, SubAssign};
use std::os::raw::c_int;

pub trait Parameter: Copy + Default + PartialEq + 'static {
    /// `la_dtype` of include/linalg_b200.h.
    const DTYPE: c_int;
    fn type_to_str() -> &'static str;
}

pub trait Real:
    Parameter + Add<Output = Self> + AddAssign + Sub<Output = Self> + SubAssign + Mul<Output = Self> + MulAssign + Div<Output = Self> + DivAssign + Neg<Output = Self>
{
    fn sqrt(self) -> Self;
    fn one() -> Self;
    fn zero() -> Self;
}

macro_rules! parameter {
    ($($ty:ty => $code:expr, $name:expr;)+) => {
        $(impl Parameter for $ty {
            const DTYPE: c_int = $code;
            fn type_to_str() -> &'static str { $name }
        })+
    };
}

// i8 computes as "uchar", exactly as the reference's table has it (src/traits.rs:36 there): same bits for + - * sum
// dot ==, an unsigned divide for `/`.
parameter! {
    u8 => 0, "uchar";   i8 => 0, "uchar";
    u16 => 2, "ushort"; i16 => 3, "short";
    u32 => 4, "uint";   i32 => 5, "int";
    u64 => 6, "ulong";  i64 => 7, "long";
    f32 => 8, "float";  f64 => 9, "double";
}
#[cfg(target_pointer_width = "64")]
parameter! { usize => 6, "ulong"; isize => 7, "long"; }
#[cfg(target_pointer_width = "32")]
parameter! { usize => 4, "uint"; isize => 5, "int"; }

macro_rules! real {
    ($($ty:ident),+) => {
        $(impl Real for $ty {
            fn sqrt(self) -> Self { $ty::sqrt(self) }
            fn one() -> Self { 1.0 }
            fn zero() -> Self { 0.0 }
        })+
    };
}
real!(f32, f64);
