//! `Vector<T>`: one device buffer of `len` elements (the reference's src/vector.rs surface over the C ABI).
//! Shape checks stay `assert_eq!` panics; operators block until the device is done; out-of-place operators
//! allocate a fresh result.  NOT COMPILED IN THE BUILD IMAGE (see Cargo.toml).
use std::fmt;
use std::fs::File;
use std::io::{Error, ErrorKind, Read, Write};
use std::marker::PhantomData;
use std::mem::size_of;
use std::ops::{Add, AddAssign, Div, DivAssign, Mul, MulAssign, Sub, SubAssign};
use std::os::raw::{c_int, c_void};
use std::ptr::null_mut;

use crate::ffi::{self, check, wait, DEFAULT_STREAM};
use crate::matrix::Matrix;
use crate::traits::{Parameter, Real};
use crate::util::{read_u64, write_u64};
use crate::Kernel;

pub struct Vector<T: Parameter> {
    ptr: *mut c_void,
    len: usize,
    marker: PhantomData<T>,
}

impl<T: Parameter> Vector<T> {
    pub fn new(default_val: T, size: usize) -> Vector<T> {
        let res = unsafe { Vector::uninitialized(size) };
        check(unsafe { ffi::la_fill(T::DTYPE, res.ptr, &default_val as *const T as *const c_void, size, DEFAULT_STREAM) });
        wait();
        res
    }

    /// Contents are undefined until written.
    pub unsafe fn uninitialized(size: usize) -> Vector<T> {
        let mut ptr = null_mut();
        check(ffi::la_malloc(&mut ptr, size * size_of::<T>()));
        Vector { ptr, len: size, marker: PhantomData }
    }

    pub fn from_vec(v: Vec<T>) -> Vector<T> {
        Vector::from_slice(&v)
    }

    pub fn from_slice(v: &[T]) -> Vector<T> {
        let res = unsafe { Vector::uninitialized(v.len()) };
        check(unsafe { ffi::la_upload(res.ptr, v.as_ptr() as *const c_void, v.len() * size_of::<T>(), DEFAULT_STREAM) });
        wait(); // `v` may be freed by the caller right after
        res
    }

    pub fn to_vec(&self) -> Vec<T> {
        let mut out = vec![T::default(); self.len];
        check(unsafe { ffi::la_download(out.as_mut_ptr() as *mut c_void, self.ptr, self.len * size_of::<T>(), DEFAULT_STREAM) });
        out
    }

    pub fn len(&self) -> usize {
        self.len
    }

    /// Raw device pointer: what `get_buffer` / `get_buf` hand out as an `ocl::Buffer` in the reference.
    pub fn device_ptr(&self) -> *mut c_void {
        self.ptr
    }

    // ---- user kernels: argument slots "C" (result or self), "A", "B"
    pub fn generate(kernel: &mut Kernel, size: usize) -> Vector<T> {
        let res = unsafe { Vector::uninitialized(size) };
        kernel.set_arg_buf_named("C", &res);
        kernel.enq(size);
        res
    }

    pub fn from_for_each2(a: &Vector<T>, b: &Vector<T>, kernel: &mut Kernel) -> Vector<T> {
        assert_eq!(a.len(), b.len());
        let res = unsafe { Vector::uninitialized(a.len()) };
        kernel.set_arg_buf_named("C", &res);
        kernel.set_arg_buf_named("A", a);
        kernel.set_arg_buf_named("B", b);
        kernel.enq(res.len());
        res
    }

    pub fn for_each_mut(&mut self, other: &Vector<T>, kernel: &mut Kernel) {
        assert_eq!(self.len(), other.len());
        kernel.set_arg_buf_named("C", self);
        kernel.set_arg_buf_named("B", other);
        kernel.enq(self.len());
    }

    pub fn map_mut(&mut self, kernel: &mut Kernel) {
        kernel.set_arg_buf_named("C", self);
        kernel.enq(self.len());
    }

    pub fn map(&self, kernel: &mut Kernel) -> Vector<T> {
        let res = unsafe { Vector::uninitialized(self.len()) };
        kernel.set_arg_buf_named("C", &res);
        kernel.set_arg_buf_named("B", self);
        kernel.enq(res.len());
        res
    }

    // ---- the built-in elementwise kernels
    fn binary(a: &Vector<T>, b: &Vector<T>, op: c_int) -> Vector<T> {
        assert_eq!(a.len(), b.len());
        let res = unsafe { Vector::uninitialized(a.len()) };
        check(unsafe { ffi::la_ew_vec_vec(T::DTYPE, op, res.ptr, a.ptr, b.ptr, a.len, DEFAULT_STREAM) });
        wait();
        res
    }

    fn binary_assign(&mut self, other: &Vector<T>, op: c_int) {
        assert_eq!(self.len(), other.len());
        check(unsafe { ffi::la_ew_assign_vec_vec(T::DTYPE, op, self.ptr, other.ptr, self.len, DEFAULT_STREAM) });
        wait();
    }

    fn scalar(&self, s: T, op: c_int) -> Vector<T> {
        let res = unsafe { Vector::uninitialized(self.len()) };
        check(unsafe { ffi::la_ew_vec_scl(T::DTYPE, op, res.ptr, self.ptr, &s as *const T as *const c_void, self.len, DEFAULT_STREAM) });
        wait();
        res
    }

    fn scalar_assign(&mut self, s: T, op: c_int) {
        check(unsafe { ffi::la_ew_assign_vec_scl(T::DTYPE, op, self.ptr, &s as *const T as *const c_void, self.len, DEFAULT_STREAM) });
        wait();
    }

    pub fn sum(&self) -> T {
        let mut out = T::default();
        check(unsafe { ffi::la_sum_vec(T::DTYPE, self.ptr, self.len, &mut out as *mut T as *mut c_void, DEFAULT_STREAM) });
        out
    }

    // ---- persistence: u64 element size, u64 count, raw elements, all native-endian (byte-compatible with the reference)
    pub(crate) fn write_file_only_data(&self, file: &mut File) -> Result<(), Error> {
        let data = self.to_vec();
        let bytes = unsafe { std::slice::from_raw_parts(data.as_ptr() as *const u8, data.len() * size_of::<T>()) };
        file.write_all(bytes)
    }

    pub(crate) fn read_file_only_data(file: &mut File, count: usize) -> Result<Vector<T>, Error> {
        let mut data = vec![T::default(); count];
        let bytes = unsafe { std::slice::from_raw_parts_mut(data.as_mut_ptr() as *mut u8, count * size_of::<T>()) };
        file.read_exact(bytes)?;
        Ok(Vector::from_vec(data))
    }

    pub fn write_to_file(&self, file: &mut File) -> Result<(), Error> {
        write_u64(file, size_of::<T>() as u64)?;
        write_u64(file, self.len() as u64)?;
        self.write_file_only_data(file)
    }

    pub fn read_from_file(file: &mut File) -> Result<Vector<T>, Error> {
        let elem_size = read_u64(file)?;
        let count = read_u64(file)?;
        if elem_size as usize != size_of::<T>() {
            return Err(Error::new(ErrorKind::InvalidData, "Elem size from buffer does not seem to match, what was expected!Missmatch in endiannes?"));
        }
        Vector::read_file_only_data(file, count as usize)
    }

    pub fn save(&self, path: &str) -> Result<(), Error> {
        let mut file = File::create(path)?;
        self.write_to_file(&mut file)
    }

    pub fn open(path: &str) -> Result<Vector<T>, Error> {
        let mut file = File::open(path)?;
        Vector::read_from_file(&mut file)
    }
}

impl<T: Real> Vector<T> {
    pub fn length(&self) -> T {
        dot(self, self).sqrt()
    }

    /// One fused launch chain on the device (dot, sqrt, reciprocal, scale): the same bits as
    /// `self *= T::one() / self.length()`.
    pub fn normalize(self) -> Self {
        check(unsafe { ffi::la_normalize_vec(T::DTYPE, self.ptr, self.ptr, self.len, DEFAULT_STREAM) });
        wait();
        self
    }

    pub fn normalized(&self) -> Self {
        let res = unsafe { Vector::uninitialized(self.len()) };
        check(unsafe { ffi::la_normalize_vec(T::DTYPE, res.ptr, self.ptr, self.len, DEFAULT_STREAM) });
        wait();
        res
    }
}

pub fn dot<T: Parameter>(a: &Vector<T>, b: &Vector<T>) -> T {
    assert_eq!(a.len(), b.len());
    let mut out = T::default();
    check(unsafe { ffi::la_dot_vec_vec(T::DTYPE, a.ptr, b.ptr, a.len, &mut out as *mut T as *mut c_void, DEFAULT_STREAM) });
    out
}

/// `M * v`: dot of every row of `other_m` with `vector` (length = column count).
pub fn mul_transpose_mat<T: Parameter>(vector: &Vector<T>, other_m: &Matrix<T>) -> Vector<T> {
    assert_eq!(vector.len(), other_m.get_col_count());
    let res = unsafe { Vector::uninitialized(other_m.get_row_count()) };
    check(unsafe {
        ffi::la_mul_vec_transpose_mat(T::DTYPE, res.ptr, vector.ptr, other_m.data().ptr, other_m.get_row_count(), other_m.get_col_count(), DEFAULT_STREAM)
    });
    wait();
    res
}

// `&a op &b` (new vector), `a op= &b` (in place), `a op &b` (consumes a, in place)
macro_rules! elementwise {
    ($Op:ident, $op:ident, $OpAssign:ident, $op_assign:ident, $code:expr) => {
        impl<'a, 'b, T: Parameter> $Op<&'b Vector<T>> for &'a Vector<T> {
            type Output = Vector<T>;
            fn $op(self, other: &'b Vector<T>) -> Vector<T> {
                Vector::binary(self, other, $code)
            }
        }
        impl<'a, T: Parameter> $OpAssign<&'a Vector<T>> for Vector<T> {
            fn $op_assign(&mut self, other: &'a Vector<T>) {
                self.binary_assign(other, $code)
            }
        }
        impl<'a, T: Parameter> $Op<&'a Vector<T>> for Vector<T> {
            type Output = Vector<T>;
            fn $op(mut self, other: &'a Vector<T>) -> Vector<T> {
                self.binary_assign(other, $code);
                self
            }
        }
    };
}
elementwise!(Add, add, AddAssign, add_assign, ffi::ADD);
elementwise!(Sub, sub, SubAssign, sub_assign, ffi::SUB);
elementwise!(Mul, mul, MulAssign, mul_assign, ffi::MUL);
elementwise!(Div, div, DivAssign, div_assign, ffi::DIV);

// `&a op s`, `a op= s`, `a op s`
macro_rules! by_scalar {
    ($Op:ident, $op:ident, $OpAssign:ident, $op_assign:ident, $code:expr) => {
        impl<'a, T: Parameter> $Op<T> for &'a Vector<T> {
            type Output = Vector<T>;
            fn $op(self, s: T) -> Vector<T> {
                self.scalar(s, $code)
            }
        }
        impl<T: Parameter> $OpAssign<T> for Vector<T> {
            fn $op_assign(&mut self, s: T) {
                self.scalar_assign(s, $code)
            }
        }
        impl<T: Parameter> $Op<T> for Vector<T> {
            type Output = Vector<T>;
            fn $op(mut self, s: T) -> Vector<T> {
                self.scalar_assign(s, $code);
                self
            }
        }
    };
}
by_scalar!(Mul, mul, MulAssign, mul_assign, ffi::MUL);
by_scalar!(Div, div, DivAssign, div_assign, ffi::DIV);

/// `v * M`: a row vector (length = row count) times the matrix.
impl<'a, 'b, T: Parameter> Mul<&'b Matrix<T>> for &'a Vector<T> {
    type Output = Vector<T>;
    fn mul(self, other_m: &'b Matrix<T>) -> Vector<T> {
        assert_eq!(self.len(), other_m.get_row_count());
        let res = unsafe { Vector::uninitialized(other_m.get_col_count()) };
        check(unsafe {
            ffi::la_mul_vec_mat(T::DTYPE, res.ptr, self.ptr, other_m.data().ptr, other_m.get_row_count(), other_m.get_col_count(), DEFAULT_STREAM)
        });
        wait();
        res
    }
}

impl<T: Parameter> PartialEq for Vector<T> {
    fn eq(&self, other: &Vector<T>) -> bool {
        if self.len() != other.len() {
            return false;
        }
        let mut equal: c_int = 0;
        check(unsafe { ffi::la_eq_vec(T::DTYPE, self.ptr, other.ptr, self.len, &mut equal, DEFAULT_STREAM) });
        equal != 0
    }
}

impl<T: Parameter + fmt::Display> fmt::Display for Vector<T> {
    fn fmt(&self, f: &mut fmt::Formatter) -> fmt::Result {
        let items: Vec<String> = self.to_vec().iter().map(|x| x.to_string()).collect();
        write!(f, "[{}]", items.join(", "))
    }
}

impl<T: Parameter + fmt::Debug> fmt::Debug for Vector<T> {
    fn fmt(&self, f: &mut fmt::Formatter) -> fmt::Result {
        write!(f, "{:?}", self.to_vec())
    }
}

impl<T: Parameter> Clone for Vector<T> {
    fn clone(&self) -> Vector<T> {
        let res = unsafe { Vector::uninitialized(self.len()) };
        check(unsafe { ffi::la_copy(res.ptr, self.ptr, self.len * size_of::<T>(), DEFAULT_STREAM) });
        wait();
        res
    }
}

impl<T: Parameter> Drop for Vector<T> {
    fn drop(&mut self) {
        unsafe { ffi::la_free(self.ptr) };
    }
}
