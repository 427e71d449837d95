//! `Matrix<T>`: a row-major `Vector<T>` of `row_count * col_count` elements plus its shape (the reference's
//! src/matrix.rs surface over the C ABI).  NOT COMPILED IN THE BUILD IMAGE (see Cargo.toml).
use std::fmt;
use std::fs::File;
use std::io::{Error, ErrorKind};
use std::mem::size_of;
use std::ops::{Add, AddAssign, Div, Mul, MulAssign, Sub, SubAssign};
use std::os::raw::c_void;

use crate::ffi::{self, check, wait, DEFAULT_STREAM};
use crate::traits::Parameter;
use crate::util::{read_u64, write_u64};
use crate::vector::Vector;
use crate::Kernel;

pub struct Matrix<T: Parameter> {
    data: Vector<T>,
    row_count: usize,
    col_count: usize,
}

impl<T: Parameter> Matrix<T> {
    pub fn new(default_value: T, row_count: usize, col_count: usize) -> Matrix<T> {
        Matrix { data: Vector::new(default_value, row_count * col_count), row_count, col_count }
    }

    pub unsafe fn uninitialized(row_count: usize, col_count: usize) -> Matrix<T> {
        Matrix { data: Vector::uninitialized(row_count * col_count), row_count, col_count }
    }

    pub fn from_vec(v: Vec<T>, row_count: usize, col_count: usize) -> Matrix<T> {
        assert_eq!(row_count * col_count, v.len());
        Matrix { data: Vector::from_vec(v), row_count, col_count }
    }

    pub fn to_vec(&self) -> Vec<T> {
        self.data.to_vec()
    }

    pub fn generate(kernel: &mut Kernel, row_count: usize, col_count: usize) -> Matrix<T> {
        Matrix { data: Vector::generate(kernel, row_count * col_count), row_count, col_count }
    }

    pub fn len(&self) -> (usize, usize) {
        (self.row_count, self.col_count)
    }

    pub fn get_row_count(&self) -> usize {
        self.row_count
    }

    pub fn get_col_count(&self) -> usize {
        self.col_count
    }

    pub fn data(&self) -> &Vector<T> {
        &self.data
    }

    /// out[c][r] = self[r][c].  Not in the reference (it only has `mul_transpose_mat`).
    pub fn transpose(&self) -> Matrix<T> {
        let res = unsafe { Matrix::uninitialized(self.col_count, self.row_count) };
        check(unsafe { ffi::la_transpose(T::DTYPE, res.data.device_ptr(), self.data.device_ptr(), self.row_count, self.col_count, DEFAULT_STREAM) });
        wait();
        res
    }

    fn same_shape(&self, other: &Matrix<T>) {
        assert_eq!(self.get_row_count(), other.get_row_count());
        assert_eq!(self.get_col_count(), other.get_col_count());
    }

    // ---- persistence: u64 element size, u64 rows, u64 columns, raw elements (byte-compatible with the reference)
    pub fn write_to_file(&self, file: &mut File) -> Result<(), Error> {
        write_u64(file, size_of::<T>() as u64)?;
        write_u64(file, self.row_count as u64)?;
        write_u64(file, self.col_count as u64)?;
        self.data.write_file_only_data(file)
    }

    pub fn read_from_file(file: &mut File) -> Result<Matrix<T>, Error> {
        let elem_size = read_u64(file)?;
        let row_count = read_u64(file)? as usize;
        let col_count = read_u64(file)? as usize;
        if elem_size as usize != size_of::<T>() {
            return Err(Error::new(ErrorKind::InvalidData, "Elem size from buffer does not seem to match, what was expected!Missmatch in endiannes?"));
        }
        Ok(Matrix { data: Vector::read_file_only_data(file, row_count * col_count)?, row_count, col_count })
    }

    pub fn save(&self, path: &str) -> Result<(), Error> {
        let mut file = File::create(path)?;
        self.write_to_file(&mut file)
    }

    pub fn open(path: &str) -> Result<Matrix<T>, Error> {
        let mut file = File::open(path)?;
        Matrix::read_from_file(&mut file)
    }
}

fn mul_helper<T: Parameter>(a: &Vector<T>, b: &Vector<T>, a_rows: usize, a_cols: usize, b_cols: usize) -> Matrix<T> {
    let res = unsafe { Matrix::uninitialized(a_rows, b_cols) };
    check(unsafe {
        ffi::la_mul_mat_mat(T::DTYPE, res.data.device_ptr(), a.device_ptr(), b.device_ptr(), a_rows, a_cols, b_cols, DEFAULT_STREAM)
    });
    wait();
    res
}

/// Outer product: `column` (n x 1) times `row` (1 x m).
pub fn mul_column_row<T: Parameter>(column: &Vector<T>, row: &Vector<T>) -> Matrix<T> {
    mul_helper(column, row, column.len(), 1, row.len())
}

/// `a * b` for row-major HOST operands, result on the host: uploads, multiply and download overlap
/// (la_mul_mat_mat_host).  Not in the reference; it replaces `Matrix::from_vec(a) * Matrix::from_vec(b)` + `to_vec()`.
pub fn mul_host<T: Parameter>(a: &[T], b: &[T], a_rows: usize, a_cols: usize, b_cols: usize) -> Vec<T> {
    assert_eq!(a.len(), a_rows * a_cols);
    assert_eq!(b.len(), a_cols * b_cols);
    let mut c = vec![T::default(); a_rows * b_cols];
    check(unsafe {
        ffi::la_mul_mat_mat_host(T::DTYPE, c.as_mut_ptr() as *mut c_void, a.as_ptr() as *const c_void, b.as_ptr() as *const c_void, a_rows, a_cols, b_cols)
    });
    c
}

impl<'a, 'b, T: Parameter> Mul<&'b Matrix<T>> for &'a Matrix<T> {
    type Output = Matrix<T>;
    fn mul(self, other: &'b Matrix<T>) -> Matrix<T> {
        assert_eq!(self.get_col_count(), other.get_row_count());
        mul_helper(&self.data, &other.data, self.row_count, self.col_count, other.col_count)
    }
}

macro_rules! matrix_elementwise {
    ($Op:ident, $op:ident, $OpAssign:ident, $op_assign:ident) => {
        impl<'a, 'b, T: Parameter> $Op<&'b Matrix<T>> for &'a Matrix<T> {
            type Output = Matrix<T>;
            fn $op(self, other: &'b Matrix<T>) -> Matrix<T> {
                self.same_shape(other);
                Matrix { data: (&self.data).$op(&other.data), row_count: self.row_count, col_count: self.col_count }
            }
        }
        impl<'a, T: Parameter> $OpAssign<&'a Matrix<T>> for Matrix<T> {
            fn $op_assign(&mut self, other: &'a Matrix<T>) {
                self.same_shape(other);
                self.data.$op_assign(&other.data);
            }
        }
        impl<'a, T: Parameter> $Op<&'a Matrix<T>> for Matrix<T> {
            type Output = Matrix<T>;
            fn $op(mut self, other: &'a Matrix<T>) -> Matrix<T> {
                self.$op_assign(other);
                self
            }
        }
    };
}
matrix_elementwise!(Add, add, AddAssign, add_assign);
matrix_elementwise!(Sub, sub, SubAssign, sub_assign);

impl<'a, T: Parameter> Mul<T> for &'a Matrix<T> {
    type Output = Matrix<T>;
    fn mul(self, s: T) -> Matrix<T> {
        Matrix { data: &self.data * s, row_count: self.row_count, col_count: self.col_count }
    }
}

impl<T: Parameter> MulAssign<T> for Matrix<T> {
    fn mul_assign(&mut self, s: T) {
        self.data *= s;
    }
}

impl<T: Parameter> Mul<T> for Matrix<T> {
    type Output = Matrix<T>;
    fn mul(mut self, s: T) -> Matrix<T> {
        self.data *= s;
        self
    }
}

impl<'a, T: Parameter> Div<T> for &'a Matrix<T> {
    type Output = Matrix<T>;
    fn div(self, s: T) -> Matrix<T> {
        Matrix { data: &self.data / s, row_count: self.row_count, col_count: self.col_count }
    }
}

impl<T: Parameter> PartialEq for Matrix<T> {
    fn eq(&self, other: &Matrix<T>) -> bool {
        self.len() == other.len() && self.data == other.data
    }
}

impl<T: Parameter + fmt::Debug> fmt::Debug for Matrix<T> {
    fn fmt(&self, f: &mut fmt::Formatter) -> fmt::Result {
        let v = self.data.to_vec();
        for r in 0..self.row_count {
            writeln!(f, "{:?}", &v[r * self.col_count..(r + 1) * self.col_count])?;
        }
        Ok(())
    }
}

impl<T: Parameter + fmt::Display> fmt::Display for Matrix<T> {
    fn fmt(&self, f: &mut fmt::Formatter) -> fmt::Result {
        let v = self.data.to_vec();
        for r in 0..self.row_count {
            let row: Vec<String> = v[r * self.col_count..(r + 1) * self.col_count].iter().map(|x| x.to_string()).collect();
            writeln!(f, "{}", row.join(" "))?;
        }
        Ok(())
    }
}

impl<T: Parameter> Clone for Matrix<T> {
    fn clone(&self) -> Matrix<T> {
        Matrix { data: self.data.clone(), row_count: self.row_count, col_count: self.col_count }
    }
}
