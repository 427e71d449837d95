//! linear_algebra over liblinalg_b200: the public surface of usbalbin/linear_algebra with the OpenCL layer
//! (`ocl` crate, JIT-compiled src/kernels.cl, the process-global mutex of src/lib.rs:24-149 there) replaced by the
//! B200 library behind include/linalg_b200.h.  Kernels are compiled ahead of time and launches are stateless, so
//! there is no `cl_data()` / `get_kernels()` and no lock; every operator still returns a finished result.
//!
//! NOT COMPILED IN THE BUILD IMAGE (no rustc there).  cpp/linalg.hpp and linear_algebra_b200/*.py are the tested
//! mirrors; this crate is the binding a maintainer of the reference would start from.
pub mod ffi;
pub mod matrix;
pub mod traits;
pub mod util;
pub mod vector;

use std::ffi::CString;
use std::os::raw::c_void;
use std::ptr::null_mut;

use traits::Parameter;
use vector::Vector;

/// Binds the calling thread to a GPU (`la_init`); optional, the first operation does it lazily for device 0.
pub fn init(device: i32) {
    ffi::check(unsafe { ffi::la_init(device) });
}

/// A user kernel with its argument slots, declared by name in signature order: what `ocl::Kernel` is to the
/// reference's `create_kernel` (src/lib.rs:156-171 there) and to `Vector::generate / for_each_mut / map_mut / ...`.
pub struct Kernel {
    raw: *mut c_void,
    names: Vec<String>,
    slots: Vec<Vec<u8>>, // slot j holds the bytes of parameter j (a device pointer or a scalar)
}

/// Text of ./src/extra_kernels.cl, where the reference looks for user kernels (src/lib.rs:173-180 there).
pub fn load_extra_src() -> Result<String, std::io::Error> {
    std::fs::read_to_string("src/extra_kernels.cl")
}

/// Kernel `kernel_name` (full name, `{T}` already substituted: "float_my_kernel") from the user's
/// src/extra_kernels.cl.  The text keeps the conventions the reference documents: `{T}` is the element type,
/// `i` is get_global_id(0), functions are prefixed `{T}_`, `IS_<TYPE>` is defined for the type being compiled.
pub fn create_kernel<T: Parameter>(kernel_name: &str) -> Kernel {
    create_kernel_from_source::<T>(kernel_name, &load_extra_src().unwrap_or_default())
}

pub fn create_kernel_from_source<T: Parameter>(kernel_name: &str, source: &str) -> Kernel {
    let src = CString::new(source).expect("kernel source contains a NUL byte");
    let name = CString::new(kernel_name).expect("kernel name contains a NUL byte");
    let mut raw = null_mut();
    let status = unsafe { ffi::la_user_kernel_create(T::DTYPE, src.as_ptr(), name.as_ptr(), &mut raw) };
    if status != ffi::LA_OK {
        panic!("Failed to create kernel. Forgot to add kernel source to kernel.cl?: {}", ffi::last_error());
    }
    Kernel { raw, names: Vec::new(), slots: Vec::new() }
}

impl Kernel {
    pub fn arg_buf_named(mut self, name: &str) -> Kernel {
        self.names.push(name.to_owned());
        self.slots.push(vec![0u8; std::mem::size_of::<*mut c_void>()]);
        self
    }

    pub fn arg_scl_named<S: Copy>(mut self, name: &str, value: S) -> Kernel {
        self.names.push(name.to_owned());
        self.slots.push(vec![0u8; std::mem::size_of::<S>()]);
        self.set_arg_scl_named(name, value);
        self
    }

    fn store(&mut self, name: &str, bytes: &[u8]) {
        let j = self.names.iter().position(|n| n == name).unwrap_or_else(|| panic!("kernel has no argument named {}", name));
        assert_eq!(self.slots[j].len(), bytes.len(), "kernel argument {} has another size", name);
        self.slots[j].copy_from_slice(bytes);
    }

    pub fn set_arg_buf_named<T: Parameter>(&mut self, name: &str, vector: &Vector<T>) {
        let p = vector.device_ptr() as usize;
        self.store(name, &p.to_ne_bytes());
    }

    pub fn set_arg_scl_named<S: Copy>(&mut self, name: &str, value: S) {
        let bytes = unsafe { std::slice::from_raw_parts(&value as *const S as *const u8, std::mem::size_of::<S>()) };
        let copy = bytes.to_vec();
        self.store(name, &copy);
    }

    /// One work-item per index below `gws`, then wait (`kernel.cmd().gws(n).enq()` + `event.wait_for()`).
    pub fn enq(&mut self, gws: usize) {
        let mut argv: Vec<*mut c_void> = self.slots.iter_mut().map(|s| s.as_mut_ptr() as *mut c_void).collect();
        ffi::check(unsafe { ffi::la_user_kernel_launch(self.raw, gws, argv.as_mut_ptr(), ffi::DEFAULT_STREAM) });
        ffi::wait();
    }
}

impl Drop for Kernel {
    fn drop(&mut self) {
        unsafe { ffi::la_user_kernel_destroy(self.raw) };
    }
}
