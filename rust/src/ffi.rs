//! The C ABI of include/linalg_b200.h, declared by hand (no bindgen in the build image).  It stands where the
//! reference keeps its OpenCL state: `cl_data`, `get_kernels`, `Kernels`, `setup_kernels` (src/lib.rs:24-335 there).
#![allow(non_camel_case_types)]
use std::ffi::CStr;
use std::os::raw::{c_char, c_int, c_void};

pub const LA_OK: c_int = 0;
pub const ADD: c_int = 0;
pub const SUB: c_int = 1;
pub const MUL: c_int = 2;
pub const DIV: c_int = 3;
pub type Stream = *mut c_void;
/// `la_exchange_fn`: all-gather of `bytes` bytes per rank between the processes of a communicator (0 = success).
pub type ExchangeFn = unsafe extern "C" fn(user: *mut c_void, send: *const c_void, recv: *mut c_void, bytes: usize) -> c_int;
pub const DEFAULT_STREAM: Stream = 0 as Stream;

extern "C" {
    pub fn la_init(device: c_int) -> c_int;
    pub fn la_last_error() -> *const c_char;
    pub fn la_malloc(dev_ptr: *mut *mut c_void, bytes: usize) -> c_int;
    pub fn la_free(dev_ptr: *mut c_void) -> c_int;
    pub fn la_upload(dst_dev: *mut c_void, src_host: *const c_void, bytes: usize, stream: Stream) -> c_int;
    pub fn la_download(dst_host: *mut c_void, src_dev: *const c_void, bytes: usize, stream: Stream) -> c_int;
    pub fn la_copy(dst_dev: *mut c_void, src_dev: *const c_void, bytes: usize, stream: Stream) -> c_int;
    pub fn la_fill(dtype: c_int, dst_dev: *mut c_void, value_host: *const c_void, n: usize, stream: Stream) -> c_int;
    pub fn la_stream_sync(stream: Stream) -> c_int;
    pub fn la_ew_vec_vec(dtype: c_int, op: c_int, c: *mut c_void, a: *const c_void, b: *const c_void, n: usize, stream: Stream) -> c_int;
    pub fn la_ew_assign_vec_vec(dtype: c_int, op: c_int, c: *mut c_void, b: *const c_void, n: usize, stream: Stream) -> c_int;
    pub fn la_ew_vec_scl(dtype: c_int, op: c_int, c: *mut c_void, a: *const c_void, scalar_host: *const c_void, n: usize, stream: Stream) -> c_int;
    pub fn la_ew_assign_vec_scl(dtype: c_int, op: c_int, c: *mut c_void, scalar_host: *const c_void, n: usize, stream: Stream) -> c_int;
    pub fn la_eq_vec(dtype: c_int, a: *const c_void, b: *const c_void, n: usize, equal_host: *mut c_int, stream: Stream) -> c_int;
    pub fn la_sum_vec(dtype: c_int, data: *const c_void, n: usize, out_host: *mut c_void, stream: Stream) -> c_int;
    pub fn la_dot_vec_vec(dtype: c_int, a: *const c_void, b: *const c_void, n: usize, out_host: *mut c_void, stream: Stream) -> c_int;
    pub fn la_normalize_vec(dtype: c_int, c: *mut c_void, a: *const c_void, n: usize, stream: Stream) -> c_int;
    pub fn la_mul_vec_mat(dtype: c_int, c: *mut c_void, a: *const c_void, b: *const c_void, rows: usize, cols: usize, stream: Stream) -> c_int;
    pub fn la_mul_vec_transpose_mat(dtype: c_int, c: *mut c_void, a: *const c_void, b: *const c_void, rows: usize, cols: usize, stream: Stream) -> c_int;
    pub fn la_mul_mat_mat(dtype: c_int, c: *mut c_void, a: *const c_void, b: *const c_void, m: usize, k: usize, n: usize, stream: Stream) -> c_int;
    pub fn la_mul_mat_mat_host(dtype: c_int, c_host: *mut c_void, a_host: *const c_void, b_host: *const c_void, m: usize, k: usize, n: usize) -> c_int;
    pub fn la_gemm_prepare_b(dtype: c_int, b: *const c_void, k: usize, n: usize, stream: Stream, handle: *mut *mut c_void) -> c_int;
    pub fn la_mul_mat_mat_prepared(c: *mut c_void, a: *const c_void, handle: *const c_void, m: usize, stream: Stream) -> c_int;
    pub fn la_gemm_prepared_free(handle: *mut c_void) -> c_int;
    // multi-GPU: one process per GPU, row-block shards (csrc/comm.cu); `exchange` all-gathers `bytes` bytes per rank
    pub fn la_comm_create(world: c_int, rank: c_int, exchange: Option<ExchangeFn>, user: *mut c_void, comm: *mut *mut c_void) -> c_int;
    pub fn la_comm_destroy(comm: *mut c_void) -> c_int;
    pub fn la_comm_info(comm: *const c_void, world: *mut c_int, rank: *mut c_int, peer_memory: *mut c_int) -> c_int;
    pub fn la_comm_barrier(comm: *mut c_void) -> c_int;
    pub fn la_partition(n: usize, world: c_int, rank: c_int, lo: *mut usize, hi: *mut usize);
    pub fn la_sum_vec_sharded(comm: *mut c_void, dtype: c_int, data_local: *const c_void, n_local: usize, out_host: *mut c_void, stream: Stream) -> c_int;
    pub fn la_dot_vec_vec_sharded(comm: *mut c_void, dtype: c_int, a_local: *const c_void, b_local: *const c_void, n_local: usize, out_host: *mut c_void, stream: Stream) -> c_int;
    pub fn la_eq_vec_sharded(comm: *mut c_void, dtype: c_int, a_local: *const c_void, b_local: *const c_void, n_local: usize, equal_host: *mut c_int, stream: Stream) -> c_int;
    pub fn la_mul_vec_mat_sharded(comm: *mut c_void, dtype: c_int, c: *mut c_void, v_local: *const c_void, m_local: *const c_void, rows_local: usize, cols: usize, stream: Stream) -> c_int;
    pub fn la_transpose_sharded(comm: *mut c_void, dtype: c_int, out_local: *mut c_void, in_local: *const c_void, rows: usize, cols: usize, stream: Stream) -> c_int;
    pub fn la_mul_mat_mat_host_sharded(comm: *mut c_void, dtype: c_int, c_host_local: *mut c_void, a_host_local: *const c_void, b_host: *const c_void, m: usize, k: usize, n: usize) -> c_int;
    pub fn la_transpose(dtype: c_int, out: *mut c_void, input: *const c_void, rows: usize, cols: usize, stream: Stream) -> c_int;
    pub fn la_user_kernel_create(dtype: c_int, source: *const c_char, kernel_name: *const c_char, kernel: *mut *mut c_void) -> c_int;
    pub fn la_user_kernel_launch(kernel: *mut c_void, global_size: usize, arg_values: *mut *mut c_void, stream: Stream) -> c_int;
    pub fn la_user_kernel_destroy(kernel: *mut c_void) -> c_int;
}

pub fn last_error() -> String {
    unsafe { CStr::from_ptr(la_last_error()).to_string_lossy().into_owned() }
}

/// Every OpenCL call of the reference is followed by `.unwrap()`; a failing library call panics the same way.
pub fn check(status: c_int) {
    if status != LA_OK {
        panic!("liblinalg_b200: status {}: {}", status, last_error());
    }
}

/// `event.wait_for()` after every launch (src/vector.rs:108 of the reference): operators return finished results.
pub fn wait() {
    check(unsafe { la_stream_sync(DEFAULT_STREAM) });
}
