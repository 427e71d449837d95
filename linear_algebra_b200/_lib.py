"""ctypes binding of liblinalg_b200.so (the C ABI declared in include/linalg_b200.h).

There is no fallback of any kind: if the shared library is missing this module raises on
import, and without a CUDA device every compute call raises LaError (LA_ERR_CUDA)."""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liblinalg_b200.so")

LA_OK, LA_ERR_ARG, LA_ERR_SHAPE, LA_ERR_DTYPE, LA_ERR_OOM, LA_ERR_CUDA, LA_ERR_UNSUPPORTED = range(7)
ADD, SUB, MUL, DIV = 0, 1, 2, 3
GEMM_SIMT, GEMM_3XTF32, GEMM_AUTO = 0, 1, 2


class LaError(RuntimeError):
    def __init__(self, status, message):
        super().__init__("liblinalg_b200: status %d: %s" % (status, message))
        self.status = status


def build():
    """Compile the library in-tree with nvcc for sm_100a (works without a GPU)."""
    import subprocess
    subprocess.run(["make", "-C", os.path.join(HERE, "csrc"), "-j8"], check=True, stdout=subprocess.DEVNULL)


if not os.path.exists(LIB_PATH):
    raise ImportError(
        "%s not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` or "
        "`make -C linear_algebra_b200/csrc` (there is no CPU fallback)" % LIB_PATH)

lib = ctypes.CDLL(LIB_PATH)

_V, _Z, _I = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int
_PV, _PI = ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_int)

# name -> (argtypes); every one returns la_status except the three noted below.
SIGNATURES = {
    "la_init": [_I],
    "la_device_count": [_PI],
    "la_device_name": [ctypes.c_char_p, _Z],
    "la_sm_count": [_PI],
    "la_malloc": [_PV, _Z],
    "la_free": [_V],
    "la_upload": [_V, _V, _Z, _V],
    "la_download": [_V, _V, _Z, _V],
    "la_copy": [_V, _V, _Z, _V],
    "la_copy_2d": [_V, _Z, _V, _Z, _Z, _Z, _V],
    "la_fill": [_I, _V, _V, _Z, _V],
    "la_synth": [_I, _V, ctypes.c_uint64, _Z, _V],
    "la_host_alloc": [_PV, _Z],
    "la_host_free": [_V],
    "la_stream_create": [_PV],
    "la_stream_destroy": [_V],
    "la_stream_sync": [_V],
    "la_event_create": [_PV],
    "la_event_destroy": [_V],
    "la_event_record": [_V, _V],
    "la_event_elapsed_ms": [_V, _V, ctypes.POINTER(ctypes.c_float)],
    "la_ew_vec_vec": [_I, _I, _V, _V, _V, _Z, _V],
    "la_ew_assign_vec_vec": [_I, _I, _V, _V, _Z, _V],
    "la_ew_vec_scl": [_I, _I, _V, _V, _V, _Z, _V],
    "la_ew_assign_vec_scl": [_I, _I, _V, _V, _Z, _V],
    "la_eq_vec": [_I, _V, _V, _Z, _PI, _V],
    "la_sum_vec": [_I, _V, _Z, _V, _V],
    "la_dot_vec_vec": [_I, _V, _V, _Z, _V, _V],
    "la_sum_vec_dev": [_I, _V, _Z, _V, _V],
    "la_dot_vec_vec_dev": [_I, _V, _V, _Z, _V, _V],
    "la_normalize_vec": [_I, _V, _V, _Z, _V],
    "la_p2p_buffer_create": [_Z, _PV, _V],
    "la_p2p_buffer_open": [_V, _PV],
    "la_p2p_buffer_close": [_V],
    "la_p2p_buffer_destroy": [_V],
    "la_vec_allreduce": [_I, _V, _Z, _PV, _Z, _PV, _I, _I, ctypes.c_uint64, _V],
    "la_sum_vec_allreduce": [_I, _V, _Z, _V, _PV, _I, _I, ctypes.c_uint64, _V],
    "la_dot_vec_vec_allreduce": [_I, _V, _V, _Z, _V, _PV, _I, _I, ctypes.c_uint64, _V],
    "la_comm_create": [_I, _I, _V, _V, _PV],
    "la_comm_destroy": [_V],
    "la_comm_info": [_V, _PI, _PI, _PI],
    "la_comm_barrier": [_V],
    "la_sum_vec_sharded": [_V, _I, _V, _Z, _V, _V],
    "la_dot_vec_vec_sharded": [_V, _I, _V, _V, _Z, _V, _V],
    "la_eq_vec_sharded": [_V, _I, _V, _V, _Z, _PI, _V],
    "la_mul_vec_mat_sharded": [_V, _I, _V, _V, _V, _Z, _Z, _V],
    "la_transpose_sharded": [_V, _I, _V, _V, _Z, _Z, _V],
    "la_mul_mat_mat_host_sharded": [_V, _I, _V, _V, _V, _Z, _Z, _Z],
    "la_mul_vec_mat": [_I, _V, _V, _V, _Z, _Z, _V],
    "la_mul_vec_transpose_mat": [_I, _V, _V, _V, _Z, _Z, _V],
    "la_mul_vec_mat_noalloc": [_I, _V, _V, _V, _Z, _Z, _V],
    "la_mul_mat_mat": [_I, _V, _V, _V, _Z, _Z, _Z, _V],
    "la_mul_mat_mat_host": [_I, _V, _V, _V, _Z, _Z, _Z],
    "la_mul_mat_mat_info": [_I, _V, _V, _V, _Z, _Z, _Z, _V, _V],
    "la_gemm_prepare_b": [_I, _V, _Z, _Z, _V, _PV],
    "la_mul_mat_mat_prepared": [_V, _V, _V, _Z, _V],
    "la_gemm_prepared_free": [_V],
    "la_set_gemm_mode": [_I],
    "la_get_gemm_mode": [_PI],
    "la_transpose": [_I, _V, _V, _Z, _Z, _V],
    "la_user_kernel_compile": [_I, ctypes.c_char_p, ctypes.POINTER(ctypes.c_size_t)],
    "la_user_kernel_create": [_I, ctypes.c_char_p, ctypes.c_char_p, _PV],
    "la_user_kernel_launch": [_V, _Z, _PV, _V],
    "la_user_kernel_destroy": [_V],
}
EXCHANGE_FN = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t)
NON_STATUS = {"la_partition": None, "la_last_error": ctypes.c_char_p, "la_version": ctypes.c_char_p, "la_dtype_size": ctypes.c_size_t,
              "la_p2p_mailbox_bytes": ctypes.c_size_t, "la_p2p_handle_bytes": ctypes.c_size_t,
              "la_user_kernel_build_log": ctypes.c_char_p, "la_user_kernel_type_name": ctypes.c_char_p}

for _name, _args in SIGNATURES.items():
    _f = getattr(lib, _name)
    _f.argtypes = _args
    _f.restype = ctypes.c_int
lib.la_partition.argtypes = [_Z, _I, _I, ctypes.POINTER(_Z), ctypes.POINTER(_Z)]
lib.la_partition.restype = None
lib.la_last_error.argtypes = []
lib.la_last_error.restype = ctypes.c_char_p
lib.la_version.argtypes = []
lib.la_version.restype = ctypes.c_char_p
lib.la_p2p_mailbox_bytes.argtypes = []
lib.la_p2p_mailbox_bytes.restype = ctypes.c_size_t
lib.la_p2p_handle_bytes.argtypes = []
lib.la_p2p_handle_bytes.restype = ctypes.c_size_t
lib.la_user_kernel_build_log.argtypes = []
lib.la_user_kernel_build_log.restype = ctypes.c_char_p
lib.la_user_kernel_type_name.argtypes = [_I]
lib.la_user_kernel_type_name.restype = ctypes.c_char_p
lib.la_dtype_size.argtypes = [_I]
lib.la_dtype_size.restype = ctypes.c_size_t
# tuning / introspection hooks (include/linalg_b200_tuning.h)
TUNING = {
    "la_gemm_set_variant": [_I], "la_gemm_tc_set_variant": [_I], "la_gemm_tc_set_persistent": [_I], "la_gemm_tc_set_chunk": [_I],
    "la_gemm_tc_set_cross": [_I], "la_gemm_tc_set_a_direct": [_I], "la_gemm_tc_set_f16": [_I], "la_gemm_tc_set_tile_n": [_I], "la_gemm_set_stream_tail": [_I], "la_gemm_set_raster": [_I],
    "la_gemm_kernel_timing": [_I], "la_gemm_kernel_time_ms": [ctypes.POINTER(ctypes.c_float), _PI],
    "la_gemm_last_engine": [], "la_gemm_last_padded": [], "la_gemm_tc_last_form": [], "la_gemm_tc_last_launches": [],
    "la_transpose_set_variant": [_I], "la_transpose_set_band": [_I], "la_transpose_set_hints": [_I], "la_transpose_set_mode": [_I],
}
for _name, _args in TUNING.items():
    _f = getattr(lib, _name)
    _f.argtypes = _args
    _f.restype = ctypes.c_int


def check(status):
    if status != LA_OK:
        raise LaError(status, lib.la_last_error().decode("utf-8", "replace"))


def call(name, *args):
    check(getattr(lib, name)(*args))
