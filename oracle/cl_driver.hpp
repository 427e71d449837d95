// oracle/cl_driver.hpp -- TEST INFRASTRUCTURE ONLY.
//
// Host-side launchers for the reference kernels compiled through cl_shim.hpp.  Each
// launcher reproduces the NDRange the reference's Rust host code enqueues (cited per
// function) and nothing else.  Included AFTER the templated kernels.cl text, so the
// reference's un-#undef'd helper macros are dropped first.
#undef i
#undef lz
#undef kernel
#undef global
#undef local

// 1-D NDRange of `gws` work-items, no local size (vector.rs:105-109, matrix.rs:258-262).
#define CL_NDRANGE(gws, ...)                                             \
    _Pragma("omp parallel for schedule(static)")                           \
    for (long g_ = 0; g_ < (long)(gws); ++g_) {                            \
        cl_wi.gid = (size_t)g_; cl_wi.lid = 0; cl_wi.wgid = (size_t)g_;    \
        cl_wi.lsz = 1; cl_wi.gsz = (size_t)(gws);                          \
        __VA_ARGS__;                                                          \
    }

// gws = groups*W, lws = W, one `local T temp[W]` (vector.rs:272-288, 343-360; lib.rs:101-122).
// Work-groups run one after another; the W work-items of a group are W OpenMP threads so
// the kernel's own barrier() calls are real barriers.
#define CL_NDRANGE_GROUPS(T, groups, W, ...)                            \
    {                                                                      \
        T* temp = new T[(W)];                                              \
        for (long wg_ = 0; wg_ < (long)(groups); ++wg_) {                  \
            _Pragma("omp parallel num_threads(W)")                         \
            {                                                              \
                cl_wi.lid = (size_t)omp_get_thread_num();                  \
                cl_wi.wgid = (size_t)wg_; cl_wi.lsz = (size_t)(W);         \
                cl_wi.gsz = (size_t)(groups) * (size_t)(W);                \
                cl_wi.gid = cl_wi.wgid * cl_wi.lsz + cl_wi.lid;            \
                __VA_ARGS__;                                                  \
            }                                                              \
        }                                                                  \
        delete[] temp;                                                     \
    }

#define REF_WRAPPERS(T)                                                                        \
extern "C" void ref_##T##_ew_vec_vec(int op, T* C, T* A, T* B, long n) {                       \
    switch (op) {                                                                              \
    case 0: CL_NDRANGE(n, T##_add_vec_vec(C, A, B)) break;                                     \
    case 1: CL_NDRANGE(n, T##_sub_vec_vec(C, A, B)) break;                                     \
    case 2: CL_NDRANGE(n, T##_mul_vec_vec(C, A, B)) break;                                     \
    case 3: CL_NDRANGE(n, T##_div_vec_vec(C, A, B)) break; }                                   \
}                                                                                              \
extern "C" void ref_##T##_ew_assign_vec_vec(int op, T* C, T* B, long n) {                      \
    switch (op) {                                                                              \
    case 0: CL_NDRANGE(n, T##_add_assign_vec_vec(C, B)) break;                                 \
    case 1: CL_NDRANGE(n, T##_sub_assign_vec_vec(C, B)) break;                                 \
    case 2: CL_NDRANGE(n, T##_mul_assign_vec_vec(C, B)) break;                                 \
    case 3: CL_NDRANGE(n, T##_div_assign_vec_vec(C, B)) break; }                               \
}                                                                                              \
extern "C" void ref_##T##_ew_vec_scl(int op, T* C, T* A, const T* B, long n) {                 \
    if (op == 2) { CL_NDRANGE(n, T##_mul_vec_scl(C, A, *B)) }                                  \
    else         { CL_NDRANGE(n, T##_div_vec_scl(C, A, *B)) }                                  \
}                                                                                              \
extern "C" void ref_##T##_ew_assign_vec_scl(int op, T* C, const T* B, long n) {                \
    if (op == 2) { CL_NDRANGE(n, T##_mul_assign_vec_scl(C, *B)) }                              \
    else         { CL_NDRANGE(n, T##_div_assign_vec_scl(C, *B)) }                              \
}                                                                                              \
extern "C" void ref_##T##_eq_vec(uchar* C, T* A, T* B, long n) {                               \
    CL_NDRANGE(n, T##_eq_vec(C, A, B))                                                         \
}                                                                                              \
extern "C" void ref_##T##_sum_vec(const T* data, T* results, int count, int groups, int W) {   \
    CL_NDRANGE_GROUPS(T, groups, W, T##_sum_vec(data, results, count, temp))                   \
}                                                                                              \
extern "C" void ref_##T##_dot_vec_vec(const T* a, const T* b, T* results, int count,           \
                                      int groups, int W) {                                     \
    CL_NDRANGE_GROUPS(T, groups, W, T##_dot_vec_vec(a, b, results, count, temp))               \
}                                                                                              \
extern "C" void ref_##T##_mul_vec_mat(T* C, T* A, T* B, int B_col_count, int A_len) {          \
    CL_NDRANGE(B_col_count, T##_mul_vec_mat(C, A, B, B_col_count, A_len))                      \
}                                                                                              \
extern "C" void ref_##T##_mul_vec_transpose_mat(T* C, T* A, T* B, int B_col_count, int A_len,  \
                                                long B_row_count) {                            \
    CL_NDRANGE(B_row_count, T##_mul_vec_transpose_mat(C, A, B, B_col_count, A_len))            \
}                                                                                              \
extern "C" void ref_##T##_mul_mat_mat(T* C, T* A, T* B, int C_col_count, int A_col_count,      \
                                      long C_row_count) {                                      \
    CL_NDRANGE(C_row_count * (long)C_col_count,                                                \
               T##_mul_mat_mat(C, A, B, C_col_count, A_col_count))                             \
}
