/* oracle/la_oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement, in plain C, of the arithmetic of usbalbin/linear_algebra's hot path: every
 * kernel of /root/reference/src/kernels.cl plus the two host-side reductions of
 * /root/reference/src/vector.rs that finish `sum` and `dot`.  Each function cites the lines it
 * follows.  It is the checker for the CUDA path; only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load it.  The product library never does.
 *
 * Parity is PINNED: tests/test_oracle.py checks this file against (1) every golden vector of the
 * reference's own src/tests.rs for this path and (2) the reference's kernels.cl itself, compiled
 * verbatim as C++ into oracle/_ref (oracle/build_ref.py), on randomized inputs for every dtype.
 *
 * Conventions
 *  - dtype / op codes are those of include/linalg_b200.h.
 *  - Integer arithmetic wraps mod 2^n (computed in the unsigned type of the same width; that is
 *    what every OpenCL device does, and it makes sum/dot/GEMM order-independent).
 *  - Floating point follows the reference's ORDER exactly: serial k-ascending dot products
 *    (kernels.cl:134-140), strided per-work-item partial sums + power-of-two tree + serial host
 *    sum for sum/dot (kernels.cl:83-124, vector.rs:290,361).  Built with -ffp-contract=off; the
 *    *_fma entry points use fma() to bracket OpenCL's permitted contraction.
 *  - `transpose` has no counterpart in the reference; its oracle is the
 *    definition out[c*R + r] = in[r*C + c].
 */
#include <stdint.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

enum { LA_U8 = 0, LA_I8, LA_U16, LA_I16, LA_U32, LA_I32, LA_U64, LA_I64, LA_F32, LA_F64 };
enum { LA_ADD = 0, LA_SUB, LA_MUL, LA_DIV };
#define WORK_GROUP_COUNT 64 /* lib.rs:114 */

/* T = element type, U = type the wrap-safe arithmetic is carried out in. */
#define FOR_EACH_TYPE(X)      \
    X(U8, u8, uint8_t, uint32_t)   \
    X(I8, i8, int8_t, uint32_t)    \
    X(U16, u16, uint16_t, uint32_t)\
    X(I16, i16, int16_t, uint32_t) \
    X(U32, u32, uint32_t, uint32_t)\
    X(I32, i32, int32_t, uint32_t) \
    X(U64, u64, uint64_t, uint64_t)\
    X(I64, i64, int64_t, uint64_t) \
    X(F32, f32, float, float)      \
    X(F64, f64, double, double)

#define DEFINE_KERNELS(TAG, N, T, U)                                                                  \
static inline T N##_add(T a, T b) { return (T)((U)a + (U)b); }                                    \
static inline T N##_sub(T a, T b) { return (T)((U)a - (U)b); }                                    \
static inline T N##_mul(T a, T b) { return (T)((U)a * (U)b); }                                    \
static inline T N##_div(T a, T b) { return (T)(a / b); }                                          \
static inline T N##_apply(int op, T a, T b) {                                                     \
    switch (op) { case LA_ADD: return N##_add(a, b); case LA_SUB: return N##_sub(a, b);           \
                  case LA_MUL: return N##_mul(a, b); default: return N##_div(a, b); }             \
}                                                                                                 \
/* kernels.cl:15-29 ({T}_add/sub/mul/div_vec_vec), one work-item per element */                   \
static void N##_ew_vec_vec(int op, T* C, const T* A, const T* B, long n) {                        \
    _Pragma("omp parallel for schedule(static)")                                                  \
    for (long g = 0; g < n; ++g) C[g] = N##_apply(op, A[g], B[g]);                                \
}                                                                                                 \
/* kernels.cl:33-47 ({T}_*_assign_vec_vec) */                                                     \
static void N##_ew_assign_vec_vec(int op, T* C, const T* B, long n) {                             \
    _Pragma("omp parallel for schedule(static)")                                                  \
    for (long g = 0; g < n; ++g) C[g] = N##_apply(op, C[g], B[g]);                                \
}                                                                                                 \
/* kernels.cl:51-57 ({T}_mul_vec_scl, {T}_div_vec_scl) */                                         \
static void N##_ew_vec_scl(int op, T* C, const T* A, T s, long n) {                               \
    _Pragma("omp parallel for schedule(static)")                                                  \
    for (long g = 0; g < n; ++g) C[g] = N##_apply(op, A[g], s);                                   \
}                                                                                                 \
/* kernels.cl:61-67 ({T}_mul_assign_vec_scl, {T}_div_assign_vec_scl) */                           \
static void N##_ew_assign_vec_scl(int op, T* C, T s, long n) {                                    \
    _Pragma("omp parallel for schedule(static)")                                                  \
    for (long g = 0; g < n; ++g) C[g] = N##_apply(op, C[g], s);                                   \
}                                                                                                 \
/* kernels.cl:71-74 ({T}_eq_vec): flag pre-set to 1 by vector.rs:658-661, cleared on mismatch */  \
static int N##_eq_vec(const T* A, const T* B, long n) {                                           \
    int flag = 1;                                                                                 \
    _Pragma("omp parallel for schedule(static) reduction(&& : flag)")                             \
    for (long g = 0; g < n; ++g) if (A[g] != B[g]) flag = 0;                                      \
    return flag;                                                                                  \
}                                                                                                 \
/* kernels.cl:83-102 ({T}_sum_vec) / :105-124 ({T}_dot_vec_vec) by loop fission, launched with   \
 * gws = 64*W, lws = W (vector.rs:282-287, lib.rs:114-115); b == NULL selects sum.  With `fm`    \
 * the multiply-add of :108 is contracted.  Writes the 64 per-group results. */                  \
static void N##_reduce_groups(const T* a, const T* b, T* results, long count, int W, int fm) {    \
    long gz = (long)WORK_GROUP_COUNT * W;                                                         \
    _Pragma("omp parallel for schedule(static)")                                                  \
    for (long wg = 0; wg < WORK_GROUP_COUNT; ++wg) {                                              \
        T* temp = (T*)malloc(sizeof(T) * (size_t)W);                                              \
        for (long lid = 0; lid < W; ++lid) {          /* :84-90 / :106-112 */                     \
            T value = 0;                                                                          \
            for (long k = wg * W + lid; k < count; k += gz) {                                     \
                if (!b)      value = N##_add(value, a[k]);                                        \
                else if (fm) value = N##_fma(a[k], b[k], value);                                  \
                else         value = N##_add(value, N##_mul(a[k], b[k]));                         \
            }                                                                                     \
            temp[lid] = value;                                                                    \
        }                                                                                         \
        for (long offset = W / 2; offset > 0; offset /= 2)   /* :93-97 / :115-119 */              \
            for (long lid = 0; lid < offset; ++lid)                                               \
                temp[lid] = N##_add(temp[lid], temp[lid + offset]);                               \
        results[wg] = temp[0];                        /* :99-101 / :121-123 */                    \
        free(temp);                                                                               \
    }                                                                                             \
}                                                                                                 \
/* Vector::sum vector.rs:256-293 and dot vector.rs:326-363: host path when len <= 64             \
 * (:268-270, :338-340), else kernel + in-order host sum of the 64 partials (:290, :361). */     \
static T N##_reduce(const T* a, const T* b, long n, int W, int fm) {                              \
    T res = 0;                                                                                    \
    if (n <= WORK_GROUP_COUNT) {                                                                  \
        for (long k = 0; k < n; ++k)                                                              \
            res = N##_add(res, b ? N##_mul(a[k], b[k]) : a[k]);                                   \
        return res;                                                                               \
    }                                                                                             \
    T partial[WORK_GROUP_COUNT];                                                                  \
    N##_reduce_groups(a, b, partial, n, W, fm);                                                   \
    for (int k = 0; k < WORK_GROUP_COUNT; ++k) res = N##_add(res, partial[k]);                    \
    return res;                                                                                   \
}                                                                                                 \
/* kernels.cl:134-140 ({T}_dot_product): serial, j ascending, accumulate in T */                  \
static inline T N##_dot_product(const T* a, long sa, const T* b, long sb, long count, int fm) {   \
    T res = 0;                                                                                    \
    for (long j = 0; j < count; ++j)                                                              \
        res = fm ? N##_fma(a[j * sa], b[j * sb], res) : N##_add(res, N##_mul(a[j * sa], b[j * sb])); \
    return res;                                                                                   \
}                                                                                                 \
/* kernels.cl:143-147 ({T}_mul_vec_mat): C[col] = A . B[:,col]; gws = cols (vector.rs:532) */      \
static void N##_mul_vec_mat(T* C, const T* A, const T* B, long rows, long cols, int fm) {         \
    _Pragma("omp parallel for schedule(static)")                                                  \
    for (long col = 0; col < cols; ++col) C[col] = N##_dot_product(A, 1, &B[col], cols, rows, fm);\
}                                                                                                 \
/* kernels.cl:150-152 ({T}_mul_vec_transpose_mat): C[r] = A . B[r,:]; gws = rows (vector.rs:565) */\
static void N##_mul_vec_transpose_mat(T* C, const T* A, const T* B, long rows, long cols, int fm) {\
    _Pragma("omp parallel for schedule(static)")                                                  \
    for (long r = 0; r < rows; ++r) C[r] = N##_dot_product(A, 1, &B[r * cols], 1, cols, fm);      \
}                                                                                                 \
/* kernels.cl:161-169 ({T}_mul_mat_mat): row = i / N, col = i % N, gws = M*N (matrix.rs:260) */    \
static void N##_mul_mat_mat(T* C, const T* A, const T* B, long M, long K, long Nc, int fm) {      \
    _Pragma("omp parallel for schedule(static)")                                                  \
    for (long g = 0; g < M * Nc; ++g) {                                                           \
        long row = g / Nc, col = g % Nc;                                                          \
        C[g] = N##_dot_product(&A[K * row], 1, &B[col], Nc, K, fm);                               \
    }                                                                                             \
}                                                                                                 \
/* transpose: not in the reference; definition out[c*R + r] = in[r*C + c] */                      \
static void N##_transpose(T* out, const T* in, long R, long Cc) {                                 \
    _Pragma("omp parallel for schedule(static)")                                                  \
    for (long r = 0; r < R; ++r) for (long c = 0; c < Cc; ++c) out[c * R + r] = in[r * Cc + c];   \
}

/* fma helper per type: integers have no contraction question. */
#define DEFINE_FMA_INT(N, T, U) static inline T N##_fma(T a, T b, T c) { return (T)((U)a * (U)b + (U)c); }
DEFINE_FMA_INT(u8, uint8_t, uint32_t) DEFINE_FMA_INT(i8, int8_t, uint32_t)
DEFINE_FMA_INT(u16, uint16_t, uint32_t) DEFINE_FMA_INT(i16, int16_t, uint32_t)
DEFINE_FMA_INT(u32, uint32_t, uint32_t) DEFINE_FMA_INT(i32, int32_t, uint32_t)
DEFINE_FMA_INT(u64, uint64_t, uint64_t) DEFINE_FMA_INT(i64, int64_t, uint64_t)
static inline float f32_fma(float a, float b, float c) { return fmaf(a, b, c); }
static inline double f64_fma(double a, double b, double c) { return fma(a, b, c); }

FOR_EACH_TYPE(DEFINE_KERNELS)

/* ------------------------------------------------------------------ exported, dtype-dispatched */

int oracle_ew_vec_vec(int dtype, int op, void* C, const void* A, const void* B, long n) {
    switch (dtype) {
#define X(TAG, N, T, U) case LA_##TAG: N##_ew_vec_vec(op, (T*)C, (const T*)A, (const T*)B, n); return 0;
        FOR_EACH_TYPE(X)
#undef X
    }
    return -1;
}
int oracle_ew_assign_vec_vec(int dtype, int op, void* C, const void* B, long n) {
    switch (dtype) {
#define X(TAG, N, T, U) case LA_##TAG: N##_ew_assign_vec_vec(op, (T*)C, (const T*)B, n); return 0;
        FOR_EACH_TYPE(X)
#undef X
    }
    return -1;
}
int oracle_ew_vec_scl(int dtype, int op, void* C, const void* A, const void* s, long n) {
    switch (dtype) {
#define X(TAG, N, T, U) case LA_##TAG: N##_ew_vec_scl(op, (T*)C, (const T*)A, *(const T*)s, n); return 0;
        FOR_EACH_TYPE(X)
#undef X
    }
    return -1;
}
int oracle_ew_assign_vec_scl(int dtype, int op, void* C, const void* s, long n) {
    switch (dtype) {
#define X(TAG, N, T, U) case LA_##TAG: N##_ew_assign_vec_scl(op, (T*)C, *(const T*)s, n); return 0;
        FOR_EACH_TYPE(X)
#undef X
    }
    return -1;
}
/* returns 1 (equal) / 0 (not equal) / -1 (bad dtype) */
int oracle_eq_vec(int dtype, const void* A, const void* B, long n) {
    switch (dtype) {
#define X(TAG, N, T, U) case LA_##TAG: return N##_eq_vec((const T*)A, (const T*)B, n);
        FOR_EACH_TYPE(X)
#undef X
    }
    return -1;
}
/* The 64 per-work-group results of the device kernel alone (b == NULL: sum_vec). */
int oracle_reduce_groups(int dtype, const void* a, const void* b, void* results64, long n, int W, int fm) {
    switch (dtype) {
#define X(TAG, N, T, U) case LA_##TAG: N##_reduce_groups((const T*)a, (const T*)b, (T*)results64, n, W, fm); return 0;
        FOR_EACH_TYPE(X)
#undef X
    }
    return -1;
}
/* Vector::sum (b == NULL) / dot, end to end, W = work-group size the device would report. */
int oracle_reduce(int dtype, const void* a, const void* b, long n, int W, int fm, void* out) {
    switch (dtype) {
#define X(TAG, N, T, U) case LA_##TAG: *(T*)out = N##_reduce((const T*)a, (const T*)b, n, W, fm); return 0;
        FOR_EACH_TYPE(X)
#undef X
    }
    return -1;
}
int oracle_mul_vec_mat(int dtype, void* C, const void* A, const void* B, long rows, long cols, int fm) {
    switch (dtype) {
#define X(TAG, N, T, U) case LA_##TAG: N##_mul_vec_mat((T*)C, (const T*)A, (const T*)B, rows, cols, fm); return 0;
        FOR_EACH_TYPE(X)
#undef X
    }
    return -1;
}
int oracle_mul_vec_transpose_mat(int dtype, void* C, const void* A, const void* B, long rows, long cols, int fm) {
    switch (dtype) {
#define X(TAG, N, T, U) case LA_##TAG: N##_mul_vec_transpose_mat((T*)C, (const T*)A, (const T*)B, rows, cols, fm); return 0;
        FOR_EACH_TYPE(X)
#undef X
    }
    return -1;
}
int oracle_mul_mat_mat(int dtype, void* C, const void* A, const void* B, long M, long K, long N, int fm) {
    switch (dtype) {
#define X(TAG, N_, T, U) case LA_##TAG: N_##_mul_mat_mat((T*)C, (const T*)A, (const T*)B, M, K, N, fm); return 0;
        FOR_EACH_TYPE(X)
#undef X
    }
    return -1;
}
int oracle_transpose(int dtype, void* out, const void* in, long R, long C) {
    switch (dtype) {
#define X(TAG, N, T, U) case LA_##TAG: N##_transpose((T*)out, (const T*)in, R, C); return 0;
        FOR_EACH_TYPE(X)
#undef X
    }
    return -1;
}

/* ------------------------------------------------ f64-accumulated "truth" for f32 error bounds
 * (e(x) = |x - x64| / sum_k |a_k b_k|).  Not reference behaviour. */
void oracle_f32_dot_truth(const float* a, long sa, const float* b, long sb, long n, double* dot, double* absdot) {
    double s = 0.0, t = 0.0;
    for (long k = 0; k < n; ++k) { double p = (double)a[k * sa] * (double)b[k * sb]; s += p; t += fabs(p); }
    *dot = s; *absdot = t;
}
/* C64[M*N] = A*B in f64, ABS[M*N] = |A|*|B| */
void oracle_f32_mul_mat_mat_truth(double* C64, double* ABS, const float* A, const float* B, long M, long K, long N) {
    #pragma omp parallel for schedule(static)
    for (long r = 0; r < M; ++r) {
        double* c = C64 + r * N; double* t = ABS + r * N;
        for (long j = 0; j < N; ++j) { c[j] = 0.0; t[j] = 0.0; }
        for (long k = 0; k < K; ++k) {
            double a = (double)A[r * K + k];
            const float* brow = B + k * N;
            for (long j = 0; j < N; ++j) { double p = a * (double)brow[j]; c[j] += p; t[j] += fabs(p); }
        }
    }
}
