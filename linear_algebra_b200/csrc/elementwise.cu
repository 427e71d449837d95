// csrc/elementwise.cu -- the elementwise family, equality and fill.
//
// Replaces (src/kernels.cl line ranges): {T}_add/sub/mul/div_vec_vec :15-29,
// {T}_*_assign_vec_vec :33-47, {T}_mul/div_vec_scl :51-57, {T}_mul/div_assign_vec_scl :61-67,
// {T}_eq_vec :71-74.  The reference runs one work-item per element with scalar accesses; here
// every thread moves EW_UNROLL 16-byte packets per operand (all loads issued before the first
// use), through streaming (evict-first) accesses, since each byte is touched exactly once.
//
// Roofline: HBM.  Algorithmic bytes per element: 3*sizeof(T) (vec op vec, either form),
// 2*sizeof(T) (vec op scalar, either form; eq).
#include "common.cuh"

namespace {

constexpr int EW_THREADS = 256;
constexpr int EW_UNROLL = 4;

// C[i] = A[i] op (SCALAR ? s : B[i]).  C may alias A (the *_assign forms): every thread reads
// its own elements before it writes them and no element is shared between threads.
template <typename T, int OP, bool SCALAR>
__global__ void __launch_bounds__(EW_THREADS)
ew_pack_kernel(T* C, const T* A, const T* B, T s, size_t n) {
    using P = Pack<T>;
    constexpr int N = P::N;
    const size_t n_pack = n / N;
    const size_t base = (size_t)blockIdx.x * (EW_THREADS * EW_UNROLL) + threadIdx.x;
    P* Cp = reinterpret_cast<P*>(C);
    const P* Ap = reinterpret_cast<const P*>(A);
    const P* Bp = reinterpret_cast<const P*>(B);

    if ((size_t)(blockIdx.x + 1) * (EW_THREADS * EW_UNROLL) <= n_pack) {
        P a[EW_UNROLL], b[EW_UNROLL];
#pragma unroll
        for (int j = 0; j < EW_UNROLL; ++j) {
            a[j] = ld_stream(Ap + base + (size_t)j * EW_THREADS);
            if (!SCALAR) b[j] = ld_stream(Bp + base + (size_t)j * EW_THREADS);
        }
#pragma unroll
        for (int j = 0; j < EW_UNROLL; ++j) {
            P c;
#pragma unroll
            for (int e = 0; e < N; ++e) c.v[e] = ApplyOp<T, OP>::apply(a[j].v[e], SCALAR ? s : b[j].v[e]);
            st_stream(Cp + base + (size_t)j * EW_THREADS, c);
        }
    } else {
#pragma unroll
        for (int j = 0; j < EW_UNROLL; ++j) {
            const size_t p = base + (size_t)j * EW_THREADS;
            if (p < n_pack) {
                P a = ld_stream(Ap + p), b, c;
                if (!SCALAR) b = ld_stream(Bp + p);
#pragma unroll
                for (int e = 0; e < N; ++e) c.v[e] = ApplyOp<T, OP>::apply(a.v[e], SCALAR ? s : b.v[e]);
                st_stream(Cp + p, c);
            }
        }
        // ragged tail: fewer than N elements after the last whole packet
        if (blockIdx.x == gridDim.x - 1) {
            const size_t i = n_pack * N + threadIdx.x;
            if (i < n) C[i] = ApplyOp<T, OP>::apply(A[i], SCALAR ? s : B[i]);
        }
    }
}

// Any alignment: one element per thread per step, still fully coalesced.
template <typename T, int OP, bool SCALAR>
__global__ void __launch_bounds__(EW_THREADS)
ew_elem_kernel(T* C, const T* A, const T* B, T s, size_t n) {
    const size_t stride = (size_t)gridDim.x * EW_THREADS;
    for (size_t i = (size_t)blockIdx.x * EW_THREADS + threadIdx.x; i < n; i += stride)
        C[i] = ApplyOp<T, OP>::apply(A[i], SCALAR ? s : B[i]);
}

template <typename T, int OP, bool SCALAR>
la_status ew_launch(T* C, const T* A, const T* B, T s, size_t n, cudaStream_t st, int sm_count) {
    if (n == 0) return LA_OK;
    const bool aligned = la_aligned16(C) && la_aligned16(A) && (SCALAR || la_aligned16(B));
    if (aligned) {
        const size_t n_pack = n / Pack<T>::N;
        // +1 block when the packets divide evenly but a ragged tail remains
        size_t blocks = (n_pack + EW_THREADS * EW_UNROLL - 1) / (EW_THREADS * EW_UNROLL);
        if (blocks * (size_t)(EW_THREADS * EW_UNROLL) == n_pack && n_pack * Pack<T>::N != n) blocks += 1;
        if (blocks == 0) blocks = 1;
        if (blocks > 0x7fffffffull) return la_fail(LA_ERR_UNSUPPORTED, "elementwise: %zu elements exceed the grid limit", n);
        ew_pack_kernel<T, OP, SCALAR><<<(unsigned)blocks, EW_THREADS, 0, st>>>(C, A, B, s, n);
    } else {
        size_t blocks = (n + EW_THREADS - 1) / EW_THREADS;
        const size_t cap = (size_t)sm_count * 32;
        if (blocks > cap) blocks = cap;
        ew_elem_kernel<T, OP, SCALAR><<<(unsigned)blocks, EW_THREADS, 0, st>>>(C, A, B, s, n);
    }
    LA_LAUNCH_CHECK();
    return LA_OK;
}

template <typename T, bool SCALAR>
la_status ew_op_dispatch(int op, T* C, const T* A, const T* B, T s, size_t n, cudaStream_t st, int sm) {
    switch (op) {
    case LA_ADD: if (SCALAR) break; return ew_launch<T, LA_ADD, SCALAR>(C, A, B, s, n, st, sm);
    case LA_SUB: if (SCALAR) break; return ew_launch<T, LA_SUB, SCALAR>(C, A, B, s, n, st, sm);
    case LA_MUL: return ew_launch<T, LA_MUL, SCALAR>(C, A, B, s, n, st, sm);
    case LA_DIV: return ew_launch<T, LA_DIV, SCALAR>(C, A, B, s, n, st, sm);
    default: return la_fail(LA_ERR_ARG, "unknown op %d", op);
    }
    // the reference has only mul/div by a scalar (kernels.cl:51-67)
    return la_fail(LA_ERR_UNSUPPORTED, "vec-scalar op %d: only LA_MUL and LA_DIV exist", op);
}

// ---------------------------------------------------------------- eq
// flag[0] becomes non-zero if any A[i] != B[i].  Integers compare packet-wise on the raw words.
template <typename T>
__global__ void __launch_bounds__(EW_THREADS)
eq_kernel(const T* A, const T* B, size_t n, unsigned* flag, bool packed) {
    using P = Pack<T>;
    constexpr int N = P::N;
    bool diff = false;
    const size_t stride = (size_t)gridDim.x * EW_THREADS;
    const size_t tid = (size_t)blockIdx.x * EW_THREADS + threadIdx.x;
    size_t done = 0;
    if (packed) {
        const size_t n_pack = n / N;
        const P* Ap = reinterpret_cast<const P*>(A);
        const P* Bp = reinterpret_cast<const P*>(B);
        size_t p = tid;
        for (; p + (EW_UNROLL - 1) * stride < n_pack; p += EW_UNROLL * stride) {
            P a[EW_UNROLL], b[EW_UNROLL];
#pragma unroll
            for (int j = 0; j < EW_UNROLL; ++j) { a[j] = ld_stream(Ap + p + j * stride); b[j] = ld_stream(Bp + p + j * stride); }
#pragma unroll
            for (int j = 0; j < EW_UNROLL; ++j)
#pragma unroll
                for (int e = 0; e < N; ++e) diff |= Arith<T>::ne(a[j].v[e], b[j].v[e]);
        }
        for (; p < n_pack; p += stride) {
            P a = ld_stream(Ap + p), b = ld_stream(Bp + p);
#pragma unroll
            for (int e = 0; e < N; ++e) diff |= Arith<T>::ne(a.v[e], b.v[e]);
        }
        done = n_pack * N;
    }
    for (size_t i = done + tid; i < n; i += stride) diff |= Arith<T>::ne(A[i], B[i]);
    if (__syncthreads_or(diff) && threadIdx.x == 0) atomicOr(flag, 1u);
}

template <typename T>
__global__ void __launch_bounds__(EW_THREADS) fill_kernel(T* dst, T value, size_t n) {
    const size_t stride = (size_t)gridDim.x * EW_THREADS;
    for (size_t i = (size_t)blockIdx.x * EW_THREADS + threadIdx.x; i < n; i += stride) dst[i] = value;
}

// Synthetic operands generated on the device: counter-based, element i
// depends only on (seed, i), so the host can reproduce any element without moving the array.
__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
template <typename T> __device__ __forceinline__ T synth_value(uint64_t bits);
template <> __device__ __forceinline__ float synth_value<float>(uint64_t bits) { return (float)(bits >> 40) * (2.0f / 16777216.0f) - 1.0f; }       // U[-1,1), 24 bits
template <> __device__ __forceinline__ double synth_value<double>(uint64_t bits) { return (double)(bits >> 11) * (2.0 / 9007199254740992.0) - 1.0; }
#define LA_SYNTH_INT(T) template <> __device__ __forceinline__ T synth_value<T>(uint64_t bits) { return (T)(bits >> (64 - 8 * sizeof(T))); }
LA_SYNTH_INT(uint8_t) LA_SYNTH_INT(int8_t) LA_SYNTH_INT(uint16_t) LA_SYNTH_INT(int16_t)
LA_SYNTH_INT(uint32_t) LA_SYNTH_INT(int32_t) LA_SYNTH_INT(uint64_t) LA_SYNTH_INT(int64_t)

template <typename T>
__global__ void __launch_bounds__(EW_THREADS) synth_kernel(T* dst, uint64_t seed, size_t n) {
    const size_t stride = (size_t)gridDim.x * EW_THREADS;
    for (size_t i = (size_t)blockIdx.x * EW_THREADS + threadIdx.x; i < n; i += stride) dst[i] = synth_value<T>(splitmix64(seed ^ (uint64_t)i));
}

// C[i] = A[i] * (1 / sqrt(*dot_dev)): the scale step of normalize / normalized (src/vector.rs:313-323) with the
// squared length taken from device memory, so length -> 1/length -> scale needs no host round trip.  sqrt and the
// divide are the correctly rounded IEEE operations the reference's host code performs (traits.rs:63-65, vector.rs:314).
template <typename T>
__global__ void __launch_bounds__(EW_THREADS)
scale_by_inv_norm_kernel(T* C, const T* A, const T* __restrict__ dot_dev, size_t n) {
    const T inv = T(1) / sqrt(*dot_dev);
    const size_t stride = (size_t)gridDim.x * EW_THREADS;
    if (la_aligned16(C) && la_aligned16(A)) {
        using P = Pack<T>;
        const size_t n_pack = n / P::N;
        for (size_t p = (size_t)blockIdx.x * EW_THREADS + threadIdx.x; p < n_pack; p += stride) {
            P a = ld_stream(reinterpret_cast<const P*>(A) + p), c;
#pragma unroll
            for (int k = 0; k < P::N; ++k) c.v[k] = a.v[k] * inv;
            st_stream(reinterpret_cast<P*>(C) + p, c);
        }
        for (size_t i = n_pack * P::N + (size_t)blockIdx.x * EW_THREADS + threadIdx.x; i < n; i += stride) C[i] = A[i] * inv;
    } else {
        for (size_t i = (size_t)blockIdx.x * EW_THREADS + threadIdx.x; i < n; i += stride) C[i] = A[i] * inv;
    }
}

la_status check_common(int dtype, size_t n) {
    if (!la_valid_dtype(dtype)) return la_fail(LA_ERR_ARG, "unknown dtype %d", dtype);
    (void)n;
    return LA_OK;
}

}  // namespace

extern "C" la_status la_ew_vec_vec(int dtype, int op, void* C, const void* A, const void* B, size_t n, void* stream) {
    LA_TRY(check_common(dtype, n));
    if (n && (!C || !A || !B)) return la_fail(LA_ERR_ARG, "la_ew_vec_vec: NULL operand");
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    LA_DISPATCH_DTYPE(dtype, T, return (ew_op_dispatch<T, false>(op, (T*)C, (const T*)A, (const T*)B, T(0), n, (cudaStream_t)stream, info->sm_count)));
    return LA_OK;
}
extern "C" la_status la_ew_assign_vec_vec(int dtype, int op, void* C, const void* B, size_t n, void* stream) {
    return la_ew_vec_vec(dtype, op, C, C, B, n, stream);
}
extern "C" la_status la_ew_vec_scl(int dtype, int op, void* C, const void* A, const void* scalar_host, size_t n, void* stream) {
    LA_TRY(check_common(dtype, n));
    if (!scalar_host || (n && (!C || !A))) return la_fail(LA_ERR_ARG, "la_ew_vec_scl: NULL operand");
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    LA_DISPATCH_DTYPE(dtype, T, return (ew_op_dispatch<T, true>(op, (T*)C, (const T*)A, (const T*)nullptr, *(const T*)scalar_host, n, (cudaStream_t)stream, info->sm_count)));
    return LA_OK;
}
extern "C" la_status la_ew_assign_vec_scl(int dtype, int op, void* C, const void* scalar_host, size_t n, void* stream) {
    return la_ew_vec_scl(dtype, op, C, C, scalar_host, n, stream);
}

extern "C" la_status la_eq_vec(int dtype, const void* A, const void* B, size_t n, int* equal_host, void* stream) {
    LA_TRY(check_common(dtype, n));
    if (!equal_host || (n && (!A || !B))) return la_fail(LA_ERR_ARG, "la_eq_vec: NULL operand");
    if (n == 0) { *equal_host = 1; return LA_OK; }
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    cudaStream_t st = (cudaStream_t)stream;
    void* scratch; LA_TRY(la_scratch(st, 0, &scratch));
    unsigned* flag = (unsigned*)scratch + 1;            // word 1 of the zeroed header (word 0: tickets)
    const bool packed = la_aligned16(A) && la_aligned16(B);
    LA_DISPATCH_DTYPE(dtype, T, {
        size_t work = packed ? (n / Pack<T>::N + 1) : n;
        size_t blocks = (work + EW_THREADS - 1) / EW_THREADS;
        const size_t cap = (size_t)info->sm_count * 8;
        if (blocks > cap) blocks = cap;
        eq_kernel<T><<<(unsigned)blocks, EW_THREADS, 0, st>>>((const T*)A, (const T*)B, n, flag, packed);
    });
    LA_LAUNCH_CHECK();
    unsigned h = 0;
    LA_CUDA(cudaMemcpyAsync(&h, flag, sizeof(h), cudaMemcpyDeviceToHost, st));
    LA_CUDA(cudaMemsetAsync(flag, 0, sizeof(unsigned), st));     // leave the header zeroed
    LA_CUDA(cudaStreamSynchronize(st));
    *equal_host = h ? 0 : 1;
    return LA_OK;
}

extern "C" la_status la_fill(int dtype, void* dst_dev, const void* value_host, size_t n, void* stream) {
    LA_TRY(check_common(dtype, n));
    if (!value_host || (n && !dst_dev)) return la_fail(LA_ERR_ARG, "la_fill: NULL operand");
    if (n == 0) return LA_OK;
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    LA_DISPATCH_DTYPE(dtype, T, {
        size_t blocks = (n + EW_THREADS - 1) / EW_THREADS;
        const size_t cap = (size_t)info->sm_count * 16;
        if (blocks > cap) blocks = cap;
        fill_kernel<T><<<(unsigned)blocks, EW_THREADS, 0, (cudaStream_t)stream>>>((T*)dst_dev, *(const T*)value_host, n);
    });
    LA_LAUNCH_CHECK();
    return LA_OK;
}

extern "C" la_status la_synth(int dtype, void* dst_dev, uint64_t seed, size_t n, void* stream) {
    LA_TRY(check_common(dtype, n));
    if (n && !dst_dev) return la_fail(LA_ERR_ARG, "la_synth: NULL operand");
    if (n == 0) return LA_OK;
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    LA_DISPATCH_DTYPE(dtype, T, {
        size_t blocks = (n + EW_THREADS - 1) / EW_THREADS;
        const size_t cap = (size_t)info->sm_count * 16;
        if (blocks > cap) blocks = cap;
        synth_kernel<T><<<(unsigned)blocks, EW_THREADS, 0, (cudaStream_t)stream>>>((T*)dst_dev, seed, n);
    });
    LA_LAUNCH_CHECK();
    return LA_OK;
}

// normalized (C != A) / normalize (C == A), vector.rs:313-323: dot(a, a), sqrt, 1/length, scale -- stream-ordered,
// no host wait.  Real types only (traits.rs:76-78).
extern "C" la_status la_normalize_vec(int dtype, void* C, const void* A, size_t n, void* stream) {
    if (dtype != LA_F32 && dtype != LA_F64) return la_fail(LA_ERR_DTYPE, "la_normalize_vec: needs a Real element type (f32 / f64)");
    if (n && (!C || !A)) return la_fail(LA_ERR_ARG, "la_normalize_vec: NULL operand");
    if (n == 0) return LA_OK;
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    cudaStream_t st = (cudaStream_t)stream;
    void* scratch; LA_TRY(la_scratch(st, LA_REDUCE_SCRATCH_BYTES, &scratch));   // sized as the reduction below sizes it: the block cannot move under `slot`
    void* slot = (char*)scratch + 32;                      // bytes 32..39 of the header (restored to zero below)
    LA_TRY(la_dot_vec_vec_dev(dtype, A, A, n, slot, stream));
    size_t blocks = (n / 4 + EW_THREADS) / EW_THREADS;
    const size_t cap = (size_t)info->sm_count * 16;
    if (blocks > cap) blocks = cap;
    if (dtype == LA_F32) scale_by_inv_norm_kernel<float><<<(unsigned)blocks, EW_THREADS, 0, st>>>((float*)C, (const float*)A, (const float*)slot, n);
    else                 scale_by_inv_norm_kernel<double><<<(unsigned)blocks, EW_THREADS, 0, st>>>((double*)C, (const double*)A, (const double*)slot, n);
    LA_LAUNCH_CHECK();
    LA_CUDA(cudaMemsetAsync(slot, 0, 8, st));
    return LA_OK;
}
