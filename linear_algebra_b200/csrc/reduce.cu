// csrc/reduce.cu -- Vector::sum and dot.
//
// Replaces {T}_sum_vec (src/kernels.cl:83-102) + the 64-partial download and host iter::sum()
// of Vector::sum (src/vector.rs:256-293), and {T}_dot_vec_vec (src/kernels.cl:105-124) + dot()
// (src/vector.rs:326-363).  The reference launches a fixed 64 work-groups of one warp each
// (src/lib.rs:114) and finishes on the host; here ONE launch does everything:
//   stage 1  grid-stride over 16-byte packets, RED_UNROLL packets per operand in flight per
//            thread, independent accumulators, warp-shuffle + shared-memory block reduction,
//            one partial per block into scratch;
//   stage 2  the last block to finish (atomic ticket) adds the partials in a fixed order.
// Grid size and every reduction shape depend only on (n, SM count), so a result is
// reproducible run to run.  Integer accumulation wraps (exact mod 2^n whatever the order);
// 8/16-bit types accumulate in 32 bits and truncate at the end.
//
// Roofline: HBM.  Algorithmic bytes per element: sizeof(T) (sum), 2*sizeof(T) (dot).
#include "common.cuh"

namespace {

constexpr int RED_THREADS = 256;
// 16-byte packets per operand in flight per thread
template <bool DOT> struct RedUnroll { static constexpr int value = 4; };
template <bool DOT> struct RedBlocksPerSm { static constexpr int value = DOT ? 4 : 6; };     // sum needs fewer registers: more blocks in flight

template <typename T, bool DOT>
__global__ void __launch_bounds__(RED_THREADS)
reduce_kernel(const T* __restrict__ a, const T* __restrict__ b, size_t n, bool packed,
              typename Arith<T>::Acc* partials, unsigned* ticket, T* out) {
    using A = typename Arith<T>::Acc;
    using P = Pack<T>;
    constexpr int N = P::N;
    constexpr int RED_UNROLL = RedUnroll<DOT>::value;
    __shared__ A smem[32];
    __shared__ bool is_last;

    const size_t stride = (size_t)gridDim.x * RED_THREADS;
    const size_t tid = (size_t)blockIdx.x * RED_THREADS + threadIdx.x;
    A acc[RED_UNROLL];
#pragma unroll
    for (int j = 0; j < RED_UNROLL; ++j) acc[j] = A(0);

    size_t done = 0;
    if (packed) {
        const size_t n_pack = n / N;
        const P* ap = reinterpret_cast<const P*>(a);
        const P* bp = reinterpret_cast<const P*>(b);
        size_t p = tid;
        for (; p + (RED_UNROLL - 1) * stride < n_pack; p += RED_UNROLL * stride) {
            P x[RED_UNROLL], y[RED_UNROLL];
#pragma unroll
            for (int j = 0; j < RED_UNROLL; ++j) {
                x[j] = ld_stream(ap + p + j * stride);
                if (DOT) y[j] = ld_stream(bp + p + j * stride);
            }
#pragma unroll
            for (int j = 0; j < RED_UNROLL; ++j)
#pragma unroll
                for (int e = 0; e < N; ++e)
                    acc[j] = DOT ? Arith<T>::fma(x[j].v[e], y[j].v[e], acc[j]) : acc[j] + Arith<T>::widen(x[j].v[e]);
        }
        for (; p < n_pack; p += stride) {
            P x = ld_stream(ap + p), y;
            if (DOT) y = ld_stream(bp + p);
#pragma unroll
            for (int e = 0; e < N; ++e)
                acc[0] = DOT ? Arith<T>::fma(x.v[e], y.v[e], acc[0]) : acc[0] + Arith<T>::widen(x.v[e]);
        }
        done = n_pack * N;
    }
    for (size_t i = done + tid; i < n; i += stride)
        acc[0] = DOT ? Arith<T>::fma(a[i], b[i], acc[0]) : acc[0] + Arith<T>::widen(a[i]);

    A v = (acc[0] + acc[1]) + (acc[2] + acc[3]);
    if (RED_UNROLL == 8) v += (acc[RED_UNROLL - 4] + acc[RED_UNROLL - 3]) + (acc[RED_UNROLL - 2] + acc[RED_UNROLL - 1]);
    v = block_sum(v, smem);

    if (gridDim.x == 1) {
        if (threadIdx.x == 0) *out = Arith<T>::from_acc(v);
        return;
    }
    if (threadIdx.x == 0) {
        partials[blockIdx.x] = v;
        __threadfence();
        is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    A t = A(0);
    for (unsigned k = threadIdx.x; k < gridDim.x; k += RED_THREADS) t += __ldcg(partials + k);
    t = block_sum(t, smem);
    if (threadIdx.x == 0) {
        *out = Arith<T>::from_acc(t);
        *ticket = 0;                 // leave the scratch header zeroed for the next launch
    }
}

// out_dev: device pointer to one T.
template <typename T, bool DOT>
la_status reduce_launch(const T* a, const T* b, size_t n, T* out_dev, cudaStream_t st, int sm_count) {
    using A = typename Arith<T>::Acc;
    if (n == 0) {                    // empty sum = 0 (Rust: `iter::sum()` of nothing)
        LA_CUDA(cudaMemsetAsync(out_dev, 0, sizeof(T), st));
        return LA_OK;
    }
    const bool packed = la_aligned16(a) && (!DOT || la_aligned16(b));
    const size_t work = packed ? (n / Pack<T>::N + 1) : n;
    constexpr int RED_UNROLL = RedUnroll<DOT>::value;
    size_t blocks = (work + (size_t)RED_THREADS * RED_UNROLL - 1) / ((size_t)RED_THREADS * RED_UNROLL);
    const size_t cap = (size_t)sm_count * RedBlocksPerSm<DOT>::value;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    static_assert(sizeof(A) <= 8, "partials are at most 8 bytes");
    if (blocks * sizeof(A) > LA_REDUCE_SCRATCH_BYTES) blocks = LA_REDUCE_SCRATCH_BYTES / sizeof(A);
    // Always the same request, so a caller that already holds a pointer into this stream's scratch header (the result slot
    // of reduce_entry, the length slot of la_normalize_vec -- both made with the same request) never sees the block move.
    void* scratch;
    LA_TRY(la_scratch(st, LA_REDUCE_SCRATCH_BYTES, &scratch));
    unsigned* ticket = (unsigned*)scratch;                                  // word 0 of the zeroed header
    A* partials = (A*)((char*)scratch + LA_SCRATCH_ZEROED);
    reduce_kernel<T, DOT><<<(unsigned)blocks, RED_THREADS, 0, st>>>(a, b, n, packed, partials, ticket, out_dev);
    LA_LAUNCH_CHECK();
    return LA_OK;
}

la_status reduce_entry(int dtype, bool dot, const void* a, const void* b, size_t n, void* out, bool out_is_host, void* stream) {
    if (!la_valid_dtype(dtype)) return la_fail(LA_ERR_ARG, "unknown dtype %d", dtype);
    if (!out || (n && (!a || (dot && !b)))) return la_fail(LA_ERR_ARG, "%s: NULL operand", dot ? "la_dot_vec_vec" : "la_sum_vec");
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    cudaStream_t st = (cudaStream_t)stream;
    void* out_dev = out;
    if (out_is_host) {
        // result slot: bytes 16..23 of the scratch header (re-zeroed below)
        void* scratch; LA_TRY(la_scratch(st, LA_REDUCE_SCRATCH_BYTES, &scratch));    // the one request reduce_launch makes too
        out_dev = (char*)scratch + 16;
    }
    LA_DISPATCH_DTYPE(dtype, T, {
        la_status s = dot ? reduce_launch<T, true>((const T*)a, (const T*)b, n, (T*)out_dev, st, info->sm_count)
                          : reduce_launch<T, false>((const T*)a, (const T*)nullptr, n, (T*)out_dev, st, info->sm_count);
        if (s != LA_OK) return s;
    });
    if (out_is_host) {
        LA_CUDA(cudaMemcpyAsync(out, out_dev, la_dtype_size(dtype), cudaMemcpyDeviceToHost, st));
        LA_CUDA(cudaMemsetAsync(out_dev, 0, 8, st));
        LA_CUDA(cudaStreamSynchronize(st));
    }
    return LA_OK;
}

}  // namespace

extern "C" la_status la_sum_vec(int dtype, const void* data, size_t n, void* out_host, void* stream) {
    return reduce_entry(dtype, false, data, nullptr, n, out_host, true, stream);
}
extern "C" la_status la_dot_vec_vec(int dtype, const void* a, const void* b, size_t n, void* out_host, void* stream) {
    return reduce_entry(dtype, true, a, b, n, out_host, true, stream);
}
extern "C" la_status la_sum_vec_dev(int dtype, const void* data, size_t n, void* out_dev, void* stream) {
    return reduce_entry(dtype, false, data, nullptr, n, out_dev, false, stream);
}
extern "C" la_status la_dot_vec_vec_dev(int dtype, const void* a, const void* b, size_t n, void* out_dev, void* stream) {
    return reduce_entry(dtype, true, a, b, n, out_dev, false, stream);
}
