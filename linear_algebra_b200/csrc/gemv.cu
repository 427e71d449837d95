// csrc/gemv.cu -- Vector x Matrix in both orientations.
//
// la_mul_vec_transpose_mat  replaces {T}_mul_vec_transpose_mat (src/kernels.cl:150-152, host
//   src/vector.rs:542-569):  C[r] = sum_j A[j] * B[r*cols + j]          ("Matrix x Vector")
//   Reference: one work-item per row, neighbouring work-items `cols` elements apart ->
//   uncoalesced.  Here: a warp owns MV_ROWS consecutive rows and sweeps them with 16-byte
//   streaming loads, re-using each packet of A (kept in L1) for all MV_ROWS rows; shuffle
//   reduction per row.  Few-row matrices use one block per row instead.
//
// la_mul_vec_mat            replaces {T}_mul_vec_mat (src/kernels.cl:143-147, host
//   src/vector.rs:508-539):  C[col] = sum_j A[j] * B[j*cols + col]
//   Reference: one work-item per column walking `rows` strided elements -> only `cols`
//   threads, latency-bound.  Here: 2-D decomposition, a thread owns one 16-byte packet of
//   columns over a strip of rows (VM_UNROLL rows in flight); per-strip partial vectors go to
//   scratch and a second kernel adds the strips in ascending order (deterministic, no atomics).
//
// Roofline: HBM.  Algorithmic bytes: sizeof(T) * (rows*cols + rows + cols).
// Accumulation is in T's accumulator type (wraps for integers; f32 stays f32), in a fixed order.
#include "common.cuh"
#include <mutex>

namespace {

// ------------------------------------------------------------------ M x v  (row dots)
constexpr int MV_THREADS = 256;
constexpr int MV_LOADS = 8;     // 16-byte matrix packets in flight per lane (MV_ROWS x MV_UNROLL)

// MV_ROWS rows per warp (4 for tall matrices; 2 or 1 when there are too few rows to fill the SMs otherwise)
template <typename T, int MV_ROWS>
__global__ void __launch_bounds__(MV_THREADS, 4)
matvec_rows_pack_kernel(T* __restrict__ C, const T* __restrict__ A, const T* __restrict__ B, size_t rows, size_t cols) {
    using Acc = typename Arith<T>::Acc;
    using P = Pack<T>;
    constexpr int N = P::N;
    constexpr int MV_UNROLL = MV_LOADS / MV_ROWS;    // column packets in flight per row per lane
    const int lane = threadIdx.x & 31;
    const size_t warp = ((size_t)blockIdx.x * MV_THREADS + threadIdx.x) >> 5;
    const size_t r0 = warp * MV_ROWS;
    if (r0 >= rows) return;
    const size_t n_pack = cols / N;                  // cols % N == 0 on this path
    const P* vp = reinterpret_cast<const P*>(A);
    const P* bp[MV_ROWS];
#pragma unroll
    for (int i = 0; i < MV_ROWS; ++i) {
        const size_t r = (r0 + i < rows) ? r0 + i : rows - 1;      // clamp: duplicates are discarded below
        bp[i] = reinterpret_cast<const P*>(B + r * cols);
    }
    Acc acc[MV_ROWS];
#pragma unroll
    for (int i = 0; i < MV_ROWS; ++i) acc[i] = Acc(0);

    size_t p = lane;
    for (; p + (MV_UNROLL - 1) * 32 < n_pack; p += MV_UNROLL * 32) {
        P v[MV_UNROLL], m[MV_UNROLL][MV_ROWS];
#pragma unroll
        for (int u = 0; u < MV_UNROLL; ++u) {
            v[u] = ld_cached(vp + p + u * 32);
#pragma unroll
            for (int i = 0; i < MV_ROWS; ++i) m[u][i] = ld_stream(bp[i] + p + u * 32);
        }
#pragma unroll
        for (int u = 0; u < MV_UNROLL; ++u)
#pragma unroll
            for (int i = 0; i < MV_ROWS; ++i)
#pragma unroll
                for (int e = 0; e < N; ++e) acc[i] = Arith<T>::fma(v[u].v[e], m[u][i].v[e], acc[i]);
    }
    for (; p < n_pack; p += 32) {
        P v = ld_cached(vp + p);
#pragma unroll
        for (int i = 0; i < MV_ROWS; ++i) {
            P m = ld_stream(bp[i] + p);
#pragma unroll
            for (int e = 0; e < N; ++e) acc[i] = Arith<T>::fma(v.v[e], m.v[e], acc[i]);
        }
    }
#pragma unroll
    for (int i = 0; i < MV_ROWS; ++i) {
        Acc s = warp_sum(acc[i]);
        if (lane == 0 && r0 + i < rows) C[r0 + i] = Arith<T>::from_acc(s);
    }
}

// One block per row; any alignment, any cols.  Used for ragged shapes and few-row matrices.
template <typename T>
__global__ void __launch_bounds__(MV_THREADS)
matvec_rows_block_kernel(T* __restrict__ C, const T* __restrict__ A, const T* __restrict__ B, size_t rows, size_t cols) {
    using Acc = typename Arith<T>::Acc;
    __shared__ Acc smem[32];
    for (size_t r = blockIdx.x; r < rows; r += gridDim.x) {
        const T* row = B + r * cols;
        Acc acc = Acc(0);
        for (size_t j = threadIdx.x; j < cols; j += MV_THREADS) acc = Arith<T>::fma(A[j], row[j], acc);
        acc = block_sum(acc, smem);
        if (threadIdx.x == 0) C[r] = Arith<T>::from_acc(acc);
    }
}

// One block per row, 16-byte packets (aligned rows): matrices with few, long rows.
template <typename T>
__global__ void __launch_bounds__(MV_THREADS)
matvec_rows_block_pack_kernel(T* __restrict__ C, const T* __restrict__ A, const T* __restrict__ B, size_t rows, size_t cols) {
    using Acc = typename Arith<T>::Acc;
    using P = Pack<T>;
    constexpr int N = P::N;
    __shared__ Acc smem[32];
    const size_t n_pack = cols / N;
    const P* vp = reinterpret_cast<const P*>(A);
    for (size_t r = blockIdx.x; r < rows; r += gridDim.x) {
        const P* bp = reinterpret_cast<const P*>(B + r * cols);
        Acc acc[4] = {Acc(0), Acc(0), Acc(0), Acc(0)};
        size_t p = threadIdx.x;
        for (; p + 3 * MV_THREADS < n_pack; p += 4 * MV_THREADS) {
            P v[4], m[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) { m[u] = ld_stream(bp + p + u * MV_THREADS); v[u] = ld_cached(vp + p + u * MV_THREADS); }
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int e = 0; e < N; ++e) acc[u] = Arith<T>::fma(v[u].v[e], m[u].v[e], acc[u]);
        }
        for (; p < n_pack; p += MV_THREADS) {
            P v = ld_cached(vp + p), m = ld_stream(bp + p);
#pragma unroll
            for (int e = 0; e < N; ++e) acc[0] = Arith<T>::fma(v.v[e], m.v[e], acc[0]);
        }
        Acc s = block_sum((acc[0] + acc[1]) + (acc[2] + acc[3]), smem);
        if (threadIdx.x == 0) C[r] = Arith<T>::from_acc(s);
    }
}

// ------------------------------------------------------------------ v x M  (column sums)
constexpr int VM_THREADS = 128;
constexpr int VM_UNROLL = 8;    // rows in flight per thread

// Thread = one packet of columns over rows [strip*rows_per_strip, ...).  Writes Acc partials to
// part[strip][cols] (or final T values straight to C when there is a single strip).
template <typename T>
__global__ void __launch_bounds__(VM_THREADS)
vecmat_pack_kernel(T* __restrict__ C, typename Arith<T>::Acc* __restrict__ part, const T* __restrict__ A,
                   const T* __restrict__ B, size_t rows, size_t cols, size_t rows_per_strip) {
    using Acc = typename Arith<T>::Acc;
    using P = Pack<T>;
    constexpr int N = P::N;
    const size_t n_pack = cols / N;
    const size_t p = (size_t)blockIdx.x * VM_THREADS + threadIdx.x;
    if (p >= n_pack) return;
    const size_t r_begin = (size_t)blockIdx.y * rows_per_strip;
    size_t r_end = r_begin + rows_per_strip;
    if (r_end > rows) r_end = rows;
    Acc acc[N];
#pragma unroll
    for (int e = 0; e < N; ++e) acc[e] = Acc(0);
    const P* bp = reinterpret_cast<const P*>(B) + p;
    size_t r = r_begin;
    for (; r + VM_UNROLL <= r_end; r += VM_UNROLL) {
        P m[VM_UNROLL]; T v[VM_UNROLL];
#pragma unroll
        for (int u = 0; u < VM_UNROLL; ++u) { m[u] = ld_stream(bp + (r + u) * n_pack); v[u] = __ldg(A + r + u); }
#pragma unroll
        for (int u = 0; u < VM_UNROLL; ++u)
#pragma unroll
            for (int e = 0; e < N; ++e) acc[e] = Arith<T>::fma(v[u], m[u].v[e], acc[e]);
    }
    for (; r < r_end; ++r) {
        P m = ld_stream(bp + r * n_pack); T v = __ldg(A + r);
#pragma unroll
        for (int e = 0; e < N; ++e) acc[e] = Arith<T>::fma(v, m.v[e], acc[e]);
    }
    if (gridDim.y == 1) {
        P out;
#pragma unroll
        for (int e = 0; e < N; ++e) out.v[e] = Arith<T>::from_acc(acc[e]);
        reinterpret_cast<P*>(C)[p] = out;
    } else {
        Acc* dst = part + (size_t)blockIdx.y * cols + p * N;
#pragma unroll
        for (int e = 0; e < N; ++e) dst[e] = acc[e];
    }
}

// Any alignment / ragged cols: thread = one column.
template <typename T>
__global__ void __launch_bounds__(VM_THREADS)
vecmat_elem_kernel(T* __restrict__ C, typename Arith<T>::Acc* __restrict__ part, const T* __restrict__ A,
                   const T* __restrict__ B, size_t rows, size_t cols, size_t rows_per_strip) {
    using Acc = typename Arith<T>::Acc;
    const size_t c = (size_t)blockIdx.x * VM_THREADS + threadIdx.x;
    if (c >= cols) return;
    const size_t r_begin = (size_t)blockIdx.y * rows_per_strip;
    size_t r_end = r_begin + rows_per_strip;
    if (r_end > rows) r_end = rows;
    Acc acc = Acc(0);
    for (size_t r = r_begin; r < r_end; ++r) acc = Arith<T>::fma(__ldg(A + r), B[r * cols + c], acc);
    if (gridDim.y == 1) C[c] = Arith<T>::from_acc(acc);
    else part[(size_t)blockIdx.y * cols + c] = acc;
}

template <typename T>
__global__ void __launch_bounds__(256)
vecmat_finish_kernel(T* __restrict__ C, const typename Arith<T>::Acc* __restrict__ part, size_t cols, int strips) {
    using Acc = typename Arith<T>::Acc;
    const size_t c = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (c >= cols) return;
    Acc acc = Acc(0);
    for (int s = 0; s < strips; ++s) acc += part[(size_t)s * cols + c];     // ascending strip order
    C[c] = Arith<T>::from_acc(acc);
}

template <typename T, int MV_ROWS>
la_status matvec_rows_launch(T* C, const T* A, const T* B, size_t rows, size_t cols, cudaStream_t st) {
    const size_t warps = (rows + MV_ROWS - 1) / MV_ROWS;
    const size_t blocks = (warps + MV_THREADS / 32 - 1) / (MV_THREADS / 32);
    if (blocks > 0x7fffffffull) return la_fail(LA_ERR_UNSUPPORTED, "mul_vec_transpose_mat: too many rows");
    matvec_rows_pack_kernel<T, MV_ROWS><<<(unsigned)blocks, MV_THREADS, 0, st>>>(C, A, B, rows, cols);
    LA_LAUNCH_CHECK();
    return LA_OK;
}

template <typename T>
la_status matvec_rows(T* C, const T* A, const T* B, size_t rows, size_t cols, cudaStream_t st, int sm) {
    const bool packed = la_aligned16(A) && la_aligned16(B) && (cols % Pack<T>::N == 0) && cols >= 32 * Pack<T>::N;
    const size_t fill = (size_t)sm * 16;                        // warps that keep every SM busy
    if (packed && rows / 4 >= fill) return matvec_rows_launch<T, 4>(C, A, B, rows, cols, st);
    if (packed && rows / 2 >= fill) return matvec_rows_launch<T, 2>(C, A, B, rows, cols, st);
    if (packed && rows >= fill)     return matvec_rows_launch<T, 1>(C, A, B, rows, cols, st);
    size_t blocks = rows;
    const size_t cap = (size_t)sm * 16;
    if (blocks > cap) blocks = cap;
    if (packed) matvec_rows_block_pack_kernel<T><<<(unsigned)blocks, MV_THREADS, 0, st>>>(C, A, B, rows, cols);
    else        matvec_rows_block_kernel<T><<<(unsigned)blocks, MV_THREADS, 0, st>>>(C, A, B, rows, cols);
    LA_LAUNCH_CHECK();
    return LA_OK;
}

template <typename T>
la_status vecmat_cols(T* C, const T* A, const T* B, size_t rows, size_t cols, cudaStream_t st, int sm) {
    using Acc = typename Arith<T>::Acc;
    const bool packed = la_aligned16(C) && la_aligned16(B) && (cols % Pack<T>::N == 0);
    const size_t units = packed ? cols / Pack<T>::N : cols;
    const size_t col_blocks = (units + VM_THREADS - 1) / VM_THREADS;
    // enough strips to put ~16 blocks on every SM, at least 4*VM_UNROLL rows each
    size_t strips = ((size_t)sm * 16 + col_blocks - 1) / col_blocks;
    const size_t max_strips = (rows + 4 * VM_UNROLL - 1) / (4 * VM_UNROLL);
    if (strips > max_strips) strips = max_strips;
    if (strips > 65535) strips = 65535;
    if (strips < 1) strips = 1;
    const size_t rows_per_strip = (rows + strips - 1) / strips;
    strips = (rows + rows_per_strip - 1) / rows_per_strip;
    if (col_blocks > 0x7fffffffull) return la_fail(LA_ERR_UNSUPPORTED, "mul_vec_mat: too many columns");
    Acc* part = nullptr;
    if (strips > 1) {
        void* scratch; LA_TRY(la_scratch(st, strips * cols * sizeof(Acc), &scratch));
        part = (Acc*)((char*)scratch + LA_SCRATCH_ZEROED);
    }
    dim3 grid((unsigned)col_blocks, (unsigned)strips);
    if (packed) vecmat_pack_kernel<T><<<grid, VM_THREADS, 0, st>>>(C, part, A, B, rows, cols, rows_per_strip);
    else        vecmat_elem_kernel<T><<<grid, VM_THREADS, 0, st>>>(C, part, A, B, rows, cols, rows_per_strip);
    LA_LAUNCH_CHECK();
    if (strips > 1) {
        vecmat_finish_kernel<T><<<(unsigned)((cols + 255) / 256), 256, 0, st>>>(C, part, cols, (int)strips);
        LA_LAUNCH_CHECK();
    }
    return LA_OK;
}

la_status check(const char* name, int dtype, void* C, const void* A, const void* B) {
    if (!la_valid_dtype(dtype)) return la_fail(LA_ERR_ARG, "%s: unknown dtype %d", name, dtype);
    if (!C || !A || !B) return la_fail(LA_ERR_ARG, "%s: NULL operand", name);
    return LA_OK;
}

}  // namespace

extern "C" la_status la_mul_vec_transpose_mat(int dtype, void* C, const void* A, const void* B, size_t rows, size_t cols, void* stream) {
    LA_TRY(check("la_mul_vec_transpose_mat", dtype, C, A, B));
    if (rows == 0) return LA_OK;
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    if (cols == 0) { LA_CUDA(cudaMemsetAsync(C, 0, rows * la_dtype_size(dtype), (cudaStream_t)stream)); return LA_OK; }
    LA_DISPATCH_DTYPE(dtype, T, return matvec_rows<T>((T*)C, (const T*)A, (const T*)B, rows, cols, (cudaStream_t)stream, info->sm_count));
    return LA_OK;
}

extern "C" la_status la_mul_vec_mat(int dtype, void* C, const void* A, const void* B, size_t rows, size_t cols, void* stream) {
    LA_TRY(check("la_mul_vec_mat", dtype, C, A, B));
    if (cols == 0) return LA_OK;
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    if (rows == 0) { LA_CUDA(cudaMemsetAsync(C, 0, cols * la_dtype_size(dtype), (cudaStream_t)stream)); return LA_OK; }
    LA_DISPATCH_DTYPE(dtype, T, return vecmat_cols<T>((T*)C, (const T*)A, (const T*)B, rows, cols, (cudaStream_t)stream, info->sm_count));
    return LA_OK;
}

// =====================================================================================
// la_mul_vec_mat_noalloc -- the same v x M computation for callers that must not allocate.
//
// Who needs it: a host program that has already handed (nearly) all device memory to its own
// allocator, so that a cudaMalloc issued from inside this library would fail; and a caller in a
// latency-critical section that cannot afford the device-wide synchronisation a cudaMalloc /
// cudaFree implies.  la_mul_vec_mat gets its per-strip partial vectors from la_scratch(), which
// grows by cudaMalloc on demand.  This variant keeps them in the fixed block below instead and
// calls no cudaMalloc, cudaFree, cudaMallocAsync or any other allocating routine, on any call.
//
// The price: the block is part of the library image, so every GPU that loads the library
// holds LA_NOALLOC_SCRATCH_BYTES of device memory until the process exits, and problems whose
// partials do not fit are refused (LA_ERR_UNSUPPORTED) rather than served more slowly.
//
// Sharing: there is ONE block per device, used by every call whatever its stream.  Two calls
// whose kernels overlap on the GPU (different streams, or different host threads) race on it
// and produce garbage; serialise them.  The block keeps whatever the previous call left in
// it.  Nothing in it has to start at zero: every partial the finishing kernel reads is
// written first by the main kernel of the same call, so no clearing pass is needed.
//
// First call on a device: checks that the device is a B200-class GPU (sm_10x; the library holds
// no code for anything else), looks up the device address of the block (cudaGetSymbolAddress)
// and the SM count, once, with that device current.  Under CUDA's lazy module loading that
// first call is also what places this file's kernels and the 8 MiB block on the device, so it
// can itself need device memory and may synchronise.  A host that is about to hand all device
// memory to its own allocator should therefore make one small warm-up call per device, per
// element type it will use, during start-up, BEFORE its allocator takes the memory and before
// starting any stream capture.  Later calls only launch kernels.
constexpr size_t LA_NOALLOC_SCRATCH_BYTES = size_t(8) << 20;      // 8 MiB per GPU
static __device__ __align__(256) unsigned char la_noalloc_scratch[LA_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr int LA_NOALLOC_MAX_DEVICES = 64;
struct NoallocDevice { bool ready; void* scratch; int sm_count; };
NoallocDevice g_noalloc[LA_NOALLOC_MAX_DEVICES];     // zero-initialised: ready == false
std::mutex g_noalloc_mu;

// In host code `la_noalloc_scratch` is only a handle for the runtime, not an address a kernel
// can use; the device address has to be asked for, per device.
la_status noalloc_setup(NoallocDevice** out) {
    int dev = -1;
    LA_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= LA_NOALLOC_MAX_DEVICES) return la_fail(LA_ERR_UNSUPPORTED, "la_mul_vec_mat_noalloc: device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    NoallocDevice& d = g_noalloc[dev];
    if (!d.ready) {
        // la_ensure_init() is bypassed on purpose (it sizes a memory pool); its device check is repeated here
        int major = 0, minor = 0;
        LA_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
        LA_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
        if (major != 10)
            return la_fail(LA_ERR_CUDA, "la_mul_vec_mat_noalloc: device %d is sm_%d%d; liblinalg_b200 carries sm_100a code only", dev, major, minor);
        void* p = nullptr;
        LA_CUDA(cudaGetSymbolAddress(&p, la_noalloc_scratch));
        if (!p) return la_fail(LA_ERR_CUDA, "la_mul_vec_mat_noalloc: cudaGetSymbolAddress returned NULL");
        int sms = 0;
        LA_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        d.scratch = p; d.sm_count = sms; d.ready = true;
    }
    *out = &d;
    return LA_OK;
}

// Same decomposition as vecmat_cols() above -- same strips, same kernels, same order of additions -- so the two entry
// points give bit-identical results for every element type, floats included (tests/test_zz_no_alloc.py asserts it),
// with the partials in the fixed block.
template <typename T>
la_status vecmat_cols_noalloc(T* C, const T* A, const T* B, size_t rows, size_t cols, cudaStream_t st, const NoallocDevice& d) {
    using Acc = typename Arith<T>::Acc;
    const int sm = d.sm_count;
    const bool packed = la_aligned16(C) && la_aligned16(B) && (cols % Pack<T>::N == 0);
    const size_t units = packed ? cols / Pack<T>::N : cols;
    const size_t col_blocks = (units + VM_THREADS - 1) / VM_THREADS;
    size_t strips = ((size_t)sm * 16 + col_blocks - 1) / col_blocks;
    const size_t max_strips = (rows + 4 * VM_UNROLL - 1) / (4 * VM_UNROLL);
    if (strips > max_strips) strips = max_strips;
    if (strips > 65535) strips = 65535;
    if (strips < 1) strips = 1;
    const size_t rows_per_strip = (rows + strips - 1) / strips;
    strips = (rows + rows_per_strip - 1) / rows_per_strip;
    if (col_blocks > 0x7fffffffull) return la_fail(LA_ERR_UNSUPPORTED, "la_mul_vec_mat_noalloc: too many columns");
    Acc* part = nullptr;
    if (strips > 1) {
        const size_t need = strips * cols * sizeof(Acc);
        if (need > LA_NOALLOC_SCRATCH_BYTES)
            return la_fail(LA_ERR_UNSUPPORTED,
                           "la_mul_vec_mat_noalloc: %zu x %zu needs %zu bytes of partials (%zu strips), the fixed block holds %zu; use la_mul_vec_mat",
                           rows, cols, need, strips, (size_t)LA_NOALLOC_SCRATCH_BYTES);
        part = (Acc*)d.scratch;
    }
    dim3 grid((unsigned)col_blocks, (unsigned)strips);
    if (packed) vecmat_pack_kernel<T><<<grid, VM_THREADS, 0, st>>>(C, part, A, B, rows, cols, rows_per_strip);
    else        vecmat_elem_kernel<T><<<grid, VM_THREADS, 0, st>>>(C, part, A, B, rows, cols, rows_per_strip);
    LA_LAUNCH_CHECK();
    if (strips > 1) {
        vecmat_finish_kernel<T><<<(unsigned)((cols + 255) / 256), 256, 0, st>>>(C, part, cols, (int)strips);
        LA_LAUNCH_CHECK();
    }
    return LA_OK;
}

}  // namespace

extern "C" la_status la_mul_vec_mat_noalloc(int dtype, void* C, const void* A, const void* B, size_t rows, size_t cols, void* stream) {
    LA_TRY(check("la_mul_vec_mat_noalloc", dtype, C, A, B));
    if (cols == 0) return LA_OK;
    NoallocDevice* d = nullptr;
    LA_TRY(noalloc_setup(&d));
    if (rows == 0) { LA_CUDA(cudaMemsetAsync(C, 0, cols * la_dtype_size(dtype), (cudaStream_t)stream)); return LA_OK; }
    LA_DISPATCH_DTYPE(dtype, T, return vecmat_cols_noalloc<T>((T*)C, (const T*)A, (const T*)B, rows, cols, (cudaStream_t)stream, *d));
    return LA_OK;
}
