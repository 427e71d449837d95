// csrc/common.cuh -- shared pieces of liblinalg_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <type_traits>
#include "../../include/linalg_b200.h"

// ---------------------------------------------------------------- errors
void la_set_error(const char* fmt, ...);
la_status la_fail(la_status code, const char* fmt, ...);
la_status la_cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define LA_CUDA(call)                                                       \
    do {                                                                    \
        cudaError_t e_ = (call);                                            \
        if (e_ != cudaSuccess) return la_cuda_fail(e_, #call, __FILE__, __LINE__); \
    } while (0)

#define LA_TRY(call)                           \
    do {                                       \
        la_status s_ = (call);                 \
        if (s_ != LA_OK) return s_;            \
    } while (0)

// Launch check: configuration errors surface here, execution errors at the next sync.
#define LA_LAUNCH_CHECK() LA_CUDA(cudaGetLastError())

// ---------------------------------------------------------------- runtime state
struct LaDeviceInfo {
    int device;
    int sm_count;
    int max_smem_optin;
};
la_status la_ensure_init(LaDeviceInfo** info);   // lazily binds the current device

// Scratch: global memory beyond a kernel's inputs and outputs (reduction partials, tickets,
// GEMV strip partials, eq flags).  One growing block per (device, stream); contents undefined
// except for the first LA_SCRATCH_ZEROED bytes, which every kernel leaves zeroed on exit
// (tickets / flags).
enum { LA_SCRATCH_ZEROED = 256 };
la_status la_scratch(cudaStream_t stream, size_t bytes, void** ptr);
// What every reduction asks la_scratch for (its per-block partials: at most 8 bytes x a few blocks per SM).  One fixed
// request, so pointers into the header taken before a reduction stay valid across it.
constexpr size_t LA_REDUCE_SCRATCH_BYTES = 64 << 10;
// Shortest f32 product LA_GEMM_AUTO hands to the tensor engine (gemm.cu, gemm_host.cu): below it the truncating
// accumulation inside a few tensor-core chains can land above the reference order's own error.
constexpr size_t LA_TC_AUTO_MIN_K = LA_GEMM_AUTO_MIN_K;      // include/linalg_b200.h
// One zero-copy flag word per (device, stream): kernels store into *dev_ptr, the host reads
// *host_ptr after waiting for them (no device-to-host copy involved).
la_status la_host_flag(cudaStream_t stream, volatile unsigned** host_ptr, unsigned** dev_ptr);
// One event (no timing) per (device, stream), for waiting on a point inside a chain of launches.
la_status la_stream_event(cudaStream_t stream, cudaEvent_t* event);

// What la_gemm_prepare_b keeps about a B that many products will share (include/linalg_b200.h): the caller's B itself (the SIMT
// and generic engines read it as it is) and, where the tensor engines apply, their pre-processed forms of it.  The planes
// are plain cudaMalloc blocks of the device the handle was made on.
struct la_gemm_prepared {
    int dtype = 0; int device = -1;
    size_t K = 0, N = 0;
    const void* B = nullptr;                 // caller's K x N matrix: must stay allocated and unchanged while the handle lives
    float* tc_hi = nullptr; float* tc_lo = nullptr; bool tc_cross_bf16 = false;    // f32: split planes of B^T (gemm_tc.cu); NULL when not applicable
    void* f16_planes = nullptr; bool f16_b_tiny = false, f16_b_sparse = false;                           // f32, f16-split form: fp16 hi | bf16 hi | bf16 lo planes of the column-scaled B^T, then the column exponents and the list of left-out elements (one block)
    uint8_t* limb_planes = nullptr;          // 8/16/32-bit integers: byte planes of B^T (gemm_limb.cu); NULL when not applicable
};

int la_gemm_raster_group(size_t pair_tiles_m);   // gemm_tc.cu: tile rows per band of the tensor engines' tile order (0 = m-major; tuning hook la_gemm_set_raster)

// ---------------------------------------------------------------- dtype traits
template <typename T> struct Arith {                       // wrap-safe arithmetic
    using U = typename std::make_unsigned<T>::type;        // same width, unsigned
    using Acc = typename std::conditional<(sizeof(T) < 4), uint32_t, U>::type;  // accumulators
    static __host__ __device__ __forceinline__ T add(T a, T b) { return (T)((U)a + (U)b); }
    static __host__ __device__ __forceinline__ T sub(T a, T b) { return (T)((U)a - (U)b); }
    static __host__ __device__ __forceinline__ T mul(T a, T b) { return (T)((U)a * (U)b); }
    static __host__ __device__ __forceinline__ T div(T a, T b) { return (T)(a / b); }
    static __host__ __device__ __forceinline__ Acc to_acc(T a) { return (Acc)(U)a; }    // zero- or sign-agnostic: mod 2^n
    static __host__ __device__ __forceinline__ Acc widen(T a) { return (Acc)a; }        // value-preserving (sign-extends)
    static __host__ __device__ __forceinline__ T from_acc(Acc a) { return (T)a; }
    static __host__ __device__ __forceinline__ Acc fma(T a, T b, Acc c) { return (Acc)a * (Acc)b + c; }
    static __host__ __device__ __forceinline__ bool ne(T a, T b) { return a != b; }
};
template <> struct Arith<float> {
    using U = float; using Acc = float;
    static __host__ __device__ __forceinline__ float add(float a, float b) { return a + b; }
    static __host__ __device__ __forceinline__ float sub(float a, float b) { return a - b; }
    static __host__ __device__ __forceinline__ float mul(float a, float b) { return a * b; }
    static __host__ __device__ __forceinline__ float div(float a, float b) { return a / b; }
    static __host__ __device__ __forceinline__ float widen(float a) { return a; }
    static __host__ __device__ __forceinline__ float from_acc(float a) { return a; }
    static __device__ __forceinline__ float fma(float a, float b, float c) { return fmaf(a, b, c); }
    static __host__ __device__ __forceinline__ bool ne(float a, float b) { return a != b; }
};
template <> struct Arith<double> {
    using U = double; using Acc = double;
    static __host__ __device__ __forceinline__ double add(double a, double b) { return a + b; }
    static __host__ __device__ __forceinline__ double sub(double a, double b) { return a - b; }
    static __host__ __device__ __forceinline__ double mul(double a, double b) { return a * b; }
    static __host__ __device__ __forceinline__ double div(double a, double b) { return a / b; }
    static __host__ __device__ __forceinline__ double widen(double a) { return a; }
    static __host__ __device__ __forceinline__ double from_acc(double a) { return a; }
    static __device__ __forceinline__ double fma(double a, double b, double c) { return ::fma(a, b, c); }
    static __host__ __device__ __forceinline__ bool ne(double a, double b) { return a != b; }
};

// 8-bit quotients through the FP32 pipe: floor((|a| + 0.5) / |b|) == floor(|a| / |b|) for integers, and the half
// keeps the operand at least 1/510 away from every integer, far outside the error of the approximate divide, so the
// truncated result is exact.  (The integer divide sequence is ~3x longer and made 8-bit `/` compute bound.)
__device__ __forceinline__ uint8_t la_div_u8(uint8_t a, uint8_t b) {
    return (uint8_t)(unsigned)__fdividef((float)a + 0.5f, (float)b);
}
__device__ __forceinline__ int8_t la_div_i8(int8_t a, int8_t b) {
    const int ia = a, ib = b;
    const int q = (int)__fdividef((float)abs(ia) + 0.5f, (float)abs(ib));
    return (int8_t)(((ia ^ ib) < 0) ? -q : q);
}

template <typename T, int OP> struct ApplyOp {
    static __device__ __forceinline__ T apply(T a, T b) {
        if (OP == LA_ADD) return Arith<T>::add(a, b);
        if (OP == LA_SUB) return Arith<T>::sub(a, b);
        if (OP == LA_MUL) return Arith<T>::mul(a, b);
        return Arith<T>::div(a, b);
    }
};
template <> struct ApplyOp<uint8_t, LA_DIV> { static __device__ __forceinline__ uint8_t apply(uint8_t a, uint8_t b) { return la_div_u8(a, b); } };
template <> struct ApplyOp<int8_t, LA_DIV> { static __device__ __forceinline__ int8_t apply(int8_t a, int8_t b) { return la_div_i8(a, b); } };

// dtype -> C++ type dispatch.  F is a generic lambda taking a value-initialised T.
#define LA_DISPATCH_DTYPE(dtype, T, ...)                                         \
    switch (dtype) {                                                             \
    case LA_U8:  { using T = uint8_t;  __VA_ARGS__; } break;                     \
    case LA_I8:  { using T = int8_t;   __VA_ARGS__; } break;                     \
    case LA_U16: { using T = uint16_t; __VA_ARGS__; } break;                     \
    case LA_I16: { using T = int16_t;  __VA_ARGS__; } break;                     \
    case LA_U32: { using T = uint32_t; __VA_ARGS__; } break;                     \
    case LA_I32: { using T = int32_t;  __VA_ARGS__; } break;                     \
    case LA_U64: { using T = uint64_t; __VA_ARGS__; } break;                     \
    case LA_I64: { using T = int64_t;  __VA_ARGS__; } break;                     \
    case LA_F32: { using T = float;    __VA_ARGS__; } break;                     \
    case LA_F64: { using T = double;   __VA_ARGS__; } break;                     \
    default: return la_fail(LA_ERR_ARG, "unknown dtype %d", (int)(dtype));       \
    }

static inline bool la_valid_dtype(int d) { return d >= LA_U8 && d <= LA_F64; }
static __host__ __device__ inline bool la_aligned16(const void* p) { return (((uintptr_t)p) & 15u) == 0; }

// ---------------------------------------------------------------- 128-bit streaming access
// One 16-byte packet of T.  Streaming (evict-first) loads/stores: every hot-path operand of the
// HBM-bound kernels is touched exactly once.
template <typename T> struct alignas(16) Pack {
    static constexpr int N = 16 / sizeof(T);
    T v[N];
};
template <typename T> __device__ __forceinline__ Pack<T> ld_stream(const Pack<T>* p) {
    uint4 r = __ldcs(reinterpret_cast<const uint4*>(p));
    return *reinterpret_cast<Pack<T>*>(&r);
}
template <typename T> __device__ __forceinline__ void st_stream(Pack<T>* p, const Pack<T>& x) {
    __stcs(reinterpret_cast<uint4*>(p), *reinterpret_cast<const uint4*>(&x));
}
template <typename T> __device__ __forceinline__ Pack<T> ld_cached(const Pack<T>* p) {
    uint4 r = __ldg(reinterpret_cast<const uint4*>(p));
    return *reinterpret_cast<Pack<T>*>(&r);
}

// ---------------------------------------------------------------- warp / block reductions
template <typename A> __device__ __forceinline__ A warp_sum(A v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
// All threads of the block must call; result valid in thread 0.  `smem` holds >= 32 A's.
// Fixed shape => deterministic floating-point result for a given block size.
template <typename A> __device__ __forceinline__ A block_sum(A v, A* smem) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    if (lane == 0) smem[warp] = v;
    __syncthreads();
    A r = A(0);
    if (warp == 0) {
        r = lane < nwarps ? smem[lane] : A(0);
        r = warp_sum(r);
    }
    __syncthreads();
    return r;
}
