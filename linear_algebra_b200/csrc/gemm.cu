// csrc/gemm.cu -- Matrix x Matrix.
//
// Replaces {T}_mul_mat_mat (src/kernels.cl:161-169) and its serial {T}_dot_product
// (src/kernels.cl:134-140), launched by mul_helper (src/matrix.rs:240-265).  The reference
// runs one work-item per C element with two global loads per multiply-add and no reuse.
//
// la_mul_mat_mat dispatches over five engines (la_mul_mat_mat_impl at the bottom of this file):
//  * gemm_tc.cu    3xTF32 tcgen05 tensor-core engine: f32, >= 512^3 multiply-adds, operands in range (LA_GEMM_AUTO)
//  * gemm_limb.cu  exact byte-limb products on the INT8 tensor cores: 8/16/32-bit integers, >= 512^3 multiply-adds
//  * gemm_tma_kernel  (below; 4- and 8-byte element types, M and N multiples of 16 bytes' worth of elements, 16-byte
//      aligned bases; what LA_GEMM_SIMT forces, what mid-sized or out-of-range f32 products get, and where f64 / 64-bit
//      integer products run):
//      TMA-fed, shared-memory-staged, register-tiled SIMT kernel.  A dedicated producer warpgroup issues
//      cp.async.bulk.tensor.2d loads of a [BK x BM] slab of A^T and a [BK x BN] slab of B into a
//      STAGES-deep ring guarded by full/empty mbarriers; 8 consumer warps (setmaxnreg 240) read both
//      operands k-major with conflict-free LDS.128, fragments double-buffered one k-step ahead across
//      slab boundaries, and keep a TM x TN accumulator tile per thread.  f32 uses the sm_100 packed FFMA2
//      (two FMAs per issued instruction, scalar-broadcast operand form); integers use IMAD and wrap.  Each
//      C element is accumulated k-ascending with one fused multiply-add per k -- the reference's order
//      (kernels.cl:136-138) with contraction.  A^T is produced once per call by the transpose kernel into
//      scratch (8 B/element of A, <1% of the GEMM time at 8192^2); edge tiles rely on TMA's out-of-bounds
//      zero fill.
//  * zero-padding path: large products with ragged sizes / unaligned bases are copied into padded, aligned
//      temporaries and run on the engines above.
//  * gemm_generic_kernel (every dtype, any shape / alignment; small problems): 64x64x16
//      shared-memory tiling with guarded scalar loads.  8/16-bit integers accumulate in 32 bits and truncate.
//
// Roofline: FP32 FMA pipe (f32), FP64 pipe (f64), integer multiply-add pipe (integers).  Work: 2*M*N*K flop.
#include "common.cuh"
#include "../../include/linalg_b200_tuning.h"
#include <cuda.h>
#include <cstdlib>
#include <mutex>

la_status la_transpose_raw32(uint32_t* out, const uint32_t* in, size_t R, size_t C, cudaStream_t st);   // transpose.cu
la_status la_transpose_raw64(uint64_t* out, const uint64_t* in, size_t R, size_t C, cudaStream_t st);

namespace {

// ------------------------------------------------------------------ generic engine
constexpr int GG_BM = 64, GG_BN = 64, GG_BK = 16, GG_THREADS = 256;

template <typename T>
__global__ void __launch_bounds__(GG_THREADS)
gemm_generic_kernel(T* __restrict__ C, const T* __restrict__ A, const T* __restrict__ B, size_t M, size_t K, size_t N, unsigned tiles_n) {
    using Acc = typename Arith<T>::Acc;
    __shared__ T As[GG_BK][GG_BM + 4];     // k-major
    __shared__ T Bs[GG_BK][GG_BN + 4];
    const size_t m0 = (size_t)(blockIdx.x / tiles_n) * GG_BM, n0 = (size_t)(blockIdx.x % tiles_n) * GG_BN;
    const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;     // 16 x 16 threads, 4 x 4 outputs each
    Acc acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = Acc(0);

    for (size_t k0 = 0; k0 < K; k0 += GG_BK) {
        // A tile: 64 rows x 16 k  -> As[k][m];  B tile: 16 k x 64 cols -> Bs[k][n]
#pragma unroll
        for (int idx = threadIdx.x; idx < GG_BM * GG_BK; idx += GG_THREADS) {
            const int m = idx / GG_BK, k = idx % GG_BK;
            As[k][m] = (m0 + m < M && k0 + k < K) ? A[(m0 + m) * K + k0 + k] : T(0);
        }
#pragma unroll
        for (int idx = threadIdx.x; idx < GG_BK * GG_BN; idx += GG_THREADS) {
            const int k = idx / GG_BN, n = idx % GG_BN;
            Bs[k][n] = (k0 + k < K && n0 + n < N) ? B[(k0 + k) * N + n0 + n] : T(0);
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < GG_BK; ++k) {
            T a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = As[k][ty * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = Bs[k][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = Arith<T>::fma(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const size_t m = m0 + ty * 4 + i, n = n0 + tx * 4 + j;
            if (m < M && n < N) C[m * N + n] = Arith<T>::from_acc(acc[i][j]);
        }
}

template <typename T>
la_status gemm_generic(T* C, const T* A, const T* B, size_t M, size_t K, size_t N, cudaStream_t st) {
    const size_t tm = (M + GG_BM - 1) / GG_BM, tn = (N + GG_BN - 1) / GG_BN;
    if (tm * tn > 0x7fffffffull) return la_fail(LA_ERR_UNSUPPORTED, "mul_mat_mat: %zu x %zu output is too large", M, N);
    gemm_generic_kernel<T><<<(unsigned)(tm * tn), GG_THREADS, 0, st>>>(C, A, B, M, K, N, (unsigned)tn);
    LA_LAUNCH_CHECK();
    return LA_OK;
}

// ------------------------------------------------------------------ TMA-fed SIMT engine
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// Bounded wait: a broken pipeline traps instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done = 0;
    for (uint32_t spin = 0; !done; ++spin) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
        if (!done && spin > (1u << 22)) __trap();
    }
}
// Non-blocking probe (test_wait never suspends the thread).
__device__ __forceinline__ uint32_t mbar_test(uint64_t* bar, uint32_t parity) {
    uint32_t done;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return done;
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int x, int y) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(x), "r"(y) : "memory");
}

template <typename T> struct Mac;       // multiply-accumulate on the 4- and 8-byte element types
template <> struct Mac<double> {
    using Acc = double;
    static __device__ __forceinline__ void mac(double& c, double a, double b) { c = fma(a, b, c); }
};
template <> struct Mac<uint64_t> {
    using Acc = uint64_t;
    static __device__ __forceinline__ void mac(uint64_t& c, uint64_t a, uint64_t b) { c += a * b; }
};
template <> struct Mac<float> {
    using Acc = float;
    static __device__ __forceinline__ void mac(float& c, float a, float b) { c = fmaf(a, b, c); }
};
template <> struct Mac<uint32_t> {
    using Acc = uint32_t;
    static __device__ __forceinline__ void mac(uint32_t& c, uint32_t a, uint32_t b) { c += a * b; }
};

// Geometry: lanes are 4 (m) x 8 (n).  With CH = 16 / sizeof(T) elements per 16-byte chunk (4 for the
// 4-byte types, 2 for the 8-byte ones), a thread's TM rows are TM/CH chunks of CH rows, 4*CH rows
// apart; its TN columns are TN/CH chunks of CH columns, 8*CH columns apart.  So an A fragment load
// is a broadcast within each quarter-warp and a B fragment load is 128 contiguous bytes per
// quarter-warp: both conflict-free LDS.128.
template <int TM_, int TN_, int WARPS_M_, int WARPS_N_, int BK_, int STAGES_, int MIN_BLOCKS_, int CONSUMER_REGS_>
struct GemmShape {
    static constexpr int TM = TM_, TN = TN_, WARPS_M = WARPS_M_, WARPS_N = WARPS_N_, BK = BK_, STAGES = STAGES_;
    static constexpr int MIN_BLOCKS = MIN_BLOCKS_, CONSUMER_REGS = CONSUMER_REGS_, PRODUCER_REGS = 24;
    static constexpr int WM = 4 * TM, WN = 8 * TN;
    static constexpr int BM = WARPS_M * WM, BN = WARPS_N * WN;
    static constexpr int CONSUMER_WARPS = WARPS_M * WARPS_N;
    static_assert(CONSUMER_WARPS % 4 == 0, "whole consumer warpgroups");
    // consumer warpgroups + one producer warpgroup (setmaxnreg moves registers between whole
    // warpgroups; a lone extra warp would put one more warp on one SM sub-partition and cap
    // everyone's registers).
    static constexpr int THREADS = (CONSUMER_WARPS + 4) * 32;
    static constexpr int STAGE_ELEMS = (BM + BN) * BK;
    static constexpr size_t smem_bytes(size_t elem_size) { return (size_t)STAGES * STAGE_ELEMS * elem_size; }
    // setmaxnreg trades registers inside the CTA's own launch-time pool: THREADS * (launch regs),
    // launch regs = 65536 / (THREADS * MIN_BLOCKS) rounded down to a multiple of 8.
    static constexpr int LAUNCH_REGS = (65536 / (THREADS * MIN_BLOCKS)) / 8 * 8;
    static_assert(128 * PRODUCER_REGS + CONSUMER_WARPS * 32 * CONSUMER_REGS <= THREADS * LAUNCH_REGS, "CTA register pool");
};

template <typename T, typename S, bool PACKED>
__global__ void __launch_bounds__(S::THREADS, S::MIN_BLOCKS)
gemm_tma_kernel(T* __restrict__ C, const __grid_constant__ CUtensorMap map_at, const __grid_constant__ CUtensorMap map_b,
                int M, int N, int K, int tiles_m) {
    constexpr int TM = S::TM, TN = S::TN, BK = S::BK, BM = S::BM, BN = S::BN, STAGES = S::STAGES;
    constexpr int CH = 16 / (int)sizeof(T);                 // elements per 16-byte chunk
    static_assert(TM % CH == 0 && TN % CH == 0, "thread tile is made of whole 16-byte chunks");
    static_assert(!PACKED || sizeof(T) == 4, "FFMA2 is the f32 path");
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ alignas(8) uint64_t full_bar[STAGES];
    __shared__ alignas(8) uint64_t empty_bar[STAGES];
    T* smem = reinterpret_cast<T*>(smem_raw);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = (blockIdx.x % tiles_m) * BM, n0 = (blockIdx.x / tiles_m) * BN;
    const int num_kb = (K + BK - 1) / BK;

    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], S::CONSUMER_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();

    if (warp >= S::CONSUMER_WARPS) {
        // ---------------- producer warpgroup: gives its registers away, one lane drives TMA
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(S::PRODUCER_REGS));
        if (warp == S::CONSUMER_WARPS && lane == 0) {
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % STAGES;
                const uint32_t ph = (uint32_t)(kb / STAGES) & 1u;
                mbar_wait(&empty_bar[s], ph ^ 1u);
                T* as = smem + (size_t)s * S::STAGE_ELEMS;
                T* bs = as + BM * BK;
                mbar_expect_tx(&full_bar[s], (uint32_t)(S::STAGE_ELEMS * sizeof(T)));
                tma_load_2d(as, &map_at, &full_bar[s], m0, kb * BK);      // A^T slab: BK rows of BM
                tma_load_2d(bs, &map_b, &full_bar[s], n0, kb * BK);       // B slab:   BK rows of BN
            }
        }
        return;
    }
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(S::CONSUMER_REGS));

    // ---------------- consumer warps
    using Acc = typename Mac<T>::Acc;
    const int wm = warp % S::WARPS_M, wn = warp / S::WARPS_M;
    const int mi = lane >> 3, ni = lane & 7;
    const int a_off = wm * S::WM + mi * CH;       // + cm*4*CH
    const int b_off = wn * S::WN + ni * CH;       // + cn*8*CH

    Acc acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = Acc(0);

    // Fragments are double-buffered in registers and loaded one k-step ahead, ACROSS slab
    // boundaries: the next slab's barrier is probed (non-blocking test_wait) half a slab early
    // and its k=0 fragments are fetched before the current slab's last k-step is multiplied, so
    // neither the barrier round trip nor the first LDS latency ever leaves the FMA pipe idle
    // (both warps of a sub-partition reach a slab boundary at the same moment).
    T fa[2][TM], fb[2][TN];
    auto load_frag = [&](int buf, const T* as, const T* bs, int k) {
#pragma unroll
        for (int cm = 0; cm < TM / CH; ++cm)
            *reinterpret_cast<uint4*>(&fa[buf][cm * CH]) = *reinterpret_cast<const uint4*>(as + k * BM + cm * 4 * CH);
#pragma unroll
        for (int cn = 0; cn < TN / CH; ++cn)
            *reinterpret_cast<uint4*>(&fb[buf][cn * CH]) = *reinterpret_cast<const uint4*>(bs + k * BN + cn * 8 * CH);
    };
    static_assert(BK % 2 == 0, "fragment double-buffering alternates per k-step");

    mbar_wait(&full_bar[0], 0);
    load_frag(0, smem + a_off, smem + BM * BK + b_off, 0);

    for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % STAGES;
        const T* as = smem + (size_t)s * S::STAGE_ELEMS + a_off;
        const T* bs = smem + (size_t)s * S::STAGE_ELEMS + BM * BK + b_off;
        const int kn = kb + 1;
        const int sn = kn % STAGES;
        const uint32_t phn = (uint32_t)(kn / STAGES) & 1u;
        const bool has_next = kn < num_kb;
        uint32_t next_ready = 0;
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            if (k == BK / 2 && has_next) next_ready = mbar_test(&full_bar[sn], phn);
            if (k + 1 < BK) {
                load_frag((k + 1) & 1, as, bs, k + 1);
            } else if (has_next) {
                if (!next_ready) mbar_wait(&full_bar[sn], phn);
                load_frag(0, smem + (size_t)sn * S::STAGE_ELEMS + a_off, smem + (size_t)sn * S::STAGE_ELEMS + BM * BK + b_off, 0);
            }
            const T* a = fa[k & 1];
            const T* b = fb[k & 1];
            if constexpr (PACKED) {
                // f32 only: two FMAs per issued FFMA2 (its scalar-broadcast operand form takes
                // the unpaired fragment).  Pairs run along the longer side of the thread tile.
                if constexpr (TM >= TN) {
#pragma unroll
                    for (int j = 0; j < TN; ++j) {
                        const float2 bb = make_float2(b[j], b[j]);
#pragma unroll
                        for (int i = 0; i < TM; i += 2) {
                            float2 c = make_float2(acc[i][j], acc[i + 1][j]);
                            c = __ffma2_rn(make_float2(a[i], a[i + 1]), bb, c);
                            acc[i][j] = c.x; acc[i + 1][j] = c.y;
                        }
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < TM; ++i) {
                        const float2 aa = make_float2(a[i], a[i]);
#pragma unroll
                        for (int j = 0; j < TN; j += 2) {
                            float2 c = make_float2(acc[i][j], acc[i][j + 1]);
                            c = __ffma2_rn(aa, make_float2(b[j], b[j + 1]), c);
                            acc[i][j] = c.x; acc[i][j + 1] = c.y;
                        }
                    }
                }
            } else {
#pragma unroll
                for (int i = 0; i < TM; ++i)
#pragma unroll
                    for (int j = 0; j < TN; ++j) Mac<T>::mac(acc[i][j], a[i], b[j]);
            }
        }
        // every fragment of slab `kb` has been consumed by the FMAs above: hand the slot back
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[s]);
    }

    // ---------------- epilogue: 16-byte stores, 128 contiguous bytes per quarter-warp
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        const int m = m0 + a_off + (i / CH) * (4 * CH) + (i % CH);
        if (m < M) {
#pragma unroll
            for (int cn = 0; cn < TN / CH; ++cn) {
                const int n = n0 + b_off + cn * 8 * CH;
                if (n < N) {   // N % CH == 0: a chunk of CH columns is wholly inside or outside
                    uint4 v;
                    memcpy(&v, &acc[i][cn * CH], 16);
                    __stcs(reinterpret_cast<uint4*>(C + (size_t)m * N + n), v);
                }
            }
        }
    }
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

la_status get_encode(EncodeTiledFn* out) {
    static EncodeTiledFn fn = nullptr;
    static std::mutex mu;
    std::lock_guard<std::mutex> lk(mu);
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        LA_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
        if (q != cudaDriverEntryPointSuccess || !p) return la_fail(LA_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
        fn = (EncodeTiledFn)p;
    }
    *out = fn;
    return LA_OK;
}

// 2-D row-major tensor of 4- or 8-byte elements: `rows` x `cols`, box `box_rows` x `box_cols`.
la_status make_map(CUtensorMap* map, const void* base, size_t rows, size_t cols, int box_rows, int box_cols, bool is_float, size_t elem_size) {
    EncodeTiledFn enc; LA_TRY(get_encode(&enc));
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)cols * elem_size};
    cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    const CUtensorMapDataType type = elem_size == 8 ? (is_float ? CU_TENSOR_MAP_DATA_TYPE_FLOAT64 : CU_TENSOR_MAP_DATA_TYPE_UINT64)
                                                    : (is_float ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_UINT32);
    CUresult r = enc(map, type, 2, const_cast<void*>(base),
                     dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return la_fail(LA_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d (rows=%zu cols=%zu box=%dx%d)", (int)r, rows, cols, box_rows, box_cols);
    return LA_OK;
}

int g_gemm_variant = 0;     // la_gemm_set_variant (tuning hook, not part of the reference surface)

template <typename T, typename S, bool PACKED>
la_status gemm_tma_launch(T* C, const T* At, const T* B, size_t M, size_t K, size_t N, cudaStream_t st) {
    CUtensorMap map_at, map_b;
    static_assert(S::smem_bytes(sizeof(T)) <= 227 * 1024, "shared-memory ring fits one SM");
    static_assert(S::BM <= 256 && S::BN <= 256, "TMA box dimension limit");
    constexpr bool is_float = std::is_floating_point<T>::value;
    LA_TRY(make_map(&map_at, At, K, M, S::BK, S::BM, is_float, sizeof(T)));
    LA_TRY(make_map(&map_b, B, K, N, S::BK, S::BN, is_float, sizeof(T)));
    auto kern = gemm_tma_kernel<T, S, PACKED>;
    constexpr size_t smem = S::smem_bytes(sizeof(T));
    LA_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const size_t tm = (M + S::BM - 1) / S::BM, tn = (N + S::BN - 1) / S::BN;
    kern<<<(unsigned)(tm * tn), S::THREADS, smem, st>>>(C, map_at, map_b, (int)M, (int)N, (int)K, (int)tm);
    LA_LAUNCH_CHECK();
    return LA_OK;
}

//                        TM  TN  WM WN  BK ST  blocks/SM  consumer regs
using Shape_16x8  = GemmShape<16,  8, 4, 2, 16, 4, 1, 240>;    // 256 x 128 block
using Shape_16x8b = GemmShape<16,  8, 2, 4, 16, 4, 1, 240>;    // 128 x 256 block
using Shape_8x16  = GemmShape< 8, 16, 4, 2, 16, 4, 1, 240>;    // 128 x 256 block
using Shape_16x8k = GemmShape<16,  8, 4, 2,  8, 8, 1, 240>;    // 256 x 128 block, BK 8
using Shape_8x8   = GemmShape< 8,  8, 4, 2, 16, 4, 2, 104>;    // 128 x 128 block, two per SM
using Shape_8x8s  = GemmShape< 8,  8, 4, 2, 16, 6, 1, 240>;    // 128 x 128 block, one per SM
using Shape_8x8w  = GemmShape< 8,  8, 6, 2, 16, 4, 1, 160>;    // 192 x 128 block, 12 consumer warps (3 per scheduler)
using Shape_8x8v  = GemmShape< 8,  8, 4, 3, 16, 4, 1, 160>;    // 128 x 192 block, 12 consumer warps
using Shape_16x8d = GemmShape<16,  8, 4, 2, 32, 3, 1, 240>;    // 256 x 128 block, BK 32

// 8-byte element types: thread tile 8 x 8 (64 accumulators = 128 registers)
using Shape64_8x8  = GemmShape< 8,  8, 4, 2, 16, 4, 1, 240>;    // 128 x 128 block, 8 consumer warps, 4 x 32 KB stages
using Shape64_8x8d = GemmShape< 8,  8, 4, 2, 32, 3, 1, 240>;    // 128 x 128 block, BK 32, 3 x 64 KB stages
using Shape64_8x8k = GemmShape< 8,  8, 4, 2,  8, 8, 1, 240>;    // 128 x 128 block, BK 8, 8 x 16 KB stages
using Shape64_8x4  = GemmShape< 8,  4, 4, 2, 16, 4, 1, 240>;    // 128 x 64 block: the 64-bit integer multiply-add needs the registers an 8 x 8 tile would take
using Shape64_4x8  = GemmShape< 4,  8, 4, 2, 16, 4, 1, 240>;    // 64 x 128 block

template <typename T>
la_status gemm_tma(T* C, const T* A, const T* B, size_t M, size_t K, size_t N, cudaStream_t st) {
    // A^T (K x M) into scratch
    void* scratch; LA_TRY(la_scratch(st, M * K * sizeof(T), &scratch));
    T* At = (T*)((char*)scratch + LA_SCRATCH_ZEROED);
    if constexpr (sizeof(T) == 8) {
        LA_TRY(la_transpose_raw64((uint64_t*)At, (const uint64_t*)A, M, K, st));
        switch (g_gemm_variant) {
        default:
        case 0:                                                                          // f64: 8 x 8 tile, BK 32 (28.7 TFLOP/s at 4096^3; generic engine 12.1)
            if (std::is_same<T, double>::value) return gemm_tma_launch<T, Shape64_8x8d, false>(C, At, B, M, K, N, st);
            return gemm_tma_launch<T, Shape64_8x4, false>(C, At, B, M, K, N, st);
        case 1: return gemm_tma_launch<T, Shape64_8x8, false>(C, At, B, M, K, N, st);
        case 2: return gemm_tma_launch<T, Shape64_8x8k, false>(C, At, B, M, K, N, st);
        case 3: return gemm_tma_launch<T, Shape64_8x4, false>(C, At, B, M, K, N, st);
        case 4: return gemm_tma_launch<T, Shape64_4x8, false>(C, At, B, M, K, N, st);
        case 5: return gemm_tma_launch<T, Shape64_8x8d, false>(C, At, B, M, K, N, st);
        case 13: return std::is_same<T, double>::value ? gemm_tma_launch<T, Shape64_8x8d, false>(C, At, B, M, K, N, st) : gemm_tma_launch<T, Shape64_8x4, false>(C, At, B, M, K, N, st);
        }
    } else {
    LA_TRY(la_transpose_raw32((uint32_t*)At, (const uint32_t*)A, M, K, st));
    constexpr bool F = std::is_same<T, float>::value;
    switch (g_gemm_variant) {
    default:
    case 0:                                                                      // 256 x 128 tile; f32: BK 32 (65.0 vs 63.5 TFLOP/s at 8192^2),
        if (F) return gemm_tma_launch<T, Shape_16x8d, F>(C, At, B, M, K, N, st);  // integers: BK 16 (35.7 vs 30.4 Top/s: the IMAD body spills at BK 32)
        return gemm_tma_launch<T, Shape_16x8, F>(C, At, B, M, K, N, st);
    case 12: return gemm_tma_launch<T, Shape_16x8, F>(C, At, B, M, K, N, st);
    case 13: return F ? gemm_tma_launch<T, Shape_16x8d, F>(C, At, B, M, K, N, st) : gemm_tma_launch<T, Shape_16x8, F>(C, At, B, M, K, N, st);   // the default shapes, at any size
    case 1: return gemm_tma_launch<T, Shape_16x8, false>(C, At, B, M, K, N, st);
    case 2: return gemm_tma_launch<T, Shape_8x8, F>(C, At, B, M, K, N, st);
    case 3: return gemm_tma_launch<T, Shape_8x8, false>(C, At, B, M, K, N, st);
    case 4: return gemm_tma_launch<T, Shape_8x16, F>(C, At, B, M, K, N, st);
    case 5: return gemm_tma_launch<T, Shape_16x8b, F>(C, At, B, M, K, N, st);
    case 6: return gemm_tma_launch<T, Shape_16x8k, F>(C, At, B, M, K, N, st);
    case 7: return gemm_tma_launch<T, Shape_8x8s, F>(C, At, B, M, K, N, st);
    case 8: return gemm_tma_launch<T, Shape_8x16, false>(C, At, B, M, K, N, st);
    case 9: return gemm_tma_launch<T, Shape_8x8w, F>(C, At, B, M, K, N, st);
    case 10: return gemm_tma_launch<T, Shape_8x8v, F>(C, At, B, M, K, N, st);
    }
    }
}

// LA_GEMM_MODE=simt|3xtf32|auto in the environment selects the f32 engine at load time.
int initial_gemm_mode() {
    const char* e = getenv("LA_GEMM_MODE");
    if (e && (e[0] == 's' || e[0] == 'S')) return LA_GEMM_SIMT;
    if (e && (e[0] == '3' || e[0] == 't' || e[0] == 'T')) return LA_GEMM_3XTF32;
    return LA_GEMM_AUTO;
}
int g_gemm_mode = initial_gemm_mode();
// What this thread's last product did (la_gemm_last_engine / la_gemm_last_padded / la_mul_mat_mat_info): thread-local, so
// callers on different threads and streams do not overwrite each other's answer.
thread_local int g_last_padded = 0;        // went through zero-padded temporaries
thread_local int g_last_engine = -1;       // 0 generic, 1 TMA SIMT, 2 3xTF32, 3 int8-limb tensor cores

}  // namespace

la_status la_gemm_3xtf32(float* C, const float* A, const float* B, size_t M, size_t K, size_t N, cudaStream_t st, bool* handled, const la_gemm_prepared* prep);   // gemm_tc.cu
la_status la_gemm_limb(int dtype, void* C, const void* A, const void* B, size_t M, size_t K, size_t N, cudaStream_t st, bool* handled, const la_gemm_prepared* prep);  // gemm_limb.cu
la_status la_tc_prepare_b(la_gemm_prepared* h, cudaStream_t st);      // gemm_tc.cu
la_status la_limb_prepare_b(la_gemm_prepared* h, cudaStream_t st);    // gemm_limb.cu

extern "C" la_status la_set_gemm_mode(int mode) {
    if (mode != LA_GEMM_SIMT && mode != LA_GEMM_3XTF32 && mode != LA_GEMM_AUTO) return la_fail(LA_ERR_ARG, "unknown GEMM mode %d", mode);
    g_gemm_mode = mode;
    return LA_OK;
}
extern "C" la_status la_get_gemm_mode(int* mode) {
    if (!mode) return la_fail(LA_ERR_ARG, "mode is NULL");
    *mode = g_gemm_mode;
    return LA_OK;
}
// Tuning hook used by tools/ and bench.py --sweep: picks the tile shape of the TMA engine
// (0 = default shape, used from the size where it beats the generic kernel; k > 0 = that shape at any size; 13 = the default
// shapes at any size).  -1 forces the generic engine.
extern "C" la_status la_gemm_set_variant(int v) { g_gemm_variant = v; return LA_OK; }
extern "C" int la_gemm_last_engine(void) { return g_last_engine; }
extern "C" int la_gemm_last_padded(void) { return g_last_padded; }

la_status la_mul_mat_mat_impl(int dtype, void* C, const void* A, const void* B, size_t M, size_t K, size_t N, void* stream, int mode, const la_gemm_prepared* prep = nullptr);

extern "C" la_status la_mul_mat_mat(int dtype, void* C, const void* A, const void* B, size_t M, size_t K, size_t N, void* stream) {
    return la_mul_mat_mat_impl(dtype, C, A, B, M, K, N, stream, g_gemm_mode);
}

extern "C" la_status la_mul_mat_mat_info(int dtype, void* C, const void* A, const void* B, size_t M, size_t K, size_t N, void* stream, la_gemm_info* info) {
    g_last_engine = -1; g_last_padded = 0;
    const la_status s = la_mul_mat_mat_impl(dtype, C, A, B, M, K, N, stream, g_gemm_mode);
    if (info) { info->engine = g_last_engine; info->padded = g_last_padded; }
    return s;
}

// ---------------------------------------------------------------- a B shared by many products
// What a product does to B before its main kernel -- the tensor engines' split / transposed planes -- depends on B alone.
// la_gemm_prepare_b does it once; la_mul_mat_mat_prepared then runs la_mul_mat_mat's engine selection with that part
// skipped: the same kernels on the same bits, so results are bit-identical to la_mul_mat_mat(A, B).  (The row-block
// sharded multi-GPU product, B replicated on every rank, is the case it is for: SURVEY.md section 8(e) "B broadcast once;
// report steady-state and cold separately".)
extern "C" la_status la_gemm_prepare_b(int dtype, const void* B, size_t K, size_t N, void* stream, la_gemm_prepared** handle) {
    if (!handle) return la_fail(LA_ERR_ARG, "la_gemm_prepare_b: handle is NULL");
    *handle = nullptr;
    if (!la_valid_dtype(dtype)) return la_fail(LA_ERR_ARG, "la_gemm_prepare_b: unknown dtype %d", dtype);
    if (!B && K && N) return la_fail(LA_ERR_ARG, "la_gemm_prepare_b: B is NULL");
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    la_gemm_prepared* h = new la_gemm_prepared();
    h->dtype = dtype; h->device = info->device; h->K = K; h->N = N; h->B = B;
    la_status s = LA_OK;
    if (K && N) {
        if (dtype == LA_F32) s = la_tc_prepare_b(h, (cudaStream_t)stream);
        else if (dtype != LA_F64 && la_dtype_size(dtype) <= 4) s = la_limb_prepare_b(h, (cudaStream_t)stream);
    }
    if (s != LA_OK) { delete h; return s; }
    *handle = h;
    return LA_OK;
}
extern "C" la_status la_gemm_prepared_free(la_gemm_prepared* handle) {
    if (!handle) return LA_OK;
    // products that use the planes may still be queued: a plain cudaFree waits for the device
    if (handle->tc_hi) cudaFree(handle->tc_hi);         // (tc_lo is the second half of the same block)
    if (handle->f16_planes) cudaFree(handle->f16_planes);       // (exponents and list live in the same block)
    if (handle->limb_planes) cudaFree(handle->limb_planes);
    (void)cudaGetLastError();
    delete handle;
    return LA_OK;
}
extern "C" la_status la_mul_mat_mat_prepared(void* C, const void* A, const la_gemm_prepared* handle, size_t M, void* stream) {
    if (!handle) return la_fail(LA_ERR_ARG, "la_mul_mat_mat_prepared: handle is NULL");
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    if (info->device != handle->device) return la_fail(LA_ERR_ARG, "la_mul_mat_mat_prepared: the handle was prepared on device %d, current device is %d", handle->device, info->device);
    return la_mul_mat_mat_impl(handle->dtype, C, A, handle->B, M, handle->K, handle->N, stream, g_gemm_mode, handle);
}

int la_gemm_current_mode() { return g_gemm_mode; }

// `mode`: LA_GEMM_SIMT / LA_GEMM_3XTF32 / LA_GEMM_AUTO for this call (gemm_host.cu forces SIMT per block).
la_status la_mul_mat_mat_impl(int dtype, void* C, const void* A, const void* B, size_t M, size_t K, size_t N, void* stream, int mode, const la_gemm_prepared* prep) {
    if (!la_valid_dtype(dtype)) return la_fail(LA_ERR_ARG, "la_mul_mat_mat: unknown dtype %d", dtype);
    if (M == 0 || N == 0) return LA_OK;
    if (!C || !A || !B) return la_fail(LA_ERR_ARG, "la_mul_mat_mat: NULL operand");
    if (C == A || C == B) return la_fail(LA_ERR_ARG, "la_mul_mat_mat: C must not alias A or B");
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    cudaStream_t st = (cudaStream_t)stream;
    if (K == 0) { LA_CUDA(cudaMemsetAsync(C, 0, M * N * la_dtype_size(dtype), st)); return LA_OK; }

    const size_t chunk = 16 / la_dtype_size(dtype);        // elements per 16 bytes: TMA row pitches and the 16-byte C stores need whole chunks
    const bool tma_ok = la_dtype_size(dtype) >= 4 && g_gemm_variant >= 0 && M % chunk == 0 && N % chunk == 0 && la_aligned16(A) && la_aligned16(B) && la_aligned16(C) &&
                        M >= 64 && N >= 64 && K >= 16 && M < (1u << 31) && N < (1u << 31) && K < (1u << 31);
    g_last_padded = 0;
    const double macs = (double)M * (double)N * (double)K;
    const bool big = macs >= 134217728.0;                                   // 512^3 multiply-adds
    // The TMA-fed SIMT kernel gives one CTA a 256 x 128 (8-byte types: 128 x 128) tile: with fewer tiles than SMs it runs at
    // the speed of one SM per tile (K x 0.135 us whatever M and N), and the generic kernel's 64 x 64 tiles fill the GPU
    // better.  Measured crossovers: f32 1536^3 (72 tiles), i32 ~1800^3, f64 ~900^3 (tools/gemm_generic_vs_tma.py); e.g.
    // 512^3 f32 84.5 -> 32.1 us, i32 145 -> 37 us.  An explicit tile-shape variant (tuning hook) always takes the TMA kernel.
    const size_t tma_tiles = la_dtype_size(dtype) == 8 ? ((M + 127) / 128) * ((N + 127) / 128) : ((M + 255) / 256) * ((N + 127) / 128);
    const bool tma_worth = g_gemm_variant > 0 || tma_tiles >= (la_dtype_size(dtype) == 8 ? 48u : 96u);
    // AUTO: the tensor engine for problems big enough to amortise its pre-pass (its floor is ~27 us: it overtakes the generic
    // kernel at 512^3, 27.4 against 32.2 us)
    // ... and only from K = LA_GEMM_AUTO_MIN_K: a shorter product is one to four accumulation chains inside the tensor cores,
    // whose FP32 accumulation truncates, and its worst error over many outputs can land slightly above the reference order's
    // own (4096 x 64 x 4096: 3.1e-7 of sum|a||b| against 3.0e-7), which the gate of DESIGN.md section 4 does not allow.  Those
    // stay on the k-ascending SIMT engine (the tensor engine gains little there anyway); LA_GEMM_3XTF32 remains the opt-in.
    if (dtype == LA_F32 && (mode == LA_GEMM_3XTF32 || (mode == LA_GEMM_AUTO && big && K >= LA_TC_AUTO_MIN_K))) {
        bool handled = false;
        LA_TRY(la_gemm_3xtf32((float*)C, (const float*)A, (const float*)B, M, K, N, st, &handled, prep));
        if (handled) { g_last_engine = 2; return LA_OK; }
    }
    // 8/16/32-bit integers: exact limb products on the INT8 tensor cores (mode SIMT keeps the IMAD engines)
    if (dtype != LA_F32 && dtype != LA_F64 && la_dtype_size(dtype) <= 4 && mode != LA_GEMM_SIMT && big) {
        bool handled = false;
        LA_TRY(la_gemm_limb(dtype, C, A, B, M, K, N, st, &handled, prep));
        if (handled) { g_last_engine = 3; return LA_OK; }
    }
    // Ragged sizes or unaligned bases of a large product: instead of dropping to the generic kernel (a 30x
    // cliff), copy the operands into zero-padded, aligned temporaries (every dimension a multiple of 16),
    // run the fast engines on those and copy C back.  Zero padding adds exact zeros to every dot product.
    const size_t es = la_dtype_size(dtype);
    const bool ragged = (M % 16) || (K % 16) || (N % 16) || !la_aligned16(A) || !la_aligned16(B) || !la_aligned16(C);   // false on the padded call
    if (ragged && big && M < (1u << 30) && N < (1u << 30) && K < (1u << 30)) {
        const size_t Mp = (M + 15) / 16 * 16, Kp = (K + 15) / 16 * 16, Np = (N + 15) / 16 * 16;
        void *Ap = nullptr, *Bp = nullptr, *Cp = nullptr;
        LA_CUDA(cudaMallocAsync(&Ap, Mp * Kp * es, st));
        LA_CUDA(cudaMallocAsync(&Bp, Kp * Np * es, st));
        LA_CUDA(cudaMallocAsync(&Cp, Mp * Np * es, st));
        LA_CUDA(cudaMemsetAsync(Ap, 0, Mp * Kp * es, st));
        LA_CUDA(cudaMemsetAsync(Bp, 0, Kp * Np * es, st));
        LA_CUDA(cudaMemcpy2DAsync(Ap, Kp * es, A, K * es, K * es, M, cudaMemcpyDeviceToDevice, st));
        LA_CUDA(cudaMemcpy2DAsync(Bp, Np * es, B, N * es, N * es, K, cudaMemcpyDeviceToDevice, st));
        const la_status ps = la_mul_mat_mat_impl(dtype, Cp, Ap, Bp, Mp, Kp, Np, stream, mode);
        if (ps == LA_OK) LA_CUDA(cudaMemcpy2DAsync(C, N * es, Cp, Np * es, N * es, M, cudaMemcpyDeviceToDevice, st));
        cudaFreeAsync(Ap, st); cudaFreeAsync(Bp, st); cudaFreeAsync(Cp, st);
        g_last_padded = 1;
        return ps;
    }
    if (tma_ok && tma_worth) {
        g_last_engine = 1;
        if (dtype == LA_F32) return gemm_tma<float>((float*)C, (const float*)A, (const float*)B, M, K, N, st);
        if (dtype == LA_F64) return gemm_tma<double>((double*)C, (const double*)A, (const double*)B, M, K, N, st);
        if (la_dtype_size(dtype) == 8) return gemm_tma<uint64_t>((uint64_t*)C, (const uint64_t*)A, (const uint64_t*)B, M, K, N, st);   // i64 == u64 mod 2^64
        return gemm_tma<uint32_t>((uint32_t*)C, (const uint32_t*)A, (const uint32_t*)B, M, K, N, st);   // i32 == u32 mod 2^32
    }
    g_last_engine = 0;
    LA_DISPATCH_DTYPE(dtype, T, return gemm_generic<T>((T*)C, (const T*)A, (const T*)B, M, K, N, st));
    return LA_OK;
}
