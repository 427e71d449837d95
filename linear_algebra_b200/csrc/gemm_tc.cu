// csrc/gemm_tc.cu -- split-f32 tensor-core engine for f32 Matrix x Matrix (LA_GEMM_3XTF32; what LA_GEMM_AUTO picks from 512^3
// multiply-adds up when K >= 256).
//
// Same contract as the SIMT engine in gemm.cu ({T}_mul_mat_mat, src/kernels.cl:161-169), other arithmetic: every f32
// operand is split x = hi + lo with an 11-bit hi, and each product is formed as hi*hi + hi*lo + lo*hi on the 5th-generation
// tensor cores with FP32 accumulation in tensor memory, in one of two operand formats:
//   tf32 form  (256 <= K < 1024; K % 8 != 0; operands outside the other form's range)   hi = tf32(x), lo = tf32(x - hi), three
//              tcgen05.mma kind::tf32 per k-step ("3xTF32"): 3 tensor time units per product
//   f16 form   (K >= 1024)   rows of A / columns of B scaled by powers of two into fp16's range, three fp16 planes (hi, hi * 2^-6,
//              (x - hi) * 2^6), three kind::f16 MMAs at twice the rate: 1.5 units per product, 6 bytes of planes per element;
//              elements far below their row's maximum are left out of the planes and added afterwards in f32; operands with
//              one-product outputs (identity, diagonal, ...) are handed to the exact engine -- see gemm_f16split_2cta_kernel
// The dropped lo*lo term is below 2^-22 relative.  Accumulation runs in chains of 32 .. 256 k inside the tensor cores
// (their FP32 accumulation truncates) and in FP32 registers across chains.  Results stay inside the f32 bar of DESIGN.md
// section 4 -- tests/test_gpu_parity.py gates both forms / every chain length / tile width on it, on random and on
// structured data; la_set_gemm_mode(LA_GEMM_SIMT) switches the engine off.
//
// Kernels in this file (la_gemm_tc_set_variant picks one of the tf32 form's; 0 is the default):
//   -  gemm_f16split_2cta_kernel<BN>      the f16 form: CTA pairs, 4-stage ring of six 8 KB planes, otherwise as the next one
//   0  gemm_3xtf32_2cta_kernel<BN, BF16X>  CTA pairs (tcgen05 cta_group::2), 256 x BN tile per pair (BN = 256 / 224 / 192 by
//                                   wave count), 3-stage TMA ring, work items = (tile, chain range) so the ragged last
//                                   wave can be cut in two along K, banded tile order above 4096 rows; BF16X = the cross
//                                   terms as bf16 MMAs on bf16(hi), bf16(x - hi) planes (2 units; a tuning setting: ~2^-20
//                                   per product, above the gate for one-product outputs)
//   5/3 gemm_3xtf32_chunked_kernel  one CTA per 128x256 tile (with / without B multicast), 64-chains  (tuning references)
//   1/2 gemm_3xtf32_kernel          single accumulation chain (tuning reference; drifts)
// Common structure:
//   pre-pass   tf32 form: split_both_kernel / split_tf32_kernel / split_transpose_kernel -- A's planes and the same planes of B^T
//              (N x K), so both operands are K-major; f16 form: f16_prepass1_kernel (A's rows; B's column maxima and sums),
//              f16_split_transpose_kernel (B^T's planes), f16_tiny_fix_kernel after the product          (all in scratch)
//   TMA warp   cp.async.bulk.tensor.2d of the slabs of a 32-wide k-block (128-byte rows / SWIZZLE_128B for tf32 planes,
//              64-byte rows / SWIZZLE_64B for 16-bit planes) into a ring guarded by full/empty mbarriers
//   MMA warp   owns tensor memory; one elected lane issues the MMAs of a k-block and tcgen05.commit's the stage back
//   epilogue   tcgen05.ld of the FP32 accumulator after every chain, FP32 adds in registers, 16-byte streaming stores to C
//              (atomicAdd of the two halves for a tile cut along K)
// Non-finite inputs: inf * (lo == 0) would be NaN, and tiny values lose their lo part; the pre-passes detect NaN / inf and
// magnitudes outside their form's range ([2^-40, 2^40] for row / column maxima of the f16 form, [2^-100, 2^100] for the tf32
// form) and the call then runs the next engine over the same C, so the tensor engine never changes the result class of a product.
// Roofline: tensor pipe, 1.5 (f16 form) or 3 (tf32 form) tf32-equivalent MMA units per useful product; the f16 form's kernel runs
// at 0.80 - 0.84 of that floor at 8192^3 (tensor pipe 82 % active).
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include "tc_common.cuh"
#include "../../include/linalg_b200_tuning.h"

la_status la_transpose_raw32(uint32_t* out, const uint32_t* in, size_t R, size_t C, cudaStream_t st);   // transpose.cu

namespace {

constexpr int TC_BM = 128;          // UMMA M (cta_group::1)
constexpr int TC_BK = 32;           // 32 f32 = 128 bytes = one swizzle row
constexpr int TC_UMMA_K = 8;        // tf32: 32 bytes of K per instruction
constexpr int TC_THREADS = 192;


// kind::tf32, FP32 accumulate, both operands K-major.
__host__ __device__ constexpr uint32_t umma_idesc_tf32(int M, int N) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}

template <int BN, int STAGES>
struct TcShape {
    static constexpr int A_TILE = TC_BM * TC_BK * 4;          // bytes
    static constexpr int B_TILE = BN * TC_BK * 4;
    static constexpr int STAGE_BYTES = 2 * A_TILE + 2 * B_TILE;
    static constexpr size_t SMEM_BYTES = (size_t)STAGES * STAGE_BYTES + 1024;   // + alignment slack
    static constexpr int TMEM_COLS = BN;                       // power of two >= 32
    static_assert(BN == 128 || BN == 256, "accumulator columns");
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(TC_THREADS, 1)
gemm_3xtf32_kernel(float* __restrict__ C, const __grid_constant__ CUtensorMap map_a_hi, const __grid_constant__ CUtensorMap map_a_lo,
                   const __grid_constant__ CUtensorMap map_b_hi, const __grid_constant__ CUtensorMap map_b_lo,
                   int M, int N, int K, int tiles_m) {
    using S = TcShape<BN, STAGES>;
    extern __shared__ uint8_t smem_raw[];
    __shared__ alignas(8) uint64_t full_bar[STAGES];
    __shared__ alignas(8) uint64_t empty_bar[STAGES];
    __shared__ alignas(8) uint64_t tmem_full_bar;
    __shared__ uint32_t tmem_base_slot;

    // 1024-byte aligned ring (the swizzle pattern is a function of the absolute address)
    const uint32_t ring = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* ring_ptr = smem_raw + (ring - smem_u32(smem_raw));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = (blockIdx.x % tiles_m) * TC_BM, n0 = (blockIdx.x / tiles_m) * BN;
    const int num_kb = (K + TC_BK - 1) / TC_BK;

    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(&tmem_full_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 5) {    // one warp allocates tensor memory and later frees it
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_slot)), "n"(S::TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = tmem_base_slot;

    if (warp == 4) {
        // ---------------- TMA producer
        if (lane == 0) {
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % STAGES;
                const uint32_t ph = (uint32_t)(kb / STAGES) & 1u;
                mbar_wait(&empty_bar[s], ph ^ 1u);
                uint8_t* st = ring_ptr + (size_t)s * S::STAGE_BYTES;
                mbar_expect_tx(&full_bar[s], (uint32_t)S::STAGE_BYTES);
                tma_load_2d(st, &map_a_hi, &full_bar[s], kb * TC_BK, m0);
                tma_load_2d(st + S::A_TILE, &map_a_lo, &full_bar[s], kb * TC_BK, m0);
                tma_load_2d(st + 2 * S::A_TILE, &map_b_hi, &full_bar[s], kb * TC_BK, n0);
                tma_load_2d(st + 2 * S::A_TILE + S::B_TILE, &map_b_lo, &full_bar[s], kb * TC_BK, n0);
            }
        }
    } else if (warp == 5) {
        // ---------------- MMA issuer (one elected lane)
        constexpr uint32_t idesc = umma_idesc_tf32(TC_BM, BN);
        for (int kb = 0; kb < num_kb; ++kb) {
            const int s = kb % STAGES;
            const uint32_t ph = (uint32_t)(kb / STAGES) & 1u;
            mbar_wait(&full_bar[s], ph);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (elect_one()) {
                const uint32_t st = ring + (uint32_t)s * S::STAGE_BYTES;
                const uint64_t a_hi = umma_desc_k_sw128(st), a_lo = umma_desc_k_sw128(st + S::A_TILE);
                const uint64_t b_hi = umma_desc_k_sw128(st + 2 * S::A_TILE), b_lo = umma_desc_k_sw128(st + 2 * S::A_TILE + S::B_TILE);
#pragma unroll
                for (int kk = 0; kk < TC_BK / TC_UMMA_K; ++kk) {
                    const uint64_t adv = (uint64_t)((kk * TC_UMMA_K * 4) >> 4);     // 32 bytes of K per step
                    // small cross terms first, then the leading term
                    umma_tf32(tmem_d, a_hi + adv, b_lo + adv, idesc, (kb | kk) != 0);
                    umma_tf32(tmem_d, a_lo + adv, b_hi + adv, idesc, 1u);
                    umma_tf32(tmem_d, a_hi + adv, b_hi + adv, idesc, 1u);
                }
                umma_commit(&empty_bar[s]);                                         // frees the stage when the MMAs retire
                if (kb == num_kb - 1) umma_commit(&tmem_full_bar);                  // accumulator complete
            }
            __syncwarp();
        }
    } else {
        // ---------------- epilogue warps 0..3: TMEM lane quadrant = warp index
        mbar_wait(&tmem_full_bar, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int row = m0 + warp * 32 + lane;
        float* crow = C + (size_t)row * N + n0;
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 32) {
            uint32_t v[32];
            const uint32_t taddr = tmem_d + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0;
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                  "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                  "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                  "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                : "r"(taddr) : "memory");
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            if (row < M) {
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const int n = n0 + c0 + q * 4;
                    if (n < N)      // N % 4 == 0: a 4-column chunk is wholly inside or outside
                        __stcs(reinterpret_cast<uint4*>(crow + c0 + q * 4), make_uint4(v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]));
                }
            }
        }
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 5) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "n"(S::TMEM_COLS) : "memory");
    }
}

// ------------------------------------------------------------------ accurate variant
// The tensor core adds into its FP32 accumulator with truncation, so a long accumulation chain
// drifts (measured: 2e-6 of sum|a||b| at K = 8192, 50x the SIMT engine).  This kernel therefore
// lets the tensor core accumulate only TC_CHUNK_K = 64 of K at a time into one of two 256-column
// accumulators in tensor memory; 8 epilogue warps drain the finished chunk with tcgen05.ld and add
// it to FP32 register accumulators with ordinary round-to-nearest FADDs while the MMA warp is
// already filling the other accumulator.  384 threads: warps 0-7 epilogue (lane quadrant =
// warp % 4, column half = warp / 4, 128 accumulators per thread, setmaxnreg 232), warp 8 TMA
// producer, warp 9 MMA issuer + tensor-memory owner.
constexpr int TC_CHUNK_KB = 2;                    // k-blocks of 32 per tensor-core accumulation chain
constexpr int TC2_BN = 256;
constexpr int TC2_STAGES = 2;
constexpr int TC2_THREADS = 384;


// CLUSTER == 2: the two CTAs of a cluster own vertically adjacent 128-row tiles of the same 256
// columns; each loads half of the B slabs and TMA-multicasts it to both, which cuts L2 -> SM
// traffic per tile by a third (the single-CTA version is L2-bandwidth bound, not tensor bound).
template <int CLUSTER>
__global__ void __launch_bounds__(TC2_THREADS, 1)
gemm_3xtf32_chunked_kernel(float* __restrict__ C, const __grid_constant__ CUtensorMap map_a_hi, const __grid_constant__ CUtensorMap map_a_lo,
                           const __grid_constant__ CUtensorMap map_b_hi, const __grid_constant__ CUtensorMap map_b_lo,
                           int M, int N, int K, int tiles_m) {
    using S = TcShape<TC2_BN, TC2_STAGES>;
    constexpr int BN = TC2_BN, STAGES = TC2_STAGES;
    extern __shared__ uint8_t smem_raw[];
    __shared__ alignas(8) uint64_t full_bar[STAGES];
    __shared__ alignas(8) uint64_t empty_bar[STAGES];
    __shared__ alignas(8) uint64_t acc_full_bar[2];      // MMA -> epilogue: chunk finished in accumulator b
    __shared__ alignas(8) uint64_t acc_empty_bar[2];     // epilogue -> MMA: accumulator b drained
    __shared__ uint32_t tmem_base_slot;

    const uint32_t ring = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* ring_ptr = smem_raw + (ring - smem_u32(smem_raw));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = (blockIdx.x % tiles_m) * TC_BM, n0 = (blockIdx.x / tiles_m) * BN;
    const int num_kb = (K + TC_BK - 1) / TC_BK;
    const int num_chunks = (num_kb + TC_CHUNK_KB - 1) / TC_CHUNK_KB;
    const uint32_t crank = (CLUSTER > 1) ? cluster_ctarank() : 0u;
    constexpr uint16_t cmask = (uint16_t)((1u << CLUSTER) - 1u);

    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], CLUSTER); }   // a slot is free when every CTA writing into it has consumed it
#pragma unroll
        for (int b = 0; b < 2; ++b) { mbar_init(&acc_full_bar[b], 1); mbar_init(&acc_empty_bar[b], 8); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 9) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_slot)), "n"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (CLUSTER > 1) cluster_sync_all();          // peers' barriers are initialised before anyone signals them
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_slot;

    if (warp >= 8) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        if (warp == 8) {
            // ---------------- TMA producer
            if (lane == 0) {
                for (int kb = 0; kb < num_kb; ++kb) {
                    const int s = kb % STAGES;
                    const uint32_t ph = (uint32_t)(kb / STAGES) & 1u;
                    mbar_wait(&empty_bar[s], ph ^ 1u);
                    uint8_t* st = ring_ptr + (size_t)s * S::STAGE_BYTES;
                    mbar_expect_tx(&full_bar[s], (uint32_t)S::STAGE_BYTES);
                    tma_load_2d(st, &map_a_hi, &full_bar[s], kb * TC_BK, m0);
                    tma_load_2d(st + S::A_TILE, &map_a_lo, &full_bar[s], kb * TC_BK, m0);
                    if (CLUSTER == 1) {
                        tma_load_2d(st + 2 * S::A_TILE, &map_b_hi, &full_bar[s], kb * TC_BK, n0);
                        tma_load_2d(st + 2 * S::A_TILE + S::B_TILE, &map_b_lo, &full_bar[s], kb * TC_BK, n0);
                    } else {
                        // my half of the B rows, delivered to both CTAs (box = BN / 2 rows)
                        const int half_rows = BN / CLUSTER, half_bytes = half_rows * TC_BK * 4;
                        const int y = n0 + (int)crank * half_rows;
                        tma_load_2d_mcast(st + 2 * S::A_TILE + crank * half_bytes, &map_b_hi, &full_bar[s], kb * TC_BK, y, cmask);
                        tma_load_2d_mcast(st + 2 * S::A_TILE + S::B_TILE + crank * half_bytes, &map_b_lo, &full_bar[s], kb * TC_BK, y, cmask);
                    }
                }
            }
        } else if (warp == 9) {
            // ---------------- MMA issuer
            constexpr uint32_t idesc = umma_idesc_tf32(TC_BM, BN);
            int kb = 0;
            for (int c = 0; c < num_chunks; ++c) {
                const int b = c & 1;
                const uint32_t use = (uint32_t)(c >> 1);
                mbar_wait(&acc_empty_bar[b], (use & 1u) ^ 1u);          // first use of each accumulator passes at once
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t tmem_d = tmem_base + (uint32_t)(b * BN);
                const int kb_end = (kb + TC_CHUNK_KB < num_kb) ? kb + TC_CHUNK_KB : num_kb;
                for (int first = 1; kb < kb_end; ++kb) {
                    const int s = kb % STAGES;
                    const uint32_t ph = (uint32_t)(kb / STAGES) & 1u;
                    mbar_wait(&full_bar[s], ph);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    if (elect_one()) {
                        const uint32_t st = ring + (uint32_t)s * S::STAGE_BYTES;
                        const uint64_t a_hi = umma_desc_k_sw128(st), a_lo = umma_desc_k_sw128(st + S::A_TILE);
                        const uint64_t b_hi = umma_desc_k_sw128(st + 2 * S::A_TILE), b_lo = umma_desc_k_sw128(st + 2 * S::A_TILE + S::B_TILE);
#pragma unroll
                        for (int kk = 0; kk < TC_BK / TC_UMMA_K; ++kk) {
                            const uint64_t adv = (uint64_t)((kk * TC_UMMA_K * 4) >> 4);
                            umma_tf32(tmem_d, a_hi + adv, b_lo + adv, idesc, (first && kk == 0) ? 0u : 1u);
                            umma_tf32(tmem_d, a_lo + adv, b_hi + adv, idesc, 1u);
                            umma_tf32(tmem_d, a_hi + adv, b_hi + adv, idesc, 1u);
                        }
                        if (CLUSTER == 1) umma_commit(&empty_bar[s]);
                        else umma_commit_mcast(&empty_bar[s], cmask);      // frees the slot in both CTAs
                        if (kb == kb_end - 1) umma_commit(&acc_full_bar[b]);
                    }
                    __syncwarp();
                    first = 0;
                }
            }
        }
    } else {
        // ---------------- epilogue warps 0..7
        asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
        const int quad = warp & 3, half = warp >> 2;
        float acc[128];
#pragma unroll
        for (int i = 0; i < 128; ++i) acc[i] = 0.f;
        for (int c = 0; c < num_chunks; ++c) {
            const int b = c & 1;
            const uint32_t use = (uint32_t)(c >> 1);
            mbar_wait(&acc_full_bar[b], use & 1u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(b * BN + half * 128);
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                uint32_t v0[32], v1[32];
                tmem_ld32(taddr + j * 64, v0);
                tmem_ld32(taddr + j * 64 + 32, v1);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    acc[j * 64 + i] += __uint_as_float(v0[i]);
                    acc[j * 64 + 32 + i] += __uint_as_float(v1[i]);
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty_bar[b]);
        }
        const int row = m0 + quad * 32 + lane;
        if (row < M) {
            float* crow = C + (size_t)row * N + n0 + half * 128;
#pragma unroll
            for (int q = 0; q < 32; ++q) {
                const int n = n0 + half * 128 + q * 4;
                if (n < N) __stcs(reinterpret_cast<float4*>(crow + q * 4), make_float4(acc[q * 4], acc[q * 4 + 1], acc[q * 4 + 2], acc[q * 4 + 3]));
            }
        }
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (CLUSTER > 1) cluster_sync_all();          // nobody leaves while a peer may still write into it
    if (warp == 9) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
    }
}

// ------------------------------------------------------------------ CTA-pair variant (cta_group::2)
// Two CTAs on the SMs of one TPC share one 256 x 256 output tile: CTA r owns rows r*128.. of it
// (its own tensor memory holds those 128 accumulator rows) and loads its 128 rows of A plus HALF
// of the B slabs (128 of the 256 columns); tcgen05.mma.cta_group::2, issued by the leader CTA
// only, reads both CTAs' shared memory.  Per CTA and 32-wide k-block only 64 KB enter shared
// memory (the single-CTA kernel needs 96 KB, which runs into the SM's ingest rate), and the ring
// gets 3 stages.  Accumulation is chunked exactly as above; the leader's MMA warp waits for the
// epilogue warps of BOTH CTAs before it reuses an accumulator.
constexpr int TC3_STAGES = 3;
constexpr int TC3_THREADS = 384;
constexpr int TC3_TILE = 128 * TC_BK * 4;                         // one 128-row slab: 16 KB
constexpr int TC3_STAGE_BYTES = 4 * TC3_TILE;                     // A_hi, A_lo, Bhalf_hi, Bhalf_lo
constexpr size_t TC3_SMEM_BYTES = (size_t)TC3_STAGES * TC3_STAGE_BYTES + 1024;

__device__ __forceinline__ void umma_tf32_2sm(uint32_t tmem_d, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}

// Work plan of the CTA-pair kernel.  Items 0 .. dp_items-1 are whole tiles.  The tiles after them (stream_tiles of them:
// the last full wave plus the ragged remainder) are cut into stream_items equal spans of accumulation chunks, one span per
// CTA pair, so the final wave is as long for every pair instead of leaving most SMs idle behind a partial wave.  A span is
// at least one tile long (stream_tiles >= stream_items), so a tile receives at most two contributions: both are added
// onto zeroed memory with atomics, and a sum of two addends does not depend on their order -- results stay reproducible.
struct TcPlan {
    int pair_tiles_m, tiles_n, raster_group;
    int num_kb, chunk_kb, num_chunks;            // k-blocks of 32; k-blocks per tensor-core accumulation chain; chains per tile
    int dp_items, stream_items, stream_tiles;
    int persist;                                 // > 0: that many CTA pairs loop over all work items (persistent grid)
};

// Segment `seg` of work item `item`: tile (linear index) and its chunk range [c_begin, c_end).  False when there is none.
__device__ __forceinline__ bool tc_segment(const TcPlan& p, int item, int seg, int& tile, int& c_begin, int& c_end) {
    if (p.persist > 0) {
        // persistent grid: CTA pair `item` walks the work items item, item + persist, item + 2 persist, ... as its segments
        // (each of those is a single segment: a whole tile or one half of a split tile), so the drain and the C stores of one
        // tile overlap the first chains of the next and tensor memory, barriers and the ring are set up once per pair
        item += seg * p.persist;
        if (item >= p.dp_items + p.stream_items) return false;
        seg = 0;
    }
    if (item < p.dp_items) {
        tile = item; c_begin = 0; c_end = p.num_chunks;
        return seg == 0;
    }
    const long long total = (long long)p.stream_tiles * p.num_chunks;
    const long long j = item - p.dp_items;
    const long long span_b = j * total / p.stream_items, span_e = (j + 1) * total / p.stream_items;
    const long long t = span_b / p.num_chunks + seg;
    const long long tb = t * p.num_chunks;
    if (tb >= span_e) return false;
    tile = p.dp_items + (int)t;
    c_begin = (int)((span_b > tb ? span_b : tb) - tb);
    c_end = (int)((span_e < tb + p.num_chunks ? span_e : tb + p.num_chunks) - tb);
    return c_begin < c_end;
}

__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N) {       // kind::f16: D = f32, A = B = bf16, both K-major
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_bf16_2sm(uint32_t tmem_d, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}

// BN: columns of the pair's tile (a multiple of 32 up to 256): 256 unless another width packs the waves better (tc_launch_2cta)
// BF16X: the two cross terms hi*lo and lo*hi as bf16 MMAs (twice the tf32 rate) on bf16(hi), bf16(x - hi) planes: the
// `lo` maps then describe the bf16(hi) planes and the `x` maps the bf16(x - hi) planes; 64-byte rows, SWIZZLE_64B.
template <int BN, bool BF16X>
__global__ void __launch_bounds__(TC3_THREADS, 1)
gemm_3xtf32_2cta_kernel(float* __restrict__ C, const __grid_constant__ CUtensorMap map_a_hi, const __grid_constant__ CUtensorMap map_a_lo,
                        const __grid_constant__ CUtensorMap map_b_hi, const __grid_constant__ CUtensorMap map_b_lo,
                        const __grid_constant__ CUtensorMap map_a_x, const __grid_constant__ CUtensorMap map_b_x,
                        int M, int N, const TcPlan plan) {
    constexpr int STAGES = TC3_STAGES;
    constexpr int HC = BN / 2;                                   // B rows held by each CTA = accumulator columns drained by each epilogue half
    static_assert(BN % 32 == 0 && BN >= 64 && BN <= 256, "pair tile width");
    extern __shared__ uint8_t smem_raw[];
    __shared__ alignas(8) uint64_t full_bar[STAGES];     // used in the leader: both CTAs' slabs have landed
    __shared__ alignas(8) uint64_t empty_bar[STAGES];    // per CTA: the pair's MMAs are done with this slot
    __shared__ alignas(8) uint64_t acc_full_bar[2];      // per CTA: chunk finished in accumulator b
    __shared__ alignas(8) uint64_t acc_empty_bar[2];     // used in the leader: 16 epilogue warps drained accumulator b
    __shared__ uint32_t tmem_base_slot;

    const uint32_t ring = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* ring_ptr = smem_raw + (ring - smem_u32(smem_raw));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t crank = cluster_ctarank();
    const int item = (int)(blockIdx.x >> 1);
    const int num_kb = plan.num_kb;

    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
#pragma unroll
        for (int b = 0; b < 2; ++b) { mbar_init(&acc_full_bar[b], 1); mbar_init(&acc_empty_bar[b], 16); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 9) {    // the same warp of both CTAs allocates the pair's tensor memory
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_slot)), "n"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    cluster_sync_all();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_slot;

    if (warp >= 8) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        if (warp == 8) {
            // ---------------- TMA producer (both CTAs)
            if (lane == 0) {
                int it = 0;                                           // k-blocks issued so far, over all segments: ring position
                int tile, cb, ce;
                for (int seg = 0; tc_segment(plan, item, seg, tile, cb, ce); ++seg) {
                    int tile_m, tile_n;
                    la_raster_tile(tile, plan.pair_tiles_m, plan.tiles_n, plan.raster_group, tile_m, tile_n);
                    const int m0 = tile_m * 256 + (int)crank * 128, n0 = tile_n * BN;
                    const int kb_end = min(ce * plan.chunk_kb, num_kb);
                    for (int kb = cb * plan.chunk_kb; kb < kb_end; ++kb, ++it) {
                        const int s = it % STAGES;
                        const uint32_t ph = (uint32_t)(it / STAGES) & 1u;
                        mbar_wait(&empty_bar[s], ph ^ 1u);
                        uint8_t* st = ring_ptr + (size_t)s * TC3_STAGE_BYTES;
                        if (crank == 0) mbar_expect_tx(&full_bar[s], 2u * (2u * TC3_TILE + 2u * HC * TC_BK * 4u));     // my slabs + the peer's
                        tma_load_2d_2sm(st, &map_a_hi, &full_bar[s], kb * TC_BK, m0);
                        tma_load_2d_2sm(st + TC3_TILE, &map_a_lo, &full_bar[s], kb * TC_BK, m0);
                        tma_load_2d_2sm(st + 2 * TC3_TILE, &map_b_hi, &full_bar[s], kb * TC_BK, n0 + (int)crank * HC);
                        tma_load_2d_2sm(st + 3 * TC3_TILE, &map_b_lo, &full_bar[s], kb * TC_BK, n0 + (int)crank * HC);
                        // (With bf16 cross terms the kernel runs at 90 % of its tensor-time floor (8192^3: 2.96 ms against
                        //  2.67) while pulling 34.4 GB per launch through L2 -> SM = 11.6 TB/s, close to what the L2 slices
                        //  deliver chip-wide.  Experiments: a 2-stage ring costs 22 %; prefetching boxes into L2 ahead of the
                        //  ring makes it 27 % SLOWER (more L2 requests); not loading A at all for every other tile -- the upper
                        //  bound of what multicasting A across a 4-CTA cluster could save -- gains only 3 %.  So the operand
                        //  path is near its limit but not the binding one; what is left is per-tile ramp-up and drain.)
                        if constexpr (BF16X) {      // the second bf16 plane of each operand, in the upper half of its 16 KB slot
                            tma_load_2d_2sm(st + TC3_TILE + TC3_TILE / 2, &map_a_x, &full_bar[s], kb * TC_BK, m0);
                            tma_load_2d_2sm(st + 3 * TC3_TILE + TC3_TILE / 2, &map_b_x, &full_bar[s], kb * TC_BK, n0 + (int)crank * HC);
                        }
                    }
                }
            }
        } else if (warp == 9 && crank == 0) {
            // ---------------- MMA issuer (leader CTA only)
            constexpr uint32_t idesc = umma_idesc_tf32(256, BN);
            int it = 0, cc = 0;                                       // k-blocks / chunks consumed so far, over all segments
            int tile, cb, ce;
            for (int seg = 0; tc_segment(plan, item, seg, tile, cb, ce); ++seg)
            for (int c = cb; c < ce; ++c, ++cc) {
                const int b = cc & 1;
                const uint32_t use = (uint32_t)(cc >> 1);
                mbar_wait(&acc_empty_bar[b], (use & 1u) ^ 1u);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t tmem_d = tmem_base + (uint32_t)(b * 256);
                int kb = c * plan.chunk_kb;
                const int kb_end = (kb + plan.chunk_kb < num_kb) ? kb + plan.chunk_kb : num_kb;
                for (int first = 1; kb < kb_end; ++kb, ++it) {
                    const int s = it % STAGES;
                    const uint32_t ph = (uint32_t)(it / STAGES) & 1u;
                    mbar_wait(&full_bar[s], ph);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    if (elect_one()) {
                        const uint32_t st = ring + (uint32_t)s * TC3_STAGE_BYTES;
                        const uint64_t a_hi = umma_desc_k_sw128(st), b_hi = umma_desc_k_sw128(st + 2 * TC3_TILE);
                        if constexpr (BF16X) {
                            constexpr uint32_t idesc16 = umma_idesc_bf16(256, BN);
                            const uint64_t a_hb = umma_desc_k_sw64(st + TC3_TILE), a_lb = umma_desc_k_sw64(st + TC3_TILE + TC3_TILE / 2);
                            const uint64_t b_hb = umma_desc_k_sw64(st + 3 * TC3_TILE), b_lb = umma_desc_k_sw64(st + 3 * TC3_TILE + TC3_TILE / 2);
#pragma unroll
                            for (int kk = 0; kk < TC_BK / 16; ++kk) {            // bf16: 16 k = 32 bytes per instruction
                                const uint64_t adv = (uint64_t)((kk * 16 * 2) >> 4);
                                umma_bf16_2sm(tmem_d, a_hb + adv, b_lb + adv, idesc16, (first && kk == 0) ? 0u : 1u);
                                umma_bf16_2sm(tmem_d, a_lb + adv, b_hb + adv, idesc16, 1u);
                            }
#pragma unroll
                            for (int kk = 0; kk < TC_BK / TC_UMMA_K; ++kk) {
                                const uint64_t adv = (uint64_t)((kk * TC_UMMA_K * 4) >> 4);
                                umma_tf32_2sm(tmem_d, a_hi + adv, b_hi + adv, idesc, 1u);
                            }
                        } else {
                            const uint64_t a_lo = umma_desc_k_sw128(st + TC3_TILE), b_lo = umma_desc_k_sw128(st + 3 * TC3_TILE);
#pragma unroll
                            for (int kk = 0; kk < TC_BK / TC_UMMA_K; ++kk) {
                                const uint64_t adv = (uint64_t)((kk * TC_UMMA_K * 4) >> 4);
                                umma_tf32_2sm(tmem_d, a_hi + adv, b_lo + adv, idesc, (first && kk == 0) ? 0u : 1u);
                                umma_tf32_2sm(tmem_d, a_lo + adv, b_hi + adv, idesc, 1u);
                                umma_tf32_2sm(tmem_d, a_hi + adv, b_hi + adv, idesc, 1u);
                            }
                        }
                        umma_commit_2sm(&empty_bar[s]);
                        if (kb == kb_end - 1) umma_commit_2sm(&acc_full_bar[b]);
                    }
                    __syncwarp();
                    first = 0;
                }
            }
        }
    } else {
        // ---------------- epilogue warps 0..7 (both CTAs; each drains its own tensor memory)
        asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
        const int quad = warp & 3, half = warp >> 2;
        float acc[HC];
        int cc = 0;
        int tile, cb, ce;
        for (int seg = 0; tc_segment(plan, item, seg, tile, cb, ce); ++seg) {
#pragma unroll
        for (int i = 0; i < HC; ++i) acc[i] = 0.f;
        for (int c = cb; c < ce; ++c, ++cc) {
            const int b = cc & 1;
            const uint32_t use = (uint32_t)(cc >> 1);
            mbar_wait(&acc_full_bar[b], use & 1u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(b * 256 + half * HC);
#pragma unroll
            for (int j = 0; j < HC / 64; ++j) {
                uint32_t v0[32], v1[32];
                tmem_ld32(taddr + j * 64, v0);
                tmem_ld32(taddr + j * 64 + 32, v1);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    acc[j * 64 + i] += __uint_as_float(v0[i]);
                    acc[j * 64 + 32 + i] += __uint_as_float(v1[i]);
                }
            }
            if constexpr (HC % 64 >= 32) {
                constexpr int o = HC / 64 * 64;
                uint32_t v0[32];
                tmem_ld32(taddr + o, v0);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int i = 0; i < 32; ++i) acc[o + i] += __uint_as_float(v0[i]);
            }
            if constexpr (HC % 32 == 16) {
                constexpr int o = HC / 32 * 32;
                uint32_t v0[16];
                tmem_ld16(taddr + o, v0);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int i = 0; i < 16; ++i) acc[o + i] += __uint_as_float(v0[i]);
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive_on_leader(&acc_empty_bar[b]);
        }
        int tile_m, tile_n;
        la_raster_tile(tile, plan.pair_tiles_m, plan.tiles_n, plan.raster_group, tile_m, tile_n);
        const int m0 = tile_m * 256 + (int)crank * 128, n0 = tile_n * BN;
        const bool whole = (cb == 0 && ce == plan.num_chunks);       // otherwise one of (at most) two partial sums of this tile
        const int row = m0 + quad * 32 + lane;
        if (row < M) {
            float* crow = C + (size_t)row * N + n0 + half * HC;
#pragma unroll
            for (int q = 0; q < HC / 4; ++q) {
                const int n = n0 + half * HC + q * 4;
                if (n < N) {
                    const float4 v = make_float4(acc[q * 4], acc[q * 4 + 1], acc[q * 4 + 2], acc[q * 4 + 3]);
                    if (whole) __stcs(reinterpret_cast<float4*>(crow + q * 4), v);
                    else atomicAdd(reinterpret_cast<float4*>(crow + q * 4), v);      // onto zeroed memory (tc_launch_2cta)
                }
            }
        }
        }
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    cluster_sync_all();
    if (warp == 9) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
    }
}

// ------------------------------------------------------------------ f16-split variant of the CTA-pair kernel
// Same structure as gemm_3xtf32_2cta_kernel, other operand format.  Every row of A and every column of B is scaled by a power of
// two so that its largest magnitude lies in [2^14, 2^15) (exact), and the scaled value is split x = hi + lo with hi = fp16(x) --
// 11 significant bits, what tf32 keeps.  Three planes per operand, all fp16: hi, hs = hi * 2^-6 and ls = fp16(lo * 2^6) (the
// residual lifted into fp16's range: 11 of its 13 bits survive, where bf16 keeps 8).  A product is three fp16 x fp16 MMAs at twice
// the tf32 rate -- hi*hi, hs*ls, ls*hs, the two powers of two cancelling exactly -- so 1.5 tf32-equivalent units per product
// instead of 2 (or 3 with tf32 cross terms), 6 bytes of planes per element instead of 8, and a per-product error of ~2^-23 |a||b|
// (bf16 cross terms: ~2^-20, fine when many comparable terms average it out, above the gate for an output that is essentially
// ONE product -- B = identity, a permutation, a diagonal).  (fp16 x bf16 in one instruction would need one plane less, but is an
// illegal instruction on this GPU.)  Neither hs nor ls may lose bits to fp16's subnormal range, which is why elements 2^-23 and
// more below their row's / column's maximum are kept out of the planes (f16_split_one).  A stage holds 32 k of the six planes (A's
// 128 rows and half of B's columns: 8 KB each, 64-byte rows, SWIZZLE_64B), four stages.  The epilogue multiplies by
// 2^(ea[row] + eb[col]).
constexpr int F16_BK = 32;                                      // k per stage
constexpr int F16_PLANE = 128 * F16_BK * 2;                     // one 128-row slab of one plane: 8 KB
constexpr int F16_STAGE_BYTES = 6 * F16_PLANE;                  // 48 KB
constexpr int F16_STAGES = 4;
static_assert((size_t)F16_STAGES * F16_STAGE_BYTES + 1024 == TC3_SMEM_BYTES, "same shared-memory budget as the tf32 pair kernel");

__host__ __device__ constexpr uint32_t umma_idesc_16(int M, int N, int a_bf16, int b_bf16) {      // kind::f16: D = f32, A / B = fp16 or bf16, K-major
    return (1u << 4) | ((uint32_t)a_bf16 << 7) | ((uint32_t)b_bf16 << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

template <int BN>
__global__ void __launch_bounds__(TC3_THREADS, 1)
gemm_f16split_2cta_kernel(float* __restrict__ C, const __grid_constant__ CUtensorMap map_a_hi, const __grid_constant__ CUtensorMap map_a_hb, const __grid_constant__ CUtensorMap map_a_lb,
                          const __grid_constant__ CUtensorMap map_b_hi, const __grid_constant__ CUtensorMap map_b_hb, const __grid_constant__ CUtensorMap map_b_lb,
                          const int* __restrict__ ea, const int* __restrict__ eb, int M, int N, const TcPlan plan) {
    constexpr int STAGES = F16_STAGES;
    constexpr int HC = BN / 2;
    static_assert(BN % 32 == 0 && BN >= 64 && BN <= 256, "pair tile width");
    extern __shared__ uint8_t smem_raw[];
    __shared__ alignas(8) uint64_t full_bar[STAGES];
    __shared__ alignas(8) uint64_t empty_bar[STAGES];
    __shared__ alignas(8) uint64_t acc_full_bar[2];
    __shared__ alignas(8) uint64_t acc_empty_bar[2];
    __shared__ uint32_t tmem_base_slot;

    const uint32_t ring = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* ring_ptr = smem_raw + (ring - smem_u32(smem_raw));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t crank = cluster_ctarank();
    const int item = (int)(blockIdx.x >> 1);
    const int num_kb = plan.num_kb;

    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
#pragma unroll
        for (int b = 0; b < 2; ++b) { mbar_init(&acc_full_bar[b], 1); mbar_init(&acc_empty_bar[b], 16); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 9) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_slot)), "n"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    cluster_sync_all();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_slot;

    if (warp >= 8) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        if (warp == 8) {
            // ---------------- TMA producer (both CTAs)
            if (lane == 0) {
                int it = 0;
                int tile, cb, ce;
                for (int seg = 0; tc_segment(plan, item, seg, tile, cb, ce); ++seg) {
                    int tile_m, tile_n;
                    la_raster_tile(tile, plan.pair_tiles_m, plan.tiles_n, plan.raster_group, tile_m, tile_n);
                    const int m0 = tile_m * 256 + (int)crank * 128, n0 = tile_n * BN;
                    const int kb_end = min(ce * plan.chunk_kb, num_kb);
                    for (int kb = cb * plan.chunk_kb; kb < kb_end; ++kb, ++it) {
                        const int s = it % STAGES;
                        const uint32_t ph = (uint32_t)(it / STAGES) & 1u;
                        mbar_wait(&empty_bar[s], ph ^ 1u);
                        uint8_t* st = ring_ptr + (size_t)s * F16_STAGE_BYTES;
                        if (crank == 0) mbar_expect_tx(&full_bar[s], 2u * (3u * F16_PLANE + 3u * HC * F16_BK * 2u));     // my slabs + the peer's
                        tma_load_2d_2sm(st, &map_a_hi, &full_bar[s], kb * F16_BK, m0);
                        tma_load_2d_2sm(st + F16_PLANE, &map_a_hb, &full_bar[s], kb * F16_BK, m0);
                        tma_load_2d_2sm(st + 2 * F16_PLANE, &map_a_lb, &full_bar[s], kb * F16_BK, m0);
                        tma_load_2d_2sm(st + 3 * F16_PLANE, &map_b_hi, &full_bar[s], kb * F16_BK, n0 + (int)crank * HC);
                        tma_load_2d_2sm(st + 4 * F16_PLANE, &map_b_hb, &full_bar[s], kb * F16_BK, n0 + (int)crank * HC);
                        tma_load_2d_2sm(st + 5 * F16_PLANE, &map_b_lb, &full_bar[s], kb * F16_BK, n0 + (int)crank * HC);
                    }
                }
            }
        } else if (warp == 9 && crank == 0) {
            // ---------------- MMA issuer (leader CTA only)
            constexpr uint32_t idesc_f16 = umma_idesc_16(256, BN, 0, 0);
            int it = 0, cc = 0;
            int tile, cb, ce;
            for (int seg = 0; tc_segment(plan, item, seg, tile, cb, ce); ++seg)
            for (int c = cb; c < ce; ++c, ++cc) {
                const int b = cc & 1;
                const uint32_t use = (uint32_t)(cc >> 1);
                mbar_wait(&acc_empty_bar[b], (use & 1u) ^ 1u);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t tmem_d = tmem_base + (uint32_t)(b * 256);
                int kb = c * plan.chunk_kb;
                const int kb_end = (kb + plan.chunk_kb < num_kb) ? kb + plan.chunk_kb : num_kb;
                for (int first = 1; kb < kb_end; ++kb, ++it) {
                    const int s = it % STAGES;
                    const uint32_t ph = (uint32_t)(it / STAGES) & 1u;
                    mbar_wait(&full_bar[s], ph);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    if (elect_one()) {
                        const uint32_t st = ring + (uint32_t)s * F16_STAGE_BYTES;
                        const uint64_t a_hi = umma_desc_k_sw64(st), a_hb = umma_desc_k_sw64(st + F16_PLANE), a_lb = umma_desc_k_sw64(st + 2 * F16_PLANE);
                        const uint64_t b_hi = umma_desc_k_sw64(st + 3 * F16_PLANE), b_hb = umma_desc_k_sw64(st + 4 * F16_PLANE), b_lb = umma_desc_k_sw64(st + 5 * F16_PLANE);
#pragma unroll
                        for (int kk = 0; kk < F16_BK / 16; ++kk) {               // 16 k = 32 bytes per instruction; same order of terms as the tf32 form
                            const uint64_t adv = (uint64_t)((kk * 16 * 2) >> 4);
                            umma_bf16_2sm(tmem_d, a_hb + adv, b_lb + adv, idesc_f16, (first && kk == 0) ? 0u : 1u);       // hs * ls
                            umma_bf16_2sm(tmem_d, a_lb + adv, b_hb + adv, idesc_f16, 1u);                                  // ls * hs
                        }
#pragma unroll
                        for (int kk = 0; kk < F16_BK / 16; ++kk) {
                            const uint64_t adv = (uint64_t)((kk * 16 * 2) >> 4);
                            umma_bf16_2sm(tmem_d, a_hi + adv, b_hi + adv, idesc_f16, 1u);
                        }
                        umma_commit_2sm(&empty_bar[s]);
                        if (kb == kb_end - 1) umma_commit_2sm(&acc_full_bar[b]);
                    }
                    __syncwarp();
                    first = 0;
                }
            }
        }
    } else {
        // ---------------- epilogue warps 0..7 (both CTAs; each drains its own tensor memory)
        asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
        const int quad = warp & 3, half = warp >> 2;
        float acc[HC];
        int cc = 0;
        int tile, cb, ce;
        for (int seg = 0; tc_segment(plan, item, seg, tile, cb, ce); ++seg) {
#pragma unroll
        for (int i = 0; i < HC; ++i) acc[i] = 0.f;
        for (int c = cb; c < ce; ++c, ++cc) {
            const int b = cc & 1;
            const uint32_t use = (uint32_t)(cc >> 1);
            mbar_wait(&acc_full_bar[b], use & 1u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(b * 256 + half * HC);
#pragma unroll
            for (int j = 0; j < HC / 64; ++j) {
                uint32_t v0[32], v1[32];
                tmem_ld32(taddr + j * 64, v0);
                tmem_ld32(taddr + j * 64 + 32, v1);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    acc[j * 64 + i] += __uint_as_float(v0[i]);
                    acc[j * 64 + 32 + i] += __uint_as_float(v1[i]);
                }
            }
            if constexpr (HC % 64 >= 32) {
                constexpr int o = HC / 64 * 64;
                uint32_t v0[32];
                tmem_ld32(taddr + o, v0);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int i = 0; i < 32; ++i) acc[o + i] += __uint_as_float(v0[i]);
            }
            if constexpr (HC % 32 == 16) {
                constexpr int o = HC / 32 * 32;
                uint32_t v0[16];
                tmem_ld16(taddr + o, v0);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int i = 0; i < 16; ++i) acc[o + i] += __uint_as_float(v0[i]);
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive_on_leader(&acc_empty_bar[b]);
        }
        int tile_m, tile_n;
        la_raster_tile(tile, plan.pair_tiles_m, plan.tiles_n, plan.raster_group, tile_m, tile_n);
        const int m0 = tile_m * 256 + (int)crank * 128, n0 = tile_n * BN;
        const bool whole = (cb == 0 && ce == plan.num_chunks);
        const int row = m0 + quad * 32 + lane;
        if (row < M) {
            const int er = __ldg(ea + row) + 127;                 // undo the scaling: x 2^(ea[row] + eb[col]), exact
            float* crow = C + (size_t)row * N + n0 + half * HC;
#pragma unroll
            for (int q = 0; q < HC / 4; ++q) {
                const int n = n0 + half * HC + q * 4;
                if (n < N) {
                    const int4 e4 = __ldg(reinterpret_cast<const int4*>(eb + n));
                    const float4 v = make_float4(acc[q * 4] * __int_as_float((er + e4.x) << 23), acc[q * 4 + 1] * __int_as_float((er + e4.y) << 23),
                                                 acc[q * 4 + 2] * __int_as_float((er + e4.z) << 23), acc[q * 4 + 3] * __int_as_float((er + e4.w) << 23));
                    if (whole) __stcs(reinterpret_cast<float4*>(crow + q * 4), v);
                    else atomicAdd(reinterpret_cast<float4*>(crow + q * 4), v);      // onto zeroed memory (tc_make_plan)
                }
            }
        }
        }
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    cluster_sync_all();
    if (warp == 9) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
    }
}

// ---- f16-split pre-pass.  Per row of A / column of B: m = max |x|; e = floor(log2 m) - 14, so that m * 2^-e lies in [2^14, 2^15)
// (fp16's largest finite value is 65504); an all-zero row gets e = 0.  The flag is raised -- and the product then goes to the
// tf32 form -- for non-finite values and for maxima outside [2^-40, 2^40] (the epilogue's 2^(ea + eb) must stay a normal float).
// A non-zero element below 2^-8 after scaling (2^-23 of its row's / column's maximum) would lose bits of hs = hi * 2^-6 or of
// ls = lo * 2^6 to fp16's subnormal range: such elements are left OUT of the planes (zeros) and noted in a list instead -- (row,
// column, value), at most F16_TINY_CAP per operand, beyond that the flag -- and their products are added afterwards in plain f32
// by f16_tiny_fix_kernel.  Uniform random data has one such element per 2^23 (8 in an 8192^2 operand); a matrix whose rows mix
// scales that far apart has many, and goes to the tf32 form.
// A second flag (word 1) asks for the exact engine: the split leaves an error of up to ~2^-21 |a||b| per PRODUCT (the dropped
// lo*lo and the rounding of ls), which the gate's 2^-22 sum|a||b| only absorbs when an output is a sum of several comparable
// products.  An output that is essentially one product -- a diagonal or one-hot operand, a row whose one outlier outweighs all the
// rest together -- is computed exactly by the reference, so a row of A / column of B whose sum of magnitudes is less than
// F16_MIN_MASS times its largest magnitude sends the product to the SIMT engine.  (A row with outliers 100x its typical element
// keeps the tensor cores: over K >= 1024 the typical elements still carry most of the mass.)
constexpr int F16_EXP_TARGET = 14;
constexpr float F16_KEEP_MIN = 0.00390625f;      // 2^-8
constexpr unsigned F16_TINY_CAP = 1024;
constexpr float F16_MIN_MASS = 4.0f;
struct F16Tiny { int row, col; float val; int pad; };
struct F16TinyList { F16Tiny* items; unsigned* count; unsigned* any_host; };      // any_host: mapped host word, set to 1 when the list is not empty

__device__ __forceinline__ int f16_exponent_of_max(float m, bool& bad) {          // m >= 0
    if (m == 0.0f) return 0;
    if (!(m >= 9.094947e-13f && m <= 1.0995116e12f)) { bad = true; return 0; }     // outside [2^-40, 2^40] or NaN / inf
    return (int)((__float_as_uint(m) >> 23) & 0xffu) - 127 - F16_EXP_TARGET;
}
// row / col: position in the operand as the product sees it (A: row i, column k; B: row k, column n)
__device__ __forceinline__ void f16_split_one(float x, float scale, int row, int col, const F16TinyList& tiny, __half& hi, __half& hs, __half& ls, bool& bad) {
    float xs = x * scale;                                         // exact: a power of two
    if (fabsf(xs) < F16_KEEP_MIN && x != 0.0f) {                  // noted, not split
        const unsigned slot = atomicAdd(tiny.count, 1u);
        if (slot < F16_TINY_CAP) { tiny.items[slot] = F16Tiny{row, col, x, 0}; if (slot == 0) *(volatile unsigned*)tiny.any_host = 1u; }
        else bad = true;
        xs = 0.0f;
    }
    hi = __float2half_rn(xs);
    const float h = __half2float(hi);
    hs = __float2half_rn(h * 0.015625f);                          // hi * 2^-6: exact (|hi| >= 2^-8, or 0)
    ls = __float2half_rn((xs - h) * 64.0f);                       // (x - hi) * 2^6: at most 2^10, and 11 bits of it survive down to |x| = 2^-8
    bad |= !(fabsf(x) <= 3.4028235e38f);                          // NaN / inf
}
// Two elements at once: packed conversions (one F2FP per pair instead of a scalar F2F per element, which is what bounded the
// pre-pass), hs = hi * 2^-6 as an exact half2 multiply.  Same bits as f16_split_one.
__device__ __forceinline__ void f16_split_two(float x0, float x1, float scale, int row0, int col0, int row1, int col1, const F16TinyList& tiny,
                                              uint32_t& hi, uint32_t& hs, uint32_t& ls, bool& bad) {
    float s0 = x0 * scale, s1 = x1 * scale;
    if ((fabsf(s0) < F16_KEEP_MIN && x0 != 0.0f) || (fabsf(s1) < F16_KEEP_MIN && x1 != 0.0f)) {      // rare: the scalar path notes them
        __half h0, h1, a0, a1, l0, l1;
        f16_split_one(x0, scale, row0, col0, tiny, h0, a0, l0, bad);
        f16_split_one(x1, scale, row1, col1, tiny, h1, a1, l1, bad);
        const __half2 ph = __halves2half2(h0, h1), pa = __halves2half2(a0, a1), pl = __halves2half2(l0, l1);
        hi = *reinterpret_cast<const uint32_t*>(&ph); hs = *reinterpret_cast<const uint32_t*>(&pa); ls = *reinterpret_cast<const uint32_t*>(&pl);
        return;
    }
    const __half2 h = __floats2half2_rn(s0, s1);
    const float2 hf = __half22float2(h);
    const __half2 a = __hmul2(h, __floats2half2_rn(0.015625f, 0.015625f));
    const __half2 l = __floats2half2_rn((s0 - hf.x) * 64.0f, (s1 - hf.y) * 64.0f);
    hi = *reinterpret_cast<const uint32_t*>(&h); hs = *reinterpret_cast<const uint32_t*>(&a); ls = *reinterpret_cast<const uint32_t*>(&l);
    bad |= !(fabsf(x0) <= 3.4028235e38f) || !(fabsf(x1) <= 3.4028235e38f);
}
struct F16PrepassA { __half* hi; __half* hb; __half* lo; int* ea; const float* A; unsigned rows; size_t K; F16TinyList tiny; };
struct F16PrepassB { unsigned* colmax_part; float* colsum_part; const float* B; size_t K, N; unsigned col_blocks, slices; size_t rows_per_slice; unsigned* tiny_count; };

// First launch of the pre-pass.  Blocks 0 .. a.rows-1: one row of A each -- pass 1 finds the row's maximum, pass 2 (the row is in
// L2 / L1 by then) scales, splits and stores.  The blocks after them (b.col_blocks x b.slices; none when B is prepared): column
// maxima of B over a slice of its rows, one partial row of maxima per slice (the bits of |x| compare like integers).
__global__ void __launch_bounds__(256)
f16_prepass1_kernel(const F16PrepassA a, const F16PrepassB b, unsigned* flag) {
    __shared__ float red[8], red_mass[8];
    __shared__ float row_scale;
    bool bad = false;
    if (blockIdx.x == 0 && threadIdx.x == 0 && b.tiny_count) *b.tiny_count = 0u;      // B's list is filled by the second launch
    if (blockIdx.x < a.rows) {
        const size_t row = blockIdx.x, K = a.K;
        const float4* src = reinterpret_cast<const float4*>(a.A + row * K);
        const int n4 = (int)(K / 4);
        float m = 0.f, mass = 0.f;
        for (int i = threadIdx.x; i < n4; i += 256) {
            const float4 v = __ldg(src + i);
            m = fmaxf(fmaxf(m, fmaxf(fabsf(v.x), fabsf(v.y))), fmaxf(fabsf(v.z), fabsf(v.w)));
            mass += (fabsf(v.x) + fabsf(v.y)) + (fabsf(v.z) + fabsf(v.w));
            if (!(fabsf(v.x) <= 3.4028235e38f) || !(fabsf(v.y) <= 3.4028235e38f) || !(fabsf(v.z) <= 3.4028235e38f) || !(fabsf(v.w) <= 3.4028235e38f)) m = __int_as_float(0x7f800000);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o)); mass += __shfl_xor_sync(0xffffffffu, mass, o); }
        if ((threadIdx.x & 31) == 0) { red[threadIdx.x >> 5] = m; red_mass[threadIdx.x >> 5] = mass; }
        __syncthreads();
        if (threadIdx.x == 0) {
            float t = red[0], tm = red_mass[0];
#pragma unroll
            for (int w = 1; w < 8; ++w) { t = fmaxf(t, red[w]); tm += red_mass[w]; }
            if (t != 0.0f && t <= 3.4028235e38f && !(tm >= F16_MIN_MASS * t)) *(volatile unsigned*)(flag + 1) = 1u;      // one element outweighs the rest: exact engine
            const int e = f16_exponent_of_max(t, bad);
            a.ea[row] = e;
            row_scale = __int_as_float((127 - e) << 23);
        }
        __syncthreads();
        const float scale = row_scale;
        uint2* hp = reinterpret_cast<uint2*>(a.hi + row * K);
        uint2* bp = reinterpret_cast<uint2*>(a.hb + row * K);
        uint2* lp = reinterpret_cast<uint2*>(a.lo + row * K);
        for (int i = threadIdx.x; i < n4; i += 256) {
            const float4 v = __ldg(src + i);
            uint2 h, b, l;
            f16_split_two(v.x, v.y, scale, (int)row, 4 * i + 0, (int)row, 4 * i + 1, a.tiny, h.x, b.x, l.x, bad);
            f16_split_two(v.z, v.w, scale, (int)row, 4 * i + 2, (int)row, 4 * i + 3, a.tiny, h.y, b.y, l.y, bad);
            __stcs(hp + i, h);
            __stcs(bp + i, b);
            __stcs(lp + i, l);
        }
    } else {
        const unsigned c = blockIdx.x - a.rows;
        const size_t c4 = (size_t)(c % b.col_blocks) * 256 + threadIdx.x;          // float4 column index
        const unsigned slice = c / b.col_blocks;
        if (c4 * 4 < b.N) {
            const size_t k0 = (size_t)slice * b.rows_per_slice, k1 = min(b.K, k0 + b.rows_per_slice);
            unsigned m0 = 0, m1 = 0, m2 = 0, m3 = 0;
            float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
            size_t k = k0;
            for (; k + 4 <= k1; k += 4) {                                             // four rows in flight
                float4 v[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) v[u] = __ldcs(reinterpret_cast<const float4*>(b.B + (k + u) * b.N) + c4);
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    m0 = max(m0, __float_as_uint(v[u].x) & 0x7fffffffu); m1 = max(m1, __float_as_uint(v[u].y) & 0x7fffffffu);
                    m2 = max(m2, __float_as_uint(v[u].z) & 0x7fffffffu); m3 = max(m3, __float_as_uint(v[u].w) & 0x7fffffffu);
                    s0 += fabsf(v[u].x); s1 += fabsf(v[u].y); s2 += fabsf(v[u].z); s3 += fabsf(v[u].w);
                }
            }
            for (; k < k1; ++k) {
                const float4 v = __ldcs(reinterpret_cast<const float4*>(b.B + k * b.N) + c4);
                m0 = max(m0, __float_as_uint(v.x) & 0x7fffffffu); m1 = max(m1, __float_as_uint(v.y) & 0x7fffffffu);
                m2 = max(m2, __float_as_uint(v.z) & 0x7fffffffu); m3 = max(m3, __float_as_uint(v.w) & 0x7fffffffu);
                s0 += fabsf(v.x); s1 += fabsf(v.y); s2 += fabsf(v.z); s3 += fabsf(v.w);
            }
            *reinterpret_cast<uint4*>(b.colmax_part + (size_t)slice * b.N + c4 * 4) = make_uint4(m0, m1, m2, m3);
            *reinterpret_cast<float4*>(b.colsum_part + (size_t)slice * b.N + c4 * 4) = make_float4(s0, s1, s2, s3);
        }
    }
    if (__syncthreads_or(bad) && threadIdx.x == 0) *(volatile unsigned*)flag = 1u;
}

// Second launch (not for a prepared B): B (K x N) -> Bt planes (N x K: fp16 hi, bf16 hi, bf16 lo) and eb[N].  64 x 64 tile through
// shared memory as split_transpose_kernel does; the column exponents come from the partial maxima of the first launch.
__global__ void __launch_bounds__(256)
f16_split_transpose_kernel(__half* __restrict__ bt_hi, __half* __restrict__ bt_hb, __half* __restrict__ bt_lo, int* __restrict__ eb,
                           const unsigned* __restrict__ colmax_part, const float* __restrict__ colsum_part, unsigned slices, const float* __restrict__ B, size_t K, size_t N, unsigned tiles_n,
                           const F16TinyList tiny, unsigned* flag) {
    __shared__ float tile[64][65];
    __shared__ float col_scale[64];
    __shared__ unsigned part_max[4][64];
    const size_t k0 = (size_t)(blockIdx.x / tiles_n) * 64, n0 = (size_t)(blockIdx.x % tiles_n) * 64;
    bool bad = false;
    float4 in[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int idx = threadIdx.x + j * 256, r = idx / 16, q = idx % 16;
        in[j] = (k0 + r < K && n0 + q * 4 < N) ? __ldcs(reinterpret_cast<const float4*>(B + (k0 + r) * N + n0 + q * 4)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    __syncwarp();                            // all four loads are issued before anything else
    {   // this tile's 64 column exponents from the partial maxima: four threads per column, a quarter of the slices each
        const int c = threadIdx.x & 63, g = threadIdx.x >> 6;
        const bool col_ok = n0 + c < N;
        unsigned m = 0;
        if (col_ok) for (unsigned sl = g; sl < slices; sl += 4) m = max(m, colmax_part[(size_t)sl * N + n0 + c]);
        part_max[g][c] = m;
        __syncthreads();
        m = max(max(part_max[0][c], part_max[1][c]), max(part_max[2][c], part_max[3][c]));
        int e = 0;
        bool col_bad = false;
        if (col_ok) e = f16_exponent_of_max(__uint_as_float(m), col_bad);      // (a NaN / inf column maximum has all exponent bits set: flagged)
        bad |= col_bad;
        const float scale_n = __int_as_float((127 - e) << 23);
        if (g == 0) { col_scale[c] = scale_n; if (k0 == 0 && col_ok) eb[n0 + c] = e; }
        if (k0 == 0) {                       // (once per column) its sum of magnitudes against its maximum
            float mass = 0.f;
            if (col_ok) for (unsigned sl = g; sl < slices; sl += 4) mass += colsum_part[(size_t)sl * N + n0 + c];
            __syncthreads();
            part_max[g][c] = __float_as_uint(mass);
            __syncthreads();
            if (g == 0 && col_ok && m != 0u && !col_bad) {
                const float total = (__uint_as_float(part_max[0][c]) + __uint_as_float(part_max[1][c])) + (__uint_as_float(part_max[2][c]) + __uint_as_float(part_max[3][c]));
                if (!(total >= F16_MIN_MASS * __uint_as_float(m))) *(volatile unsigned*)(flag + 1) = 1u;
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int idx = threadIdx.x + j * 256, r = idx / 16, q = idx % 16;
        tile[r][q * 4 + 0] = in[j].x; tile[r][q * 4 + 1] = in[j].y; tile[r][q * 4 + 2] = in[j].z; tile[r][q * 4 + 3] = in[j].w;
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int idx = threadIdx.x + j * 256, cc = idx / 16, q = idx % 16;      // Bt row n0+cc, columns k0 + q*4 ..
        const float scale = col_scale[cc];
        const int kk = (int)k0 + q * 4, nn = (int)n0 + cc;
        uint2 h, b, l;
        f16_split_two(tile[q * 4 + 0][cc], tile[q * 4 + 1][cc], scale, kk + 0, nn, kk + 1, nn, tiny, h.x, b.x, l.x, bad);
        f16_split_two(tile[q * 4 + 2][cc], tile[q * 4 + 3][cc], scale, kk + 2, nn, kk + 3, nn, tiny, h.y, b.y, l.y, bad);
        if (n0 + cc < N && k0 + q * 4 < K) {
            const size_t at = (n0 + cc) * K + k0 + q * 4;
            __stcs(reinterpret_cast<uint2*>(bt_hi + at), h);
            __stcs(reinterpret_cast<uint2*>(bt_hb + at), b);
            __stcs(reinterpret_cast<uint2*>(bt_lo + at), l);
        }
    }
    if (__syncthreads_or(bad) && threadIdx.x == 0) *(volatile unsigned*)flag = 1u;
}

// The products of the elements the planes left out, added in plain f32 after the product kernel (only launched when the list is
// not empty; F16_TINY_CAP x ceil(outputs / 1024) blocks).  Block (e, y) serves 1024 outputs of the row (ROWS) or column of list entry e -- if e is that row's first entry in
// the list, otherwise it leaves at once -- and applies ALL of the row's entries in ascending k:
//   ROWS:  C[i, :] += a * B[k, :]  for every noted (i, k, a) of A
//   else:  C[:, n] += A[:, k] * b  for every noted (k, n, b) of B, skipping the A elements that are noted themselves (their
//          products with all of B, noted or not, were added by the row pass, which runs first).
// One owner per output and a fixed order inside it: reproducible, whatever order the atomics filled the list in.
template <bool ROWS>
__global__ void __launch_bounds__(256)
f16_tiny_fix_kernel(float* __restrict__ C, const float* __restrict__ A, const float* __restrict__ B, const int* __restrict__ ea, size_t M, size_t K, size_t N,
                    const F16Tiny* __restrict__ items, const unsigned* __restrict__ count_ptr) {
    __shared__ int sel_k[64];
    __shared__ float sel_v[64];
    __shared__ int n_sel, n_match;
    const unsigned count = min(*count_ptr, F16_TINY_CAP);
    if (blockIdx.x >= count) return;
    const int mine = ROWS ? items[blockIdx.x].row : items[blockIdx.x].col;
    bool earlier = false;
    for (unsigned e = threadIdx.x; e < blockIdx.x; e += 256) earlier |= (ROWS ? items[e].row : items[e].col) == mine;
    if (__syncthreads_or(earlier)) return;
    if (threadIdx.x == 0) { n_sel = 0; n_match = 0; }
    __syncthreads();
    for (unsigned e = blockIdx.x + threadIdx.x; e < count; e += 256) {          // this row's entries: usually one or two
        const F16Tiny t = items[e];
        if ((ROWS ? t.row : t.col) != mine) continue;
        atomicAdd(&n_match, 1);
        const int slot = atomicAdd(&n_sel, 1);
        if (slot < 64) { sel_k[slot] = ROWS ? t.col : t.row; sel_v[slot] = t.val; }
    }
    __syncthreads();
    const bool many = n_match > 64;                               // (a row with more than 64 noted elements: rounds of 64, picked by k, by one thread)
    int last_k = -1;
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) {
            int n = many ? 0 : n_sel;
            if (many) {
                for (unsigned e = blockIdx.x; e < count; ++e) {
                    const F16Tiny t = items[e];
                    if ((ROWS ? t.row : t.col) != mine) continue;
                    const int k = ROWS ? t.col : t.row;
                    if (k <= last_k || (n == 64 && k > sel_k[63])) continue;
                    int y = (n < 64 ? n : 63) - 1;
                    for (; y >= 0 && sel_k[y] > k; --y) { sel_k[y + 1] = sel_k[y]; sel_v[y + 1] = sel_v[y]; }
                    sel_k[y + 1] = k; sel_v[y + 1] = t.val;
                    if (n < 64) ++n;
                }
            } else {
                for (int x = 1; x < n; ++x) {                    // insertion sort by k
                    const int kx = sel_k[x]; const float vx = sel_v[x];
                    int y = x - 1;
                    for (; y >= 0 && sel_k[y] > kx; --y) { sel_k[y + 1] = sel_k[y]; sel_v[y + 1] = sel_v[y]; }
                    sel_k[y + 1] = kx; sel_v[y + 1] = vx;
                }
            }
            n_sel = n;
        }
        __syncthreads();
        const int ns = n_sel;
        if (ns == 0) break;
        // this block's 1024 outputs of the row / column (blockIdx.y), four per thread, their loads issued together
        const size_t o0 = (size_t)blockIdx.y * 1024 + threadIdx.x;
        float acc[4];
        if (ROWS) {
#pragma unroll
            for (int u = 0; u < 4; ++u) { const size_t j = o0 + 256 * u; acc[u] = j < N ? C[(size_t)mine * N + j] : 0.f; }
            for (int x = 0; x < ns; ++x) {
                float bv[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) { const size_t j = o0 + 256 * u; bv[u] = j < N ? B[(size_t)sel_k[x] * N + j] : 0.f; }
#pragma unroll
                for (int u = 0; u < 4; ++u) acc[u] = fmaf(sel_v[x], bv[u], acc[u]);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) { const size_t j = o0 + 256 * u; if (j < N) C[(size_t)mine * N + j] = acc[u]; }
        } else {
            float undo[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) { const size_t i = o0 + 256 * u; acc[u] = i < M ? C[i * N + mine] : 0.f; undo[u] = i < M ? __int_as_float((127 - ea[i]) << 23) : 0.f; }
            for (int x = 0; x < ns; ++x) {
                float av[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) { const size_t i = o0 + 256 * u; av[u] = i < M ? A[i * K + sel_k[x]] : 0.f; }
#pragma unroll
                for (int u = 0; u < 4; ++u) if (!(fabsf(av[u] * undo[u]) < F16_KEEP_MIN && av[u] != 0.0f)) acc[u] = fmaf(av[u], sel_v[x], acc[u]);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) { const size_t i = o0 + 256 * u; if (i < M) C[i * N + mine] = acc[u]; }
        }
        last_k = sel_k[ns - 1];
        if (!many || ns < 64) break;
    }
}

// ------------------------------------------------------------------ pre-pass
// x -> hi = tf32(x) (round to nearest), lo = tf32(x - hi).  The split is only faithful for finite
// values away from the ends of the exponent range (an infinity would turn into NaN through
// inf * 0, a tiny value would lose its lo part to underflow), so the pre-pass also raises *flag
// when it sees anything else and the caller then leaves the problem to the SIMT engine.
__device__ __forceinline__ void split_one(float x, float& hi, float& lo, bool& bad) {
    uint32_t hb, lb;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hb) : "f"(x));
    hi = __uint_as_float(hb);
    const float d = x - hi;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(lb) : "f"(d));
    lo = __uint_as_float(lb);
    const float ax = fabsf(x);
    bad |= !(ax <= 1.2676506e30f) || (ax < 7.8886091e-31f && ax != 0.0f);      // outside [2^-100, 2^100] (or NaN)
}

// "A direct" (experiment, la_gemm_tc_set_a_direct): the tensor cores take a tf32 operand from the upper 19 bits of an f32 word, so
// the hi plane of a K-major operand can be the operand itself -- no hi plane is written -- provided lo is taken against the
// TRUNCATED value: hi = x with its low 13 mantissa bits cleared, lo = x - hi (exact, 0 <= lo/x < 2^-10).
__device__ __forceinline__ float trunc_tf32(float x) { return __uint_as_float(__float_as_uint(x) & 0xFFFFE000u); }
__device__ __forceinline__ bool out_of_split_range(float x) {
    const float ax = fabsf(x);
    return !(ax <= 1.2676506e30f) || (ax < 7.8886091e-31f && ax != 0.0f);      // outside [2^-100, 2^100] (or NaN)
}

// Cross-term planes of the tf32 + 2 x bf16 engine: bf16(hi) and bf16(x - hi), four values = 8 bytes each.
__device__ __forceinline__ uint2 pack_bf16x4(float a, float b, float c, float d) {
    const __nv_bfloat162 p0 = __floats2bfloat162_rn(a, b), p1 = __floats2bfloat162_rn(c, d);
    return make_uint2(*reinterpret_cast<const uint32_t*>(&p0), *reinterpret_cast<const uint32_t*>(&p1));
}
// BF16X: the `lo` plane (4 bytes per element) holds two bf16 planes instead, [bf16(hi) | bf16(x - hi)], n elements each.
template <bool BF16X>
__global__ void __launch_bounds__(256)
split_tf32_kernel(float* __restrict__ hi, float* __restrict__ lo, const float* __restrict__ x, size_t n4, unsigned* flag, int direct) {
    const size_t stride = (size_t)gridDim.x * 256;
    bool bad = false;
    uint2* hb = reinterpret_cast<uint2*>(lo);
    uint2* lb = hb + n4;                                  // n4 groups of 4 bf16 = the first plane
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += stride) {
        const float4 v = __ldcs(reinterpret_cast<const float4*>(x) + i);
        float4 h, l;
        if (BF16X && direct) {
            h = make_float4(trunc_tf32(v.x), trunc_tf32(v.y), trunc_tf32(v.z), trunc_tf32(v.w));
            bad |= out_of_split_range(v.x) | out_of_split_range(v.y) | out_of_split_range(v.z) | out_of_split_range(v.w);
        } else {
            split_one(v.x, h.x, l.x, bad); split_one(v.y, h.y, l.y, bad); split_one(v.z, h.z, l.z, bad); split_one(v.w, h.w, l.w, bad);
            __stcs(reinterpret_cast<float4*>(hi) + i, h);
        }
        if constexpr (BF16X) {
            __stcs(hb + i, pack_bf16x4(h.x, h.y, h.z, h.w));
            __stcs(lb + i, pack_bf16x4(v.x - h.x, v.y - h.y, v.z - h.z, v.w - h.w));
        } else {
            __stcs(reinterpret_cast<float4*>(lo) + i, l);
        }
    }
    if (__syncthreads_or(bad) && threadIdx.x == 0) *(volatile unsigned*)flag = 1u;     // flag may live in mapped host memory: plain store
}

// B (K x N) -> Bt_hi, Bt_lo (N x K) in one pass: 64 x 64 tile through shared memory, 16-byte
// loads along B's rows, 16-byte stores along Bt's rows.  K % 4 == 0 and N % 4 == 0.
template <bool BF16X>
__global__ void __launch_bounds__(256)
split_transpose_kernel(float* __restrict__ bt_hi, float* __restrict__ bt_lo, const float* __restrict__ B, size_t K, size_t N,
                       unsigned tiles_n, unsigned* flag) {
    __shared__ float tile[64][65];
    const size_t k0 = (size_t)(blockIdx.x / tiles_n) * 64, n0 = (size_t)(blockIdx.x % tiles_n) * 64;
    float4 in[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int idx = threadIdx.x + j * 256, r = idx / 16, q = idx % 16;
        in[j] = (k0 + r < K && n0 + q * 4 < N) ? __ldcs(reinterpret_cast<const float4*>(B + (k0 + r) * N + n0 + q * 4)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    __syncwarp();                            // all four loads are issued before the first shared-memory store
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int idx = threadIdx.x + j * 256, r = idx / 16, q = idx % 16;
        tile[r][q * 4 + 0] = in[j].x; tile[r][q * 4 + 1] = in[j].y; tile[r][q * 4 + 2] = in[j].z; tile[r][q * 4 + 3] = in[j].w;
    }
    __syncthreads();
    bool bad = false;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int idx = threadIdx.x + j * 256, cc = idx / 16, q = idx % 16;      // Bt row n0+cc, columns k0 + q*4 ..
        float4 h, l;
        split_one(tile[q * 4 + 0][cc], h.x, l.x, bad); split_one(tile[q * 4 + 1][cc], h.y, l.y, bad);
        split_one(tile[q * 4 + 2][cc], h.z, l.z, bad); split_one(tile[q * 4 + 3][cc], h.w, l.w, bad);
        if (n0 + cc < N && k0 + q * 4 < K) {
            const size_t at = (n0 + cc) * K + k0 + q * 4;
            __stcs(reinterpret_cast<float4*>(bt_hi + at), h);
            if constexpr (BF16X) {
                uint2* hb = reinterpret_cast<uint2*>(bt_lo);
                __stcs(hb + at / 4, pack_bf16x4(h.x, h.y, h.z, h.w));
                __stcs(hb + (N * K + at) / 4, pack_bf16x4(tile[q * 4 + 0][cc] - h.x, tile[q * 4 + 1][cc] - h.y, tile[q * 4 + 2][cc] - h.z, tile[q * 4 + 3][cc] - h.w));
            } else {
                __stcs(reinterpret_cast<float4*>(bt_lo + at), l);
            }
        }
    }
    if (__syncthreads_or(bad) && threadIdx.x == 0) *(volatile unsigned*)flag = 1u;     // flag may live in mapped host memory: plain store
}


// Both pre-passes as ONE launch: blocks 0 .. b_tiles-1 each turn and split one 64 x 64 tile of B, the blocks after them
// split A in a grid-stride loop.  One launch less per product, and the A blocks fill the SMs the B tiles' last wave leaves idle.
// The planes are bit for bit what split_tf32_kernel and split_transpose_kernel write.
template <bool BF16X>
__global__ void __launch_bounds__(256)
split_both_kernel(float* __restrict__ a_hi, float* __restrict__ a_lo, const float* __restrict__ A, size_t a_n4,
                  float* __restrict__ bt_hi, float* __restrict__ bt_lo, const float* __restrict__ B, size_t K, size_t N, unsigned tiles_n, unsigned b_tiles,
                  unsigned* flag, int direct) {
    __shared__ float tile[64][65];
    bool bad = false;
    if (blockIdx.x < b_tiles) {
        const size_t k0 = (size_t)(blockIdx.x / tiles_n) * 64, n0 = (size_t)(blockIdx.x % tiles_n) * 64;
        float4 in[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int idx = threadIdx.x + j * 256, r = idx / 16, q = idx % 16;
            in[j] = (k0 + r < K && n0 + q * 4 < N) ? __ldcs(reinterpret_cast<const float4*>(B + (k0 + r) * N + n0 + q * 4)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        __syncwarp();                        // all four loads are issued before the first shared-memory store (ptxas otherwise sinks the last one below it)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int idx = threadIdx.x + j * 256, r = idx / 16, q = idx % 16;
            tile[r][q * 4 + 0] = in[j].x; tile[r][q * 4 + 1] = in[j].y; tile[r][q * 4 + 2] = in[j].z; tile[r][q * 4 + 3] = in[j].w;
        }
        __syncthreads();
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int idx = threadIdx.x + j * 256, cc = idx / 16, q = idx % 16;      // Bt row n0+cc, columns k0 + q*4 ..
            float4 h, l;
            split_one(tile[q * 4 + 0][cc], h.x, l.x, bad); split_one(tile[q * 4 + 1][cc], h.y, l.y, bad);
            split_one(tile[q * 4 + 2][cc], h.z, l.z, bad); split_one(tile[q * 4 + 3][cc], h.w, l.w, bad);
            if (n0 + cc < N && k0 + q * 4 < K) {
                const size_t at = (n0 + cc) * K + k0 + q * 4;
                __stcs(reinterpret_cast<float4*>(bt_hi + at), h);
                if constexpr (BF16X) {
                    uint2* hb = reinterpret_cast<uint2*>(bt_lo);
                    __stcs(hb + at / 4, pack_bf16x4(h.x, h.y, h.z, h.w));
                    __stcs(hb + (N * K + at) / 4, pack_bf16x4(tile[q * 4 + 0][cc] - h.x, tile[q * 4 + 1][cc] - h.y, tile[q * 4 + 2][cc] - h.z, tile[q * 4 + 3][cc] - h.w));
                } else {
                    __stcs(reinterpret_cast<float4*>(bt_lo + at), l);
                }
            }
        }
    } else {
        const size_t stride = (size_t)(gridDim.x - b_tiles) * 256;
        uint2* hb = reinterpret_cast<uint2*>(a_lo);
        uint2* lb = hb + a_n4;
        for (size_t i = (size_t)(blockIdx.x - b_tiles) * 256 + threadIdx.x; i < a_n4; i += stride) {
            const float4 v = __ldcs(reinterpret_cast<const float4*>(A) + i);
            float4 h, l;
            if (BF16X && direct) {
                h = make_float4(trunc_tf32(v.x), trunc_tf32(v.y), trunc_tf32(v.z), trunc_tf32(v.w));
                bad |= out_of_split_range(v.x) | out_of_split_range(v.y) | out_of_split_range(v.z) | out_of_split_range(v.w);
            } else {
                split_one(v.x, h.x, l.x, bad); split_one(v.y, h.y, l.y, bad); split_one(v.z, h.z, l.z, bad); split_one(v.w, h.w, l.w, bad);
                __stcs(reinterpret_cast<float4*>(a_hi) + i, h);
            }
            if constexpr (BF16X) {
                __stcs(hb + i, pack_bf16x4(h.x, h.y, h.z, h.w));
                __stcs(lb + i, pack_bf16x4(v.x - h.x, v.y - h.y, v.z - h.z, v.w - h.w));
            } else {
                __stcs(reinterpret_cast<float4*>(a_lo) + i, l);
            }
        }
    }
    if (__syncthreads_or(bad) && threadIdx.x == 0) *(volatile unsigned*)flag = 1u;
}

// rows x K row-major f32, box = box_rows x 32 (128 bytes), 128-byte swizzle
la_status tc_make_map(CUtensorMap* map, const float* base, size_t rows, size_t K, int box_rows) {
    EncodeTiledFn enc; LA_TRY(tc_get_encode(&enc));
    cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)K * 4};
    cuuint32_t box[2] = {(cuuint32_t)TC_BK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return la_fail(LA_ERR_CUDA, "cuTensorMapEncodeTiled (3xTF32) failed with CUresult %d", (int)r);
    return LA_OK;
}

// rows x K row-major bf16, box = box_rows x 32 (64 bytes), 64-byte swizzle
la_status tc_make_map_bf16(CUtensorMap* map, const void* base, size_t rows, size_t K, int box_rows) {
    EncodeTiledFn enc; LA_TRY(tc_get_encode(&enc));
    cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)K * 2};
    cuuint32_t box[2] = {(cuuint32_t)TC_BK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return la_fail(LA_ERR_CUDA, "cuTensorMapEncodeTiled (bf16 cross planes) failed with CUresult %d", (int)r);
    return LA_OK;
}

// rows x K row-major 16-bit plane (fp16 or bf16), box = box_rows x 32 (64 bytes), 64-byte swizzle: the f16-split engine's operands
la_status tc_make_map_16(CUtensorMap* map, const void* base, bool bf16, size_t rows, size_t K, int box_rows) {
    EncodeTiledFn enc; LA_TRY(tc_get_encode(&enc));
    cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)K * 2};
    cuuint32_t box[2] = {(cuuint32_t)F16_BK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return la_fail(LA_ERR_CUDA, "cuTensorMapEncodeTiled (f16-split planes) failed with CUresult %d", (int)r);
    return LA_OK;
}

template <int BN, int STAGES>
la_status tc_launch(float* C, const float* a_hi, const float* a_lo, const float* b_hi, const float* b_lo, size_t M, size_t K, size_t N, cudaStream_t st) {
    using S = TcShape<BN, STAGES>;
    CUtensorMap ma_hi, ma_lo, mb_hi, mb_lo;
    LA_TRY(tc_make_map(&ma_hi, a_hi, M, K, 128)); LA_TRY(tc_make_map(&ma_lo, a_lo, M, K, 128));
    // a TMA box is at most 256 rows: B_TILE of BN rows in one request
    LA_TRY(tc_make_map(&mb_hi, b_hi, N, K, BN)); LA_TRY(tc_make_map(&mb_lo, b_lo, N, K, BN));
    auto kern = gemm_3xtf32_kernel<BN, STAGES>;
    LA_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::SMEM_BYTES));
    const size_t tm = (M + TC_BM - 1) / TC_BM, tn = (N + BN - 1) / BN;
    kern<<<(unsigned)(tm * tn), TC_THREADS, S::SMEM_BYTES, st>>>(C, ma_hi, ma_lo, mb_hi, mb_lo, (int)M, (int)N, (int)K, (int)tm);
    LA_LAUNCH_CHECK();
    return LA_OK;
}

template <int CLUSTER>
la_status tc_launch_chunked(float* C, const float* a_hi, const float* a_lo, const float* b_hi, const float* b_lo, size_t M, size_t K, size_t N, cudaStream_t st) {
    using S = TcShape<TC2_BN, TC2_STAGES>;
    CUtensorMap ma_hi, ma_lo, mb_hi, mb_lo;
    LA_TRY(tc_make_map(&ma_hi, a_hi, M, K, 128)); LA_TRY(tc_make_map(&ma_lo, a_lo, M, K, 128));
    LA_TRY(tc_make_map(&mb_hi, b_hi, N, K, TC2_BN / CLUSTER)); LA_TRY(tc_make_map(&mb_lo, b_lo, N, K, TC2_BN / CLUSTER));
    auto kern = gemm_3xtf32_chunked_kernel<CLUSTER>;
    LA_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::SMEM_BYTES));
    size_t tm = (M + TC_BM - 1) / TC_BM;
    const size_t tn = (N + TC2_BN - 1) / TC2_BN;
    tm = (tm + CLUSTER - 1) / CLUSTER * CLUSTER;      // whole clusters along M; surplus tiles are all out of bounds
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(tm * tn)); cfg.blockDim = dim3(TC2_THREADS); cfg.dynamicSmemBytes = S::SMEM_BYTES; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CLUSTER; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    LA_CUDA(cudaLaunchKernelEx(&cfg, kern, C, ma_hi, ma_lo, mb_hi, mb_lo, (int)M, (int)N, (int)K, (int)tm));
    return LA_OK;
}

int g_tc_stream = -1;           // la_gemm_set_stream_tail: -1 by the gain model below, 0 never, 1 whenever possible

// Kernel-only timing for bench.py's roofline line: with la_gemm_kernel_timing(1) every launch of the pair kernel is bracketed by
// two events on its own stream (up to 64 launches are kept); la_gemm_kernel_time_ms() waits for them and returns their sum.
int g_tc_timing = 0;
int g_tc_timed = 0;
cudaEvent_t g_tc_ev[64][2];
bool g_tc_ev_made = false;

int g_tc_persist = 0;           // la_gemm_tc_set_persistent: 0 one CTA pair per work item, 1 persistent grid of sm_count / 2 pairs
int g_tc_chunk = 0;             // la_gemm_tc_set_chunk: k-blocks of 32 per accumulation chain (0 = by K)
int g_tc_bn = 0;                // la_gemm_tc_set_tile_n: 0 = pick by wave count, else 256 / 224 / 192

// Wave-times (in units of a 256-wide tile's) that `tiles` tiles of width bn take on `pairs` CTA pairs, with the half-length
// last wave the split gives when the remainder fits half of the GPU.
static double tc_cost(long long tiles, long long pairs, int bn, bool can_split) {
    const long long waves = tiles / pairs, rest = tiles % pairs;
    const double last = rest == 0 ? 0.0 : (can_split && 2 * rest <= pairs) ? 0.5 : 1.0;
    return ((double)waves + last) * (double)bn / 256.0;
}

// Accumulation chains (k-blocks of 32 summed inside the tensor cores before the FP32 registers take over).  The tensor
// cores' accumulation truncates, and a lone chain is where that shows most (K = 64 as one chain, 16M outputs: 6.5e-7 of
// sum|a||b| against a bar of 5.9e-7), so short products use chains of 32; over long products the chain errors average
// out (K = 8192: 4.8e-8 with 64-chains, 6.2e-8 with 128, 9.6e-8 with 256, bar 5.4e-7) and longer chains halve the
// accumulator drains that otherwise bound the kernel once the cross terms run as bf16 (8192^3: 3.63 -> 3.30 -> 3.28 ms).
// Round 2, measured over 16M outputs against the unrelaxed gate e <= max(E_reference_order, 2^-22) (tools/short_k_probe.py): with
// 64-chains K = 256 comes out at 2.9e-7 against the reference order's 3.1 - 3.7e-7 -- inside, but by so little that a
// smaller problem (512 x 256 x 384) fell 1 % outside by chance; with 32-chains and tf32 cross terms it is 1.4e-7, K = 384
// 1.2e-7; K = 512 with 64-chains and bf16 cross terms 2.0 - 2.3e-7 (0.83 of the bar on a 1024^2 output), K = 1024 1.5e-7
// (profiles/r2_short_k_probe.txt).  So products below K = 1024 use chains of 32 (and tf32 cross terms, la_tc_cross_bf16).
int tc_chunk_kb(size_t K) { return g_tc_chunk > 0 ? g_tc_chunk : K < 1024 ? 1 : K >= 8192 ? 8 : K >= 2048 ? 4 : TC_CHUNK_KB; }

// Work plan of one product on the CTA-pair kernels (this engine's and the f16-split engine's): tiles of 256 x BN, accumulation
// chains of `chunk_kb` k-blocks (num_kb k-blocks in all, of the calling engine's k-block size), the K-split of a ragged last
// wave -- and the zeroing of the part of C that then receives two partial sums.
la_status tc_make_plan(float* C, size_t M, size_t N, int BN, int num_kb, int chunk_kb, cudaStream_t st, long long pairs, TcPlan* out, long long* out_items) {
    const size_t ptm = (M + 255) / 256, tn = (N + BN - 1) / BN;
    const long long tiles = (long long)(ptm * tn);
    TcPlan plan = {};
    plan.pair_tiles_m = (int)ptm; plan.tiles_n = (int)tn; plan.raster_group = la_gemm_raster_group(ptm);
    plan.num_kb = num_kb;
    plan.chunk_kb = chunk_kb;
    plan.num_chunks = (plan.num_kb + plan.chunk_kb - 1) / plan.chunk_kb;
    plan.dp_items = (int)tiles; plan.stream_items = 0; plan.stream_tiles = 0;
    // Ragged last wave: `waves` full waves and `rest` tiles more.  When the remainder fills at most half of the GPU, each of
    // its tiles is cut in two along K (two work items of half the length), so the last wave takes half a wave-time: 4096^3
    // is 3.46 waves of tiles -> 3.5 instead of 4.  Both halves of all tiles start together and advance in step, like the
    // whole tiles do, which keeps the operand slabs shared through L2.  (Cutting the last full wave plus the remainder into
    // equal spans would balance any remainder, but the spans then start at staggered K positions, the CTAs stop sharing
    // slabs in L2 and the kernel turns HBM-bound: measured 11 % slower at 2048 x 8192 x 8192.)
    const long long rest = tiles % pairs;
    if (rest > 0 && 2 * rest <= pairs && plan.num_chunks >= 2 && g_tc_stream != 0) {
        plan.stream_tiles = (int)rest;
        plan.stream_items = (int)(2 * rest);
        plan.dp_items = (int)(tiles - rest);
        // zero the tiles that receive two partial sums (whole tiles inside the zeroed rectangle simply overwrite it later).
        // m-major order: the last tile columns.  Banded order: the last tile columns of the last band.
        if (plan.raster_group <= 0 || plan.raster_group >= plan.pair_tiles_m) {
            const size_t first_col = ((size_t)plan.dp_items / ptm) * BN;
            LA_CUDA(cudaMemset2DAsync(C + first_col, N * sizeof(float), 0, (N - first_col) * sizeof(float), M, st));
        } else {
            const size_t bands = (ptm + plan.raster_group - 1) / plan.raster_group;
            const size_t first_m = (bands - 1) * plan.raster_group, band_rows = ptm - first_m;      // tile rows of the last band
            if ((size_t)rest <= band_rows * tn) {
                const size_t cols = ((size_t)rest + band_rows - 1) / band_rows;                      // tile columns holding the split tiles
                const size_t first_col = (tn - cols) * BN, first_row = first_m * 256;
                LA_CUDA(cudaMemset2DAsync(C + first_row * N + first_col, N * sizeof(float), 0, (N - first_col) * sizeof(float), M - first_row, st));
            } else {
                LA_CUDA(cudaMemsetAsync(C, 0, M * N * sizeof(float), st));
            }
        }
    }
    long long items = plan.dp_items + plan.stream_items;
    plan.persist = 0;
    if (g_tc_persist != 0 && items > pairs) { plan.persist = (int)pairs; items = pairs; }
    *out = plan; *out_items = items;
    return LA_OK;
}
void tc_pair_launch_config(cudaLaunchConfig_t* cfg, cudaLaunchAttribute* attr, long long items, cudaStream_t st) {
    *cfg = cudaLaunchConfig_t{};
    cfg->gridDim = dim3((unsigned)(2 * items)); cfg->blockDim = dim3(TC3_THREADS); cfg->dynamicSmemBytes = TC3_SMEM_BYTES; cfg->stream = st;
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg->attrs = attr; cfg->numAttrs = 1;
}
// la_gemm_kernel_timing: events around the launch of the product kernel, on its stream
la_status tc_timing_begin(cudaStream_t st) {
    if (!(g_tc_timing && g_tc_timed < 64)) return LA_OK;
    if (!g_tc_ev_made) { for (auto& e : g_tc_ev) { LA_CUDA(cudaEventCreate(&e[0])); LA_CUDA(cudaEventCreate(&e[1])); } g_tc_ev_made = true; }
    LA_CUDA(cudaEventRecord(g_tc_ev[g_tc_timed][0], st));
    return LA_OK;
}
la_status tc_timing_end(cudaStream_t st) {
    if (!(g_tc_timing && g_tc_timed < 64)) return LA_OK;
    LA_CUDA(cudaEventRecord(g_tc_ev[g_tc_timed][1], st)); ++g_tc_timed;
    return LA_OK;
}

template <int BN, bool BF16X>
la_status tc_launch_2cta_bn(float* C, const float* a_hi, const float* a_lo, const float* b_hi, const float* b_lo, size_t M, size_t K, size_t N, cudaStream_t st, long long pairs) {
    CUtensorMap ma_hi, ma_lo, mb_hi, mb_lo, ma_x, mb_x;
    LA_TRY(tc_make_map(&ma_hi, a_hi, M, K, 128)); LA_TRY(tc_make_map(&mb_hi, b_hi, N, K, BN / 2));
    if (BF16X) {        // the lo planes hold [bf16(hi) | bf16(x - hi)]
        const __nv_bfloat16* a16 = reinterpret_cast<const __nv_bfloat16*>(a_lo); const __nv_bfloat16* b16 = reinterpret_cast<const __nv_bfloat16*>(b_lo);
        LA_TRY(tc_make_map_bf16(&ma_lo, a16, M, K, 128)); LA_TRY(tc_make_map_bf16(&ma_x, a16 + M * K, M, K, 128));
        LA_TRY(tc_make_map_bf16(&mb_lo, b16, N, K, BN / 2)); LA_TRY(tc_make_map_bf16(&mb_x, b16 + N * K, N, K, BN / 2));
    } else {
        LA_TRY(tc_make_map(&ma_lo, a_lo, M, K, 128)); LA_TRY(tc_make_map(&mb_lo, b_lo, N, K, BN / 2));
        ma_x = ma_lo; mb_x = mb_lo;
    }
    TcPlan plan; long long items;
    LA_TRY(tc_make_plan(C, M, N, BN, (int)((K + TC_BK - 1) / TC_BK), tc_chunk_kb(K), st, pairs, &plan, &items));
    auto kern = gemm_3xtf32_2cta_kernel<BN, BF16X>;
    LA_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC3_SMEM_BYTES));
    cudaLaunchConfig_t cfg; cudaLaunchAttribute attr[1];
    tc_pair_launch_config(&cfg, attr, items, st);
    LA_TRY(tc_timing_begin(st));
    LA_CUDA(cudaLaunchKernelEx(&cfg, kern, C, ma_hi, ma_lo, mb_hi, mb_lo, ma_x, mb_x, (int)M, (int)N, plan));
    LA_TRY(tc_timing_end(st));
    return LA_OK;
}

// Cross terms in bf16?  hi*lo and lo*hi carry 2^-11 of a product's weight, so bf16's 8 significant bits leave an error of
// about 2^-20 of |a||b| per term with random sign: after a sum of K terms, 2^-20 * 1.3 / sqrt(K) of sum|a||b| (simulated:
// K = 8192 max 4e-8 over 150k outputs, K = 1024 1.2e-7) -- against the reference's own ~2.4e-7.  Two bf16 MMAs cost what
// one tf32 MMA does, so the product needs 2 tensor-core time units instead of 3 (8192^3: 3.97 -> 3.61 ms, and 3.28 ms with
// the longer accumulation chains the faster MMAs call for).  Measured on the GPU (tools/acc_probe.py) the error is the same
// as with tf32 cross terms within 20 % at every K tried (K = 256: 2.4e-7 vs 2.2e-7, bar 4.6e-7; K = 8192: 4.8e-8 vs 4.0e-8):
// the tensor cores' truncating accumulation dominates both.  With 16M outputs, K = 256: 2.9e-7 (tf32 cross terms 2.5e-7, the
// reference order itself 3.6e-7), and 4096 x 512 x 4096 runs 0.124 -> 0.113 ms.  Used from K = 1024 (round 2: at K = 256 the
// bf16 form sits at 0.8 - 0.9 of the reference order's error, at K = 512 at 0.6 - 0.85: too close to the unrelaxed gate); shorter products keep three
// tf32 MMAs.
int g_tc_variant = 0;
int g_tc_cross = -1;            // la_gemm_tc_set_cross: -1 by K, 0 tf32 cross terms, 1 bf16 cross terms whenever K % 8 == 0
int g_tc_adirect = 0;           // la_gemm_tc_set_a_direct: A's hi plane is A itself (tensor cores truncate), see trunc_tf32
bool la_tc_cross_bf16(size_t K);
bool tc_a_direct(size_t K) { return g_tc_adirect != 0 && g_tc_variant == 0 && la_tc_cross_bf16(K); }
// Round 2, end: bf16 cross terms are a tuning setting only (la_gemm_tc_set_cross(1)).  Their ~2^-20 |a||b| per product is above the
// gate for an output that is essentially one product (B = identity at K >= 1024: 9.5e-7 of sum|a||b| against a bar of 2.4e-7), and
// the f16-split form now does the fast path's job at 1.5 units with ~2^-23 per product; this form keeps tf32 cross terms at every K.
bool la_tc_cross_bf16(size_t K) {
    if (K % 8 != 0 || g_tc_variant != 0) return false;
    return g_tc_cross > 0;
}

int tc_pick_bn(size_t M, size_t K, size_t N, long long pairs) {
    if (g_tc_bn) return g_tc_bn;
    const long long ptm = (long long)((M + 255) / 256);
    const bool can_split = g_tc_stream != 0 && K > 32;
    int bn = 256;
    double best = tc_cost(ptm * (long long)((N + 255) / 256), pairs, 256, can_split);
    for (int cand : {224, 192}) {
        const double c = tc_cost(ptm * (long long)((N + cand - 1) / cand), pairs, cand, can_split);
        if (c < 0.90 * best) { best = c; bn = cand; }       // a 224-wide tile costs ~7 % more per flop (8192^3: 4.86 -> 5.23 ms at equal wave count): it must buy more than that
    }
    return bn;
}
// Tile width.  256 x 256 per CTA pair unless a narrower tile packs the waves better: the 8-GPU shard of the 8192^3 product,
// 1024 x 8192 x 8192, is 128 tiles = 1.73 waves -> 2 wave-times at width 256, but 148 tiles = exactly 2 waves of 0.875 at 224
// (measured 0.666 -> 0.645 ms).
la_status tc_launch_2cta(float* C, const float* a_hi, const float* a_lo, const float* b_hi, const float* b_lo, size_t M, size_t K, size_t N, cudaStream_t st) {
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    const long long pairs = info->sm_count / 2;                 // CTA pairs in flight (one per SM pair)
    const int bn = tc_pick_bn(M, K, N, pairs);
    if (la_tc_cross_bf16(K)) {
        switch (bn) {
        case 224: return tc_launch_2cta_bn<224, true>(C, a_hi, a_lo, b_hi, b_lo, M, K, N, st, pairs);
        case 192: return tc_launch_2cta_bn<192, true>(C, a_hi, a_lo, b_hi, b_lo, M, K, N, st, pairs);
        default:  return tc_launch_2cta_bn<256, true>(C, a_hi, a_lo, b_hi, b_lo, M, K, N, st, pairs);
        }
    }
    switch (bn) {
    case 224: return tc_launch_2cta_bn<224, false>(C, a_hi, a_lo, b_hi, b_lo, M, K, N, st, pairs);
    case 192: return tc_launch_2cta_bn<192, false>(C, a_hi, a_lo, b_hi, b_lo, M, K, N, st, pairs);
    default:  return tc_launch_2cta_bn<256, false>(C, a_hi, a_lo, b_hi, b_lo, M, K, N, st, pairs);
    }
}


// ---- f16-split engine: launch.  Chains: the tf32 form's (tc_chunk_kb).
int g_tc_f16 = -1;              // la_gemm_tc_set_f16: -1 by K, 0 never, 1 wherever the shapes allow
thread_local int g_tc_last_form = -1;       // la_gemm_tc_last_form: 0 tf32 form, 1 f16-split form (this thread's last tensor-engine product)
thread_local int g_tc_last_launches = 0;    // la_gemm_tc_last_launches: kernels that product launched
struct F16Planes { const __half* hi; const __half* hb; const __half* lb; };      // hi = fp16(x), hs = hi * 2^-11, ls = fp16((x - hi) * 2^11) of a scaled operand
template <int BN>
la_status f16_launch_bn(float* C, F16Planes a, F16Planes b, const int* ea, const int* eb, size_t M, size_t K, size_t N, cudaStream_t st, long long pairs) {
    CUtensorMap ma_hi, ma_hb, ma_lb, mb_hi, mb_hb, mb_lb;
    LA_TRY(tc_make_map_16(&ma_hi, a.hi, false, M, K, 128)); LA_TRY(tc_make_map_16(&ma_hb, a.hb, false, M, K, 128)); LA_TRY(tc_make_map_16(&ma_lb, a.lb, false, M, K, 128));
    LA_TRY(tc_make_map_16(&mb_hi, b.hi, false, N, K, BN / 2)); LA_TRY(tc_make_map_16(&mb_hb, b.hb, false, N, K, BN / 2)); LA_TRY(tc_make_map_16(&mb_lb, b.lb, false, N, K, BN / 2));
    TcPlan plan; long long items;
    // chains as in the tf32 form from K = 1024; below (tuning only) 64 k: measured 1.3 - 1.6e-7 at K = 256 over 16M outputs against a bar of 3.6 - 4.3e-7
    const int chunk = (K < 1024 && g_tc_chunk <= 0) ? 2 : tc_chunk_kb(K);
    LA_TRY(tc_make_plan(C, M, N, BN, (int)((K + F16_BK - 1) / F16_BK), chunk, st, pairs, &plan, &items));
    auto kern = gemm_f16split_2cta_kernel<BN>;
    LA_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC3_SMEM_BYTES));
    cudaLaunchConfig_t cfg; cudaLaunchAttribute attr[1];
    tc_pair_launch_config(&cfg, attr, items, st);
    LA_TRY(tc_timing_begin(st));
    LA_CUDA(cudaLaunchKernelEx(&cfg, kern, C, ma_hi, ma_hb, ma_lb, mb_hi, mb_hb, mb_lb, ea, eb, (int)M, (int)N, plan));
    LA_TRY(tc_timing_end(st));
    return LA_OK;
}
la_status f16_launch(float* C, F16Planes a, F16Planes b, const int* ea, const int* eb, size_t M, size_t K, size_t N, cudaStream_t st) {
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    const long long pairs = info->sm_count / 2;
    switch (tc_pick_bn(M, K, N, pairs)) {
    case 224: return f16_launch_bn<224>(C, a, b, ea, eb, M, K, N, st, pairs);
    case 192: return f16_launch_bn<192>(C, a, b, ea, eb, M, K, N, st, pairs);
    default:  return f16_launch_bn<256>(C, a, b, ea, eb, M, K, N, st, pairs);
    }
}
// the pre-pass of one operand
// planes of an operand with `elems` elements in one block: fp16 hi | bf16 hi | bf16 lo
F16Planes f16_planes(void* base, size_t elems) {
    F16Planes p; p.hi = (const __half*)base; p.hb = (const __half*)((char*)base + elems * 2); p.lb = (const __half*)((char*)base + elems * 4);
    return p;
}
// Scratch of one operand's pre-pass beyond its planes: exponents, the list of left-out elements, and (B) the partial column maxima.
struct F16Side { int* e; F16Tiny* items; unsigned* count; unsigned* colmax_part; float* colsum_part; };
constexpr unsigned F16_MAX_SLICES = 32;
size_t f16_side_bytes(size_t n, bool with_colmax) {              // n = M (A) or N (B)
    const size_t n4 = (n + 3) / 4 * 4;
    return n4 * sizeof(int) + F16_TINY_CAP * sizeof(F16Tiny) + 16 + (with_colmax ? 2 * (size_t)F16_MAX_SLICES * n4 * sizeof(unsigned) : 0);
}
F16Side f16_side(char* base, size_t n) {
    const size_t n4 = (n + 3) / 4 * 4;
    F16Side s; s.e = (int*)base; s.items = (F16Tiny*)(base + n4 * sizeof(int)); s.count = (unsigned*)(base + n4 * sizeof(int) + F16_TINY_CAP * sizeof(F16Tiny));
    s.colmax_part = (unsigned*)((char*)s.count + 16);
    s.colsum_part = (float*)(s.colmax_part + (size_t)F16_MAX_SLICES * n4);
    return s;
}
// The pre-pass: A's rows and (unless B is prepared: `B == nullptr`) B's column maxima in one launch, then B's transposing split.
// any_host: two mapped host words, [0] for A's list, [1] for B's.
la_status f16_prepass(F16Planes a, F16Side sa, const float* A, size_t M, F16Planes bt, F16Side sb, const float* B, size_t K, size_t N,
                      unsigned* flag, unsigned* any_dev, cudaStream_t st) {
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    F16PrepassA pa = {};
    F16PrepassB pb = {};
    if (A) {
        LA_CUDA(cudaMemsetAsync(sa.count, 0, sizeof(unsigned), st));
        pa.hi = const_cast<__half*>(a.hi); pa.hb = const_cast<__half*>(a.hb); pa.lo = const_cast<__half*>(a.lb); pa.ea = sa.e; pa.A = A; pa.rows = (unsigned)M; pa.K = K;
        pa.tiny = F16TinyList{sa.items, sa.count, any_dev};
    }
    unsigned b_blocks = 0;
    if (B) {
        pb.tiny_count = sb.count;
        pb.colmax_part = sb.colmax_part; pb.colsum_part = sb.colsum_part; pb.B = B; pb.K = K; pb.N = N;
        pb.col_blocks = (unsigned)((N / 4 + 255) / 256);
        size_t slices = std::max<size_t>(1, (size_t)info->sm_count * 8 / pb.col_blocks);
        slices = std::min<size_t>(std::min<size_t>(slices, F16_MAX_SLICES), K);
        pb.rows_per_slice = (K + slices - 1) / slices;
        pb.slices = (unsigned)((K + pb.rows_per_slice - 1) / pb.rows_per_slice);
        b_blocks = pb.col_blocks * pb.slices;
    }
    f16_prepass1_kernel<<<pa.rows + b_blocks, 256, 0, st>>>(pa, pb, flag);
    LA_LAUNCH_CHECK();
    if (B) {
        const size_t tk = (K + 63) / 64, tn = (N + 63) / 64;
        if (tk * tn > 0x7fffffffull) return la_fail(LA_ERR_UNSUPPORTED, "f16-split pre-pass: B is too large");
        f16_split_transpose_kernel<<<(unsigned)(tk * tn), 256, 0, st>>>(const_cast<__half*>(bt.hi), const_cast<__half*>(bt.hb), const_cast<__half*>(bt.lb), sb.e,
                                                                        sb.colmax_part, sb.colsum_part, pb.slices, B, K, N, (unsigned)tn, F16TinyList{sb.items, sb.count, any_dev + 1}, flag);
        LA_LAUNCH_CHECK();
    }
    return LA_OK;
}
la_status f16_tiny_fix(float* C, const float* A, const float* B, const int* ea, size_t M, size_t K, size_t N, const F16Side& sa, bool a_any, const F16Tiny* b_items, const unsigned* b_count, bool b_any, cudaStream_t st) {
    if (a_any) { f16_tiny_fix_kernel<true><<<dim3(F16_TINY_CAP, (unsigned)((N + 1023) / 1024)), 256, 0, st>>>(C, A, B, ea, M, K, N, sa.items, sa.count); LA_LAUNCH_CHECK(); }
    if (b_any) { f16_tiny_fix_kernel<false><<<dim3(F16_TINY_CAP, (unsigned)((M + 1023) / 1024)), 256, 0, st>>>(C, A, B, ea, M, K, N, b_items, b_count); LA_LAUNCH_CHECK(); }
    return LA_OK;
}

}  // namespace

bool la_tc_a_direct(size_t K) { return tc_a_direct(K); }      // also asked by gemm_host.cu

extern "C" la_status la_gemm_tc_set_variant(int v) { g_tc_variant = v; return LA_OK; }     // tuning hook
extern "C" int la_gemm_tc_last_form(void) { return g_tc_last_form; }
extern "C" int la_gemm_tc_last_launches(void) { return g_tc_last_launches; }
extern "C" la_status la_gemm_tc_set_f16(int v) { g_tc_f16 = v; return LA_OK; }             // tuning hook: f16-split form (-1 by K, 0 never, 1 wherever the shapes allow)
// Tile order of the tensor engines.  In m-major order one wave of 74 CTA pairs spans every tile row, so each wave streams
// all of A's planes from HBM: 8.75 GB of DRAM reads for the 1 GB of operand planes at 8192^3 (ncu), 3.5 TB/s demanded at
// M = 16384, where the kernel then waits on HBM (16384^3 f32 42.6 ms).  Bands of 8 tile rows keep a wave on ~17 row/column
// slabs whatever the size: 4.57 GB at 8192^3, and -- the engines being power-limited -- less DRAM traffic is also time:
// fresh-process runs (tools/fresh_gemm.py: 3 warm-ups + 5 steps, as bench.py times) f32 8192^3 4.117 -> 3.969 ms,
// i32 3.498 -> 3.274 ms, 16384^3 f32 42.6 -> 36.5 ms; no difference at 4096^3 (16 tile rows; 4096 x 8192 x 8192: 2.108 -> 2.092 ms), below that m-major is better (2048 x 8192 x 8192: 1.110 vs 1.138 ms).
static int initial_raster_group() { const char* e = getenv("LA_GEMM_RASTER"); return e ? atoi(e) : -1; }
static int g_raster_group = initial_raster_group();     // -1: by size; 0: m-major; g > 0: bands of g tile rows (LA_GEMM_RASTER in the environment)
int la_gemm_raster_group(size_t pair_tiles_m) { return g_raster_group >= 0 ? g_raster_group : (pair_tiles_m >= 16 ? 8 : 0); }
extern "C" la_status la_gemm_kernel_timing(int on) { g_tc_timing = on; g_tc_timed = 0; return LA_OK; }      // bench hook (not in the public header)
extern "C" la_status la_gemm_kernel_time_ms(float* total_ms, int* launches) {
    float sum = 0.f;
    for (int j = 0; j < g_tc_timed; ++j) {
        float ms = 0.f;
        LA_CUDA(cudaEventSynchronize(g_tc_ev[j][1]));
        LA_CUDA(cudaEventElapsedTime(&ms, g_tc_ev[j][0], g_tc_ev[j][1]));
        sum += ms;
    }
    if (total_ms) *total_ms = sum;
    if (launches) *launches = g_tc_timed;
    g_tc_timed = 0;
    return LA_OK;
}
extern "C" la_status la_gemm_tc_set_persistent(int on) { g_tc_persist = on; return LA_OK; }   // tuning hook
extern "C" la_status la_gemm_tc_set_chunk(int kb) { g_tc_chunk = kb; return LA_OK; }       // tuning hook
extern "C" la_status la_gemm_tc_set_a_direct(int on) { g_tc_adirect = on; return LA_OK; }   // tuning hook (experiment): A read in place as its own hi plane
extern "C" la_status la_gemm_tc_set_cross(int mode) { g_tc_cross = mode; return LA_OK; }   // tuning hook: -1 by K, 0 tf32 cross terms, 1 bf16
extern "C" la_status la_gemm_tc_set_tile_n(int bn) { g_tc_bn = (bn == 224 || bn == 192 || bn == 256) ? bn : 0; return LA_OK; }   // tuning hook
extern "C" la_status la_gemm_set_stream_tail(int mode) { g_tc_stream = mode; return LA_OK; }   // tuning hook: -1 auto, 0 off, 1 on
extern "C" la_status la_gemm_set_raster(int group) { g_raster_group = group; return LA_OK; }   // tuning hook

// Pieces shared by la_gemm_3xtf32 and the host-streaming GEMM (gemm_host.cu).
la_status la_tc_split_a(float* a_hi, float* a_lo, const float* A, size_t M, size_t K, unsigned* flag, cudaStream_t st) {
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    if (la_tc_cross_bf16(K)) split_tf32_kernel<true><<<(unsigned)info->sm_count * 8, 256, 0, st>>>(a_hi, a_lo, A, M * K / 4, flag, la_tc_a_direct(K) ? 1 : 0);
    else                     split_tf32_kernel<false><<<(unsigned)info->sm_count * 8, 256, 0, st>>>(a_hi, a_lo, A, M * K / 4, flag, 0);
    LA_LAUNCH_CHECK();
    return LA_OK;
}
la_status la_tc_split_bt(float* bt_hi, float* bt_lo, const float* B, size_t K, size_t N, unsigned* flag, cudaStream_t st) {
    const size_t tk = (K + 63) / 64, tn = (N + 63) / 64;
    if (tk * tn > 0x7fffffffull) return la_fail(LA_ERR_UNSUPPORTED, "3xTF32 pre-pass: B is too large");
    if (la_tc_cross_bf16(K)) split_transpose_kernel<true><<<(unsigned)(tk * tn), 256, 0, st>>>(bt_hi, bt_lo, B, K, N, (unsigned)tn, flag);
    else                     split_transpose_kernel<false><<<(unsigned)(tk * tn), 256, 0, st>>>(bt_hi, bt_lo, B, K, N, (unsigned)tn, flag);
    LA_LAUNCH_CHECK();
    return LA_OK;
}
// A and B in one launch.
la_status la_tc_split_both(float* a_hi, float* a_lo, const float* A, size_t M, float* bt_hi, float* bt_lo, const float* B, size_t K, size_t N, unsigned* flag, cudaStream_t st) {
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    const size_t tk = (K + 63) / 64, tn = (N + 63) / 64;
    const size_t a_n4 = M * K / 4;
    size_t a_blocks = (a_n4 + 255) / 256;
    if (a_blocks > (size_t)info->sm_count * 8) a_blocks = (size_t)info->sm_count * 8;
    if (tk * tn + a_blocks > 0x7fffffffull) return la_fail(LA_ERR_UNSUPPORTED, "3xTF32 pre-pass: B is too large");
    const unsigned grid = (unsigned)(tk * tn + a_blocks);
    if (la_tc_cross_bf16(K)) split_both_kernel<true><<<grid, 256, 0, st>>>(a_hi, a_lo, A, a_n4, bt_hi, bt_lo, B, K, N, (unsigned)tn, (unsigned)(tk * tn), flag, la_tc_a_direct(K) ? 1 : 0);
    else                     split_both_kernel<false><<<grid, 256, 0, st>>>(a_hi, a_lo, A, a_n4, bt_hi, bt_lo, B, K, N, (unsigned)tn, (unsigned)(tk * tn), flag, 0);
    LA_LAUNCH_CHECK();
    return LA_OK;
}

// TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's memory hung
// the GPU.  The one kernel with such loads (variant 5, B planes) is therefore only used when both planes are ordinary
// device allocations of the current device; anything else takes the same kernel without multicast (variant 3).
static bool tc_planes_are_local(const void* p0, const void* p1) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) return false;
    for (const void* p : {p0, p1}) {
        cudaPointerAttributes attr;
        if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) { (void)cudaGetLastError(); return false; }
        if (attr.type != cudaMemoryTypeDevice || attr.device != dev) return false;
    }
    return true;
}

la_status la_tc_run(float* C, const float* a_hi, const float* a_lo, const float* bt_hi, const float* bt_lo, size_t M, size_t K, size_t N, cudaStream_t st) {
    if (g_tc_variant == 5 && !tc_planes_are_local(bt_hi, bt_lo)) return tc_launch_chunked<1>(C, a_hi, a_lo, bt_hi, bt_lo, M, K, N, st);
    switch (g_tc_variant) {
    default:
    case 0: return tc_launch_2cta(C, a_hi, a_lo, bt_hi, bt_lo, M, K, N, st);         // CTA pairs, tcgen05 cta_group::2 (default)
    case 5: return tc_launch_chunked<2>(C, a_hi, a_lo, bt_hi, bt_lo, M, K, N, st);   // one CTA per tile, B multicast over CTA pairs
    case 3: return tc_launch_chunked<1>(C, a_hi, a_lo, bt_hi, bt_lo, M, K, N, st);   // same, no clusters
    case 1: return tc_launch<256, 2>(C, a_hi, a_lo, bt_hi, bt_lo, M, K, N, st);      // single chain (tuning reference; less accurate)
    case 2: return tc_launch<128, 3>(C, a_hi, a_lo, bt_hi, bt_lo, M, K, N, st);
    }
}
bool la_tc_shape_ok(const void* A, const void* B, const void* C, size_t M, size_t K, size_t N) {
    return !(M < 128 || N < 128 || K < 32 || K % 4 != 0 || N % 4 != 0 || M % 4 != 0 || !la_aligned16(A) || !la_aligned16(B) || !la_aligned16(C) ||
             M >= (1u << 31) || N >= (1u << 31) || K >= (1u << 31));
}

// The f16-split form of the engine: used from K = 1024 (below that its longer pre-pass -- row / column maxima first -- eats what
// its faster product kernel gains: 4096 x 512 x 4096 cold 0.129 ms against the tf32 form's 0.125), K a multiple of 8 (16-byte row
// pitch of the 16-bit planes), default pair kernel.  Operands outside its range (see the pre-pass) raise the flag and the
// product is redone by the tf32 form; operands with one-product outputs raise the other flag and the caller's exact engine runs.
static bool f16_wanted(size_t K) {
    if (g_tc_f16 == 0 || g_tc_variant != 0 || K % 8 != 0 || K < LA_TC_AUTO_MIN_K || tc_a_direct(K)) return false;      // (below 256 a product is a chain or two: the SIMT engine's)
    return g_tc_f16 > 0 || K >= 1024;
}
static la_status gemm_f16split(float* C, const float* A, const float* B, size_t M, size_t K, size_t N, cudaStream_t st, bool* handled, bool* exact, const la_gemm_prepared* prep) {
    *handled = false; *exact = false;
    const bool b_ready = prep && prep->f16_planes;
    if (prep && !b_ready) return LA_OK;                      // this B did not pass the f16 range test when it was prepared
    if (b_ready && prep->f16_b_sparse) { *exact = true; return LA_OK; }
    if (M > 65535ull * 1024 || N > 65535ull * 1024) return LA_OK;      // (the fix-up kernels' grid: the tf32 form takes such shapes)
    const size_t a_elems = M * K, b_elems = b_ready ? 0 : K * N;
    void* scratch; LA_TRY(la_scratch(st, (a_elems + b_elems) * 6 + f16_side_bytes(M, false) + (b_ready ? 0 : f16_side_bytes(N, true)), &scratch));   // planes | A side | B side
    volatile unsigned* flag_host; unsigned* flag;
    LA_TRY(la_host_flag(st, &flag_host, &flag));
    char* p = (char*)scratch + LA_SCRATCH_ZEROED;
    const F16Planes a = f16_planes(p, a_elems); p += a_elems * 6;
    F16Planes b = f16_planes(p, b_elems); p += b_elems * 6;
    const F16Side sa = f16_side(p, M); p += f16_side_bytes(M, false);
    F16Side sb = f16_side(p, N);
    if (b_ready) { b = f16_planes(prep->f16_planes, K * N); sb = f16_side((char*)prep->f16_planes + K * N * 6, N); }
    flag_host[0] = 0u; flag_host[1] = 0u; flag_host[2] = 0u; flag_host[3] = 0u;     // [0] out of range; [1] exact engine wanted; [2] / [3]: A's / B's list of left-out elements is not empty
    LA_TRY(f16_prepass(a, sa, A, M, b, sb, b_ready ? nullptr : B, K, N, flag, flag + 2, st));
    cudaEvent_t split_done; LA_TRY(la_stream_event(st, &split_done));
    LA_CUDA(cudaEventRecord(split_done, st));
    LA_TRY(f16_launch(C, a, b, sa.e, sb.e, M, K, N, st));
    LA_CUDA(cudaEventSynchronize(split_done));
    if (flag_host[1]) { *exact = true; return LA_OK; }      // (the SIMT engine next in the stream overwrites whatever this launch left in C)
    if (flag_host[0]) return LA_OK;                          // the tf32 form next in the stream overwrites it
    const bool a_any = flag_host[2] != 0u, b_any = b_ready ? prep->f16_b_tiny : flag_host[3] != 0u;
    if (a_any || b_any) LA_TRY(f16_tiny_fix(C, A, B, sa.e, M, K, N, sa, a_any, sb.items, sb.count, b_any, st));
    g_tc_last_form = 1; g_tc_last_launches = (b_ready ? 2 : 3) + (a_any ? 1 : 0) + (b_any ? 1 : 0);      // pre-pass (1 or 2), product, fix-ups
    *handled = true;
    return LA_OK;
}

// The f16-split form in pieces, for the streamed host-buffer product (gemm_host.cu): B's block once, then per row block of A its
// pre-pass, the product and the fix-ups -- the latter launched unconditionally (they leave at once when the lists are empty), so
// that the pipeline never waits for the host to read a word.  `flag`: [0] out of range, [1] exact engine wanted; read at the end.
bool la_f16_wanted(size_t K) { return f16_wanted(K); }
size_t la_f16_b_bytes(size_t K, size_t N) { return K * N * 6 + f16_side_bytes(N, true); }
size_t la_f16_a_bytes(size_t rows, size_t K) { return rows * K * 6 + f16_side_bytes(rows, false); }
la_status la_f16_split_b(void* b_block, const float* B, size_t K, size_t N, unsigned* flag, cudaStream_t st) {
    return f16_prepass(F16Planes{}, F16Side{}, nullptr, 0, f16_planes(b_block, K * N), f16_side((char*)b_block + K * N * 6, N), B, K, N, flag, flag + 2, st);
}
la_status la_f16_block(float* C, const float* A, const float* B, size_t rows, size_t K, size_t N, void* a_block, void* b_block, unsigned* flag, cudaStream_t st) {
    const F16Planes a = f16_planes(a_block, rows * K), b = f16_planes(b_block, K * N);
    const F16Side sa = f16_side((char*)a_block + rows * K * 6, rows), sb = f16_side((char*)b_block + K * N * 6, N);
    LA_TRY(f16_prepass(a, sa, A, rows, F16Planes{}, F16Side{}, nullptr, K, N, flag, flag + 2, st));
    LA_TRY(f16_launch(C, a, b, sa.e, sb.e, rows, K, N, st));
    return f16_tiny_fix(C, A, B, sa.e, rows, K, N, sa, true, sb.items, sb.count, true, st);
}

// `prep`: planes of B^T made earlier by la_tc_prepare_b (la_gemm_prepare_b), or NULL to split B here.
la_status la_gemm_3xtf32(float* C, const float* A, const float* B, size_t M, size_t K, size_t N, cudaStream_t st, bool* handled, const la_gemm_prepared* prep) {
    *handled = false;
    if (!la_tc_shape_ok(A, B, C, M, K, N)) return LA_OK;      // the SIMT engine takes it
    if (f16_wanted(K)) {
        bool exact = false;
        LA_TRY(gemm_f16split(C, A, B, M, K, N, st, handled, &exact, prep));
        if (*handled || exact) return LA_OK;                  // exact: an operand with one-product outputs -- the SIMT engine's (the caller's next choice)
    }
    if (prep && (!prep->tc_hi || prep->tc_cross_bf16 != la_tc_cross_bf16(K))) prep = nullptr;     // no planes, or made under other tuning settings
    const size_t a_elems = M * K, b_elems = prep ? 0 : K * N;
    void* scratch; LA_TRY(la_scratch(st, (2 * a_elems + 2 * b_elems) * sizeof(float), &scratch));
    volatile unsigned* flag_host; unsigned* flag;               // zero-copy flag word of this stream
    LA_TRY(la_host_flag(st, &flag_host, &flag));
    float* a_hi = (float*)((char*)scratch + LA_SCRATCH_ZEROED);
    float* a_lo = a_hi + a_elems;
    const float* b_hi = prep ? prep->tc_hi : a_lo + a_elems;    // B^T (N x K), split
    const float* b_lo = prep ? prep->tc_lo : b_hi + b_elems;
    *flag_host = 0u;                                            // no kernel of this stream that writes it is in flight (each call waits below)
    if (prep) LA_TRY(la_tc_split_a(a_hi, a_lo, A, M, K, flag, st));
    else      LA_TRY(la_tc_split_both(a_hi, a_lo, A, M, const_cast<float*>(b_hi), const_cast<float*>(b_lo), B, K, N, flag, st));
    // Were all operands inside the range the split is faithful for?  The product is launched straight behind the pre-pass, so
    // the GPU never idles while the host waits for the pre-pass (only) and reads the word.  If it is set -- inf / NaN / extreme
    // exponents -- the caller runs the plain f32 engine next in the stream, which overwrites whatever this launch left in C.
    cudaEvent_t split_done; LA_TRY(la_stream_event(st, &split_done));
    LA_CUDA(cudaEventRecord(split_done, st));
    LA_TRY(la_tc_run(C, la_tc_a_direct(K) ? A : a_hi, a_lo, b_hi, b_lo, M, K, N, st));
    LA_CUDA(cudaEventSynchronize(split_done));
    if (*flag_host) return LA_OK;
    g_tc_last_form = 0; g_tc_last_launches = 2;                  // pre-pass, product
    *handled = true;
    return LA_OK;
}

// la_gemm_prepare_b, f32 part: the split planes of B^T, made once.  No planes when B holds values the split cannot represent
// or the shape is outside the engine's: products with this handle then take the other engines.
la_status la_tc_prepare_b(la_gemm_prepared* h, cudaStream_t st) {
    const size_t K = h->K, N = h->N;
    if (N < 128 || K < 32 || K % 4 != 0 || N % 4 != 0 || !la_aligned16(h->B) || N >= (1u << 31) || K >= (1u << 31)) return LA_OK;
    volatile unsigned* flag_host; unsigned* flag;
    LA_TRY(la_host_flag(st, &flag_host, &flag));
    // The f16-split form's planes where it applies and B is inside its range: 6 bytes per element (+ exponents, list).  A product
    // whose A then turns out to be outside that range is done by the tf32 form from B itself, exactly as without a handle.
    if (f16_wanted(K)) {
        char* f16 = nullptr;
        LA_CUDA(cudaMalloc(&f16, K * N * 6 + f16_side_bytes(N, true)));
        flag_host[0] = 0u; flag_host[1] = 0u; flag_host[3] = 0u;
        la_status s = f16_prepass(F16Planes{}, F16Side{}, nullptr, 0, f16_planes(f16, K * N), f16_side(f16 + K * N * 6, N), (const float*)h->B, K, N, flag, flag + 2, st);
        if (s == LA_OK && cudaStreamSynchronize(st) != cudaSuccess) s = la_fail(LA_ERR_CUDA, "la_gemm_prepare_b: the f16 split of B failed");
        if (s != LA_OK) { cudaFree(f16); return s; }
        if (!flag_host[0]) { h->f16_planes = f16; h->f16_b_tiny = flag_host[3] != 0u; h->f16_b_sparse = flag_host[1] != 0u; return LA_OK; }
        cudaFree(f16); flag_host[0] = 0u;                       // out of the f16 form's range: the tf32 form's planes below
    }
    float* planes = nullptr;
    LA_CUDA(cudaMalloc(&planes, 2 * K * N * sizeof(float)));
    *flag_host = 0u;
    la_status s = la_tc_split_bt(planes, planes + K * N, (const float*)h->B, K, N, flag, st);
    if (s == LA_OK && cudaStreamSynchronize(st) != cudaSuccess) s = la_fail(LA_ERR_CUDA, "la_gemm_prepare_b: the split of B failed");
    if (s != LA_OK || *flag_host) { cudaFree(planes); return s; }        // out-of-range B: no planes, not an error
    h->tc_hi = planes; h->tc_lo = planes + K * N; h->tc_cross_bf16 = la_tc_cross_bf16(K);
    return LA_OK;
}
