// csrc/gemm_tc.cu -- split-f32 tensor-core engine for f32 Matrix x Matrix (LA_GEMM_3XTF32; what LA_GEMM_AUTO picks from 512^3
// multiply-adds up when K >= 256).
//
// Same contract as the SIMT engine in gemm.cu ({T}_mul_mat_mat, src/kernels.cl:161-169), other arithmetic: every f32
// operand is split x = hi + lo with hi = tf32(x) (round to nearest), and each product is formed as hi*hi + hi*lo + lo*hi
// on the 5th-generation tensor cores with FP32 accumulation in tensor memory:
//   K < 1024    three tcgen05.mma kind::tf32 per k-step ("3xTF32"), lo = tf32(x - hi);
//   K >= 1024   hi*hi as kind::tf32, the two cross terms (2^-11 of the weight) as kind::f16 bf16 MMAs at twice the rate
//               on the planes bf16(hi) and bf16(x - hi): 2 tensor time units per product instead of 3.
// The dropped lo*lo term is below 2^-22 relative.  Accumulation runs in chains of 32 .. 256 k inside the tensor cores
// (their FP32 accumulation truncates) and in FP32 registers across chains.  Results stay inside the f32 bar of DESIGN.md
// section 4 -- tests/test_gpu_parity.py gates every cross-term type / chain length / tile width on it, on random and on
// structured data; la_set_gemm_mode(LA_GEMM_SIMT) switches the engine off.
//
// Kernels in this file (la_gemm_tc_set_variant picks one; 0 is the default):
//   0  gemm_3xtf32_2cta_kernel<BN, BF16X>  CTA pairs (tcgen05 cta_group::2), 256 x BN tile per pair (BN = 256 / 224 / 192 by
//                                   wave count), 3-stage TMA ring, work items = (tile, chain range) so the ragged last
//                                   wave can be cut in two along K, banded tile order above 4096 rows
//   5/3 gemm_3xtf32_chunked_kernel  one CTA per 128x256 tile (with / without B multicast), 64-chains  (tuning references)
//   1/2 gemm_3xtf32_kernel          single accumulation chain (tuning reference; drifts)
// Common structure:
//   pre-pass   split_tf32_kernel       A -> A_hi + (A_lo | bf16(A_hi), bf16(A - A_hi))                 (scratch)
//              split_transpose_kernel  B -> the same planes of B^T (N x K), so both operands are K-major    (scratch)
//   TMA warp   cp.async.bulk.tensor.2d of the slabs of a 32-wide k-block (128-byte rows / SWIZZLE_128B for tf32 planes,
//              64-byte rows / SWIZZLE_64B for bf16 planes) into a ring guarded by full/empty mbarriers
//   MMA warp   owns tensor memory; one elected lane issues the MMAs of a k-block and tcgen05.commit's the stage back
//   epilogue   tcgen05.ld of the FP32 accumulator after every chain, FP32 adds in registers, 16-byte streaming stores to C
//              (atomicAdd of the two halves for a tile cut along K)
// Non-finite inputs: inf * (lo == 0) would be NaN, and tiny values lose their lo part; the pre-pass detects anything
// outside [2^-100, 2^100] (or NaN) and the call then runs the SIMT engine over the same C, so the tensor engine never
// changes the result class of a product.
// Roofline: tensor pipe, 3 (K < 1024) or 2 tf32-equivalent MMA units per useful product; in the second form the kernel is
// at 90 % of that floor with the L2 -> SM path (11.6 TB/s of operand loads at 8192^3) close to its own limit, see the
// note in the producer loop.
#include <cuda_bf16.h>
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
    auto kern = gemm_3xtf32_2cta_kernel<BN, BF16X>;
    LA_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC3_SMEM_BYTES));
    const size_t ptm = (M + 255) / 256, tn = (N + BN - 1) / BN;
    const long long tiles = (long long)(ptm * tn);

    TcPlan plan = {};
    plan.pair_tiles_m = (int)ptm; plan.tiles_n = (int)tn; plan.raster_group = la_gemm_raster_group(ptm);
    plan.num_kb = (int)((K + TC_BK - 1) / TC_BK);
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
    plan.chunk_kb = g_tc_chunk > 0 ? g_tc_chunk : K < 1024 ? 1 : K >= 8192 ? 8 : K >= 2048 ? 4 : TC_CHUNK_KB;
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
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(2 * items)); cfg.blockDim = dim3(TC3_THREADS); cfg.dynamicSmemBytes = TC3_SMEM_BYTES; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    const bool timed = g_tc_timing && g_tc_timed < 64;
    if (timed) {
        if (!g_tc_ev_made) { for (auto& e : g_tc_ev) { LA_CUDA(cudaEventCreate(&e[0])); LA_CUDA(cudaEventCreate(&e[1])); } g_tc_ev_made = true; }
        LA_CUDA(cudaEventRecord(g_tc_ev[g_tc_timed][0], st));
    }
    LA_CUDA(cudaLaunchKernelEx(&cfg, kern, C, ma_hi, ma_lo, mb_hi, mb_lo, ma_x, mb_x, (int)M, (int)N, plan));
    if (timed) { LA_CUDA(cudaEventRecord(g_tc_ev[g_tc_timed][1], st)); ++g_tc_timed; }
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
bool la_tc_cross_bf16(size_t K) {
    if (K % 8 != 0 || g_tc_cross == 0 || g_tc_variant != 0) return false;
    return g_tc_cross > 0 || K >= 1024;
}

// Tile width.  256 x 256 per CTA pair unless a narrower tile packs the waves better: the 8-GPU shard of the 8192^3 product,
// 1024 x 8192 x 8192, is 128 tiles = 1.73 waves -> 2 wave-times at width 256, but 148 tiles = exactly 2 waves of 0.875 at 224
// (measured 0.666 -> 0.645 ms).
la_status tc_launch_2cta(float* C, const float* a_hi, const float* a_lo, const float* b_hi, const float* b_lo, size_t M, size_t K, size_t N, cudaStream_t st) {
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    const long long pairs = info->sm_count / 2;                 // CTA pairs in flight (one per SM pair)
    const long long ptm = (long long)((M + 255) / 256);
    const bool can_split = g_tc_stream != 0 && K > 32;
    int bn = 256;
    if (g_tc_bn) bn = g_tc_bn;
    else {
        double best = tc_cost(ptm * (long long)((N + 255) / 256), pairs, 256, can_split);
        for (int cand : {224, 192}) {
            const double c = tc_cost(ptm * (long long)((N + cand - 1) / cand), pairs, cand, can_split);
            if (c < 0.90 * best) { best = c; bn = cand; }       // a 224-wide tile costs ~7 % more per flop (8192^3: 4.86 -> 5.23 ms at equal wave count): it must buy more than that
        }
    }
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


}  // namespace

bool la_tc_a_direct(size_t K) { return tc_a_direct(K); }      // also asked by gemm_host.cu

extern "C" la_status la_gemm_tc_set_variant(int v) { g_tc_variant = v; return LA_OK; }     // tuning hook
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

// `prep`: planes of B^T made earlier by la_tc_prepare_b (la_gemm_prepare_b), or NULL to split B here.
la_status la_gemm_3xtf32(float* C, const float* A, const float* B, size_t M, size_t K, size_t N, cudaStream_t st, bool* handled, const la_gemm_prepared* prep) {
    *handled = false;
    if (!la_tc_shape_ok(A, B, C, M, K, N)) return LA_OK;      // the SIMT engine takes it
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
    *handled = true;
    return LA_OK;
}

// la_gemm_prepare_b, f32 part: the split planes of B^T, made once.  No planes when B holds values the split cannot represent
// or the shape is outside the engine's: products with this handle then take the other engines.
la_status la_tc_prepare_b(la_gemm_prepared* h, cudaStream_t st) {
    const size_t K = h->K, N = h->N;
    if (N < 128 || K < 32 || K % 4 != 0 || N % 4 != 0 || !la_aligned16(h->B) || N >= (1u << 31) || K >= (1u << 31)) return LA_OK;
    float* planes = nullptr;
    LA_CUDA(cudaMalloc(&planes, 2 * K * N * sizeof(float)));
    volatile unsigned* flag_host; unsigned* flag;
    la_status s = la_host_flag(st, &flag_host, &flag);
    if (s == LA_OK) { *flag_host = 0u; s = la_tc_split_bt(planes, planes + K * N, (const float*)h->B, K, N, flag, st); }
    if (s == LA_OK && cudaStreamSynchronize(st) != cudaSuccess) s = la_fail(LA_ERR_CUDA, "la_gemm_prepare_b: the split of B failed");
    if (s != LA_OK || *flag_host) { cudaFree(planes); return s; }        // out-of-range B: no planes, not an error
    h->tc_hi = planes; h->tc_lo = planes + K * N; h->tc_cross_bf16 = la_tc_cross_bf16(K);
    return LA_OK;
}
