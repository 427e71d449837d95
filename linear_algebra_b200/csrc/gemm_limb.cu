// csrc/gemm_limb.cu -- exact integer Matrix x Matrix on the INT8 tensor cores.
//
// Same contract as the integer path of gemm.cu ({T}_mul_mat_mat, src/kernels.cl:161-169: results
// mod 2^n, any summation order gives the same bits), other arithmetic.  An n-bit element
// (n = 8 L, L = 1, 2, 4 for 8-, 16-, 32-bit types, signed or unsigned: two's complement) is the sum
// of L unsigned byte "limbs", x = sum_i x_i 2^(8 i).  Modulo 2^n
//     x * y  =  sum over i + j < L of  2^(8 (i + j)) x_i y_j ,
// so C = A x B mod 2^n needs L (L + 1) / 2 byte-matrix products A_i x B_j, accumulated per shift
// class s = i + j.  Each is a tcgen05.mma kind::i8 (u8 x u8 -> s32).  The s32 accumulators may wrap:
// only bits below 2^n of  P_0 + 2^8 P_1 + 2^16 P_2 + 2^24 P_3  survive, and wrapping keeps exactly
// those.  10 MMAs per 32-bit product still run ~8x faster than the half-rate IMAD pipe.
//
//   pre-pass   limb_split_kernel            A (M x K)  -> A_0 .. A_{L-1}   (M x K bytes each)
//              limb_split_transpose_kernel  B (K x N)  -> Bt_0 .. Bt_{L-1} (N x K bytes each), K-major
//   kernel     CTA pairs (cta_group::2) own a 256 x 128 tile; per 128-byte k-block each CTA TMA-loads
//              its 128 rows of every A limb and 64 of the 128 rows of every Bt limb (128-byte swizzle);
//              the leader's MMA warp issues the L(L+1)/2 MMAs per 32-wide k-step into L accumulators of
//              128 columns in tensor memory; 8 epilogue warps per CTA combine them with shifts and store.
// Shapes: K % 16 == 0 (TMA row pitch), N * sizeof(T) % 16 == 0, 16-byte aligned bases; anything else
// (and 64-bit types) stays on the SIMT / generic engines.  Roofline: tensor pipe (int8).
#include "tc_common.cuh"

namespace {

constexpr int LB_BK = 128;            // bytes of K per k-block = one 128-byte swizzle row = 128 byte-limbs
constexpr int LB_UMMA_K = 32;         // 8-bit operands: 32 elements (32 bytes) of K per instruction
constexpr int LB_BN = 128;            // tile columns (accumulator columns per shift class)
constexpr int LB_THREADS = 384;
constexpr int LB_A_TILE = 128 * LB_BK;        // 16 KB: 128 rows of one A limb
constexpr int LB_B_TILE = 64 * LB_BK;         // 8 KB:  this CTA's 64 of the 128 Bt rows of one limb

template <int L> struct LimbShape {
    static constexpr int STAGE_BYTES = L * (LB_A_TILE + LB_B_TILE);
    static constexpr int STAGES = (L == 4) ? 2 : (L == 2 ? 4 : 6);
    static constexpr size_t SMEM_BYTES = (size_t)STAGES * STAGE_BYTES + 1024;
    static constexpr int TMEM_COLS = L * LB_BN;            // 128, 256, 512: powers of two
};

// kind::i8: unsigned 8-bit A and B, s32 accumulator, both operands K-major, no saturation.
__host__ __device__ constexpr uint32_t umma_idesc_u8(int M, int N) {
    return (2u << 4) | (0u << 7) | (0u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_i8_2sm(uint32_t tmem_d, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}

template <typename OutT, int L>
__global__ void __launch_bounds__(LB_THREADS, 1)
gemm_limb_2cta_kernel(OutT* __restrict__ C, const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                      int M, int N, int K, int pair_tiles_m, int tiles_n, int raster_group) {
    using S = LimbShape<L>;
    constexpr int STAGES = S::STAGES;
    extern __shared__ uint8_t smem_raw[];
    __shared__ alignas(8) uint64_t full_bar[STAGES];     // leader: both CTAs' slabs have landed
    __shared__ alignas(8) uint64_t empty_bar[STAGES];    // per CTA: the pair's MMAs are done with this slot
    __shared__ alignas(8) uint64_t acc_full_bar;         // per CTA: all accumulators complete
    __shared__ uint32_t tmem_base_slot;

    const uint32_t ring = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* ring_ptr = smem_raw + (ring - smem_u32(smem_raw));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t crank = cluster_ctarank();
    int tile_m, tile_n;
    la_raster_tile((int)(blockIdx.x >> 1), pair_tiles_m, tiles_n, raster_group, tile_m, tile_n);
    const int m0 = tile_m * 256 + (int)crank * 128;
    const int n0 = tile_n * LB_BN;
    const int num_kb = (K + LB_BK - 1) / LB_BK;

    if (threadIdx.x == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(&acc_full_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 9) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_slot)), "n"(S::TMEM_COLS) : "memory");
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
            // ---------------- TMA producer (both CTAs).  The limb planes are stacked along the row
            // coordinate of one tensor map: limb i of A starts at row i * M, of Bt at row i * N.
            if (lane == 0) {
                for (int kb = 0; kb < num_kb; ++kb) {
                    const int s = kb % STAGES;
                    const uint32_t ph = (uint32_t)(kb / STAGES) & 1u;
                    mbar_wait(&empty_bar[s], ph ^ 1u);
                    uint8_t* st = ring_ptr + (size_t)s * S::STAGE_BYTES;
                    if (crank == 0) mbar_expect_tx(&full_bar[s], 2u * S::STAGE_BYTES);
#pragma unroll
                    for (int i = 0; i < L; ++i) {
                        tma_load_2d_2sm(st + i * LB_A_TILE, &map_a, &full_bar[s], kb * LB_BK, i * M + m0);
                        tma_load_2d_2sm(st + L * LB_A_TILE + i * LB_B_TILE, &map_b, &full_bar[s], kb * LB_BK, i * N + n0 + (int)crank * 64);
                    }
                }
            }
        } else if (warp == 9 && crank == 0) {
            // ---------------- MMA issuer (leader CTA only)
            constexpr uint32_t idesc = umma_idesc_u8(256, LB_BN);
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % STAGES;
                const uint32_t ph = (uint32_t)(kb / STAGES) & 1u;
                mbar_wait(&full_bar[s], ph);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                if (elect_one()) {
                    const uint32_t st = ring + (uint32_t)s * S::STAGE_BYTES;
#pragma unroll
                    for (int kk = 0; kk < LB_BK / LB_UMMA_K; ++kk) {
                        const uint64_t adv = (uint64_t)((kk * LB_UMMA_K) >> 4);      // 32 bytes of K per step
#pragma unroll
                        for (int i = 0; i < L; ++i)
#pragma unroll
                            for (int j = 0; j + i < L; ++j) {
                                // shift class i + j accumulates in columns (i + j) * 128 ..; the first product of a
                                // class at the very first k-step overwrites (j == 0 or i == 0 comes first per class)
                                const uint32_t first = (kb == 0 && kk == 0 && i == 0) ? 0u : 1u;
                                umma_i8_2sm(tmem_base + (uint32_t)((i + j) * LB_BN),
                                            umma_desc_k_sw128(st + i * LB_A_TILE) + adv,
                                            umma_desc_k_sw128(st + L * LB_A_TILE + j * LB_B_TILE) + adv, idesc, first);
                            }
                    }
                    umma_commit_2sm(&empty_bar[s]);
                    if (kb == num_kb - 1) umma_commit_2sm(&acc_full_bar);
                }
                __syncwarp();
            }
        }
    } else {
        // ---------------- epilogue warps 0..7: lane quadrant = warp % 4, column half = warp / 4
        asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
        const int quad = warp & 3, half = warp >> 2;
        mbar_wait(&acc_full_bar, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int row = m0 + quad * 32 + lane;
        constexpr int PER16 = 16 / (int)sizeof(OutT);       // elements per 16-byte store
#pragma unroll 1
        for (int c0 = half * 64; c0 < half * 64 + 64; c0 += 32) {
            uint32_t p[L][32];
#pragma unroll
            for (int s = 0; s < L; ++s) tmem_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(s * LB_BN + c0), p[s]);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            OutT out[32];
#pragma unroll
            for (int e = 0; e < 32; ++e) {
                uint32_t v = p[0][e];
#pragma unroll
                for (int s = 1; s < L; ++s) v += p[s][e] << (8 * s);
                out[e] = (OutT)v;
            }
            if (row < M) {
                OutT* crow = C + (size_t)row * N + n0 + c0;
#pragma unroll
                for (int q = 0; q < 32 / PER16; ++q) {
                    const int n = n0 + c0 + q * PER16;
                    if (n < N) {        // N * sizeof(T) % 16 == 0: a 16-byte chunk is wholly inside or outside
                        uint4 v;
                        memcpy(&v, &out[q * PER16], 16);
                        __stcs(reinterpret_cast<uint4*>(crow + q * PER16), v);
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
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(S::TMEM_COLS) : "memory");
    }
}

// ------------------------------------------------------------------ pre-pass
// A (rows x K elements of U) -> L byte planes, plane i at planes + i * rows * K.  Each thread turns 16
// consecutive elements into one 16-byte store per plane.  K % 16 == 0.
template <typename U, int L>
__global__ void __launch_bounds__(256)
limb_split_kernel(uint8_t* __restrict__ planes, const U* __restrict__ x, size_t n16, size_t plane_bytes) {
    const size_t stride = (size_t)gridDim.x * 256;
    for (size_t g = (size_t)blockIdx.x * 256 + threadIdx.x; g < n16; g += stride) {
        U v[16];
        constexpr int PACKS = (int)sizeof(U);          // 16 elements = sizeof(U) packets of 16 bytes
#pragma unroll
        for (int p = 0; p < PACKS; ++p) *reinterpret_cast<uint4*>(&v[p * (16 / sizeof(U))]) = __ldcs(reinterpret_cast<const uint4*>(x + g * 16) + p);
#pragma unroll
        for (int i = 0; i < L; ++i) {
            alignas(16) uint8_t b[16];
#pragma unroll
            for (int e = 0; e < 16; ++e) b[e] = (uint8_t)(v[e] >> (8 * i));
            __stcs(reinterpret_cast<uint4*>(planes + (size_t)i * plane_bytes + g * 16), *reinterpret_cast<const uint4*>(b));
        }
    }
}

// B (K x N elements of U) -> L byte planes of B^T (N x K bytes each).  A block turns a tile of 64 k-rows x 256 bytes of B
// (64 / 128 / 256 columns for 4- / 2- / 1-byte elements): four 16-byte streaming loads per thread, all issued before the first
// shared-memory store; the tile sits in shared memory with the 16-byte vector v of row r at position v ^ 2 (r / 16), which makes
// the column gathers below bank-conflict free; then each work item gathers 16 consecutive k of one column and stores 16 bytes
// per plane (four lanes cover 64 contiguous bytes of a plane row).  K % 16 == 0, N * sizeof(U) % 16 == 0, 16-byte aligned B.
// (Round 2: the first version read B with one 1/2/4-byte load per thread and loop trip -- 0.20 / 0.14 / 0.12 ms for 8192^2
//  u32 / u16 / u8 against 0.08 / 0.04 / 0.02 ms for the A split of the same bytes.)
template <typename U, int L>
__global__ void __launch_bounds__(256)
limb_split_transpose_kernel(uint8_t* __restrict__ planes, const U* __restrict__ B, size_t K, size_t N, unsigned tiles_n, size_t plane_bytes) {
    constexpr int ES = (int)sizeof(U), TN = 256 / ES, EPV = 16 / ES;          // tile columns; elements per 16-byte vector
    __shared__ __align__(16) uint8_t tile[64 * 256];
    const size_t k0 = (size_t)(blockIdx.x / tiles_n) * 64, n0 = (size_t)(blockIdx.x % tiles_n) * TN;
    uint4 in[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int idx = threadIdx.x + j * 256, r = idx / 16, v = idx % 16;
        in[j] = (k0 + r < K && n0 + (size_t)v * EPV < N) ? __ldcs(reinterpret_cast<const uint4*>(B + (k0 + r) * N + n0 + (size_t)v * EPV)) : make_uint4(0, 0, 0, 0);
    }
    __syncwarp();                            // all four loads are issued before the first shared-memory store
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int idx = threadIdx.x + j * 256, r = idx / 16, v = idx % 16;
        *reinterpret_cast<uint4*>(tile + r * 256 + ((v ^ (2 * (r >> 4))) << 4)) = in[j];
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < TN / 64; ++j) {
        const int idx = threadIdx.x + j * 256, cc = idx / 4, q = idx % 4;    // B^T row n0 + cc, k0 + 16 q .. + 15
        if (n0 + cc < N && k0 + (size_t)q * 16 < K) {
            const int off = cc * ES, vec = off >> 4, within = off & 15;
            U x[16];
#pragma unroll
            for (int e = 0; e < 16; ++e) x[e] = *reinterpret_cast<const U*>(tile + (q * 16 + e) * 256 + ((vec ^ (2 * q)) << 4) + within);
#pragma unroll
            for (int i = 0; i < L; ++i) {
                alignas(16) uint8_t b[16];
#pragma unroll
                for (int e = 0; e < 16; ++e) b[e] = (uint8_t)(x[e] >> (8 * i));
                __stcs(reinterpret_cast<uint4*>(planes + (size_t)i * plane_bytes + (n0 + cc) * K + k0 + (size_t)q * 16), *reinterpret_cast<const uint4*>(b));
            }
        }
    }
}

// All L planes of an operand as ONE 2-D byte tensor: (L * rows) x K, box = box_rows x 128 bytes, 128-byte swizzle.
la_status limb_make_map(CUtensorMap* map, const uint8_t* base, size_t total_rows, size_t K, int box_rows) {
    EncodeTiledFn enc; LA_TRY(tc_get_encode(&enc));
    cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)total_rows};
    cuuint64_t strides[1] = {(cuuint64_t)K};
    cuuint32_t box[2] = {(cuuint32_t)LB_BK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<uint8_t*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return la_fail(LA_ERR_CUDA, "cuTensorMapEncodeTiled (limb engine) failed with CUresult %d", (int)r);
    return LA_OK;
}

// B -> the L byte planes of B^T (the part of the pre-pass that la_gemm_prepare_b can do once)
template <typename U, int L>
la_status limb_split_b(uint8_t* b_planes, const U* B, size_t K, size_t N, cudaStream_t st) {
    constexpr size_t TN = 256 / sizeof(U);
    const size_t tk = (K + 63) / 64, tn = (N + TN - 1) / TN;
    if (tk * tn > 0x7fffffffull) return la_fail(LA_ERR_UNSUPPORTED, "limb engine pre-pass: B is too large");
    limb_split_transpose_kernel<U, L><<<(unsigned)(tk * tn), 256, 0, st>>>(b_planes, B, K, N, (unsigned)tn, N * K);
    LA_LAUNCH_CHECK();
    return LA_OK;
}

template <typename U, int L>
la_status limb_gemm(U* C, const U* A, const U* B, size_t M, size_t K, size_t N, cudaStream_t st, int sm_count, const uint8_t* prep_planes) {
    using S = LimbShape<L>;
    const size_t a_plane = M * K, b_plane = N * K;
    void* scratch; LA_TRY(la_scratch(st, (size_t)L * (a_plane + (prep_planes ? 0 : b_plane)), &scratch));
    uint8_t* a_planes = (uint8_t*)scratch + LA_SCRATCH_ZEROED;
    const uint8_t* b_planes = prep_planes ? prep_planes : a_planes + (size_t)L * a_plane;          // a_plane % 16 == 0 keeps this 16-byte aligned
    limb_split_kernel<U, L><<<(unsigned)sm_count * 8, 256, 0, st>>>(a_planes, A, a_plane / 16, a_plane);
    LA_LAUNCH_CHECK();
    if (!prep_planes) LA_TRY((limb_split_b<U, L>(const_cast<uint8_t*>(b_planes), B, K, N, st)));
    // Within a plane, rows past M (or N) would run into the next plane instead of reading zeros; they only
    // feed output rows / columns the epilogue never stores, so that is harmless.
    CUtensorMap map_a, map_b;
    LA_TRY(limb_make_map(&map_a, a_planes, (size_t)L * M, K, 128));
    LA_TRY(limb_make_map(&map_b, b_planes, (size_t)L * N, K, 64));
    auto kern = gemm_limb_2cta_kernel<U, L>;
    LA_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::SMEM_BYTES));
    const size_t ptm = (M + 255) / 256, tnn = (N + LB_BN - 1) / LB_BN;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(2 * ptm * tnn)); cfg.blockDim = dim3(LB_THREADS); cfg.dynamicSmemBytes = S::SMEM_BYTES; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    LA_CUDA(cudaLaunchKernelEx(&cfg, kern, C, map_a, map_b, (int)M, (int)N, (int)K, (int)ptm, (int)tnn, la_gemm_raster_group(ptm)));
    return LA_OK;
}

// the conditions on B alone (la_gemm_prepare_b) and on the whole product
bool limb_b_ok(int dtype, const void* B, size_t K, size_t N) {
    const size_t es = la_dtype_size(dtype);
    return !(es > 4 || dtype == LA_F32 || N < 64 || K < 128 || K % 16 != 0 || (N * es) % 16 != 0 || !la_aligned16(B) || 4 * N >= (1u << 31) || K >= (1u << 31));
}

}  // namespace

// Integer GEMM on the INT8 tensor cores when the shape allows; *handled says whether it ran.  `prep`: byte planes of B^T made
// earlier by la_limb_prepare_b, or NULL.
la_status la_gemm_limb(int dtype, void* C, const void* A, const void* B, size_t M, size_t K, size_t N, cudaStream_t st, bool* handled, const la_gemm_prepared* prep) {
    *handled = false;
    const size_t es = la_dtype_size(dtype);
    if (!limb_b_ok(dtype, B, K, N) || M < 128 || !la_aligned16(A) || !la_aligned16(C) || 4 * M >= (1u << 31)) return LA_OK;
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    const uint8_t* pp = prep ? prep->limb_planes : nullptr;
    la_status s;
    switch (es) {
    case 1: s = limb_gemm<uint8_t, 1>((uint8_t*)C, (const uint8_t*)A, (const uint8_t*)B, M, K, N, st, info->sm_count, pp); break;
    case 2: s = limb_gemm<uint16_t, 2>((uint16_t*)C, (const uint16_t*)A, (const uint16_t*)B, M, K, N, st, info->sm_count, pp); break;
    default: s = limb_gemm<uint32_t, 4>((uint32_t*)C, (const uint32_t*)A, (const uint32_t*)B, M, K, N, st, info->sm_count, pp); break;
    }
    if (s == LA_OK) *handled = true;
    return s;
}

// la_gemm_prepare_b, integer part: the byte planes of B^T, made once (no planes when the shape is outside the engine's).
la_status la_limb_prepare_b(la_gemm_prepared* h, cudaStream_t st) {
    if (!limb_b_ok(h->dtype, h->B, h->K, h->N)) return LA_OK;
    const size_t es = la_dtype_size(h->dtype);
    uint8_t* planes = nullptr;
    LA_CUDA(cudaMalloc(&planes, es * h->K * h->N));
    la_status s;
    switch (es) {
    case 1: s = limb_split_b<uint8_t, 1>(planes, (const uint8_t*)h->B, h->K, h->N, st); break;
    case 2: s = limb_split_b<uint16_t, 2>(planes, (const uint16_t*)h->B, h->K, h->N, st); break;
    default: s = limb_split_b<uint32_t, 4>(planes, (const uint32_t*)h->B, h->K, h->N, st); break;
    }
    if (s == LA_OK && cudaStreamSynchronize(st) != cudaSuccess) s = la_fail(LA_ERR_CUDA, "la_gemm_prepare_b: the byte-plane split of B failed");
    if (s != LA_OK) { cudaFree(planes); return s; }
    h->limb_planes = planes;
    return LA_OK;
}
