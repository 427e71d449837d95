// csrc/transpose.cu -- out[c*rows + r] = in[r*cols + c].
//
// Not in the reference (it only has {T}_mul_vec_transpose_mat, src/kernels.cl:150-152); named
// by the north star and bit-exact for every dtype by definition.  Shared-memory tiled: a block
// reads a TILE x TILE tile with 16-byte loads along the input rows, and writes it back with
// 16-byte stores along the output rows, so both global streams are coalesced 128-bit traffic.
// Tiles are padded by one element per row against bank conflicts.  Elements are moved as opaque
// 1/2/4/8-byte words, so signed/unsigned/float share kernels.
//
// Roofline: HBM.  Algorithmic bytes per element: 2*sizeof(T).
//
// Two more kernels for 4- and 8-byte elements with 16-byte-aligned rows move the tiles with TMA (transpose_tma_kernel: TMA
// loads and TMA stores, SASS UTMALDG / UTMASTG; transpose_tma_ldst_kernel: TMA loads, st.global stores).  Measured on the
// B200 they are SLOWER than the LDG/STG kernel (profiles/r2_transpose_tma.txt), so they are reachable through
// la_transpose_set_variant only; the default stays the kernel below.
#include "tc_common.cuh"
#include "../../include/linalg_b200_tuning.h"

namespace {

constexpr int TR_THREADS = 256;

int g_tr_variant = 0;      // la_transpose_set_variant (tuning hook), see transpose_launch
int g_tr_band = -1;        // la_transpose_set_band (tuning hook): tile rows per band of the TMA kernels' tile walk; -1 = default
int g_tr_hints = 0;        // la_transpose_set_hints (tuning hook): bit 0 evict-first L2 hint on the TMA loads, bit 1 on the TMA stores
int g_tr_mode = 0;         // la_transpose_set_mode (measurement only): 1 loads only, 2 stores only

// ------------------------------------------------------------------ TMA transpose
// Unit of work: a square sub-tile of T x T elements, T = 128 bytes / sizeof(E) (32 x 32 words of 4 bytes, 16 x 16 of 8):
// one TMA box whose rows are exactly one 128-byte shared-memory row, SWIZZLE_128B on the way in and on the way out.
// Every warp is a pipeline of its own -- no block-wide barrier anywhere:
//   lane 0   keeps IN_STAGES box loads in flight into the warp's private ring (one mbarrier per stage),
//   32 lanes turn a landed box: each lane moves an n x n block of elements (n = 16 bytes / sizeof(E)) with n 16-byte
//            shared-memory loads and n 16-byte stores; for 4- and 8-byte elements the turn inside the block is register
//            renaming, no instruction,
//   lane 0   stores the turned box with one TMA store (bulk async-group) and re-arms the input stage.
// Bank conflicts: with the 128-byte swizzle the 16-byte chunk q of row r sits at position q ^ (r & 7).  A quarter-warp
// (the unit a 16-byte shared-memory access is served in) takes lanes i = 0..7 with input chunk column cc = i and row
// block rb = i ^ d (d fixed per quarter-warp): reading row n*rb + j it touches positions i ^ (n*rb + j & 7), writing
// output row n*cc + j' it touches rb ^ (n*cc + j' & 7) -- both are bijections of i for n = 2, 4, 8, 16, so all eight lanes
// hit different 16-byte bank groups on both sides.
// The warps of a block form a WR x WC arrangement of neighbouring sub-tiles; blocks walk the tiles persistently.
// Edges: TMA zero-fills the part of a load box outside the matrix and clips a store box, so R and C need only make
// the row pitches multiples of 16 bytes.
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, const void* src, int x, int y) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(map), "r"(smem_u32(src)), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ void tma_store_2d_hint(const CUtensorMap* map, const void* src, int x, int y, uint64_t policy) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group.L2::cache_hint [%0, {%2, %3}], [%1], %4;"
                 ::"l"(map), "r"(smem_u32(src)), "r"(x), "r"(y), "l"(policy) : "memory");
}
__device__ __forceinline__ void tma_load_2d_hint(void* dst, const CUtensorMap* map, uint64_t* bar, int x, int y, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;"
        ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(x), "r"(y), "l"(policy) : "memory");
}
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint32_t word_of(const uint4& v, int j) { return j == 0 ? v.x : j == 1 ? v.y : j == 2 ? v.z : v.w; }   // j is a constant after unrolling
template <int N> __device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}

// the n x n turn of one lane's block: v[j] = 16 bytes of input row j; returns the 16 bytes of output row j
template <int ES> __device__ __forceinline__ uint4 turned_row(const uint4* v, int j) {
    if constexpr (ES == 4) return make_uint4(word_of(v[0], j), word_of(v[1], j), word_of(v[2], j), word_of(v[3], j));
    else { static_assert(ES == 8, "4- or 8-byte elements"); return j == 0 ? make_uint4(v[0].x, v[0].y, v[1].x, v[1].y) : make_uint4(v[0].z, v[0].w, v[1].z, v[1].w); }
}

// tile index -> (tile row, tile column): along the input rows, or in bands of `band` tile rows walked down the rows first
__device__ __forceinline__ void tr_tile(unsigned long long t, unsigned tiles_r, unsigned tiles_c, unsigned band, unsigned long long& tr, unsigned long long& tc) {
    if (band == 0) { tr = t / tiles_c; tc = t - tr * tiles_c; return; }
    const unsigned long long per_band = (unsigned long long)band * tiles_c, b = t / per_band, i = t - b * per_band;
    const unsigned long long first = b * band, rows = (tiles_r - first < band) ? tiles_r - first : band;
    tc = i / rows; tr = first + (i - tc * rows);
}

template <int ES, int WR, int WC, int IN_STAGES, int OUT_STAGES>
struct TrTma {
    static constexpr int WARPS = WR * WC;
    static constexpr int THREADS = WARPS * 32;
    static constexpr int N16 = 16 / ES;                 // elements per 16-byte chunk = side of a lane's block
    static constexpr int T = 8 * N16;                   // sub-tile side in elements (128 bytes)
    static constexpr int SUB_BYTES = T * 128;
    static constexpr size_t SMEM_BYTES = (size_t)WARPS * (IN_STAGES + OUT_STAGES) * SUB_BYTES + 1024;
};

// MODE 0 is the transpose.  1 (loads only) and 2 (stores only, of whatever the shared memory holds) exist to measure the two
// halves of the data path on their own; they are not reachable from la_transpose without the measurement hook.
template <int ES, int WR, int WC, int IN_STAGES, int OUT_STAGES, int MODE>
__global__ void __launch_bounds__(WR * WC * 32)
transpose_tma_kernel(const __grid_constant__ CUtensorMap map_in, const __grid_constant__ CUtensorMap map_out,
                     unsigned R, unsigned C, unsigned tiles_r, unsigned tiles_c, unsigned long long n_tiles, unsigned band, unsigned hints) {
    using S = TrTma<ES, WR, WC, IN_STAGES, OUT_STAGES>;
    constexpr int N16 = S::N16, T = S::T, SUB = S::SUB_BYTES;
    extern __shared__ uint8_t smem_raw[];
    __shared__ alignas(8) uint64_t full_bar[S::WARPS][IN_STAGES];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;                 // the swizzle is a function of the absolute address
    uint8_t* in_ring = smem_raw + (base - smem_u32(smem_raw)) + (size_t)warp * (IN_STAGES + OUT_STAGES) * SUB;
    uint8_t* out_ring = in_ring + (size_t)IN_STAGES * SUB;
    uint64_t* bars = full_bar[warp];
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < IN_STAGES; ++s) mbar_init(&bars[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncwarp();

    const unsigned wr = warp / WC, wc = warp % WC;
    const uint64_t policy = l2_evict_first_policy();
    // the block's k-th tile is blockIdx.x + k * gridDim.x; this warp's sub-tile of it starts at (r0, c0)
    auto origin = [&](unsigned long long t, unsigned& r0, unsigned& c0) -> bool {
        unsigned long long tr, tc;
        tr_tile(t, tiles_r, tiles_c, band, tr, tc);
        const unsigned long long r = (tr * WR + wr) * T, c = (tc * WC + wc) * T;
        r0 = (unsigned)r; c0 = (unsigned)c;
        return r < R && c < C;                                             // a sub-tile wholly outside the matrix is skipped
    };
    unsigned long long t_load = blockIdx.x, t_turn = blockIdx.x;
    unsigned n_loaded = 0, n_turned = 0;
    auto load_next = [&]() {                                                  // lane 0: the next sub-tile that exists, if any
        unsigned r0, c0;
        while (t_load < n_tiles) {
            const bool ok = origin(t_load, r0, c0);
            t_load += gridDim.x;
            if (ok) {
                const int s = (int)(n_loaded % IN_STAGES);
                mbar_expect_tx(&bars[s], (uint32_t)SUB);
                if (hints & 1u) tma_load_2d_hint(in_ring + (size_t)s * SUB, &map_in, &bars[s], (int)c0, (int)r0, policy);
                else            tma_load_2d(in_ring + (size_t)s * SUB, &map_in, &bars[s], (int)c0, (int)r0);
                ++n_loaded;
                return;
            }
        }
    };
    if (lane == 0 && MODE != 2) {
#pragma unroll 1
        for (int s = 0; s < IN_STAGES; ++s) load_next();
    }
    // lane -> block of the sub-tile, two rounds of 32 blocks: chunk column cc, row block rb (see the bank note above)
    const int li = lane & 7, ld = lane >> 3;
#pragma unroll 1
    for (; t_turn < n_tiles; t_turn += gridDim.x) {
        unsigned r0, c0;
        if (!origin(t_turn, r0, c0)) continue;
        const int s = (int)(n_turned % IN_STAGES), o = (int)(n_turned % OUT_STAGES);
        const uint32_t ph = (n_turned / IN_STAGES) & 1u;
        if (lane == 0 && MODE != 1) bulk_wait_read<OUT_STAGES - 1>();       // the store that last read out_ring[o] is done with it
        __syncwarp();
        if (MODE != 2) mbar_wait(&bars[s], ph);
        const uint8_t* src = in_ring + (size_t)s * SUB;
        uint8_t* dst = out_ring + (size_t)o * SUB;
        if (MODE == 0)
#pragma unroll
        for (int round = 0; round < 2; ++round) {
            const int cc = li, rb = li ^ (ld + 4 * round);
            uint4 v[N16];
#pragma unroll
            for (int j = 0; j < N16; ++j) {
                const int r = N16 * rb + j;
                v[j] = *reinterpret_cast<const uint4*>(src + r * 128 + ((cc ^ (r & 7)) << 4));
            }
#pragma unroll
            for (int j = 0; j < N16; ++j) {
                const int c = N16 * cc + j;
                *reinterpret_cast<uint4*>(dst + c * 128 + ((rb ^ (c & 7)) << 4)) = turned_row<ES>(v, j);
            }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");        // my writes, before the TMA store reads them
        __syncwarp();
        if (lane == 0) {
            if (MODE != 1) {
                if (hints & 2u) tma_store_2d_hint(&map_out, dst, (int)r0, (int)c0, policy);
                else            tma_store_2d(&map_out, dst, (int)r0, (int)c0);
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
            if (MODE != 2) load_next();                                      // every lane is done reading in_ring[s]
        }
        ++n_turned;
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// rows x cols matrix of ES-byte words, box = box_rows (default T) x T, 128-byte swizzle
la_status tr_make_map(CUtensorMap* map, const void* base, size_t rows, size_t cols, int es, int box_rows = 0) {
    EncodeTiledFn enc; LA_TRY(tc_get_encode(&enc));
    const cuuint32_t t = (cuuint32_t)(128 / es);
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)cols * es};
    cuuint32_t box[2] = {t, box_rows > 0 ? (cuuint32_t)box_rows : t};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, es == 4 ? CU_TENSOR_MAP_DATA_TYPE_UINT32 : CU_TENSOR_MAP_DATA_TYPE_UINT64, 2, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return la_fail(LA_ERR_CUDA, "cuTensorMapEncodeTiled (transpose, %zu x %zu) failed with CUresult %d", rows, cols, (int)r);
    return LA_OK;
}

template <int ES, int WR, int WC, int IN_STAGES, int OUT_STAGES>
la_status transpose_tma_launch(void* out, const void* in, size_t R, size_t C, cudaStream_t st, int ctas_per_sm) {
    using S = TrTma<ES, WR, WC, IN_STAGES, OUT_STAGES>;
    static_assert(S::SMEM_BYTES <= 227 * 1024, "shared memory per block");
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    CUtensorMap map_in, map_out;
    LA_TRY(tr_make_map(&map_in, in, R, C, ES));
    LA_TRY(tr_make_map(&map_out, out, C, R, ES));
    auto kern = g_tr_mode == 1 ? transpose_tma_kernel<ES, WR, WC, IN_STAGES, OUT_STAGES, 1> : g_tr_mode == 2 ? transpose_tma_kernel<ES, WR, WC, IN_STAGES, OUT_STAGES, 2>
                                                                                                              : transpose_tma_kernel<ES, WR, WC, IN_STAGES, OUT_STAGES, 0>;
    LA_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::SMEM_BYTES));
    const size_t tiles_r = (R + (size_t)WR * S::T - 1) / ((size_t)WR * S::T), tiles_c = (C + (size_t)WC * S::T - 1) / ((size_t)WC * S::T);
    const unsigned long long n_tiles = (unsigned long long)tiles_r * tiles_c;
    unsigned long long grid = (unsigned long long)info->sm_count * ctas_per_sm;
    if (grid > n_tiles) grid = n_tiles;
    const unsigned band = g_tr_band >= 0 ? (unsigned)g_tr_band : 0u;
    kern<<<(unsigned)grid, S::THREADS, S::SMEM_BYTES, st>>>(map_in, map_out, (unsigned)R, (unsigned)C, (unsigned)tiles_r, (unsigned)tiles_c, n_tiles, band, (unsigned)g_tr_hints);
    LA_LAUNCH_CHECK();
    return LA_OK;
}

// TROWS x TCOLS input tile (dynamic shared memory, padded rows).
template <typename E, int TROWS, int TCOLS>
__global__ void __launch_bounds__(TR_THREADS)
transpose_pack_kernel(E* __restrict__ out, const E* __restrict__ in, size_t R, size_t C, unsigned tiles_x, unsigned tiles_y, unsigned band, int no_turn) {
    using P = Pack<E>;
    constexpr int V = P::N;               // elements per 16-byte packet
    constexpr int VPR_IN = TCOLS / V;     // packets per input tile row
    constexpr int VPR_OUT = TROWS / V;    // packets per output tile row
    constexpr int PITCH = TCOLS + 1;
    extern __shared__ __align__(16) uint8_t tile_raw[];
    E* tile = reinterpret_cast<E*>(tile_raw);                  // tile[r * PITCH + c]
    unsigned long long tr, tc;
    tr_tile(blockIdx.x, tiles_y, tiles_x, band, tr, tc);       // band == 0: along the input rows (the default)
    const size_t r0 = (size_t)tr * TROWS, c0 = (size_t)tc * TCOLS;

    // R % V == 0 and C % V == 0 on this path: a packet is either wholly inside or wholly outside.
    constexpr int LOADS = TROWS * VPR_IN / TR_THREADS, STORES = TCOLS * VPR_OUT / TR_THREADS;
    static_assert(TROWS * VPR_IN % TR_THREADS == 0 && TCOLS * VPR_OUT % TR_THREADS == 0, "whole packets per thread");
    // all global loads first (LOADS independent 16-byte requests in flight per thread) ...
    P in_p[LOADS];
#pragma unroll
    for (int j = 0; j < LOADS; ++j) {
        const int idx = threadIdx.x + j * TR_THREADS;
        const int r = idx / VPR_IN, q = idx % VPR_IN;
        if (r0 + r < R && c0 + (size_t)q * V < C)
            in_p[j] = ld_stream(reinterpret_cast<const P*>(in + (r0 + r) * C + c0 + (size_t)q * V));
    }
    // Every load is issued before the first shared-memory store.  Left to itself the compiler sinks the later loads below the
    // first stores (which wait for the first load): one or two loads in flight per thread instead of LOADS, 88.6 instead of
    // 83.8 us for the i32 8192^2 transpose (6.06 -> 6.40 TB/s; ncu, one launch: 85.8 -> 80 us).
    asm volatile("" ::: "memory");
    if (no_turn) {
        // MEASUREMENT ONLY (la_transpose_set_mode(3)): the same global reads and the same global writes -- every tile lands where
        // its transpose belongs -- without the turn through shared memory (the output is then wrong inside each tile).  What this
        // runs at is what the memory system gives this access pattern; the difference to the real kernel is the cost of the turn.
        if constexpr (TROWS == TCOLS) {
#pragma unroll
            for (int j = 0; j < LOADS; ++j) {
                const int idx = threadIdx.x + j * TR_THREADS;
                const int r = idx / VPR_IN, q = idx % VPR_IN;
                if (c0 + r < C && r0 + (size_t)q * V < R) st_stream(reinterpret_cast<P*>(out + (c0 + r) * R + r0 + (size_t)q * V), in_p[j]);
            }
        }
        return;
    }
    // ... then scatter them into the tile
#pragma unroll
    for (int j = 0; j < LOADS; ++j) {
        const int idx = threadIdx.x + j * TR_THREADS;
        const int r = idx / VPR_IN, q = idx % VPR_IN;
#pragma unroll
        for (int e = 0; e < V; ++e) tile[r * PITCH + q * V + e] = in_p[j].v[e];
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < STORES; ++j) {
        const int idx = threadIdx.x + j * TR_THREADS;
        const int cc = idx / VPR_OUT, q = idx % VPR_OUT;      // output row c0+cc, output columns r0 + q*V ...
        P p;
#pragma unroll
        for (int e = 0; e < V; ++e) p.v[e] = tile[(q * V + e) * PITCH + cc];
        if (c0 + cc < C && r0 + (size_t)q * V < R)
            st_stream(reinterpret_cast<P*>(out + (c0 + cc) * R + r0 + (size_t)q * V), p);
    }
}

// Persistent form of the tile kernel: a fixed grid walks the tiles, the 16-byte loads of the NEXT tile are issued before the
// stores of the current one (two shared-memory buffers, one barrier per tile), so every thread always has loads in flight and
// there is no per-tile block launch.
template <typename E, int T>
__global__ void __launch_bounds__(TR_THREADS)
transpose_pack_persistent_kernel(E* __restrict__ out, const E* __restrict__ in, size_t R, size_t C, unsigned tiles_x, unsigned tiles) {
    using P = Pack<E>;
    constexpr int V = P::N, VPR = T / V, PITCH = T + 1;
    constexpr int LOADS = T * VPR / TR_THREADS;
    static_assert(T * VPR % TR_THREADS == 0, "whole packets per thread");
    extern __shared__ __align__(16) uint8_t tile_raw[];
    E* buf[2] = {reinterpret_cast<E*>(tile_raw), reinterpret_cast<E*>(tile_raw) + T * PITCH};
    P in_p[LOADS];
    auto fetch = [&](unsigned t) {
        const size_t r0 = (size_t)(t / tiles_x) * T, c0 = (size_t)(t % tiles_x) * T;
#pragma unroll
        for (int j = 0; j < LOADS; ++j) {
            const int idx = threadIdx.x + j * TR_THREADS, r = idx / VPR, q = idx % VPR;
            if (r0 + r < R && c0 + (size_t)q * V < C) in_p[j] = ld_stream(reinterpret_cast<const P*>(in + (r0 + r) * C + c0 + (size_t)q * V));
        }
    };
    unsigned t = blockIdx.x;
    if (t < tiles) fetch(t);
    for (int b = 0; t < tiles; t += gridDim.x, b ^= 1) {
        E* tile = buf[b];
#pragma unroll
        for (int j = 0; j < LOADS; ++j) {
            const int idx = threadIdx.x + j * TR_THREADS, r = idx / VPR, q = idx % VPR;
#pragma unroll
            for (int e = 0; e < V; ++e) tile[r * PITCH + q * V + e] = in_p[j].v[e];
        }
        __syncthreads();            // also orders this tile's writes after the reads of two tiles ago (same buffer)
        const size_t r0 = (size_t)(t / tiles_x) * T, c0 = (size_t)(t % tiles_x) * T;
        if (t + gridDim.x < tiles) fetch(t + gridDim.x);
#pragma unroll
        for (int j = 0; j < LOADS; ++j) {
            const int idx = threadIdx.x + j * TR_THREADS, cc = idx / VPR, q = idx % VPR;
            P p;
#pragma unroll
            for (int e = 0; e < V; ++e) p.v[e] = tile[(q * V + e) * PITCH + cc];
            if (c0 + cc < C && r0 + (size_t)q * V < R) st_stream(reinterpret_cast<P*>(out + (c0 + cc) * R + r0 + (size_t)q * V), p);
        }
    }
}

// 1- and 2-byte elements: the tile lives in shared memory as 32-bit words (G = 4 / sizeof(E) elements each), so the
// shared-memory traffic is word-sized instead of byte-sized; the G x G sub-blocks are transposed in registers with
// byte permutes.  Tile: TR = 128 / sizeof(E) input rows x 32 words (128 bytes) -> 32 G output rows of 128 bytes.
template <int G> __device__ __forceinline__ void subword_transpose(const uint32_t* in, uint32_t* out);
template <> __device__ __forceinline__ void subword_transpose<4>(const uint32_t* in, uint32_t* out) {      // 4 x 4 bytes
    const uint32_t t0 = __byte_perm(in[0], in[1], 0x5140), t1 = __byte_perm(in[0], in[1], 0x7362);       // a0 b0 a1 b1 | a2 b2 a3 b3
    const uint32_t t2 = __byte_perm(in[2], in[3], 0x5140), t3 = __byte_perm(in[2], in[3], 0x7362);       // c0 d0 c1 d1 | c2 d2 c3 d3
    out[0] = __byte_perm(t0, t2, 0x5410); out[1] = __byte_perm(t0, t2, 0x7632);                          // a0 b0 c0 d0 | a1 b1 c1 d1
    out[2] = __byte_perm(t1, t3, 0x5410); out[3] = __byte_perm(t1, t3, 0x7632);
}
template <> __device__ __forceinline__ void subword_transpose<2>(const uint32_t* in, uint32_t* out) {      // 2 x 2 halves
    out[0] = __byte_perm(in[0], in[1], 0x5410); out[1] = __byte_perm(in[0], in[1], 0x7632);
}

template <typename E>
__global__ void __launch_bounds__(TR_THREADS)
transpose_subword_kernel(E* __restrict__ out, const E* __restrict__ in, size_t R, size_t C, unsigned tiles_x) {
    constexpr int G = 4 / (int)sizeof(E);            // elements per word
    constexpr int TR = 128 / (int)sizeof(E);          // input rows per tile
    constexpr int TW = 32;                            // input words per tile row
    constexpr int TC = TW * G;                        // input columns (= output rows) per tile
    __shared__ uint32_t tile[TR][TW + 1];
    const size_t r0 = (size_t)(blockIdx.x / tiles_x) * TR, c0 = (size_t)(blockIdx.x % tiles_x) * TC;
    // load: TR rows x 8 packets of 16 bytes; R and C are multiples of 16 / sizeof(E) on this path
    constexpr int LOADS = TR * 8 / TR_THREADS;
    uint4 p[LOADS];
#pragma unroll
    for (int j = 0; j < LOADS; ++j) {
        const int idx = threadIdx.x + j * TR_THREADS, r = idx / 8, q = idx % 8;
        p[j] = (r0 + r < R && c0 + (size_t)q * 4 * G < C) ? __ldcs(reinterpret_cast<const uint4*>(in + (r0 + r) * C + c0 + (size_t)q * 4 * G)) : make_uint4(0, 0, 0, 0);
    }
    __syncwarp();                            // all loads are issued before the first shared-memory store (see transpose_pack_kernel)
#pragma unroll
    for (int j = 0; j < LOADS; ++j) {
        const int idx = threadIdx.x + j * TR_THREADS, r = idx / 8, q = idx % 8;
        tile[r][q * 4 + 0] = p[j].x; tile[r][q * 4 + 1] = p[j].y; tile[r][q * 4 + 2] = p[j].z; tile[r][q * 4 + 3] = p[j].w;
    }
    __syncthreads();
    // thread (w, jj): word column w, input rows 4 G jj .. 4 G jj + 4 G - 1 -> G output rows, 16 bytes each
    const int jj = threadIdx.x % 8, w = threadIdx.x / 8;
    uint32_t o[G][4];
#pragma unroll
    for (int blk = 0; blk < 4; ++blk) {
        uint32_t a[G], t[G];
#pragma unroll
        for (int g = 0; g < G; ++g) a[g] = tile[4 * G * jj + blk * G + g][w];
        subword_transpose<G>(a, t);
#pragma unroll
        for (int g = 0; g < G; ++g) o[g][blk] = t[g];
    }
#pragma unroll
    for (int g = 0; g < G; ++g) {
        const size_t orow = c0 + (size_t)w * G + g, ocol = r0 + (size_t)jj * 4 * G;
        if (orow < C && ocol < R) __stcs(reinterpret_cast<uint4*>(out + orow * R + ocol), make_uint4(o[g][0], o[g][1], o[g][2], o[g][3]));
    }
}

// Any shape / alignment: 32x32 tile, scalar accesses (coalesced on both sides).
template <typename E>
__global__ void __launch_bounds__(TR_THREADS)
transpose_elem_kernel(E* __restrict__ out, const E* __restrict__ in, size_t R, size_t C, unsigned tiles_x) {
    __shared__ E tile[32][33];
    const size_t r0 = (size_t)(blockIdx.x / tiles_x) * 32, c0 = (size_t)(blockIdx.x % tiles_x) * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;     // 32 x 8
#pragma unroll
    for (int j = ty; j < 32; j += 8)
        if (r0 + j < R && c0 + tx < C) tile[j][tx] = in[(r0 + j) * C + c0 + tx];
    __syncthreads();
#pragma unroll
    for (int j = ty; j < 32; j += 8)
        if (c0 + j < C && r0 + tx < R) out[(c0 + j) * R + r0 + tx] = tile[tx][j];
}

                           // la_transpose_set_variant (tuning hook): 4-byte tile shape; measured at i32 8192^2 / 32768^2:
                           // 64x64 6.06 / 6.51 TB/s (default), 64x32 6.26 / 6.38, 32x64 6.17 / 6.05, 128x128 4.73 / 5.11;
                           // 3..8: persistent pipelined walk with that many blocks per SM: 5.6 - 5.9 / 5.9 - 6.05 (64 registers
                           // per thread halve the occupancy; plain blocks at 8 per SM keep more loads in flight)


// ------------------------------------------------------------------ TMA loads, thread stores
// Measured on the B200 (tools/tma_bw.cu, 1 GiB, writes only): TMA stores of 128-byte-wide boxes reach 6.2 TB/s (fewer with
// more warps issuing), st.global.cs in the pattern of a transposing block (a warp writes whole 256-byte runs) 7.4 TB/s; reads
// only: TMA boxes 7.1 TB/s, ld.global.cs 6.7.  HBM moves one direction at a time, so a transpose takes about the sum of the
// two.  This kernel therefore reads with TMA and writes with st.global.cs:
//   lane 0   TMA-loads boxes of VR*T rows x 128 bytes (T = 128 / sizeof(E): VR square sub-tiles stacked) into the warp's ring,
//   32 lanes turn the box inside shared memory exactly as above (same lane map, conflict-free on both sides) into VR panels of
//            T output rows x 128 bytes,
//   32 lanes read the panels row-wise (a quarter-warp = one 128-byte panel row: conflict-free) and store them with
//            st.global.cs: a warp instruction writes 32 / (8 VR) output rows of 128 VR contiguous bytes each.
// No bulk groups, no proxy fence on the way out (the stores are ordinary), one output buffer per warp.
template <int ES, int VR, int WR, int WC, int IN_STAGES>
struct TrLdSt {
    static constexpr int WARPS = WR * WC, THREADS = WARPS * 32;
    static constexpr int N16 = 16 / ES, T = 8 * N16;
    static constexpr int BOX_ROWS = VR * T;                       // input rows per box
    static constexpr int BOX_BYTES = BOX_ROWS * 128;
    static constexpr size_t SMEM_BYTES = (size_t)WARPS * (IN_STAGES + 1) * BOX_BYTES + 1024;
};

template <int ES, int VR, int WR, int WC, int IN_STAGES>
__global__ void __launch_bounds__(WR * WC * 32)
transpose_tma_ldst_kernel(const __grid_constant__ CUtensorMap map_in, uint8_t* __restrict__ out,
                          unsigned R, unsigned C, unsigned tiles_r, unsigned tiles_c, unsigned long long n_tiles, unsigned band) {
    using S = TrLdSt<ES, VR, WR, WC, IN_STAGES>;
    constexpr int N16 = S::N16, T = S::T, BOX = S::BOX_BYTES, PANEL = T * 128;
    extern __shared__ uint8_t smem_raw[];
    __shared__ alignas(8) uint64_t full_bar[S::WARPS][IN_STAGES];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* in_ring = smem_raw + (base - smem_u32(smem_raw)) + (size_t)warp * (IN_STAGES + 1) * BOX;
    uint8_t* turned = in_ring + (size_t)IN_STAGES * BOX;
    uint64_t* bars = full_bar[warp];
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < IN_STAGES; ++s) mbar_init(&bars[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncwarp();

    const unsigned wr = warp / WC, wc = warp % WC;
    auto origin = [&](unsigned long long t, unsigned& r0, unsigned& c0) -> bool {
        unsigned long long tr, tc;
        tr_tile(t, tiles_r, tiles_c, band, tr, tc);
        const unsigned long long r = (tr * WR + wr) * S::BOX_ROWS, c = (tc * WC + wc) * T;
        r0 = (unsigned)r; c0 = (unsigned)c;
        return r < R && c < C;
    };
    unsigned long long t_load = blockIdx.x, t_turn = blockIdx.x;
    unsigned n_loaded = 0, n_turned = 0;
    auto load_next = [&]() {
        unsigned r0, c0;
        while (t_load < n_tiles) {
            const bool ok = origin(t_load, r0, c0);
            t_load += gridDim.x;
            if (ok) {
                const int s = (int)(n_loaded % IN_STAGES);
                mbar_expect_tx(&bars[s], (uint32_t)BOX);
                tma_load_2d(in_ring + (size_t)s * BOX, &map_in, &bars[s], (int)c0, (int)r0);
                ++n_loaded;
                return;
            }
        }
    };
    if (lane == 0) {
#pragma unroll 1
        for (int s = 0; s < IN_STAGES; ++s) load_next();
    }
    const int li = lane & 7, ld = lane >> 3;
    const size_t out_pitch = (size_t)R * ES;                                  // bytes per output row
#pragma unroll 1
    for (; t_turn < n_tiles; t_turn += gridDim.x) {
        unsigned r0, c0;
        if (!origin(t_turn, r0, c0)) continue;
        const int s = (int)(n_turned % IN_STAGES);
        const uint32_t ph = (n_turned / IN_STAGES) & 1u;
        mbar_wait(&bars[s], ph);
        const uint8_t* src = in_ring + (size_t)s * BOX;
        // turn: per panel 8 row blocks x 8 chunk columns = 2 rounds of 32 lane blocks; row block rb = li ^ d inside the panel
#pragma unroll
        for (int round = 0; round < 2 * VR; ++round) {
            const int d = ld + 4 * (round & 1), panel = round >> 1;
            const int cc = li, rb = li ^ d;
            uint4 v[N16];
#pragma unroll
            for (int j = 0; j < N16; ++j) {
                const int r = N16 * rb + j;                                     // row inside the panel's T input rows
                v[j] = *reinterpret_cast<const uint4*>(src + (panel * T + r) * 128 + ((cc ^ (r & 7)) << 4));
            }
#pragma unroll
            for (int j = 0; j < N16; ++j) {
                const int c = N16 * cc + j;
                *reinterpret_cast<uint4*>(turned + panel * PANEL + c * 128 + ((rb ^ (c & 7)) << 4)) = turned_row<ES>(v, j);
            }
        }
        __syncwarp();
        if (lane == 0) load_next();                                           // every lane is done reading in_ring[s]
        // out rows c0 .. c0 + T - 1, each VR * 128 bytes starting at column r0: 8 VR lanes per row
        constexpr int LANES_PER_ROW = 8 * VR, ROWS_PER_INSTR = 32 / LANES_PER_ROW;
        const int q16 = lane % LANES_PER_ROW, panel = q16 >> 3, q = q16 & 7;
        const bool col_ok = r0 + (unsigned)(q16 * N16) < R;                   // R * ES % 16 == 0: a chunk is wholly inside or outside
#pragma unroll
        for (int k = 0; k < T / ROWS_PER_INSTR; ++k) {
            const int c = k * ROWS_PER_INSTR + lane / LANES_PER_ROW;
            const uint4 w = *reinterpret_cast<const uint4*>(turned + panel * PANEL + c * 128 + ((q ^ (c & 7)) << 4));
            if (col_ok && c0 + (unsigned)c < C)
                __stcs(reinterpret_cast<uint4*>(out + (size_t)(c0 + c) * out_pitch + (size_t)r0 * ES + (size_t)q16 * 16), w);
        }
        __syncwarp();                                                         // `turned` is free for the next box
        ++n_turned;
    }
}

template <int ES, int VR, int WR, int WC, int IN_STAGES>
la_status transpose_tma_ldst_launch(void* out, const void* in, size_t R, size_t C, cudaStream_t st, int ctas_per_sm) {
    using S = TrLdSt<ES, VR, WR, WC, IN_STAGES>;
    static_assert(S::SMEM_BYTES * 1 <= 227 * 1024, "shared memory per block");
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    CUtensorMap map_in;
    LA_TRY(tr_make_map(&map_in, in, R, C, ES, S::BOX_ROWS));
    auto kern = transpose_tma_ldst_kernel<ES, VR, WR, WC, IN_STAGES>;
    LA_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::SMEM_BYTES));
    const size_t tile_rows = (size_t)WR * S::BOX_ROWS, tile_cols = (size_t)WC * S::T;
    const size_t tiles_r = (R + tile_rows - 1) / tile_rows, tiles_c = (C + tile_cols - 1) / tile_cols;
    const unsigned long long n_tiles = (unsigned long long)tiles_r * tiles_c;
    unsigned long long grid = (unsigned long long)info->sm_count * ctas_per_sm;
    if (grid > n_tiles) grid = n_tiles;
    const unsigned band = g_tr_band >= 0 ? (unsigned)g_tr_band : 0u;
    kern<<<(unsigned)grid, S::THREADS, S::SMEM_BYTES, st>>>(map_in, (uint8_t*)out, (unsigned)R, (unsigned)C, (unsigned)tiles_r, (unsigned)tiles_c, n_tiles, band);
    LA_LAUNCH_CHECK();
    return LA_OK;
}

template <int ES>
la_status transpose_tma_variant(int variant, void* out, const void* in, size_t R, size_t C, cudaStream_t st) {
    switch (variant) {
    default:
    // TMA loads + TMA stores: <ES, WR, WC, IN_STAGES, OUT_STAGES>, blocks per SM
    case 20: return transpose_tma_launch<ES, 2, 4, 4, 2>(out, in, R, C, st, 1);     // 8 warps: 64 x 128 words per block, 192 KB, one block per SM
    case 21: return transpose_tma_launch<ES, 2, 2, 4, 2>(out, in, R, C, st, 2);     // 4 warps, two blocks per SM
    case 22: return transpose_tma_launch<ES, 4, 4, 2, 1>(out, in, R, C, st, 1);     // 16 warps, shallow rings
    case 23: return transpose_tma_launch<ES, 4, 2, 4, 2>(out, in, R, C, st, 1);     // 128 x 64 words per block
    // TMA loads + st.global.cs stores: <ES, VR, WR, WC, IN_STAGES>
    case 30: return transpose_tma_ldst_launch<ES, 2, 2, 4, 2>(out, in, R, C, st, 1);    // 8 warps, 256-byte output runs
    case 31: return transpose_tma_ldst_launch<ES, 1, 2, 4, 4>(out, in, R, C, st, 1);    // 128-byte output runs
    case 32: return transpose_tma_ldst_launch<ES, 2, 1, 4, 2>(out, in, R, C, st, 2);    // 4 warps, 96 KB: two blocks per SM
    case 33: return transpose_tma_ldst_launch<ES, 1, 4, 4, 1>(out, in, R, C, st, 1);    // 16 warps, single-stage rings
    case 34: return transpose_tma_ldst_launch<ES, 1, 2, 4, 2>(out, in, R, C, st, 2);    // 8 warps x 12 KB = 96 KB: two blocks per SM (16 warps per SM)
    }
}

template <typename E, int TROWS, int TCOLS>
la_status transpose_pack_launch(E* out, const E* in, size_t R, size_t C, cudaStream_t st) {
    const size_t gx = (C + TCOLS - 1) / TCOLS, gy = (R + TROWS - 1) / TROWS;
    // 1-D grid of tiles (row-major over the input), so neither dimension hits the 65535 limit
    if (gx * gy > 0x7fffffffull || gx > 0xffffffffull) return la_fail(LA_ERR_UNSUPPORTED, "transpose: %zu x %zu is too large", R, C);
    auto kern = transpose_pack_kernel<E, TROWS, TCOLS>;
    constexpr size_t smem = (size_t)TROWS * (TCOLS + 1) * sizeof(E);
    if (smem > 48 * 1024) LA_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)(gx * gy), TR_THREADS, smem, st>>>(out, in, R, C, (unsigned)gx, (unsigned)gy, g_tr_band > 0 ? (unsigned)g_tr_band : 0u, g_tr_mode == 3 ? 1 : 0);
    LA_LAUNCH_CHECK();
    return LA_OK;
}

template <typename E, int TILE>
la_status transpose_launch(E* out, const E* in, size_t R, size_t C, cudaStream_t st) {
    constexpr int V = Pack<E>::N;
    const bool packed = la_aligned16(out) && la_aligned16(in) && R % V == 0 && C % V == 0;
    if constexpr (sizeof(E) >= 4) {
        // the TMA kernels (variants 20.. / 30..): whole rows must start on 16-byte boundaries and the coordinates are 32-bit
        if (g_tr_variant >= 20 && packed && R < (1u << 31) && C < (1u << 31)) return transpose_tma_variant<(int)sizeof(E)>(g_tr_variant, out, in, R, C, st);
    }
    if (packed) {
        if constexpr (sizeof(E) < 4) {
            constexpr int TRs = 128 / (int)sizeof(E), TCs = 32 * (4 / (int)sizeof(E));
            const size_t gx = (C + TCs - 1) / TCs, gy = (R + TRs - 1) / TRs;
            if (gx * gy > 0x7fffffffull || gx > 0xffffffffull) return la_fail(LA_ERR_UNSUPPORTED, "transpose: %zu x %zu is too large", R, C);
            if (g_tr_variant != 9) {       // 9 = the generic tile kernel (tuning reference)
                transpose_subword_kernel<E><<<(unsigned)(gx * gy), TR_THREADS, 0, st>>>(out, in, R, C, (unsigned)gx);
                LA_LAUNCH_CHECK();
                return LA_OK;
            }
        }
        if constexpr (sizeof(E) == 4) {
            if (g_tr_variant >= 3 && g_tr_variant <= 8) {          // persistent tile walk, g_tr_variant blocks per SM
                const size_t gx = (C + 63) / 64, gy = (R + 63) / 64;
                if (gx * gy > 0x7fffffffull) return la_fail(LA_ERR_UNSUPPORTED, "transpose: %zu x %zu is too large", R, C);
                LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
                auto kern = transpose_pack_persistent_kernel<E, 64>;
                constexpr size_t smem = 2 * (size_t)64 * 65 * sizeof(E);
                size_t grid = (size_t)info->sm_count * g_tr_variant;
                if (grid > gx * gy) grid = gx * gy;
                kern<<<(unsigned)grid, TR_THREADS, smem, st>>>(out, in, R, C, (unsigned)gx, (unsigned)(gx * gy));
                LA_LAUNCH_CHECK();
                return LA_OK;
            }
            switch (g_tr_variant) {
            case 1: return transpose_pack_launch<E, 32, 64>(out, in, R, C, st);
            case 2: return transpose_pack_launch<E, 64, 32>(out, in, R, C, st);
            default: break;
            }
        }
        return transpose_pack_launch<E, TILE, TILE>(out, in, R, C, st);
    }
    const size_t gx = (C + 31) / 32, gy = (R + 31) / 32;
    if (gx * gy > 0x7fffffffull || gx > 0xffffffffull) return la_fail(LA_ERR_UNSUPPORTED, "transpose: %zu x %zu is too large", R, C);
    transpose_elem_kernel<E><<<(unsigned)(gx * gy), TR_THREADS, 0, st>>>(out, in, R, C, (unsigned)gx);
    LA_LAUNCH_CHECK();
    return LA_OK;
}

}  // namespace

// used by gemm.cu to stage A^T
la_status la_transpose_raw32(uint32_t* out, const uint32_t* in, size_t R, size_t C, cudaStream_t st) {
    return transpose_launch<uint32_t, 64>(out, in, R, C, st);
}

la_status la_transpose_raw64(uint64_t* out, const uint64_t* in, size_t R, size_t C, cudaStream_t st) {
    return transpose_launch<uint64_t, 32>(out, in, R, C, st);
}

extern "C" la_status la_transpose_set_variant(int v) { g_tr_variant = v; return LA_OK; }      // tuning hook
extern "C" la_status la_transpose_set_band(int rows) { g_tr_band = rows; return LA_OK; }       // tuning hook
extern "C" la_status la_transpose_set_hints(int bits) { g_tr_hints = bits; return LA_OK; }     // tuning hook
extern "C" la_status la_transpose_set_mode(int mode) { g_tr_mode = mode; return LA_OK; }       // measurement hook (1 / 2: half of the data path only)

extern "C" la_status la_transpose(int dtype, void* out, const void* in, size_t rows, size_t cols, void* stream) {
    if (!la_valid_dtype(dtype)) return la_fail(LA_ERR_ARG, "la_transpose: unknown dtype %d", dtype);
    if (rows == 0 || cols == 0) return LA_OK;
    if (!out || !in) return la_fail(LA_ERR_ARG, "la_transpose: NULL operand");
    if (out == in) return la_fail(LA_ERR_ARG, "la_transpose: out must not alias in");
    LaDeviceInfo* info; LA_TRY(la_ensure_init(&info));
    cudaStream_t st = (cudaStream_t)stream;
    switch (la_dtype_size(dtype)) {
    case 1: return transpose_launch<uint8_t, 128>((uint8_t*)out, (const uint8_t*)in, rows, cols, st);
    case 2: return transpose_launch<uint16_t, 64>((uint16_t*)out, (const uint16_t*)in, rows, cols, st);
    case 4: return transpose_launch<uint32_t, 64>((uint32_t*)out, (const uint32_t*)in, rows, cols, st);
    default: return transpose_launch<uint64_t, 32>((uint64_t*)out, (const uint64_t*)in, rows, cols, st);
    }
}
