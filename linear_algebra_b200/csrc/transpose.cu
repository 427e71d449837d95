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
#include "common.cuh"
#include "../../include/linalg_b200_tuning.h"

namespace {

constexpr int TR_THREADS = 256;

// TROWS x TCOLS input tile (dynamic shared memory, padded rows).
template <typename E, int TROWS, int TCOLS>
__global__ void __launch_bounds__(TR_THREADS)
transpose_pack_kernel(E* __restrict__ out, const E* __restrict__ in, size_t R, size_t C, unsigned tiles_x) {
    using P = Pack<E>;
    constexpr int V = P::N;               // elements per 16-byte packet
    constexpr int VPR_IN = TCOLS / V;     // packets per input tile row
    constexpr int VPR_OUT = TROWS / V;    // packets per output tile row
    constexpr int PITCH = TCOLS + 1;
    extern __shared__ __align__(16) uint8_t tile_raw[];
    E* tile = reinterpret_cast<E*>(tile_raw);                  // tile[r * PITCH + c]
    const size_t r0 = (size_t)(blockIdx.x / tiles_x) * TROWS, c0 = (size_t)(blockIdx.x % tiles_x) * TCOLS;

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

int g_tr_variant = 0;      // la_transpose_set_variant (tuning hook): 4-byte tile shape; measured at i32 8192^2 / 32768^2:
                           // 64x64 6.06 / 6.51 TB/s (default), 64x32 6.26 / 6.38, 32x64 6.17 / 6.05, 128x128 4.73 / 5.11;
                           // 3..8: persistent pipelined walk with that many blocks per SM: 5.6 - 5.9 / 5.9 - 6.05 (64 registers
                           // per thread halve the occupancy; plain blocks at 8 per SM keep more loads in flight)

template <typename E, int TROWS, int TCOLS>
la_status transpose_pack_launch(E* out, const E* in, size_t R, size_t C, cudaStream_t st) {
    const size_t gx = (C + TCOLS - 1) / TCOLS, gy = (R + TROWS - 1) / TROWS;
    // 1-D grid of tiles (row-major over the input), so neither dimension hits the 65535 limit
    if (gx * gy > 0x7fffffffull || gx > 0xffffffffull) return la_fail(LA_ERR_UNSUPPORTED, "transpose: %zu x %zu is too large", R, C);
    auto kern = transpose_pack_kernel<E, TROWS, TCOLS>;
    constexpr size_t smem = (size_t)TROWS * (TCOLS + 1) * sizeof(E);
    if (smem > 48 * 1024) LA_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)(gx * gy), TR_THREADS, smem, st>>>(out, in, R, C, (unsigned)gx);
    LA_LAUNCH_CHECK();
    return LA_OK;
}

template <typename E, int TILE>
la_status transpose_launch(E* out, const E* in, size_t R, size_t C, cudaStream_t st) {
    constexpr int V = Pack<E>::N;
    const bool packed = la_aligned16(out) && la_aligned16(in) && R % V == 0 && C % V == 0;
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
