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

namespace {

constexpr int TR_THREADS = 256;

template <typename E, int TILE>
__global__ void __launch_bounds__(TR_THREADS)
transpose_pack_kernel(E* __restrict__ out, const E* __restrict__ in, size_t R, size_t C, unsigned tiles_x) {
    using P = Pack<E>;
    constexpr int V = P::N;               // elements per 16-byte packet
    constexpr int VPR = TILE / V;         // packets per tile row
    __shared__ E tile[TILE][TILE + 1];
    const size_t r0 = (size_t)(blockIdx.x / tiles_x) * TILE, c0 = (size_t)(blockIdx.x % tiles_x) * TILE;

    // R % V == 0 and C % V == 0 on this path: a packet is either wholly inside or wholly outside.
    constexpr int PER_THREAD = TILE * VPR / TR_THREADS;
    static_assert(TILE * VPR % TR_THREADS == 0, "whole packets per thread");
    // all global loads first (PER_THREAD independent 16-byte requests in flight per thread) ...
    P in_p[PER_THREAD];
#pragma unroll
    for (int j = 0; j < PER_THREAD; ++j) {
        const int idx = threadIdx.x + j * TR_THREADS;
        const int r = idx / VPR, q = idx % VPR;
        if (r0 + r < R && c0 + (size_t)q * V < C)
            in_p[j] = ld_stream(reinterpret_cast<const P*>(in + (r0 + r) * C + c0 + (size_t)q * V));
    }
    // ... then scatter them into the tile
#pragma unroll
    for (int j = 0; j < PER_THREAD; ++j) {
        const int idx = threadIdx.x + j * TR_THREADS;
        const int r = idx / VPR, q = idx % VPR;
#pragma unroll
        for (int e = 0; e < V; ++e) tile[r][q * V + e] = in_p[j].v[e];
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < PER_THREAD; ++j) {
        const int idx = threadIdx.x + j * TR_THREADS;
        const int cc = idx / VPR, q = idx % VPR;      // output row c0+cc, output columns r0 + q*V ...
        P p;
#pragma unroll
        for (int e = 0; e < V; ++e) p.v[e] = tile[q * V + e][cc];
        if (c0 + cc < C && r0 + (size_t)q * V < R)
            st_stream(reinterpret_cast<P*>(out + (c0 + cc) * R + r0 + (size_t)q * V), p);
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

template <typename E, int TILE>
la_status transpose_launch(E* out, const E* in, size_t R, size_t C, cudaStream_t st) {
    constexpr int V = Pack<E>::N;
    const bool packed = la_aligned16(out) && la_aligned16(in) && R % V == 0 && C % V == 0;
    const int tile = packed ? TILE : 32;
    const size_t gx = (C + tile - 1) / tile, gy = (R + tile - 1) / tile;
    // 1-D grid of tiles (row-major over the input), so neither dimension hits the 65535 limit
    if (gx * gy > 0x7fffffffull || gx > 0xffffffffull) return la_fail(LA_ERR_UNSUPPORTED, "transpose: %zu x %zu is too large", R, C);
    const unsigned grid = (unsigned)(gx * gy);
    if (packed) transpose_pack_kernel<E, TILE><<<grid, TR_THREADS, 0, st>>>(out, in, R, C, (unsigned)gx);
    else        transpose_elem_kernel<E><<<grid, TR_THREADS, 0, st>>>(out, in, R, C, (unsigned)gx);
    LA_LAUNCH_CHECK();
    return LA_OK;
}

}  // namespace

// used by gemm.cu to stage A^T
la_status la_transpose_raw32(uint32_t* out, const uint32_t* in, size_t R, size_t C, cudaStream_t st) {
    return transpose_launch<uint32_t, 64>(out, in, R, C, st);
}

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
