// tools/tma_bw.cu -- what do the two halves of a tile-moving kernel reach on their own?  (measurement tool, not product)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/build/tma_bw tools/tma_bw.cu -lcuda
// Writes / reads a 16384 x 16384 matrix of 4-byte words (1 GiB, 64 KB row pitch) in tile patterns:
//   plain   st.global.cs / ld.global.cs, 16 bytes per lane, linear
//   tile    st.global.cs in the pattern of a transposing block: a warp instruction covers 2 rows x 256 bytes
//   tma     cp.async.bulk.tensor.2d boxes of BOX_ROWS x 128 bytes (SWIZZLE_128B) from / into shared memory, W warps per
//           block each issuing its own boxes, D bulk groups (stores) or ring stages (loads) outstanding per warp
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

constexpr size_t ROWS = 16384, COLS = 16384;      // words

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void plain_write(uint4* out, size_t n16) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) __stcs(out + i, make_uint4(1, 2, 3, (unsigned)i));
}
__global__ void plain_read(const uint4* in, size_t n16, unsigned* sink) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    unsigned acc = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) { const uint4 v = __ldcs(in + i); acc ^= v.x ^ v.y ^ v.z ^ v.w; }
    if (acc == 0x12345678u) *sink = acc;
}
// one block = one 64 x 64-word tile: 64 rows x 256 bytes, 16 bytes per lane, 4 stores per thread
__global__ void tile_write(uint32_t* out, unsigned tiles_x) {
    const size_t r0 = (size_t)(blockIdx.x / tiles_x) * 64, c0 = (size_t)(blockIdx.x % tiles_x) * 64;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int idx = threadIdx.x + j * 256, r = idx / 16, q = idx % 16;
        __stcs(reinterpret_cast<uint4*>(out + (r0 + r) * COLS + c0 + q * 4), make_uint4(1, 2, 3, idx));
    }
}

template <int BOX_ROWS, int DEPTH>
__global__ void tma_store_kernel(const __grid_constant__ CUtensorMap map, unsigned boxes_x, unsigned long long n_boxes) {
    extern __shared__ uint8_t smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, warps = blockDim.x >> 5;
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t buf = base + warp * BOX_ROWS * 128;
    if (lane == 0) {
        for (unsigned long long b = (unsigned long long)blockIdx.x * warps + warp; b < n_boxes; b += (unsigned long long)gridDim.x * warps) {
            const int x = (int)(b % boxes_x) * 32, y = (int)(b / boxes_x) * BOX_ROWS;
            asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(&map), "r"(buf), "r"(x), "r"(y) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(DEPTH - 1) : "memory");
        }
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
}

template <int BOX_ROWS, int DEPTH>
__global__ void tma_load_kernel(const __grid_constant__ CUtensorMap map, unsigned boxes_x, unsigned long long n_boxes) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ alignas(8) uint64_t bars[32][DEPTH];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, warps = blockDim.x >> 5;
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t ring = base + warp * DEPTH * BOX_ROWS * 128;
    if (lane == 0) {
        for (int s = 0; s < DEPTH; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bars[warp][s])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        unsigned long long issued = 0, done = 0;
        const unsigned long long first = (unsigned long long)blockIdx.x * warps + warp, step = (unsigned long long)gridDim.x * warps;
        const unsigned long long mine = first < n_boxes ? (n_boxes - first + step - 1) / step : 0;
        while (done < mine) {
            while (issued < mine && issued < done + DEPTH) {
                const unsigned long long b = first + issued * step;
                const int s = (int)(issued % DEPTH);
                const int x = (int)(b % boxes_x) * 32, y = (int)(b / boxes_x) * BOX_ROWS;
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bars[warp][s])), "r"(BOX_ROWS * 128) : "memory");
                asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                             ::"r"(ring + s * BOX_ROWS * 128), "l"(&map), "r"(smem_u32(&bars[warp][s])), "r"(x), "r"(y) : "memory");
                ++issued;
            }
            const int s = (int)(done % DEPTH);
            const uint32_t ph = (uint32_t)(done / DEPTH) & 1u;
            uint32_t ok = 0;
            while (!ok)
                asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(&bars[warp][s])), "r"(ph) : "memory");
            ++done;
        }
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static CUtensorMap make_map(void* base, int box_rows) {
    static EncodeTiledFn enc = nullptr;
    if (!enc) { cudaDriverEntryPointQueryResult q; void* p; CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q)); enc = (EncodeTiledFn)p; }
    CUtensorMap m;
    cuuint64_t dims[2] = {COLS, ROWS}; cuuint64_t strides[1] = {COLS * 4};
    cuuint32_t box[2] = {32, (cuuint32_t)box_rows}; cuuint32_t es[2] = {1, 1};
    CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); exit(1); }
    return m;
}

template <typename F> static float time_ms(F f, int reps = 10) {
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    for (int i = 0; i < 2; ++i) f();
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(a));
    for (int i = 0; i < reps; ++i) f();
    CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    CK(cudaGetLastError());
    float ms; CK(cudaEventElapsedTime(&ms, a, b));
    return ms / reps;
}

template <int BOX_ROWS, int DEPTH>
static void run_tma(void* buf, int warps, int sms, int ctas_per_sm) {
    const double bytes = (double)ROWS * COLS * 4;
    CUtensorMap m = make_map(buf, BOX_ROWS);
    const unsigned boxes_x = COLS / 32; const unsigned long long n_boxes = (unsigned long long)boxes_x * (ROWS / BOX_ROWS);
    const size_t smem_st = (size_t)warps * BOX_ROWS * 128 + 1024, smem_ld = (size_t)warps * DEPTH * BOX_ROWS * 128 + 1024;
    if (smem_st <= 227 * 1024 / ctas_per_sm) {
        CK(cudaFuncSetAttribute(tma_store_kernel<BOX_ROWS, DEPTH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_st));
        const float ms = time_ms([&] { tma_store_kernel<BOX_ROWS, DEPTH><<<sms * ctas_per_sm, warps * 32, smem_st>>>(m, boxes_x, n_boxes); });
        printf("tma STORE box %3d x 128 B  depth %d  %2d warps x %d blocks/SM on %3d SMs: %.4f ms  %6.0f GB/s\n", BOX_ROWS, DEPTH, warps, ctas_per_sm, sms, ms, bytes / ms / 1e6);
    }
    if (smem_ld <= 227 * 1024 / ctas_per_sm) {
        CK(cudaFuncSetAttribute(tma_load_kernel<BOX_ROWS, DEPTH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_ld));
        const float ms = time_ms([&] { tma_load_kernel<BOX_ROWS, DEPTH><<<sms * ctas_per_sm, warps * 32, smem_ld>>>(m, boxes_x, n_boxes); });
        printf("tma LOAD  box %3d x 128 B  depth %d  %2d warps x %d blocks/SM on %3d SMs: %.4f ms  %6.0f GB/s\n", BOX_ROWS, DEPTH, warps, ctas_per_sm, sms, ms, bytes / ms / 1e6);
    }
    fflush(stdout);
}

int main() {
    void* buf; CK(cudaMalloc(&buf, ROWS * COLS * 4));
    unsigned* sink; CK(cudaMalloc(&sink, 4));
    CK(cudaMemset(buf, 1, ROWS * COLS * 4));
    int sms; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const double bytes = (double)ROWS * COLS * 4;
    const size_t n16 = ROWS * COLS / 4;
    float ms = time_ms([&] { plain_write<<<sms * 16, 256>>>((uint4*)buf, n16); });
    printf("plain st.global.cs linear write: %.4f ms  %6.0f GB/s\n", ms, bytes / ms / 1e6);
    ms = time_ms([&] { plain_read<<<sms * 16, 256>>>((const uint4*)buf, n16, sink); });
    printf("plain ld.global.cs linear read : %.4f ms  %6.0f GB/s\n", ms, bytes / ms / 1e6);
    ms = time_ms([&] { CK(cudaMemsetAsync(buf, 0, ROWS * COLS * 4)); });
    printf("cudaMemsetAsync                : %.4f ms  %6.0f GB/s\n", ms, bytes / ms / 1e6);
    ms = time_ms([&] { tile_write<<<(unsigned)((ROWS / 64) * (COLS / 64)), 256>>>((uint32_t*)buf, (unsigned)(COLS / 64)); });
    printf("64 x 64-word tile write (st.cs): %.4f ms  %6.0f GB/s\n", ms, bytes / ms / 1e6);
    run_tma<32, 2>(buf, 8, sms, 1);
    run_tma<32, 4>(buf, 8, sms, 1);
    run_tma<32, 4>(buf, 16, sms, 1);
    run_tma<64, 2>(buf, 8, sms, 1);
    run_tma<128, 2>(buf, 4, sms, 1);
    run_tma<256, 2>(buf, 2, sms, 1);
    run_tma<32, 4>(buf, 8, sms / 2, 1);      // half of the SMs: is the limit per SM or chip-wide?
    run_tma<8, 8>(buf, 8, sms, 1);
    return 0;
}
