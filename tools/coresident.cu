// tools/coresident.cu -- can a block of a small kernel share an SM with a CTA of a kernel that holds 144 x 384 registers and 194 KB
// of shared memory?  Kernel X (one CTA per SM, clusters of 2 like the f32 tensor engine's pair kernel) spins until kernel Y, launched
// on another stream, has set a word -- or gives up after 50 ms.  Reported per configuration: whether Y ran while X was resident.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/coresident tools/coresident.cu && /tmp/coresident
#include <cstdio>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s failed: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

template <int BIG_REGS>
__global__ void __maxnreg__(BIG_REGS) big_kernel(volatile unsigned* word, unsigned* seen, long long limit_clocks, float* sink, int rounds) {
    extern __shared__ unsigned char smem[];
    // keep ~BIG_REGS - 16 values live so that the kernel really is allocated BIG_REGS registers per thread
    constexpr int LIVE = BIG_REGS - 10;
    float acc[LIVE];
#pragma unroll
    for (int i = 0; i < LIVE; ++i) acc[i] = (float)(threadIdx.x + i);
    if (threadIdx.x == 0) {
        smem[0] = 1;
        const long long t0 = clock64();
        unsigned v = 0;
        while ((v = *word) < gridDim.x && clock64() - t0 < limit_clocks) __nanosleep(200);      // until EVERY small block has started
        if (v >= gridDim.x) atomicAdd(seen, 1u);
        atomicMax(seen + 1, v);
    }
    __syncthreads();
    for (int r = 0; r < rounds; ++r) {
#pragma unroll
        for (int i = 0; i < LIVE; ++i) acc[i] = acc[i] * 1.0001f + acc[(i + 1) % LIVE];
    }
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < LIVE; ++i) sum += acc[i];
    if (sum == 12345.678f) sink[threadIdx.x] = sum;
}

template <int REGS>
__global__ void __maxnreg__(REGS) small_kernel(unsigned* word, unsigned* ran, float* sink, int rounds) {
    __shared__ float tile[64][65];
    constexpr int LIVE = REGS - 4;
    float acc[LIVE];
#pragma unroll
    for (int i = 0; i < LIVE; ++i) acc[i] = (float)(threadIdx.x * 3 + i);
    tile[threadIdx.x % 64][threadIdx.x / 64] = (float)threadIdx.x;
    __syncthreads();
    if (threadIdx.x == 0) { atomicAdd(ran, 1u); atomicAdd(word, (unsigned)tile[1][0]); }     // tile[1][0] = 1
    for (int r = 0; r < rounds; ++r) {
#pragma unroll
        for (int i = 0; i < LIVE; ++i) acc[i] = acc[i] * 1.0001f + acc[(i + 1) % LIVE];
    }
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < LIVE; ++i) sum += acc[i];
    if (sum == 12345.678f) sink[threadIdx.x] = sum;
}

template <int BIG_REGS, int REGS>
static int run(const char* what, int big_threads, size_t big_smem, int small_threads, int sms) {
    unsigned *word, *seen, *ran; float* sink;
    CK(cudaMalloc(&word, 4)); CK(cudaMalloc(&seen, 8)); CK(cudaMalloc(&ran, 4)); CK(cudaMalloc(&sink, 4096));
    CK(cudaMemset(word, 0, 4)); CK(cudaMemset(seen, 0, 8)); CK(cudaMemset(ran, 0, 4));
    cudaStream_t s1, s2;
    CK(cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking)); CK(cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking));
    CK(cudaFuncSetAttribute(big_kernel<BIG_REGS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)big_smem));
    CK(cudaFuncSetAttribute(big_kernel<BIG_REGS>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    CK(cudaFuncSetAttribute(small_kernel<REGS>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    cudaFuncAttributes a, b; CK(cudaFuncGetAttributes(&a, big_kernel<BIG_REGS>)); CK(cudaFuncGetAttributes(&b, small_kernel<REGS>));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(sms); cfg.blockDim = dim3(big_threads); cfg.dynamicSmemBytes = big_smem; cfg.stream = s1;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension; attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    const long long limit = 100000000LL;       // ~50 ms
    CK(cudaLaunchKernelEx(&cfg, big_kernel<BIG_REGS>, (volatile unsigned*)word, seen, limit, sink, 1));
    small_kernel<REGS><<<sms, small_threads, 0, s2>>>(word, ran, sink, 1);
    CK(cudaDeviceSynchronize());
    unsigned h_seen2[2] = {0, 0}, h_ran = 0;
    CK(cudaMemcpy(h_seen2, seen, 8, cudaMemcpyDeviceToHost));
    const unsigned h_seen = h_seen2[0]; CK(cudaMemcpy(&h_ran, ran, 4, cudaMemcpyDeviceToHost));
    printf("%-40s big %3d regs x %d thr = %6d, small %2d regs x %d thr = %6d, sum %6d, smem %3zu KB: big CTAs that saw ALL small blocks start: %3u of %d; small blocks started while the big CTAs waited: %3u\n", what, a.numRegs, big_threads,
           a.numRegs * big_threads, b.numRegs, small_threads, b.numRegs * small_threads, a.numRegs * big_threads + b.numRegs * small_threads, big_smem / 1024, h_seen, sms, h_seen2[1]);
    cudaFree(word); cudaFree(seen); cudaFree(ran); cudaFree(sink); cudaStreamDestroy(s1); cudaStreamDestroy(s2);
    return 0;
}

int main() {
    int sms = 0; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const size_t smem = 3 * 65536 + 1024;
    // the small kernel is launched AFTER the big one: it only sets the word in time if its blocks share the SMs with the big CTAs
    run<144, 40>("144 + 40", 384, smem, 256, sms);
    run<144, 32>("144 + 32", 384, smem, 256, sms);
    run<144, 40>("144 + 40, small 128 thr", 384, smem, 128, sms);
    run<136, 40>("136 + 40", 384, smem, 256, sms);
    run<136, 32>("136 + 32", 384, smem, 256, sms);
    run<128, 40>("128 + 40", 384, smem, 256, sms);
    run<128, 32>("128 + 32", 384, smem, 256, sms);
    run<120, 40>("120 + 40", 384, smem, 256, sms);
    run<112, 40>("112 + 40", 384, smem, 256, sms);
    run<96, 40>("96 + 40", 384, smem, 256, sms);
    run<144, 40>("144 + 40, big 160 KB", 384, 160 * 1024, 256, sms);
    run<144, 40>("144 + 40, big 128 KB", 384, 128 * 1024, 256, sms);
    run<144, 40>("144 + 40, big 96 KB", 384, 96 * 1024, 256, sms);
    run<128, 40>("128 + 40, big 128 KB", 384, 128 * 1024, 256, sms);
    run<96, 40>("96 + 40, big 96 KB", 384, 96 * 1024, 256, sms);
    return 0;
}
