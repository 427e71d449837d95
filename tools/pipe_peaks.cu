// Measured pipe peaks on this GPU from register-resident loops: DMMA.8x8x4 (mma.sync m8n8k4 f64), DFMA, FFMA and the packed
// FFMA2 (the rates the SIMT GEMM engines are held against in DESIGN.md).
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(256) dmma_loop(double* out, int iters) {
    double acc[16][2];
    for (int j = 0; j < 16; ++j) { acc[j][0] = threadIdx.x; acc[j][1] = j; }
    double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int j = 0; j < 16; ++j)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(acc[j][0]), "+d"(acc[j][1]) : "d"(a), "d"(b));
    double s = 0; for (int j = 0; j < 16; ++j) s += acc[j][0] + acc[j][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void __launch_bounds__(256) dfma_loop(double* out, int iters) {
    double acc[32];
    for (int j = 0; j < 32; ++j) acc[j] = threadIdx.x + j;
    double a = 1.0 + threadIdx.x * 1e-9, b = 1e-9;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int j = 0; j < 32; ++j) acc[j] = fma(acc[j], a, b);
    double s = 0; for (int j = 0; j < 32; ++j) s += acc[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void __launch_bounds__(256) ffma_loop(float* out, int iters) {
    float acc[32];
    for (int j = 0; j < 32; ++j) acc[j] = threadIdx.x + j;
    float a = 1.0f + threadIdx.x * 1e-7f, b = 1e-7f;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int j = 0; j < 32; ++j) acc[j] = fmaf(acc[j], a, b);
    float s = 0; for (int j = 0; j < 32; ++j) s += acc[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void __launch_bounds__(256) ffma2_loop(float* out, int iters) {
    float2 acc[16];
    for (int j = 0; j < 16; ++j) acc[j] = make_float2(threadIdx.x + j, j);
    const float2 a = make_float2(1.0f + threadIdx.x * 1e-7f, 1.0f - threadIdx.x * 1e-7f), b = make_float2(1e-7f, 2e-7f);
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int j = 0; j < 16; ++j) acc[j] = __ffma2_rn(acc[j], a, b);
    float s = 0; for (int j = 0; j < 16; ++j) s += acc[j].x + acc[j].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    double* out; cudaMalloc(&out, (size_t)sms * 8 * 256 * 8);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 4096;
    for (int blocks_per_sm : {2, 4, 8}) {
        dmma_loop<<<sms * blocks_per_sm, 256>>>(out, 16); cudaDeviceSynchronize();
        cudaEventRecord(e0); dmma_loop<<<sms * blocks_per_sm, 256>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        double flops = (double)sms * blocks_per_sm * 8 /*warps*/ * iters * 16.0 * (8 * 8 * 4 * 2);
        printf("DMMA %d blocks/SM: %.3f ms  %.1f TFLOP/s\n", blocks_per_sm, ms, flops / ms / 1e9);
        dfma_loop<<<sms * blocks_per_sm, 256>>>(out, 16); cudaDeviceSynchronize();
        cudaEventRecord(e0); dfma_loop<<<sms * blocks_per_sm, 256>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        flops = (double)sms * blocks_per_sm * 256 * iters * 32.0 * 2;
        printf("DFMA %d blocks/SM: %.3f ms  %.1f TFLOP/s\n", blocks_per_sm, ms, flops / ms / 1e9);
        float* fout = reinterpret_cast<float*>(out);
        ffma_loop<<<sms * blocks_per_sm, 256>>>(fout, 16); cudaDeviceSynchronize();
        cudaEventRecord(e0); ffma_loop<<<sms * blocks_per_sm, 256>>>(fout, iters * 4); cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        flops = (double)sms * blocks_per_sm * 256 * (iters * 4.0) * 32.0 * 2;
        printf("FFMA %d blocks/SM: %.3f ms  %.1f TFLOP/s\n", blocks_per_sm, ms, flops / ms / 1e9);
        ffma2_loop<<<sms * blocks_per_sm, 256>>>(fout, 16); cudaDeviceSynchronize();
        cudaEventRecord(e0); ffma2_loop<<<sms * blocks_per_sm, 256>>>(fout, iters * 4); cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        flops = (double)sms * blocks_per_sm * 256 * (iters * 4.0) * 32.0 * 2;
        printf("FFMA2 %d blocks/SM: %.3f ms  %.1f TFLOP/s\n", blocks_per_sm, ms, flops / ms / 1e9);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
