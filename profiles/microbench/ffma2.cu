// Micro-benchmark: issue rate of scalar FFMA vs packed FFMA2 (fma.rn.f32x2) on sm_100a.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 ffma2.cu -o ffma2 ; run on the B200.
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) {
    u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d;
}
__global__ void k_scalar(float* out, float a, float b, int iters) {
    float x[16];
#pragma unroll
    for (int i = 0; i < 16; i++) x[i] = threadIdx.x + i;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 16; i++) x[i] = fmaf(x[i], a, b);
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_packed(float* out, float a, float b, int iters) {
    u64 x[8];
    float2 aa = {a, a}, bb = {b, b};
    u64 A = *reinterpret_cast<u64*>(&aa), B = *reinterpret_cast<u64*>(&bb);
#pragma unroll
    for (int i = 0; i < 8; i++) { float2 v = {threadIdx.x + 2.f * i, threadIdx.x + 2.f * i + 1}; x[i] = *reinterpret_cast<u64*>(&v); }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) x[i] = fma2(x[i], A, B);
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) { float2 v = *reinterpret_cast<float2*>(&x[i]); s += v.x + v.y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
    const int blocks = 148 * 8, threads = 256, iters = 4096;
    float* out; cudaMalloc(&out, blocks * threads * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int rep = 0; rep < 3; rep++) {
        float ms;
        cudaEventRecord(e0); k_scalar<<<blocks, threads>>>(out, 1.0001f, 0.5f, iters); cudaEventRecord(e1);
        cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
        double fl = 2.0 * 16 * iters * (double)blocks * threads;
        printf("scalar FFMA : %.3f ms  %.1f TFLOP/s\n", ms, fl / ms / 1e9);
        cudaEventRecord(e0); k_packed<<<blocks, threads>>>(out, 1.0001f, 0.5f, iters); cudaEventRecord(e1);
        cudaEventSynchronize(e1); cudaEventElapsedTime(&ms, e0, e1);
        printf("packed FFMA2: %.3f ms  %.1f TFLOP/s\n", ms, fl / ms / 1e9);
    }
    return 0;
}
