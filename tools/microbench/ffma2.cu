// Microbenchmark: throughput of scalar FFMA vs packed FFMA2 (fma.rn.f32x2) on sm_100a.
// Build+run on the GPU box: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ffma2 tools/microbench/ffma2.cu && /tmp/ffma2
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) {
    unsigned long long ra = *reinterpret_cast<unsigned long long*>(&a), rb = *reinterpret_cast<unsigned long long*>(&b), rc = *reinterpret_cast<unsigned long long*>(&c), rd;
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    return *reinterpret_cast<float2*>(&rd);
}
template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, float a, float b, int iters) {
    float2 r[8];
    for (int j = 0; j < 8; j++) r[j] = make_float2(threadIdx.x * 0.001f + j, threadIdx.x * 0.002f + j);
    const float2 va = make_float2(a, a), vb = make_float2(b, b);
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (MODE == 0) { r[j].x = fmaf(r[j].x, a, b); r[j].y = fmaf(r[j].y, a, b); }
            else r[j] = fma2(r[j], va, vb);
        }
    }
    float s = 0;
    for (int j = 0; j < 8; j++) s += r[j].x + r[j].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
    float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 20000;
    for (int mode = 0; mode < 2; mode++) {
        for (int rep = 0; rep < 3; rep++) {
            cudaEventRecord(e0);
            if (mode == 0) k<0><<<148 * 8, 256>>>(d, 0.999f, 0.001f, iters); else k<1><<<148 * 8, 256>>>(d, 0.999f, 0.001f, iters);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            double fma = 148.0 * 8 * 256 * (double)iters * 16;
            if (rep == 2) printf("%s: %.3f ms, %.1f TFLOP/s (2 flop per FMA)\n", mode == 0 ? "FFMA (scalar)" : "FFMA2 (packed)", ms, 2 * fma / ms * 1e-9);
        }
    }
    return 0;
}
