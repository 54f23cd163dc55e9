// Microbenchmark: scalar FP32 (FMUL/FADD/FFMA) vs packed f32x2 (sm_100a FMUL2/FADD2/FFMA2) issue rates,
// alone and interleaved with integer ALU work.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ float2 mul2(float2 a, float2 b) {
    unsigned long long r, x = *reinterpret_cast<unsigned long long*>(&a), y = *reinterpret_cast<unsigned long long*>(&b);
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(x), "l"(y));
    return *reinterpret_cast<float2*>(&r);
}
__device__ __forceinline__ float2 add2(float2 a, float2 b) {
    unsigned long long r, x = *reinterpret_cast<unsigned long long*>(&a), y = *reinterpret_cast<unsigned long long*>(&b);
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(x), "l"(y));
    return *reinterpret_cast<float2*>(&r);
}
__global__ void k_scalar(float* out, int iters, float m, float c) {
    float a[16];
#pragma unroll
    for (int i = 0; i < 16; i++) a[i] = threadIdx.x * 1e-3f + i;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 16; i++) a[i] = __fadd_rn(__fmul_rn(a[i], m), c);
    }
    float s = 0; for (int i = 0; i < 16; i++) s += a[i];
    if (s == 1.2345f) out[0] = s;
}
__global__ void k_packed(float* out, int iters, float m, float c) {
    float2 a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = make_float2(threadIdx.x * 1e-3f + i, threadIdx.x * 1e-3f + i + 8);
    float2 mm = make_float2(m, m), cc = make_float2(c, c);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) a[i] = add2(mul2(a[i], mm), cc);
    }
    float s = 0; for (int i = 0; i < 8; i++) s += a[i].x + a[i].y;
    if (s == 1.2345f) out[0] = s;
}
__global__ void k_scalar_int(float* out, int iters, float m, float c) {
    float a[16]; unsigned x[8];
#pragma unroll
    for (int i = 0; i < 16; i++) a[i] = threadIdx.x * 1e-3f + i;
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = threadIdx.x + i;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 16; i++) a[i] = __fadd_rn(__fmul_rn(a[i], m), c);
#pragma unroll
        for (int i = 0; i < 8; i++) { x[i] ^= x[i] << 13; x[i] ^= x[i] >> 17; }
    }
    float s = 0; for (int i = 0; i < 16; i++) s += a[i];
    unsigned y = 0; for (int i = 0; i < 8; i++) y ^= x[i];
    if (s == 1.2345f || y == 0x12345u) out[0] = s;
}
__global__ void k_packed_int(float* out, int iters, float m, float c) {
    float2 a[8]; unsigned x[8];
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = make_float2(threadIdx.x * 1e-3f + i, threadIdx.x * 1e-3f + i + 8);
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = threadIdx.x + i;
    float2 mm = make_float2(m, m), cc = make_float2(c, c);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) a[i] = add2(mul2(a[i], mm), cc);
#pragma unroll
        for (int i = 0; i < 8; i++) { x[i] ^= x[i] << 13; x[i] ^= x[i] >> 17; }
    }
    float s = 0; for (int i = 0; i < 8; i++) s += a[i].x + a[i].y;
    unsigned y = 0; for (int i = 0; i < 8; i++) y ^= x[i];
    if (s == 1.2345f || y == 0x12345u) out[0] = s;
}
template <class K> float run(K k, float* d, int iters) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 1e9;
    for (int r = 0; r < 4; r++) {
        cudaEventRecord(a); k<<<148 * 8, 256>>>(d, iters, 0.999f, 1e-3f); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); if (r && ms < best) best = ms;
    }
    return best;
}
int main() {
    float* d; cudaMalloc(&d, 4);
    const int iters = 4096;
    double flop = 2.0 * 16 * iters * 256.0 * 148 * 8;
    float t1 = run(k_scalar, d, iters), t2 = run(k_packed, d, iters), t3 = run(k_scalar_int, d, iters), t4 = run(k_packed_int, d, iters);
    printf("scalar mul+add : %.3f ms  %.1f TFLOP/s\n", t1, flop / t1 / 1e9);
    printf("packed mul2+add2: %.3f ms  %.1f TFLOP/s\n", t2, flop / t2 / 1e9);
    printf("scalar + int   : %.3f ms\n", t3);
    printf("packed + int   : %.3f ms\n", t4);
    return cudaGetLastError() != cudaSuccess;
}
