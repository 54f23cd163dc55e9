// Does ptxas keep mul.rn.f32x2 + add.rn.f32x2 unfused?  Compares against scalar __fmul_rn/__fadd_rn.
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
__global__ void chk(unsigned* bad, unsigned* fusedlike) {
    unsigned s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    unsigned nb = 0, nf = 0;
    for (int i = 0; i < 1000; i++) {
        float v[6];
        for (int k = 0; k < 6; k++) {
            s ^= s << 13; s ^= s >> 17; s ^= s << 5;
            v[k] = __uint_as_float((s & 0x007fffffu) | 0x3f800000u) - 1.5f;
        }
        float2 r = add2(mul2(make_float2(v[0], v[1]), make_float2(v[2], v[3])), make_float2(v[4], v[5]));
        float e0 = __fadd_rn(__fmul_rn(v[0], v[2]), v[4]), e1 = __fadd_rn(__fmul_rn(v[1], v[3]), v[5]);
        float f0 = __fmaf_rn(v[0], v[2], v[4]), f1 = __fmaf_rn(v[1], v[3], v[5]);
        if (r.x != e0) nb++;
        if (r.y != e1) nb++;
        if (r.x == f0 && f0 != e0) nf++;
        if (r.y == f1 && f1 != e1) nf++;
    }
    atomicAdd(bad, nb);
    atomicAdd(fusedlike, nf);
}
int main() {
    unsigned *d, h[2] = {0, 0};
    cudaMalloc(&d, 8);
    cudaMemset(d, 0, 8);
    chk<<<64, 256>>>(d, d + 1);
    cudaMemcpy(h, d, 8, cudaMemcpyDeviceToHost);
    printf("mismatches vs unfused: %u ; of which equal to fused: %u (of %u)\n", h[0], h[1], 64 * 256 * 2000);
    return 0;
}
