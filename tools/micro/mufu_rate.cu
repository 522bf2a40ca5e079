// Micro-benchmark: MUFU throughput per SM for ex2.approx.f32, ex2.approx.f16x2, tanh.approx.f32/f16x2.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mufu_rate mufu_rate.cu && ./mufu_rate
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(unsigned * out, int iters) {
    unsigned x[8];
    for (int i = 0; i < 8; ++i) x[i] = 0x3c003800u + threadIdx.x + i;
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+r"(x[i]));
            if (MODE == 1) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(x[i]));
            if (MODE == 2) asm volatile("tanh.approx.f32 %0, %0;" : "+r"(x[i]));
            if (MODE == 3) asm volatile("tanh.approx.f16x2 %0, %0;" : "+r"(x[i]));
            if (MODE == 4) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+r"(x[i]));
        }
    }
    long long t1 = clock64();
    unsigned s = 0;
    for (int i = 0; i < 8; ++i) s ^= x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = (unsigned) (t1 - t0);
}
int main() {
    unsigned * d; cudaMalloc(&d, 1 << 20);
    const char * names[] = {"ex2.f32", "ex2.f16x2", "tanh.f32", "tanh.f16x2", "rcp.f32"};
    for (int warps = 4; warps <= 16; warps *= 2) {
        for (int m = 0; m < 5; ++m) {
            const int iters = 2000;
            if (m == 0) k<0><<<1, 32 * warps>>>(d, iters);
            if (m == 1) k<1><<<1, 32 * warps>>>(d, iters);
            if (m == 2) k<2><<<1, 32 * warps>>>(d, iters);
            if (m == 3) k<3><<<1, 32 * warps>>>(d, iters);
            if (m == 4) k<4><<<1, 32 * warps>>>(d, iters);
            cudaDeviceSynchronize();
            unsigned cyc; cudaMemcpy(&cyc, d, 4, cudaMemcpyDeviceToHost);
            const double ops = (double) iters * 8 * 32 * warps * (m == 1 || m == 3 ? 2 : 1);
            printf("%-10s warps=%2d  %.2f results/clk/SM\n", names[m], warps, ops / cyc);
        }
    }
    return 0;
}
