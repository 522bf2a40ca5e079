// Micro-benchmark: does the f32 -> f16x2 pack (F2FP.F16.F32.PACK_AB) share a pipe with MUFU.EX2?
// Per iteration each thread issues NE ex2.approx.f32 and NC cvt.rn.f16x2.f32 (independent chains); if the two
// sit on different pipes the time is max(), on the same pipe it is the sum.  Also: an integer-pipe f32->f16 pack.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o xu_mix xu_mix.cu && ./xu_mix
#include <cstdio>
#include <cuda_runtime.h>
template <int NE, int NC, int NI>
__global__ void k(unsigned * out, int iters) {
    float e[8]; unsigned c[8]; unsigned q[8];
    for (int i = 0; i < 8; ++i) { e[i] = -0.001f * (threadIdx.x + i); c[i] = 0x3c003800u + threadIdx.x + i; q[i] = 0x3f000000u + threadIdx.x * 8 + i; }
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (i < NE) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(e[i]));
            if (i < NC) asm volatile("{ .reg .f32 a; mov.b32 a, %0; cvt.rn.f16x2.f32 %0, a, a; }" : "+r"(c[i]));
            if (i < NI) {
                // positive normal f32 in [2^-14, 65504] -> f16 bits with round-to-nearest (ties away), on the integer pipe
                unsigned a = q[i] + 0x00001000u - 0x38000000u, b = (q[i] ^ 0x00400000u) + 0x00001000u - 0x38000000u;
                asm volatile("prmt.b32 %0, %1, %2, 0x7632;" : "=r"(q[i]) : "r"(a << 3), "r"(b << 3));
                q[i] |= 0x3f000000u;
            }
        }
    }
    long long t1 = clock64();
    unsigned s = 0;
    for (int i = 0; i < 8; ++i) s ^= __float_as_uint(e[i]) ^ c[i] ^ q[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = (unsigned) (t1 - t0);
}
template <int NE, int NC, int NI> void run(unsigned * d, int warps) {
    const int iters = 2000;
    k<NE, NC, NI><<<1, 32 * warps>>>(d, iters);
    cudaDeviceSynchronize();
    unsigned cyc; cudaMemcpy(&cyc, d, 4, cudaMemcpyDeviceToHost);
    printf("warps=%2d  ex2=%d cvt=%d intpack=%d per iter: %.2f clk per warp-iteration-SM  (ex2 %.1f/clk/SM, cvt pairs %.1f/clk/SM, intpack %.1f/clk/SM)\n",
           warps, NE, NC, NI, (double) cyc / iters, (double) iters * NE * 32 * warps / cyc, (double) iters * NC * 32 * warps / cyc,
           (double) iters * NI * 32 * warps / cyc);
}
int main() {
    unsigned * d; cudaMalloc(&d, 1 << 20);
    for (int warps = 8; warps <= 16; warps *= 2) {
        run<8, 0, 0>(d, warps); run<0, 8, 0>(d, warps); run<8, 8, 0>(d, warps); run<8, 4, 0>(d, warps);
        run<0, 0, 8>(d, warps); run<8, 0, 4>(d, warps);
    }
    return 0;
}
