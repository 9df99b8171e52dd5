// Microbenchmark (GPU box): issue rate per SM of the instructions the dp4a GEMV is made of -- IDP.4A alone, LOP3 alone,
// the 1:1 LOP3 + IDP.4A mix of block_dot_q4_q8_fast, and that mix plus one LDS.128 per 8 -- to see which pipe bounds
// the consumers of gemv_q4_0_set.cuh / gemv_q4_0_fused.cuh.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o dp4a_rate dp4a_rate.cu ; not part of the product.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ int dp4a_us(unsigned a, unsigned b, int c) { int d; asm volatile("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }
__device__ __forceinline__ unsigned lop_and(unsigned a, unsigned b) { unsigned d; asm volatile("and.b32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }

template <int MODE>   // 0: 16 IDP.4A; 1: 16 LOP3; 2: 8 LOP3 + 8 IDP.4A (dependent pairs); 3: mode 2 + one LDS.128
__global__ void k(int * out, long long * cyc, int iters, unsigned seed) {
    __shared__ uint4 sm[1024];
    sm[threadIdx.x] = make_uint4(seed, seed * 3, seed * 5, seed * 7);
    int acc[16]; unsigned q[16];
    for (int i = 0; i < 16; i++) { acc[i] = i + threadIdx.x; q[i] = seed * (i + 1) + threadIdx.x; }
    const unsigned y = seed ^ 0x01020304u;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
        if (MODE == 0) {
#pragma unroll
            for (int i = 0; i < 16; i++) acc[i] = dp4a_us(q[i], y, acc[i]);
        } else if (MODE == 1) {
#pragma unroll
            for (int i = 0; i < 16; i++) q[i] = lop_and(q[i], y + i);
        } else {
            if (MODE == 3) { const uint4 v = sm[(threadIdx.x + it) & 1023]; q[0] ^= v.x; q[1] ^= v.y; q[2] ^= v.z; q[3] ^= v.w; }
#pragma unroll
            for (int i = 0; i < 8; i++) { const unsigned m = lop_and(q[i], 0x0F0F0F0Fu + i); acc[i] = dp4a_us(m, y, acc[i]); }
        }
    }
    long long t1 = clock64();
    int s = 0; for (int i = 0; i < 16; i++) s += acc[i] + (int)q[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

int main() {
    int * out; long long * cyc; cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
    const int iters = 20000;
    const char * names[4] = {"16 IDP.4A/iter", "16 LOP3/iter", "8 LOP3 + 8 IDP.4A/iter", "8 LOP3 + 8 IDP.4A + 1 LDS.128/iter"};
    for (int mode = 0; mode < 4; mode++) for (int threads = 128; threads <= 1024; threads *= 2) {
        for (int rep = 0; rep < 2; rep++) {
            if (mode == 0) k<0><<<148, threads>>>(out, cyc, iters, 12345u);
            if (mode == 1) k<1><<<148, threads>>>(out, cyc, iters, 12345u);
            if (mode == 2) k<2><<<148, threads>>>(out, cyc, iters, 12345u);
            if (mode == 3) k<3><<<148, threads>>>(out, cyc, iters, 12345u);
            cudaDeviceSynchronize();
        }
        long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        double c = 0; for (int i = 0; i < 148; i++) c += h[i]; c /= 148;
        const double instr = (double)iters * 16 * threads / 32;   // warp instructions per SM (16 per thread per iteration)
        printf("%-38s threads/SM %4d : %6.2f warp-instr / clk / SM  (%5.1f lanes/clk)\n", names[mode], threads, instr / c, 32 * instr / c);
    }
    return 0;
}
