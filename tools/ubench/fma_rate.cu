// Microbenchmark (GPU box): issue rate of fma.rn.f32x2 (FFMA2) vs scalar FFMA vs a mix, per SM.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o fma_rate fma_rate.cu ; not part of the product.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d;
}
__device__ __forceinline__ float fma1(float a, float b, float c) {
    float d; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d;
}

template <int MODE>   // 0: 8 FFMA2 per iter; 1: 16 FFMA per iter; 2: 4 FFMA2 + 8 FFMA per iter; 3: FFMA2 with denormal multiplicand
__global__ void k(float * out, long long * cyc, int iters, float seed) {
    uint64_t a2[8]; float a1[16];
    for (int i = 0; i < 8; i++) { float lo = seed + i + threadIdx.x, hi = lo * 0.5f; asm("mov.b64 %0, {%1,%2};" : "=l"(a2[i]) : "f"(lo), "f"(hi)); }
    for (int i = 0; i < 16; i++) a1[i] = seed * i + threadIdx.x;
    uint64_t m2, c2; float mf = 1.0000001f, cf = 1e-9f;
    if (MODE == 3) mf = __int_as_float(40000);
    asm("mov.b64 %0, {%1,%1};" : "=l"(m2) : "f"(mf)); asm("mov.b64 %0, {%1,%1};" : "=l"(c2) : "f"(cf));
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
        if (MODE == 0 || MODE == 3) {
#pragma unroll
            for (int i = 0; i < 8; i++) a2[i] = fma2(a2[i], m2, c2);
        } else if (MODE == 1) {
#pragma unroll
            for (int i = 0; i < 16; i++) a1[i] = fma1(a1[i], mf, cf);
        } else {
#pragma unroll
            for (int i = 0; i < 4; i++) { a2[i] = fma2(a2[i], m2, c2); a1[2 * i] = fma1(a1[2 * i], mf, cf); a1[2 * i + 1] = fma1(a1[2 * i + 1], mf, cf); }
        }
    }
    long long t1 = clock64();
    float s = 0; for (int i = 0; i < 8; i++) { float lo, hi; asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a2[i])); s += lo + hi; }
    for (int i = 0; i < 16; i++) s += a1[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

int main() {
    float * out; long long * cyc; cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
    const int iters = 20000;
    const char * names[4] = {"8 FFMA2/iter", "16 FFMA/iter", "4 FFMA2 + 8 FFMA/iter", "8 FFMA2/iter, denormal multiplicand"};
    for (int mode = 0; mode < 4; mode++) for (int threads = 128; threads <= 1024; threads *= 2) {
        for (int rep = 0; rep < 2; rep++) {
            if (mode == 0) k<0><<<148, threads>>>(out, cyc, iters, 1.5f);
            if (mode == 1) k<1><<<148, threads>>>(out, cyc, iters, 1.5f);
            if (mode == 2) k<2><<<148, threads>>>(out, cyc, iters, 1.5f);
            if (mode == 3) k<3><<<148, threads>>>(out, cyc, iters, 1.5f);
            cudaDeviceSynchronize();
        }
        long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        double c = 0; for (int i = 0; i < 148; i++) c += h[i]; c /= 148;
        const double fmas = (double)iters * 16 * threads;   // every mode does 16 fp32 FMAs per thread per iteration
        printf("%-38s threads/SM %4d : %7.1f cycles/iter/warp-set, %6.1f fp32 FMA / clk / SM\n", names[mode], threads, c / iters, fmas / c);
    }
    return 0;
}
